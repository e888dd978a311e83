// ll_episode.cu — K1: the fused episode kernel.  One persistent launch per generation evaluates every
// (member, layout) episode of this GPU's shard: MLP forward -> argmax action -> lander update -> step penalty,
// until touchdown or max_steps, then the terminal penalty and the per-member fitness.
//
// Replaces (reference file:line, azimonti/lunar-lander):
//   NNEngineTrainer<T>::train_generation / sim_task          src/nn_engine_trainer.cpp:443-539
//   GameLogicCpp<T>::reset/apply_action/update_physics/check_landing_crash/get_state/update
//                                                              src/game_logic.cpp:124-310
//   calculate_step_penalty / calculate_terminal_penalty(3-arg) src/game_logic.cpp:312-320, 352-388
//   nn::ANN_MLP_GA<T>::feedforward (ma-libs, call site)        src/nn_engine_trainer.cpp:501-502
//
// Mapping.  A warp owns a tile of 32/G members, G = lanes per member (32 when n_layouts >= 32, else the next
// power of two >= n_layouts).  Lane = one episode slot: its lander state, step counter and penalty accumulator
// live in registers for the whole episode.  The tile's member rows are staged ONCE into shared memory by TMA
// bulk copies (cp.async.bulk + mbarrier) in the k-major device layout, so that inside the dot products all lanes
// of a member read the same 16-byte weight vector (a shared-memory broadcast) and each lane reads only its own
// activation column [k][lane] (bank = lane, conflict free).  Lanes whose episode ended pull the member's next
// layout from a shared-memory counter; warps pull tiles from a global counter (episodes last 9..max_steps steps).
// Per-layout results are kept in shared memory and summed in layout order by one lane per member, in double,
// exactly like the reference's accumulators (src/nn_engine_trainer.cpp:477-530).
#include <stdio.h>

#include <algorithm>
#include <stdlib.h>
#include <string.h>

#include "ll_ops.cuh"

// default choice between the two tensor-core kernels (measured on B200): up to 8 layouts fill one 8-slot
// N tile of the transposed kernel exactly (configs[2]: 5.09e9 vs 3.48e9 lander-steps/s; configs[1]: 6.4e8 vs 6.0e8); with 50
// layouts the 16-row M-side kernel keeps the edge (configs[3]: 85.8 % of the FMA roofline vs 76.8 % / 69.0 % for one / two
// N tiles per warp)
#ifndef LL_MMAT_DEFAULT
#define LL_MMAT_DEFAULT(n_layouts) ((n_layouts) <= 8)
#endif

namespace
{
    struct ep_dev_args
    {
        const void* params;
        const void* layouts;
        int n_layouts, max_steps;
        unsigned long long n_members;
        double *fitness, *lr, *steps;
        int* ep_steps;
        double* ep_fit;
        unsigned int* queue;
        ll_step_rec* trace;
        int trace_cap;
        ll_episode_result* trace_result;
        // geometry
        int G, nslots;            // episode slots per member, members per warp tile
        int chunks;               // warps a member's layouts are dealt to (chunk c flies layouts c, c + chunks, ...); > 1 needs ep_steps / ep_fit
        unsigned int* part_count; // [n_members] arrival counters of the chunks, zero before the launch
        int Q;                    // lanes per episode slot (the slot's output neurons are dealt to its Q lanes); G * nslots * Q <= 32
        unsigned int n_tiles;
        int slot_stride;          // elements between member rows in shared memory
        int width_a, width_b;     // rows of the two activation buffers: a0 holds the inputs and the outputs of the odd layers, a1 the even ones
        int off_w, off_act, off_fit, off_meta, warp_bytes; // byte offsets inside a warp's shared-memory slice
        int act_pitch;                                     // row pitch of the activation buffers, elements
    };

    template <typename T, int N> struct vec_loader;
    template <int N> struct vec_loader<float, N>
    {
        static __device__ __forceinline__ void load(const float* p, float* v)
        {
#pragma unroll
            for (int i = 0; i < N; i += 4)
            {
                const float4 q = *reinterpret_cast<const float4*>(p + i);
                v[i] = q.x, v[i + 1] = q.y, v[i + 2] = q.z, v[i + 3] = q.w;
            }
        }
    };
    template <int N> struct vec_loader<double, N>
    {
        static __device__ __forceinline__ void load(const double* p, double* v)
        {
#pragma unroll
            for (int i = 0; i < N; i += 2)
            {
                const double2 q = *reinterpret_cast<const double2*>(p + i);
                v[i] = q.x, v[i + 1] = q.y;
            }
        }
    };

    template <> struct vec_loader<float, 1>
    {
        static __device__ __forceinline__ void load(const float* p, float* v) { v[0] = p[0]; }
    };
    template <> struct vec_loader<double, 1>
    {
        static __device__ __forceinline__ void load(const double* p, double* v) { v[0] = p[0]; }
    };

    // TILE consecutive output neurons of one layer for this lane's episode.
    //   w   : &Wt[0][o0] of the lane's member (shared memory), row pitch out_pad
    //   ain : this lane's activation column of the previous layer, element k at ain[k * P] (P: row pitch of the activation
    //         buffers in elements, 32 = one column per lane)
    template <typename T, int MODE, int TILE>
    __device__ __forceinline__ void layer_tile(const T* __restrict__ w, const T* __restrict__ b, const T* __restrict__ ain,
                                               T* __restrict__ aout, int in, int out_pad, int act, int P)
    {
        using O = ll_ops<T, MODE>;
        T acc[TILE];
#pragma unroll
        for (int i = 0; i < TILE; ++i) acc[i] = T(0);
#pragma unroll 2
        for (int k = 0; k < in; ++k)
        {
            const T a = ain[k * P];
            T wv[TILE];
            vec_loader<T, TILE>::load(w + k * out_pad, wv);
#pragma unroll
            for (int i = 0; i < TILE; ++i) acc[i] = O::madd(acc[i], wv[i], a);
        }
        T bv[TILE];
        vec_loader<T, TILE>::load(b, bv);
#pragma unroll
        for (int i = 0; i < TILE; ++i) acc[i] = O::add(acc[i], bv[i]);
        O::template act_n<TILE>(acc, act);
#pragma unroll
        for (int i = 0; i < TILE; ++i) aout[i * P] = acc[i];
    }

    // Whole forward pass; returns the index of the first maximum of the output layer.
    // Q > 1: the episode's column is shared by Q lanes (sub = 0..Q-1), each taking every Q-th tile of 4 output neurons; a neuron's
    // own arithmetic (sequential-k multiply-add, bias last, activation) is untouched, so the result is bit-identical to Q == 1.
    // Every lane of the warp must call this together when Q > 1 (warp barriers between layers).
    // WIDE: compiled for Q >= 16 only (one neuron per lane) — its own instantiation so that the register tiles of the other paths do
    // not set the kernel's register count (small populations want many resident warps).
    template <typename T, int MODE, bool WIDE = false>
    __device__ __forceinline__ int forward(const ll_topo& tp, const T* __restrict__ wrow, T* a0, T* a1, int sub = 0, int Q = 1, int P = 32)
    {
        constexpr int BIG = sizeof(T) == 4 ? 16 : 8;
        T* ain  = a0;
        T* aout = a1;
        for (int l = 0; l < tp.n_layers; ++l)
        {
            const int in = tp.in[l], op = tp.out_pad[l];
            const T* w = wrow + tp.w_off[l];
            const T* b = wrow + tp.b_off[l];
            if constexpr (WIDE)
            { // one neuron per lane (small populations: a whole half-warp or warp flies ONE episode, the layer's neurons side by side)
                for (int o = sub; o < tp.out[l]; o += Q) layer_tile<T, MODE, 1>(w + o, b + o, ain, aout + o * P, in, op, tp.activation, P);
                __syncwarp();
            }
            else if (Q > 1)
            {
                for (int o = sub * 4; o < op; o += Q * 4) layer_tile<T, MODE, 4>(w + o, b + o, ain, aout + o * P, in, op, tp.activation, P);
                __syncwarp();
            }
            else
            {
                int o = 0;
                for (; o + BIG <= op; o += BIG) layer_tile<T, MODE, BIG>(w + o, b + o, ain, aout + o * P, in, op, tp.activation, P);
                if (BIG == 16 && o + 8 <= op)
                {
                    layer_tile<T, MODE, 8>(w + o, b + o, ain, aout + o * P, in, op, tp.activation, P);
                    o += 8;
                }
                if (o + 4 <= op)
                {
                    layer_tile<T, MODE, 4>(w + o, b + o, ain, aout + o * P, in, op, tp.activation, P);
                    o += 4;
                }
            }
            T* t = ain;
            ain  = aout;
            aout = t;
        }
        const int n_out = tp.out[tp.n_layers - 1];
        int best        = 0;
        T vbest         = ain[0];
        for (int j = 1; j < n_out; ++j)
        {
            const T v = ain[j * P];
            if (v > vbest) vbest = v, best = j;
        }
        if (Q > 1) __syncwarp(); // the outputs are read before any lane overwrites the buffers with the next step's state
        return best;
    }

    template <typename T> __device__ __forceinline__ T clampT(T v, T lo, T hi) { return v < lo ? lo : (hi < v ? hi : v); }
    __device__ __forceinline__ float absT(float v) { return fabsf(v); }
    __device__ __forceinline__ double absT(double v) { return fabs(v); }

    // get_state(): src/game_logic.cpp:275-293
    template <typename T, int MODE>
    __device__ __forceinline__ void write_state(const ll_env<T>& e, T x, T y, T vx, T vy, T fuel, T cx, T* col, int P = 32)
    {
        using O = ll_ops<T, MODE>;
        T s0, s1, s2, s3, s4;
        const T dx = O::sub(x, cx), dy = O::sub(y, e.pad_y);
        if (O::fast)
        {
            s0 = O::mul(vx, e.inv_max_vx), s1 = O::mul(vy, e.inv_max_vy), s2 = O::mul(dx, e.inv_width);
            s3 = O::mul(dy, e.inv_height), s4 = O::mul(fuel, e.inv_max_fuel);
        }
        else
        {
            s0 = O::div(vx, e.max_vx), s1 = O::div(vy, e.max_vy), s2 = O::div(dx, e.width);
            s3 = O::div(dy, e.height), s4 = O::div(fuel, e.max_fuel);
        }
        col[0 * P] = clampT(s0, T(-2), T(2));
        col[1 * P] = clampT(s1, T(-2), T(2));
        col[2 * P] = clampT(s2, T(-5), T(5));
        col[3 * P] = clampT(s3, T(-5), T(5));
        col[4 * P] = s4;
    }

    // RP: the activation buffers have a run-time row pitch (Q > 1 and wide layers, see the launcher); otherwise the pitch is the
    // constant 32 and a slot's column is its first lane, which keeps the address arithmetic of the inner loops in immediates.
    template <typename T, int MODE, bool RP, bool WIDE = false>
    __global__ void __launch_bounds__(128, WIDE ? 6 : 1) ll_episode_kernel(const __grid_constant__ ll_topo tp, const __grid_constant__ ll_env<T> env,
                                                                           const __grid_constant__ ep_dev_args a)
    {
        using O = ll_ops<T, MODE>;
        extern __shared__ __align__(128) unsigned char smem_raw[];
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        unsigned char* ws = smem_raw + (size_t)warp * a.warp_bytes;
        uint64_t* bar     = reinterpret_cast<uint64_t*>(ws);
        int* next_layout  = reinterpret_cast<int*>(ws + 16);
        T* wsm            = reinterpret_cast<T*>(ws + a.off_w);
        const int Q = a.Q, sub = lane % Q, elane = lane / Q; // Q lanes share one episode slot (and its activation column)
        // activation buffers: row pitch P elements, the slot's column = elane (P = 32 and column = lane when Q == 1)
        const int P       = RP ? a.act_pitch : 32;
        T* act0           = reinterpret_cast<T*>(ws + a.off_act) + (RP ? elane : lane - sub);
        T* act1           = act0 + a.width_a * P;
        double* res_fit   = reinterpret_cast<double*>(ws + a.off_fit);
        int* res_meta     = reinterpret_cast<int*>(ws + a.off_meta);

        const int G = a.G, L = a.n_layouts;
        const int slot = elane / G, j = elane - slot * G;
        const unsigned gmask = (Q >= 32 ? 0xffffffffu : ((1u << Q) - 1u)) << (lane - sub);
        const T* wrow  = wsm + (size_t)min(slot, a.nslots - 1) * a.slot_stride; // surplus lane groups (slot >= nslots) idle
        const ll_layout<T>* layouts = static_cast<const ll_layout<T>*>(a.layouts);
        const uint32_t row_bytes    = (uint32_t)tp.p_dev * sizeof(T);

        if (lane == 0) ll_mbar_init(bar, 1);
        __syncwarp();
        uint32_t phase = 0;

        for (;;)
        {
            unsigned int tile = 0;
            if (lane == 0) tile = atomicAdd(a.queue, 1u);
            tile = __shfl_sync(0xffffffffu, tile, 0);
            const int C = a.chunks;
            if (tile >= a.n_tiles * (unsigned int)C) break;
            const int chunk = (int)(tile % (unsigned int)C);
            tile /= (unsigned int)C;
            const int Lc = (L - chunk + C - 1) / C; // this chunk's layouts: chunk + i * C, i < Lc
            const unsigned long long m0 = (unsigned long long)tile * a.nslots;
            const int nvalid            = (int)min((unsigned long long)a.nslots, a.n_members - m0);
            const unsigned long long member = m0 + slot;
            const bool valid                = slot < nvalid;

            // ---- stage the tile's member rows: TMA bulk copies, completion on the warp's mbarrier
            if (lane == 0) ll_mbar_expect_tx(bar, row_bytes * (uint32_t)nvalid);
            __syncwarp();
            if (lane < nvalid)
                ll_bulk_g2s(wsm + (size_t)lane * a.slot_stride, static_cast<const T*>(a.params) + (m0 + lane) * (size_t)tp.p_dev, row_bytes, bar);
            if (j == 0 && sub == 0) next_layout[slot] = min(G, Lc);
            __syncwarp();
            ll_mbar_wait(bar, phase);
            phase ^= 1u;

            // ---- per-lane episode state (registers)
            T x = T(0), y = T(0), vx = T(0), vy = T(0), fuel = T(0), cx = T(0), lpad = T(0), lpad_r = T(0);
            int dir = 0, layout = -1, steps = 0;
            double acc  = 0.0;
            bool active = false;
            auto begin_episode = [&](int l) {
                const ll_layout<T> lay = layouts[l];
                layout = l, dir = lay.dir, cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
                x = lay.init_x, y = env.y0, vx = env.vx0, vy = env.vy0, fuel = env.max_fuel;
                steps = 0, acc = 0.0;
            };
            if (valid && j < Lc)
            {
                begin_episode(chunk + j * C);
                active = true;
            }
            const bool tracer = a.trace != nullptr && tile == 0 && chunk == 0 && lane == 0;

            while (__any_sync(0xffffffffu, active))
            {
                write_state<T, MODE>(env, x, y, vx, vy, fuel, cx, act0, P);
                const int action = forward<T, MODE, WIDE>(tp, wrow, act0, act1, sub, Q, P);
                if (active)
                {
                    // apply_action: src/game_logic.cpp:156-183
                    if (fuel > T(0))
                    {
                        if (action == 1) vy = O::sub(vy, T(0.3)), fuel = O::sub(fuel, T(1.0));
                        else if (action == 2) vx = O::sub(vx, T(0.05)), fuel = O::sub(fuel, T(0.5));
                        else if (action == 3) vx = O::add(vx, T(0.05)), fuel = O::sub(fuel, T(0.5));
                        fuel = (T(0) < fuel) ? fuel : T(0);
                    }
                    // update_physics: :185-195
                    vy = O::add(vy, env.ay);
                    vx = O::mul(vx, env.one_minus_mu_x);
                    vy = O::mul(vy, env.one_minus_mu_y);
                    x  = O::add(x, vx);
                    y  = O::add(y, vy);
                    // check_landing_crash: :197-239 (skipped on the very first step)
                    bool landed = false, success = false;
                    if (steps != 0 && y >= env.ground)
                    {
                        y      = env.ground;
                        landed = true;
                        const bool on_pad = (x >= lpad) && (O::add(x, env.lander_w) <= lpad_r);
                        const bool safe   = (absT(vx) < env.max_vx) && (absT(vy) < env.max_vy);
                        success           = on_pad && safe;
                        vx = T(0), vy = T(0);
                    }
                    // calculate_step_penalty on the post-update state: :312-320, accumulated in double (trainer :507-508)
                    const T dist = O::add(absT(O::sub(x, cx)), absT(O::sub(y, env.pad_y)));
                    T sp         = O::add(T(0), O::mul(dist, env.sp_dist));
                    if (action > 0) sp = O::add(sp, env.sp_action);
                    acc = __dadd_rn(acc, (double)sp);
                    if (tracer && steps < a.trace_cap)
                    {
                        ll_step_rec r;
                        r.action = (double)action;
                        write_state<T, MODE>(env, x, y, vx, vy, fuel, cx, act0, P);
                        for (int i = 0; i < 5; ++i) r.state[i] = (double)act0[i * P];
                        r.x = (double)x, r.y = (double)y, r.vx = (double)vx, r.vy = (double)vy, r.fuel = (double)fuel;
                        a.trace[steps] = r;
                    }
                    ++steps;
                    if (landed || steps >= a.max_steps)
                    {
                        // calculate_terminal_penalty(steps, direction, bonus): :352-388
                        const bool crashed = landed && !success;
                        T term = O::add(T(0), O::mul((T)steps, env.tp_steps));
                        term   = O::add(term, O::mul(dist, env.tp_dist));
                        if (success)
                        {
                            term = O::sub(term, env.tp_landed_bonus);
                            term = O::sub(term, O::mul(fuel, env.tp_fuel_bonus));
                        }
                        else if (crashed)
                        {
                            term = O::add(term, env.tp_crashed);
                            term = O::add(term, O::mul(O::sqrt(O::add(O::mul(vx, vx), O::mul(vy, vy))), env.tp_crash_v_mag));
                        }
                        else if (fuel <= T(0) && !landed) { term = O::add(term, env.tp_no_fuel); }
                        if (!success) term = O::add(term, O::mul(O::sqrt(O::add(O::mul(vx, vx), O::mul(vy, vy))), T(1.0)));
                        const double fit = __dadd_rn((double)term, acc);
                        const int meta   = steps | (landed ? (1 << 28) : 0) | (crashed ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
                        const int ridx   = slot * L + layout;
                        if (sub == 0) res_fit[ridx] = fit, res_meta[ridx] = meta;
                        if (a.ep_steps && sub == 0)
                        {
                            a.ep_steps[member * (size_t)L + layout] = meta;
                            a.ep_fit[member * (size_t)L + layout]   = fit;
                        }
                        if (tracer && a.trace_result)
                        {
                            ll_episode_result r;
                            r.steps = steps, r.flags = (landed ? 1 : 0) | (crashed ? 2 : 0) | (success ? 4 : 0);
                            r.step_penalty_sum = acc, r.terminal_penalty = (double)term;
                            r.bonus_lr[0] = r.bonus_lr[1] = 0.0;
                            if (success) r.bonus_lr[dir] = (double)O::add(T(0), env.tp_landed_lr_bonus);
                            *a.trace_result = r;
                        }
                        int nl = 0; // the slot's Q lanes hold identical state and arrive here together
                        if (sub == 0) nl = atomicAdd(&next_layout[slot], 1);
                        if (Q > 1) nl = __shfl_sync(gmask, nl, lane - sub);
                        if (nl < Lc) begin_episode(chunk + nl * C);
                        else active = false;
                    }
                }
                __syncwarp();
            }

            // ---- per-member totals in layout order: src/nn_engine_trainer.cpp:518-530
            if (C > 1) __threadfence(); // this lane's per-episode results are visible device-wide before the arrival is counted
            __syncwarp();
            bool finalize = valid && j == 0 && sub == 0;
            if (finalize && C > 1)
            { // the member's other layouts were flown by other warps: the last chunk to arrive sums all of them, in layout order,
              // from the per-episode arrays in global memory
                __threadfence();
                finalize = atomicAdd(a.part_count + member, 1u) == (unsigned int)C - 1u;
                if (finalize)
                {
                    __threadfence();
                    a.part_count[member] = 0u;
                }
            }
            if (finalize)
            {
                double total = 0.0, total_steps = 0.0;
                T bonus0 = T(0), bonus1 = T(0);
                const volatile int* g_meta   = a.ep_steps + member * (size_t)L;
                const volatile double* g_fit = a.ep_fit + member * (size_t)L;
                for (int l = 0; l < L; ++l)
                {
                    const int meta = C > 1 ? g_meta[l] : res_meta[slot * L + l];
                    total          = __dadd_rn(total, C > 1 ? g_fit[l] : res_fit[slot * L + l]);
                    total_steps    = __dadd_rn(total_steps, (double)(meta & 0x0fffffff));
                    if (meta & (1 << 30))
                    {
                        if (layouts[l].dir == 0) bonus0 = O::add(bonus0, env.tp_landed_lr_bonus);
                        else bonus1 = O::add(bonus1, env.tp_landed_lr_bonus);
                    }
                }
                const double lr = (double)((bonus1 < bonus0) ? bonus1 : bonus0);
                a.fitness[member] = __dsub_rn(total, lr);
                a.lr[member]      = lr;
                a.steps[member]   = total_steps;
            }
            __syncwarp(); // all lanes are done with this tile's shared memory before it is overwritten
        }
    }

    // Single forward pass on a given state (ll_feedforward_index): one warp, lane 0's answer.
    template <typename T, int MODE>
    __global__ void ll_feedforward_kernel(const __grid_constant__ ll_topo tp, const T* __restrict__ row, const T* __restrict__ state,
                                          unsigned long long* action)
    {
        extern __shared__ __align__(128) unsigned char smem_raw[];
        T* wsm  = reinterpret_cast<T*>(smem_raw);
        T* act0 = wsm + tp.p_dev + threadIdx.x;
        T* act1 = act0 + tp.max_width * 32;
        for (int i = threadIdx.x; i < tp.p_dev; i += 32) wsm[i] = row[i];
        __syncwarp();
        for (int i = 0; i < tp.in[0]; ++i) act0[i * 32] = state[i];
        const int best = forward<T, MODE>(tp, wsm, act0, act1);
        if (threadIdx.x == 0) *action = (unsigned long long)best;
    }

    template <typename T, int MODE> int launch_t(ll_ctx* ctx, const ll_episode_args& h)
    {
        const ll_topo& tp = ctx->topo;
        ep_dev_args a;
        a.params = h.params, a.layouts = h.layouts, a.n_layouts = h.n_layouts, a.max_steps = h.max_steps;
        a.n_members = h.n_members, a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps;
        a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit, a.queue = h.queue;
        a.trace = h.trace, a.trace_cap = h.trace_cap, a.trace_result = h.trace_result;
        const int L = h.n_layouts;
        // Geometry of a warp tile: Q lanes per episode slot, G slots per member, nslots members.  Q = 1 (a lane walks all neurons of
        // its episode) unless lanes would idle or the population is too small to fill the GPU (both below).  LL_EXACT_Q forces a value.
        auto geometry = [&](int q, int* g_out, int* nslots_out) {
            const int E = 32 / q;
            int G       = 1;
            while (G < L && G < E) G <<= 1;
            int nslots = E / G;
            // a warp never stages more than ~48 KB of member rows: with few layouts and wide networks the tile holds
            // fewer members than the warp has lane groups, and the surplus lanes idle
            while (nslots > 1 && (size_t)nslots * tp.p_dev * sizeof(T) > (size_t)48 << 10) nslots >>= 1;
            *g_out = G, *nslots_out = nslots;
            return (unsigned long long)((h.n_members + nslots - 1) / nslots);
        };
        int Q = 1;
        if (h.trace == nullptr)
        {
            // few layouts and a wide network: the 48 KB staging cap leaves lane groups without a member (16_64_32 fp64 with 8
            // layouts: 8 busy lanes of 32) - give the idle lanes to the busy slots, Q lanes per slot, as far as the widest layer has
            // tiles of 4 neurons to deal.  The number of slots per member does not change, so no extra rounds.
            int g1 = 0, n1 = 0;
            geometry(1, &g1, &n1);
            int widest = 4;
            for (int l = 0; l < tp.n_layers; ++l) widest = std::max(widest, tp.out_pad[l]);
            while (Q * 2 * g1 * n1 <= 32 && Q * 2 <= widest / 4 && Q < 8) Q <<= 1;
            // small population (the reference's own scale): a generation lasts as long as one lane's chain of steps, each a few
            // thousand fp64 instructions of ONE warp - deal the neurons of a slot to up to 4 lanes (and its layouts to more warps,
            // below) while the GPU has warps to spare.  configs[0], 100 x 50: 46.5 / 31.9 / 25.4 / 26.6 ms for Q = 1 / 2 / 4 / 8.
            // Smaller still (configs[0]: 100 x 50 = 5 000 episodes for 148 SMs): ONE neuron per lane, Q = 16 lanes per episode (its own
            // instantiation at 76 registers, 24 warps per SM), when all the (member, layout pair) work items are resident at once:
            // 25.4 -> 18.6 ms per generation (Q = 32: 20.8 - more work items than warp slots).
            if (Q == 1 && h.ep_steps && h.ep_fit)
                for (int q = 16; q > 1; q >>= 1)
                {
                    if (q == 8) continue; // 8 lanes x tiles of 4 neurons: measured no better than 4 (26.6 vs 25.4 ms)
                    int g = 0, n = 0;
                    const unsigned long long tiles = geometry(q, &g, &n);
                    const unsigned long long ch    = L > 32 / q ? (unsigned long long)((L + 32 / q - 1) / (32 / q)) : 1ull;
                    unsigned long long warps_per_sm = 8;
                    if (q == 16) warps_per_sm = std::min<unsigned long long>(24, ((size_t)200 << 10) / ((size_t)tp.p_dev * sizeof(T) + 4096));
                    if ((q == 16 ? widest >= 16 : q <= widest / 4) && tiles * ch <= (unsigned long long)ctx->sm_count * warps_per_sm)
                    {
                        Q = q;
                        break;
                    }
                }
            const char* force = getenv("LL_EXACT_Q");
            if (force && *force)
            {
                const int f = atoi(force);
                if (f == 1 || f == 2 || f == 4 || f == 8 || ((f == 16 || f == 32) && h.ep_steps && h.ep_fit)) Q = f;
            }
        }
        a.Q = Q;
        // a member with more layouts than a warp has lanes needs several rounds; when the population is too small to fill the GPU
        // anyway, its layouts are dealt to ceil(L / 32) warps instead (one episode per lane, one round)
        int chunks = 1;
        {
            int g = 0, n = 0;
            if (h.trace == nullptr && h.ep_steps && h.ep_fit && L > 32 / Q && geometry(Q, &g, &n) < (unsigned long long)ctx->sm_count * 8ull)
                chunks = (L + 32 / Q - 1) / (32 / Q);
            const char* force = getenv("LL_EXACT_CHUNKS");
            if (force && *force && h.trace == nullptr && h.ep_steps && h.ep_fit) chunks = std::max(1, std::min(atoi(force), L));
        }
        a.chunks = chunks, a.part_count = nullptr;
        const int L_eff = (L + chunks - 1) / chunks; // layouts per chunk (the geometry is sized for these)
        {
            const int E = 32 / Q;
            int G       = 1;
            while (G < L_eff && G < E) G <<= 1;
            int nslots = E / G;
            while (nslots > 1 && (size_t)nslots * tp.p_dev * sizeof(T) > (size_t)48 << 10) nslots >>= 1;
            a.G = G, a.nslots = nslots;
        }
        if (chunks > 1)
        {
            const size_t count_bytes = (size_t)h.n_members * sizeof(unsigned int);
            if (ctx->part_cap < count_bytes)
            {
                if (ctx->d_part) cudaFree(ctx->d_part);
                ctx->d_part = nullptr, ctx->part_cap = 0;
                LL_CUDA(ctx, cudaMalloc(&ctx->d_part, count_bytes));
                ctx->part_cap = count_bytes;
            }
            a.part_count = static_cast<unsigned int*>(ctx->d_part);
            LL_CUDA(ctx, cudaMemsetAsync(a.part_count, 0, count_bytes, ctx->stream));
        }
        a.n_tiles = (unsigned int)((h.n_members + a.nslots - 1) / a.nslots);
        // member rows in shared memory: pitch == 4 words (mod 32) so that 16-byte reads of the same element by
        // lanes of different members fall into different bank groups
        int stride_words = tp.p_dev * (int)(sizeof(T) / 4);
        while (a.nslots > 1 && (stride_words % 32) != 4) stride_words += 4;
        a.slot_stride = stride_words / (int)(sizeof(T) / 4);
        auto up       = [](int v) { return (v + 127) & ~127; };
        a.off_w       = 256;
        a.off_act     = up(a.off_w + a.nslots * a.slot_stride * (int)sizeof(T));
        // the two activation buffers alternate layer by layer: each is as wide as the widest vector it ever holds (5-32-128-32-4:
        // 128 + 32 rows instead of 2 x 128 - with a 69 KB fp64 row that is the difference between one and two warps per SM)
        a.width_a = tp.in[0], a.width_b = 8;
        for (int l = 0; l < tp.n_layers; ++l)
        {
            int& w = (l & 1) ? a.width_a : a.width_b;
            w      = std::max(w, tp.out_pad[l]);
        }
        a.width_a = std::max(a.width_a, 8);
        // Row pitch of the activation buffers.  Q == 1: one column per lane, pitch 32.  Q > 1: E = 32 / Q columns; the Q lanes of a
        // slot store rows 4 apart (their tiles of 4 neurons), so the pitch is the smallest P >= E with 4 P = E modulo one sweep of
        // the banks (128 B) - the Q row segments of a store then tile the banks instead of piling onto the same ones (pitch 32 and
        // columns Q apart: 8-way conflicts on every activation store of 16_64_32 fp64, ncu).
        a.act_pitch = 32;
        int widest_out = 4;
        for (int l = 0; l < tp.n_layers; ++l) widest_out = std::max(widest_out, tp.out_pad[l]);
        const bool rp = (Q > 1 && widest_out >= 32) || Q >= 16; // narrow networks store little: the constant pitch is faster there (16_16: 6 %)
        if (Q >= 16) a.act_pitch = 32 / Q; // neuron per lane: the lanes of a slot store consecutive rows, E = 1 or 2 columns
        else if (rp)
        {
            const int E = 32 / Q, W = 128 / (int)sizeof(T);
            int P = E;
            while ((4 * P) % W != E % W) ++P;
            a.act_pitch = P;
        }
        a.off_fit = up(a.off_act + (a.width_a + a.width_b) * a.act_pitch * (int)sizeof(T));
        const int R   = std::max(32, a.nslots * L); // per-layout results of every member of the tile (indexed slot * L + layout)
        a.off_meta    = a.off_fit + R * 8;
        a.warp_bytes  = up(a.off_meta + R * 4);

        int max_smem = 0;
        LL_CUDA(ctx, cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device));
        if (a.warp_bytes > max_smem)
            return ll_fail(ctx, LL_EINVAL, "topology needs " + std::to_string(a.warp_bytes) + " B of shared memory per warp, device allows " + std::to_string(max_smem));
        int wpb = 4;
        while (wpb > 1 && wpb * a.warp_bytes > max_smem / 2) wpb >>= 1; // prefer >= 2 blocks per SM
        const size_t smem = (size_t)wpb * a.warp_bytes;
        auto kern         = Q >= 16 ? ll_episode_kernel<T, MODE, true, true> : (rp ? ll_episode_kernel<T, MODE, true> : ll_episode_kernel<T, MODE, false>);
        LL_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;
        LL_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, wpb * 32, smem));
        if (per_sm < 1) return ll_fail(ctx, LL_EINVAL, "episode kernel does not fit on an SM");
        unsigned int blocks_needed = (a.n_tiles * (unsigned int)chunks + wpb - 1) / wpb;
        unsigned int grid          = (unsigned int)(per_sm * ctx->sm_count);
        if (grid > blocks_needed) grid = blocks_needed;
        if (grid < 1) grid = 1;
        ll_env<T> env;
        ll_make_env<T>(ctx->env_cfg, MODE == LL_MODE_FAST, &env);
        LL_CUDA(ctx, cudaMemsetAsync(h.queue, 0, sizeof(unsigned int), ctx->stream));
        kern<<<grid, wpb * 32, smem, ctx->stream>>>(tp, env, a);
        LL_CUDA(ctx, cudaGetLastError());
        return LL_OK;
    }

    template <typename T, int MODE> int launch_ff(ll_ctx* ctx, const void* row, const void* state, unsigned long long* action)
    {
        const ll_topo& tp = ctx->topo;
        const size_t smem = ((size_t)tp.p_dev + 2 * (size_t)tp.max_width * 32) * sizeof(T);
        auto kern         = ll_feedforward_kernel<T, MODE>;
        LL_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<1, 32, smem, ctx->stream>>>(tp, static_cast<const T*>(row), static_cast<const T*>(state), action);
        LL_CUDA(ctx, cudaGetLastError());
        return LL_OK;
    }
} // namespace

// which tensor-core kernel: slots on the N side (ll_episode_mmat.cu) or on the M side (ll_episode_mma.cu) of the MMA
static bool ll_use_mmat(const ll_ctx* ctx, int n_layouts)
{
    if (!ll_mmat_applicable(ctx, n_layouts)) return false;
    const char* force = getenv("LL_FAST_KERNEL");
    if (force && !strcmp(force, "mmat")) return true;
    if (force && *force) return false;
    return LL_MMAT_DEFAULT(n_layouts);
}

// the Blackwell tensor path (ll_episode_tc05.cu): tcgen05.mma, accumulators and activations in TMEM
static bool ll_use_tc05(const ll_ctx* ctx, int n_layouts)
{
    if (!ll_tc05_applicable(ctx, n_layouts)) return false;
    const char* force = getenv("LL_FAST_KERNEL");
    if (force && !strcmp(force, "tc05")) return true;
    if (force && *force) return false;
    return LL_TC05_DEFAULT(ctx, n_layouts);
}

// warp per member, neurons across lanes (ll_episode_warp.cu): no weight reuse to exploit with one layout per member
static bool ll_use_warp(const ll_ctx* ctx, int n_layouts)
{
    if (!ll_warp_applicable(ctx, n_layouts)) return false;
    const char* force = getenv("LL_FAST_KERNEL");
    if (force && !strcmp(force, "warp")) return true;
    if (force && *force) return false;
    if (n_layouts <= LL_WARP_MAX_LAYOUTS) return true;
    // small populations with many layouts (configs[4]: 100 x 50): one warp per episode, 3.2 ms against 6.6 ms on the mma.sync kernel
    return ll_warp_per_episode(ctx, ctx->pop_local, n_layouts, ctx->d_ep_steps && ctx->d_ep_fit);
}

int ll_launch_episode(ll_ctx* ctx, const ll_episode_args& a)
{
    if (ctx->mode == LL_MODE_EXACT && ll_exactw_applicable(ctx, a)) return ll_launch_episode_exactw(ctx, a); // small populations
    if (ctx->precision == 64) return launch_t<double, LL_MODE_EXACT>(ctx, a);
    if (ctx->mode == LL_MODE_FAST)
    { // the team kernel needs weight reuse across layouts; traces and tiny layout counts stay lane-per-episode
        if (a.trace == nullptr && ll_use_warp(ctx, a.n_layouts)) return ll_launch_episode_warp(ctx, a);
        if (a.trace == nullptr && ll_use_tc05(ctx, a.n_layouts)) return ll_launch_episode_tc05(ctx, a);
        if (a.trace == nullptr && ll_use_mmat(ctx, a.n_layouts)) return ll_launch_episode_mmat(ctx, a);
        if (a.trace == nullptr && ll_mma_applicable(ctx, a.n_layouts)) return ll_launch_episode_mma(ctx, a);
        if (a.trace == nullptr && ll_fast_applicable(ctx, a.n_layouts)) return ll_launch_episode_fast(ctx, a);
        return launch_t<float, LL_MODE_FAST>(ctx, a);
    }
    return launch_t<float, LL_MODE_EXACT>(ctx, a);
}

int ll_launch_feedforward(ll_ctx* ctx, const void* member_row, const void* d_state, unsigned long long* d_action)
{
    if (ctx->precision == 64) return launch_ff<double, LL_MODE_EXACT>(ctx, member_row, d_state, d_action);
    if (ctx->mode == LL_MODE_FAST) return launch_ff<float, LL_MODE_FAST>(ctx, member_row, d_state, d_action);
    return launch_ff<float, LL_MODE_EXACT>(ctx, member_row, d_state, d_action);
}

const char* ll_episode_kernel_name(const ll_ctx* ctx, int n_layouts)
{
    if (ctx->mode == LL_MODE_EXACT && n_layouts > 0)
    { // what ll_evaluate will pick (per-episode arrays present, no trace)
        ll_episode_args probe{};
        probe.n_layouts = n_layouts, probe.n_members = ctx->pop_local;
        probe.ep_steps = ctx->d_ep_steps, probe.ep_fit = ctx->d_ep_fit;
        if (ll_exactw_applicable(ctx, probe)) return ctx->precision == 64 ? "ll_episode_exactw_kernel<double> (warp per episode)" : "ll_episode_exactw_kernel<float> (warp per episode)";
    }
    if (ctx->precision == 64) return "ll_episode_kernel<double,exact>";
    if (ctx->mode == LL_MODE_FAST && ll_use_warp(ctx, n_layouts))
        return ll_warp_per_episode(ctx, ctx->pop_local, n_layouts, ctx->d_ep_steps && ctx->d_ep_fit) ? "ll_episode_warp_ep_kernel (warp per episode, FFMA + smem)"
                                                                                                    : "ll_episode_warp_kernel (warp per member, FFMA + smem)";
    if (ctx->mode == LL_MODE_FAST && ll_use_tc05(ctx, n_layouts)) return "ll_episode_tc05_kernel (tcgen05 + TMEM; stragglers: mma.sync)";
    if (ctx->mode == LL_MODE_FAST && ll_use_mmat(ctx, n_layouts)) return "ll_episode_mmat_kernel (3 x f16-split mma.sync, slots on N)";
    if (ctx->mode == LL_MODE_FAST && ll_mma_applicable(ctx, n_layouts)) return "ll_episode_mma_kernel (3 x f16-split mma.sync)";
    if (ctx->mode == LL_MODE_FAST) return ll_fast_applicable(ctx, n_layouts) ? "ll_episode_fast_kernel (slot-pool GEMM)" : "ll_episode_kernel<float,fast>";
    return "ll_episode_kernel<float,exact>";
}
