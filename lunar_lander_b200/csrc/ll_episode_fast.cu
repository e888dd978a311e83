// ll_episode_fast.cu — K1 fast (fp32, LL_MODE_FAST): the throughput variant of the fused episode kernel.
//
// Replaces the same reference code as ll_episode.cu (train_generation / sim_task src/nn_engine_trainer.cpp:443-539,
// GameLogicCpp<T> src/game_logic.cpp:124-388, the ma-libs feedforward call :501-502), with free summation order,
// fused multiply-add and tanh.approx.f32 — judged on outcome rates, not on bit patterns (tests/test_gpu_parity.py).
//
// Mapping.  A lane-per-episode GEMV fetches one shared-memory word per FMA; shared memory returns 32
// words/clk/SM against 128 FMA lanes/clk/SM, i.e. <= 25 % of the FP32 pipe.  All L layouts of a member share its
// weights, so a forward pass is a small GEMM D[slots x out] = A[slots x in] * W^T[in x out]: each thread computes a
// register tile of E episode slots x 8 outputs (E*8 FFMA per E+8 words), K-deep layers are split over k-slices that
// live in one warp and are reduced by shuffles.
//
// Episodes last 9..max_steps steps, so a member alone leaves most of its slots idle for most of its life.  A CTA
// therefore keeps M members resident (their rows staged once each by a TMA bulk copy, cp.async.bulk + mbarrier) and owns
// a POOL of G slot groups (E slots each) that the members share: a group belongs to one member at a time, slots pull
// that member's next layout when their episode ends, groups drained by a finishing member go to the oldest member that
// still has layouts to start, and a member whose episodes are all done hands its buffer to the next member of the
// global queue.  Per step: env phase (one thread per slot: argmax -> lander update -> penalties -> next state / next
// layout), then one phase per layer over the tiles of all live groups, separated by CTA barriers.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "ll_env.cuh"

namespace
{
    constexpr int FO       = 8;   // outputs per thread tile
    constexpr int MAX_M    = 8;   // resident members per CTA
    constexpr int MAX_G    = 32;  // slot groups per CTA
    constexpr int MAX_RES  = 256; // layouts whose per-episode results are kept in shared memory
    constexpr int MAX_INIT = MAX_G * 8;

    struct fast_layer
    {
        int in, out, out_pad, w_off, b_off; // topology (element offsets inside a member row)
        int nog, ks;                        // output groups of 8, k-slices per tile
        int in_off, out_off;                // element offsets of the activation buffers
        int last;
    };

    struct fast_args
    {
        const float* params;
        const ll_layout<float>* layouts;
        int n_layouts, max_steps;
        unsigned long long n_members;
        double *fitness, *lr, *steps;
        int* ep_steps;
        double* ep_fit;
        unsigned int* queue;
        unsigned long long* prof;
        int n_layers, act, p_dev;
        int M, G, S, threads;                           // resident members, slot groups, slots = G * E, CTA size
        int w_stride;                                   // floats between weight buffers
        int off_ctl, off_w, off_act, off_slot, off_res; // byte offsets inside the CTA's shared memory
        int res_in_smem;
        fast_layer layer[LL_MAX_LAYERS];
    };

    enum
    {
        ST_EMPTY   = 0,
        ST_LOADING = 1,
        ST_ACTIVE  = 2
    };

    struct pool_ctl
    {
        uint64_t bar[MAX_M];
        unsigned long long member[MAX_M];
        int status[MAX_M], next_layout[MAX_M], inflight[MAX_M], parity[MAX_M];
        unsigned int age[MAX_M];
        unsigned int age_counter;
        int dirty, n_g, queue_done, exit_flag, n_init;
        int gmem[MAX_G];  // group -> member buffer, -1 = free
        int glist[MAX_G]; // groups that hold at least one live slot
        int init_slot[MAX_INIT], init_layout[MAX_INIT];
    };

    struct slot_view
    {
        float *x, *y, *vx, *vy, *fuel, *cx, *lpad, *lpad_r;
        int *steps, *layout, *active;
        double* acc;
        __device__ __forceinline__ slot_view(unsigned char* base, int S)
        {
            acc    = reinterpret_cast<double*>(base);
            x      = reinterpret_cast<float*>(acc + S);
            y = x + S, vx = y + S, vy = vx + S, fuel = vy + S, cx = fuel + S, lpad = cx + S, lpad_r = lpad + S;
            steps  = reinterpret_cast<int*>(lpad_r + S);
            layout = steps + S, active = layout + S;
        }
        static __host__ __device__ int bytes(int S) { return S * (8 + 8 * 4 + 3 * 4); }
    };

    __device__ __forceinline__ float fast_act(float z, int act) { return ll_ops<float, LL_MODE_FAST>::act(z, act); }

    // One (episode group, output group, k-slice) register tile: acc[e][j] = sum_k in[k][g*E + e] * W[k][col(j)], where the
    // tile's 8 outputs are columns go*4..go*4+3 and half+go*4..half+go*4+3 (half = out_pad/2): the 8 lanes of a quarter-warp
    // then read 128 contiguous bytes per LDS.128 (no bank conflict) when consecutive lanes hold consecutive output groups.
    // Pointer-increment form: per k two row-pointer bumps, two LDS.128 (weights), E LDS.32 (activations), E*8 FFMA.
    template <int E>
    __device__ __forceinline__ void gemm_tile(const float* __restrict__ wp, const float* __restrict__ ap, int klen, int out_pad, int S, int half,
                                              float (&acc)[E][FO])
    {
#pragma unroll
        for (int e = 0; e < E; ++e)
#pragma unroll
            for (int j = 0; j < FO; ++j) acc[e][j] = 0.f;
#pragma unroll 4
        for (int k = 0; k < klen; ++k)
        {
            const float4 w0 = *reinterpret_cast<const float4*>(wp);
            const float4 w1 = *reinterpret_cast<const float4*>(wp + half);
            float av[E];
#pragma unroll
            for (int e = 0; e < E; ++e) av[e] = ap[e];
            wp += out_pad, ap += S;
#pragma unroll
            for (int e = 0; e < E; ++e)
            {
                acc[e][0] = fmaf(av[e], w0.x, acc[e][0]);
                acc[e][1] = fmaf(av[e], w0.y, acc[e][1]);
                acc[e][2] = fmaf(av[e], w0.z, acc[e][2]);
                acc[e][3] = fmaf(av[e], w0.w, acc[e][3]);
                acc[e][4] = fmaf(av[e], w1.x, acc[e][4]);
                acc[e][5] = fmaf(av[e], w1.y, acc[e][5]);
                acc[e][6] = fmaf(av[e], w1.z, acc[e][6]);
                acc[e][7] = fmaf(av[e], w1.w, acc[e][7]);
            }
        }
    }

    // Split-K reduce-scatter inside a warp.  The KS lanes of a tile sit TPW = 32/KS lanes apart (so a quarter-warp reads one
    // weight row: no bank conflicts); after log2(KS) exchange rounds lane slice `ksi` owns the 8/KS consecutive outputs
    // starting at j0, fully summed.  res[e][0..8/KS) receives them.
    template <int E, int KS>
    __device__ __forceinline__ int reduce_scatter(float (&acc)[E][FO], int ksi, unsigned mask, float (&res)[E][FO / KS])
    {
        constexpr int TPW = 32 / KS;
        if constexpr (KS == 1)
        {
#pragma unroll
            for (int e = 0; e < E; ++e)
#pragma unroll
                for (int j = 0; j < FO; ++j) res[e][j] = acc[e][j];
            return 0;
        }
        else
        {
            const bool b0 = ksi & 1;
            float r1[E][4];
#pragma unroll
            for (int e = 0; e < E; ++e)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                {
                    const float send = b0 ? acc[e][j] : acc[e][j + 4];
                    const float keep = b0 ? acc[e][j + 4] : acc[e][j];
                    r1[e][j]         = keep + __shfl_xor_sync(mask, send, TPW);
                }
            int j0 = b0 ? 4 : 0;
            if constexpr (KS == 2)
            {
#pragma unroll
                for (int e = 0; e < E; ++e)
#pragma unroll
                    for (int j = 0; j < 4; ++j) res[e][j] = r1[e][j];
                return j0;
            }
            else
            {
                const bool b1 = ksi & 2;
#pragma unroll
                for (int e = 0; e < E; ++e)
#pragma unroll
                    for (int j = 0; j < 2; ++j)
                    {
                        const float send = b1 ? r1[e][j] : r1[e][j + 2];
                        const float keep = b1 ? r1[e][j + 2] : r1[e][j];
                        res[e][j]        = keep + __shfl_xor_sync(mask, send, 2 * TPW);
                    }
                return j0 + (b1 ? 2 : 0);
            }
        }
    }


    // One layer over the tiles of all live groups.  tile -> (live group gi, output group go), go fastest so that a
    // quarter-warp shares the slot group (activation broadcast) and walks consecutive weight columns.  KS k-slices of a
    // tile (rows k = KS*i + ksi) are reduced by shuffles; the owners add the bias, apply the activation and store
    // out[o][slot].  The output layer stores into the Z buffer the env phase takes its argmax from.
    template <int E, int KS>
    __device__ __forceinline__ void layer_phase(const fast_args& a, const fast_layer& ly, const pool_ctl* ctl, const float* __restrict__ W0,
                                                float* __restrict__ ACT, int tid, int n_g)
    {
        constexpr int TPW = 32 / KS;
        const int S = a.S;
        const float* in = ACT + ly.in_off;
        float* outb     = ACT + ly.out_off;
        const int tiles = n_g * ly.nog;
        int tile, stride, ksi = 0;
        unsigned mask = 0xffffffffu;
        if constexpr (KS == 1) tile = tid, stride = a.threads;
        else
        {
            const int lane = tid & 31, q = lane & (TPW - 1);
            ksi    = lane / TPW;
            tile   = (tid >> 5) * TPW + q;
            stride = (a.threads >> 5) * TPW;
            mask   = 0;
#pragma unroll
            for (int i = 0; i < KS; ++i) mask |= 1u << (i * TPW + q);
        }
        const int half = ly.out_pad >> 1;
        for (; tile < tiles; tile += stride)
        {
            const int gi = tile / ly.nog, go = tile - gi * ly.nog;
            const int g  = ctl->glist[gi];
            const float* W = W0 + (size_t)ctl->gmem[g] * a.w_stride;
            const int kl   = (ly.in - ksi + KS - 1) / KS;
            float acc[E][FO];
            gemm_tile<E>(W + ly.w_off + ksi * ly.out_pad + go * 4, in + ksi * S + g * E, kl, KS * ly.out_pad, KS * S, half, acc);
            float res[E][FO / KS];
            const int j0 = reduce_scatter<E, KS>(acc, ksi, mask, res);
            if constexpr (KS == 1)
            { // owns both halves: columns go*4 + j and half + go*4 + j
#pragma unroll
                for (int j = 0; j < FO; ++j)
                {
                    const int o = go * 4 + (j & 3) + ((j & 4) ? half : 0);
                    if (o < ly.out)
                    {
                        const float bj = W[ly.b_off + o];
#pragma unroll
                        for (int e = 0; e < E; ++e) outb[o * S + g * E + e] = fast_act(res[e][j] + bj, a.act);
                    }
                }
            }
            else
            {
                const int o0   = go * 4 + (j0 & 3) + ((j0 & 4) ? half : 0); // first owned column (the owned ones are consecutive)
                const float* b = W + ly.b_off + o0;
                float* out     = outb + o0 * S + g * E;
#pragma unroll
                for (int j = 0; j < FO / KS; ++j)
                    if (o0 + j < ly.out)
                    {
                        const float bj = b[j];
#pragma unroll
                        for (int e = 0; e < E; ++e) out[j * S + e] = fast_act(res[e][j] + bj, a.act);
                    }
            }
        }
    }

    template <int E>
    __device__ __forceinline__ void layer_dispatch(const fast_args& a, const fast_layer& ly, const pool_ctl* ctl, const float* __restrict__ W0,
                                                   float* __restrict__ ACT, int tid, int n_g)
    {
        if (ly.ks == 4) layer_phase<E, 4>(a, ly, ctl, W0, ACT, tid, n_g);
        else if (ly.ks == 2) layer_phase<E, 2>(a, ly, ctl, W0, ACT, tid, n_g);
        else layer_phase<E, 1>(a, ly, ctl, W0, ACT, tid, n_g);
    }

    // begin an episode of layout l in slot s (reset(spad_x1, lpad_x1), src/game_logic.cpp:143-154) and publish its state
    __device__ __forceinline__ void slot_begin(const fast_args& a, const ll_env<float>& env, slot_view& sv, float* A0, int s, int l)
    {
        const ll_layout<float> lay = a.layouts[l];
        sv.layout[s] = l;
        sv.cx[s] = lay.center_x, sv.lpad[s] = lay.lpad_x1, sv.lpad_r[s] = lay.lpad_right;
        ll_lander<float> p{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
        sv.x[s] = p.x, sv.y[s] = p.y, sv.vx[s] = p.vx, sv.vy[s] = p.vy, sv.fuel[s] = p.fuel;
        sv.steps[s] = 0, sv.acc[s] = 0.0, sv.active[s] = 1;
        ll_get_state<float, LL_MODE_FAST>(env, p, lay.center_x, A0 + s, a.S);
    }

    // First layer (5 inputs) for one slot by its quad of lanes: lane j computes the 8-output chunks j, j+4, ... from the
    // state held in registers and stores the activations A1[o][slot].
    __device__ __forceinline__ void first_layer_quad(const fast_args& a, const fast_layer& l0, const float* __restrict__ W, float* __restrict__ ACT,
                                                     const float (&st)[5], int s, int j)
    {
        float* out = ACT + l0.out_off + s;
        for (int c = j; c < l0.nog; c += 4)
        {
            const float* wp = W + l0.w_off + c * FO;
            const float4 b0 = *reinterpret_cast<const float4*>(W + l0.b_off + c * FO);
            const float4 b1 = *reinterpret_cast<const float4*>(W + l0.b_off + c * FO + 4);
            float acc[FO]   = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int k = 0; k < 5; ++k)
            {
                const float4 w0 = *reinterpret_cast<const float4*>(wp + k * l0.out_pad);
                const float4 w1 = *reinterpret_cast<const float4*>(wp + k * l0.out_pad + 4);
                acc[0] = fmaf(st[k], w0.x, acc[0]), acc[1] = fmaf(st[k], w0.y, acc[1]), acc[2] = fmaf(st[k], w0.z, acc[2]);
                acc[3] = fmaf(st[k], w0.w, acc[3]), acc[4] = fmaf(st[k], w1.x, acc[4]), acc[5] = fmaf(st[k], w1.y, acc[5]);
                acc[6] = fmaf(st[k], w1.z, acc[6]), acc[7] = fmaf(st[k], w1.w, acc[7]);
            }
#pragma unroll
            for (int i = 0; i < FO; ++i)
                if (c * FO + i < l0.out) out[(c * FO + i) * a.S] = fast_act(acc[i], a.act);
        }
    }

    // The light phase of one live slot, executed by a quad of consecutive lanes (j = lane & 3):
    //   output layer (k split over the quad, butterfly reduce) -> argmax -> lander update -> penalties -> episode end
    //   handling (results, next layout or retirement) -> next state -> first layer.
    // The lander arithmetic is done redundantly by the four lanes (same inputs, same results); lane 0 owns the writes.
    __device__ __forceinline__ void light_slot(const fast_args& a, const fast_layer& l0, const fast_layer& ll, const ll_env<float>& env, slot_view& sv,
                                               const float* __restrict__ W, float* __restrict__ ACT, pool_ctl* ctl, double* res_fit, int* res_meta,
                                               int mb, int s, int j, unsigned qmask)
    {
        // ---- output layer: z[o] = act(b[o] + sum_k A[k][s] * W[k][o]), out <= 8
        float z[FO] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        {
            const float* ap = ACT + ll.in_off + s;
            const float* wp = W + ll.w_off;
            for (int k = j; k < ll.in; k += 4)
            {
                const float av  = ap[k * a.S];
                const float4 w0 = *reinterpret_cast<const float4*>(wp + k * ll.out_pad);
                z[0] = fmaf(av, w0.x, z[0]), z[1] = fmaf(av, w0.y, z[1]), z[2] = fmaf(av, w0.z, z[2]), z[3] = fmaf(av, w0.w, z[3]);
                if (ll.out > 4)
                {
                    const float4 w1 = *reinterpret_cast<const float4*>(wp + k * ll.out_pad + 4);
                    z[4] = fmaf(av, w1.x, z[4]), z[5] = fmaf(av, w1.y, z[5]), z[6] = fmaf(av, w1.z, z[6]), z[7] = fmaf(av, w1.w, z[7]);
                }
            }
        }
        int action  = 0;
        float vbest = 0.f;
#pragma unroll
        for (int o = 0; o < FO; ++o)
            if (o < ll.out)
            {
                float v = z[o];
                v += __shfl_xor_sync(qmask, v, 1);
                v += __shfl_xor_sync(qmask, v, 2);
                v = fast_act(v + W[ll.b_off + o], a.act);
                if (o == 0 || v > vbest) vbest = v, action = o;
            }
        // ---- lander update (src/game_logic.cpp:295-320 through ll_env.cuh)
        ll_lander<float> p{sv.x[s], sv.y[s], sv.vx[s], sv.vy[s], sv.fuel[s]};
        float cx  = sv.cx[s];
        int steps = sv.steps[s];
        bool landed, success;
        float dist;
        const float sp   = ll_env_step<float, LL_MODE_FAST>(env, p, action, steps == 0, cx, sv.lpad[s], sv.lpad_r[s], &landed, &success, &dist);
        const double acc = sv.acc[s] + (double)sp;
        ++steps;
        if (landed || steps >= a.max_steps)
        {
            int nl = 0;
            if (j == 0)
            {
                const float term = ll_terminal_penalty<float, LL_MODE_FAST>(env, p, steps, landed, success, dist);
                const double fit = (double)term + acc;
                const int layout = sv.layout[s];
                const int meta   = steps | (landed ? (1 << 28) : 0) | ((landed && !success) ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
                if (a.res_in_smem) res_fit[mb * a.n_layouts + layout] = fit, res_meta[mb * a.n_layouts + layout] = meta;
                if (a.ep_steps)
                {
                    const unsigned long long member = ctl->member[mb];
                    a.ep_steps[member * (size_t)a.n_layouts + layout] = meta;
                    a.ep_fit[member * (size_t)a.n_layouts + layout]   = fit;
                }
                nl = atomicAdd(&ctl->next_layout[mb], 1);
            }
            nl = __shfl_sync(qmask, nl, 0, 4);
            if (nl >= a.n_layouts)
            { // retired: the manager repacks / hands the group over before the next forward pass
                if (j == 0)
                {
                    sv.active[s] = 0;
                    atomicSub(&ctl->inflight[mb], 1);
                    ctl->dirty = 1;
                }
                return;
            }
            const ll_layout<float> lay = a.layouts[nl]; // reset(spad_x1, lpad_x1): src/game_logic.cpp:143-154
            p  = ll_lander<float>{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
            cx = lay.center_x;
            if (j == 0)
            {
                sv.layout[s] = nl, sv.cx[s] = cx, sv.lpad[s] = lay.lpad_x1, sv.lpad_r[s] = lay.lpad_right;
                sv.x[s] = p.x, sv.y[s] = p.y, sv.vx[s] = p.vx, sv.vy[s] = p.vy, sv.fuel[s] = p.fuel;
                sv.steps[s] = 0, sv.acc[s] = 0.0;
            }
        }
        else if (j == 0)
        {
            sv.x[s] = p.x, sv.y[s] = p.y, sv.vx[s] = p.vx, sv.vy[s] = p.vy, sv.fuel[s] = p.fuel;
            sv.steps[s] = steps, sv.acc[s] = acc;
        }
        float st[5];
        ll_get_state<float, LL_MODE_FAST>(env, p, cx, st, 1);
        if (j == 0)
        {
            float* A0 = ACT + l0.in_off + s;
#pragma unroll
            for (int i = 0; i < 5; ++i) A0[i * a.S] = st[i];
        }
        first_layer_quad(a, l0, W, ACT, st, s, j);
    }

    // per-member totals in layout order (src/nn_engine_trainer.cpp:518-530)
    __device__ void member_finalize(const fast_args& a, const ll_env<float>& env, const double* res_fit, const int* res_meta, int mb,
                                    unsigned long long member)
    {
        double total = 0.0, total_steps = 0.0;
        float bonus0 = 0.f, bonus1 = 0.f;
        for (int l = 0; l < a.n_layouts; ++l)
        {
            const int meta   = a.res_in_smem ? res_meta[mb * a.n_layouts + l] : a.ep_steps[member * (size_t)a.n_layouts + l];
            const double fit = a.res_in_smem ? res_fit[mb * a.n_layouts + l] : a.ep_fit[member * (size_t)a.n_layouts + l];
            total += fit;
            total_steps += (double)(meta & 0x0fffffff);
            if (meta & (1 << 30))
            {
                if (a.layouts[l].dir == 0) bonus0 += env.tp_landed_lr_bonus;
                else bonus1 += env.tp_landed_lr_bonus;
            }
        }
        const double lr = (double)(bonus1 < bonus0 ? bonus1 : bonus0);
        a.fitness[member] = total - lr;
        a.lr[member]      = lr;
        a.steps[member]   = total_steps;
    }

    __device__ __forceinline__ void slot_move(const fast_args& a, slot_view& sv, float* A0, int dst, int src)
    {
        sv.acc[dst] = sv.acc[src];
        sv.x[dst] = sv.x[src], sv.y[dst] = sv.y[src], sv.vx[dst] = sv.vx[src], sv.vy[dst] = sv.vy[src], sv.fuel[dst] = sv.fuel[src];
        sv.cx[dst] = sv.cx[src], sv.lpad[dst] = sv.lpad[src], sv.lpad_r[dst] = sv.lpad_r[src];
        sv.steps[dst] = sv.steps[src], sv.layout[dst] = sv.layout[src];
        sv.active[dst] = 1, sv.active[src] = 0;
        for (int q = 0; q < a.layer[0].in; ++q) A0[q * a.S + dst] = A0[q * a.S + src];
    }

    // The pool manager (thread 0, between two barriers, only after an episode ended without a successor): finalize
    // members whose episodes are all done, repack each member's live slots into the fewest groups, free drained groups,
    // pull new members into empty buffers (TMA), give free groups to the oldest member that still has layouts to start.
    template <int E>
    __device__ void pool_manage(const fast_args& a, const ll_env<float>& env, pool_ctl* ctl, slot_view& sv, float* A0, float* W0, const double* res_fit,
                                const int* res_meta)
    {
        const int L = a.n_layouts, M = a.M, G = a.G;
        ctl->dirty = 0, ctl->n_init = 0;
        for (int mb = 0; mb < M; ++mb)
            if (ctl->status[mb] == ST_ACTIVE && ctl->inflight[mb] == 0 && ctl->next_layout[mb] >= L)
            {
                member_finalize(a, env, res_fit, res_meta, mb, ctl->member[mb]);
                ctl->status[mb] = ST_EMPTY;
            }
        for (int mb = 0; mb < M; ++mb)
        {
            if (ctl->status[mb] == ST_EMPTY) continue;
            int gl[MAX_G], ng = 0, live = 0;
            for (int g = 0; g < G; ++g)
                if (ctl->gmem[g] == mb)
                {
                    gl[ng++] = g;
                    for (int e = 0; e < E; ++e) live += sv.active[g * E + e];
                }
            if (ng > 0 && live <= E * (ng - 1))
            { // holes at the front are filled from the back: the member's last groups drain
                int i = 0, j = ng * E - 1;
                for (;;)
                {
                    while (i <= j && sv.active[gl[i / E] * E + i % E]) ++i;
                    while (j >= i && !sv.active[gl[j / E] * E + j % E]) --j;
                    if (i >= j) break;
                    slot_move(a, sv, A0, gl[i / E] * E + i % E, gl[j / E] * E + j % E);
                }
            }
        }
        for (int g = 0; g < G; ++g)
            if (ctl->gmem[g] >= 0)
            {
                int live = 0;
                for (int e = 0; e < E; ++e) live += sv.active[g * E + e];
                if (!live) ctl->gmem[g] = -1;
            }
        for (int mb = 0; mb < M; ++mb)
            if (ctl->status[mb] == ST_EMPTY && !ctl->queue_done)
            {
                const unsigned long long m = (unsigned long long)atomicAdd(a.queue, 1u);
                if (m < a.n_members)
                {
                    ctl->status[mb] = ST_LOADING, ctl->member[mb] = m, ctl->next_layout[mb] = 0, ctl->inflight[mb] = 0;
                    ctl->age[mb] = ++ctl->age_counter;
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    ll_mbar_expect_tx(&ctl->bar[mb], (uint32_t)a.p_dev * 4u);
                    ll_bulk_g2s(W0 + (size_t)mb * a.w_stride, a.params + m * (size_t)a.p_dev, (uint32_t)a.p_dev * 4u, &ctl->bar[mb]);
                }
                else ctl->queue_done = 1;
            }
        for (int g = 0; g < G; ++g)
        {
            if (ctl->gmem[g] >= 0) continue;
            int best = -1;
            for (int mb = 0; mb < M; ++mb)
                if (ctl->status[mb] != ST_EMPTY && ctl->next_layout[mb] < L && (best < 0 || ctl->age[mb] < ctl->age[best])) best = mb;
            if (best < 0) break;
            ctl->gmem[g] = best;
            for (int e = 0; e < E; ++e)
            {
                const int s = g * E + e, nl = ctl->next_layout[best];
                if (nl < L)
                {
                    ctl->next_layout[best] = nl + 1, ctl->inflight[best] += 1;
                    ctl->init_slot[ctl->n_init] = s, ctl->init_layout[ctl->n_init] = nl, ctl->n_init += 1;
                }
                else sv.active[s] = 0;
            }
        }
        int n_g = 0, busy = 0;
        for (int g = 0; g < G; ++g)
            if (ctl->gmem[g] >= 0) ctl->glist[n_g++] = g;
        ctl->n_g = n_g;
        for (int mb = 0; mb < M; ++mb) busy += ctl->status[mb] != ST_EMPTY;
        ctl->exit_flag = (busy == 0 && ctl->queue_done) ? 1 : 0;
    }

    // PROF: thread 0 accumulates clock64() per phase (env, manage, layers 0..) and the live-group count, and adds them
    // to a.prof at exit — a development aid behind LL_FAST_PROF, compiled out of the production instantiation.
    template <int E, bool PROF>
    __global__ void __launch_bounds__(384) ll_episode_fast_kernel(const __grid_constant__ ll_env<float> env, const __grid_constant__ fast_args a)
    {
        extern __shared__ __align__(128) unsigned char smem_raw[];
        __shared__ fast_layer s_layer[LL_MAX_LAYERS]; // indexed reads of kernel parameters are slow: keep the table in shared memory
        const int tid = threadIdx.x, threads = a.threads;
        pool_ctl* ctl = reinterpret_cast<pool_ctl*>(smem_raw + a.off_ctl);
        float* W0     = reinterpret_cast<float*>(smem_raw + a.off_w);
        float* ACT    = reinterpret_cast<float*>(smem_raw + a.off_act);
        slot_view sv(smem_raw + a.off_slot, a.S);
        double* res_fit = reinterpret_cast<double*>(smem_raw + a.off_res);
        int* res_meta   = reinterpret_cast<int*>(res_fit + (a.res_in_smem ? a.M * a.n_layouts : 0));
        float* A0       = ACT + a.layer[0].in_off;
        const int n_layers = a.n_layers;

        for (int i = tid; i < n_layers; i += threads) s_layer[i] = a.layer[i];
        for (int i = tid; i < (a.off_slot - a.off_act) / 4; i += threads) ACT[i] = 0.f;
        for (int i = tid; i < a.S; i += threads) sv.active[i] = 0;
        if (tid == 0)
        {
            for (int mb = 0; mb < a.M; ++mb)
            {
                ll_mbar_init(&ctl->bar[mb], 1);
                ctl->status[mb] = ST_EMPTY, ctl->parity[mb] = 0, ctl->inflight[mb] = 0, ctl->next_layout[mb] = 0, ctl->age[mb] = 0;
            }
            for (int g = 0; g < a.G; ++g) ctl->gmem[g] = -1;
            ctl->age_counter = 0, ctl->dirty = 1, ctl->n_g = 0, ctl->queue_done = 0, ctl->exit_flag = 0, ctl->n_init = 0;
        }
        __syncthreads();

        long long pc[4 + LL_MAX_LAYERS] = {0}, t0 = 0;
        if (PROF && tid == 0) t0 = clock64();
#define LL_PROF_MARK(i)                                                                                              \
    if (PROF && tid == 0)                                                                                            \
    {                                                                                                                \
        const long long t1 = clock64();                                                                              \
        pc[i] += t1 - t0, t0 = t1;                                                                                   \
    }
        for (;;)
        {
            int n_g = ctl->n_g;
            if (PROF && tid == 0) pc[2] += 1, pc[3] += n_g;
            // ---- light phase: a quad of lanes per slot of the live groups (output layer -> env -> first layer)
            for (int i = tid; i < n_g * E * 4; i += threads)
            {
                const int q = i >> 2, g = ctl->glist[q / E], s = g * E + q % E, mb = ctl->gmem[g];
                if (sv.active[s])
                    light_slot(a, s_layer[0], s_layer[n_layers - 1], env, sv, W0 + (size_t)mb * a.w_stride, ACT, ctl, res_fit, res_meta, mb, s, tid & 3,
                               0xFu << (tid & 28));
            }
            const int ev = __syncthreads_or(ctl->dirty != 0);
            LL_PROF_MARK(0)
            if (ev)
            {
                // ---- rare path: some episode ended without a successor
                if (tid == 0) pool_manage<E>(a, env, ctl, sv, A0, W0, res_fit, res_meta);
                __syncthreads();
                if (ctl->exit_flag) break;
                for (int mb = 0; mb < a.M; ++mb)
                    if (ctl->status[mb] == ST_LOADING) ll_mbar_wait(&ctl->bar[mb], (uint32_t)ctl->parity[mb]);
                for (int i = tid; i < ctl->n_init; i += threads) slot_begin(a, env, sv, A0, ctl->init_slot[i], ctl->init_layout[i]);
                __syncthreads();
                if (tid == 0)
                    for (int mb = 0; mb < a.M; ++mb)
                        if (ctl->status[mb] == ST_LOADING) ctl->status[mb] = ST_ACTIVE, ctl->parity[mb] ^= 1;
                n_g = ctl->n_g;
                // slots may have moved or (re)started: rebuild the first layer's activations of every live slot
                for (int i = tid; i < n_g * E * 4; i += threads)
                {
                    const int q = i >> 2, g = ctl->glist[q / E], s = g * E + q % E;
                    if (sv.active[s])
                    {
                        float st[5];
#pragma unroll
                        for (int k = 0; k < 5; ++k) st[k] = A0[k * a.S + s];
                        first_layer_quad(a, s_layer[0], W0 + (size_t)ctl->gmem[g] * a.w_stride, ACT, st, s, tid & 3);
                    }
                }
                __syncthreads();
                LL_PROF_MARK(1)
            }
            // ---- forward pass: one phase per hidden-to-hidden layer (the first and the output layer belong to the light phase)
            for (int l = 1; l < n_layers - 1; ++l)
            {
                const fast_layer ly = s_layer[l];
                layer_dispatch<E>(a, ly, ctl, W0, ACT, tid, n_g);
                __syncthreads();
                LL_PROF_MARK(4 + l)
            }
        }
#undef LL_PROF_MARK
        if (PROF && tid == 0 && a.prof)
            for (int i = 0; i < 4 + n_layers; ++i) atomicAdd(&a.prof[i], (unsigned long long)pc[i]);
    }

    int pick_E(int L)
    { // slots per group (= episode rows of a register tile): the candidate with the best utilisation of a member's groups
        const int cand[5] = {8, 5, 4, 2, 1};
        int best = 1;
        double best_u = -1.0;
        for (int E : cand)
        {
            const int groups = (L + E - 1) / E;
            const double u   = (double)L / ((double)groups * E);
            if (u > best_u + 1e-9) best_u = u, best = E;
        }
        return best;
    }

    int env_int(const char* name, int fallback)
    {
        const char* v = getenv(name);
        return v && *v ? atoi(v) : fallback;
    }

    template <int E> int launch_fast_t(ll_ctx* ctx, const ll_episode_args& h)
    {
        const ll_topo& tp = ctx->topo;
        fast_args a{};
        a.params = static_cast<const float*>(h.params), a.layouts = static_cast<const ll_layout<float>*>(h.layouts);
        a.n_layouts = h.n_layouts, a.max_steps = h.max_steps, a.n_members = h.n_members;
        a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps, a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit, a.queue = h.queue;
        a.n_layers = tp.n_layers, a.act = tp.activation, a.p_dev = tp.p_dev;
        const int L   = h.n_layouts;
        a.res_in_smem = L <= MAX_RES ? 1 : 0;
        if (!a.res_in_smem && !h.ep_steps) return ll_fail(ctx, LL_EINVAL, "fast episode kernel: more than 256 layouts need the per-episode buffers");
        a.w_stride = (tp.p_dev + 31) & ~31;

        int max_smem = 0;
        LL_CUDA(ctx, cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device));
        int nog_max = 1, floats_per_slot = tp.out[tp.n_layers - 1], even_w = 0, odd_w = 0;
        for (int l = 0; l < tp.n_layers; ++l)
        {
            nog_max = std::max(nog_max, tp.out_pad[l] / FO);
            ((l & 1) ? odd_w : even_w) = std::max((l & 1) ? odd_w : even_w, tp.in[l]);
        }
        floats_per_slot += even_w + odd_w;
        const int gpm         = (L + E - 1) / E; // groups one member can keep busy
        const int group_bytes = E * (floats_per_slot * 4 + slot_view::bytes(1));
        const int member_bytes = a.w_stride * 4 + (a.res_in_smem ? L * 12 : 0);
        const int fixed_bytes  = (int)sizeof(pool_ctl) + 1024;
        const int g_cap        = std::min(MAX_G, 384 / nog_max > 0 ? 384 / nog_max : 1);
        // geometry: with `ctas` CTAs per SM, as many resident members as leave room for a pool of >= one member's groups
        int M = 0, G = 0;
        for (int ctas = 2; ctas >= 1 && M == 0; --ctas)
        {
            const int budget = max_smem / ctas - 1024;
            for (int m = std::min<long long>(MAX_M, (long long)h.n_members); m >= 1; --m)
            {
                const int room = budget - fixed_bytes - m * member_bytes;
                if (room < 0) continue;
                int g = std::min({room / group_bytes, g_cap, m * gpm});
                if (g >= std::min(gpm, g_cap) && (m >= 2 || ctas == 1))
                {
                    M = m, G = g;
                    break;
                }
            }
        }
        if (M == 0) return ll_fail(ctx, LL_EINVAL, "fast episode kernel: topology does not fit in shared memory");
        M = env_int("LL_FAST_M", M), G = env_int("LL_FAST_G", G); // tuning knobs for experiments
        if (M < 1 || M > MAX_M || G < 1 || G > MAX_G) return ll_fail(ctx, LL_EINVAL, "fast episode kernel: bad LL_FAST_M / LL_FAST_G");
        a.M = M, a.G = G, a.S = G * E;
        int threads = std::min(384, G * nog_max);
        threads     = std::max(32, (threads + 31) & ~31);
        a.threads   = threads;

        const int even = even_w * a.S, odd = odd_w * a.S;
        for (int l = 0; l < tp.n_layers; ++l)
        {
            fast_layer& ly = a.layer[l];
            ly.in = tp.in[l], ly.out = tp.out[l], ly.out_pad = tp.out_pad[l], ly.w_off = tp.w_off[l], ly.b_off = tp.b_off[l];
            ly.last = l == tp.n_layers - 1;
            ly.nog  = tp.out_pad[l] / FO;
            int ks  = 1; // k-slices per tile: fill the CTA's threads, keep >= 8 rows per slice
            while (ks < 4 && G * ly.nog * ks * 2 <= threads && ly.in / (ks * 2) >= 8) ks *= 2;
            ly.ks      = ks;
            ly.in_off  = (l & 1) ? even : 0;
            ly.out_off = ly.last ? even + odd : ((l & 1) ? 0 : even);
        }
        const int act_elems = even + odd + tp.out[tp.n_layers - 1] * a.S;
        auto up             = [](int v) { return (v + 127) & ~127; };
        a.off_ctl           = 0;
        a.off_w             = up((int)sizeof(pool_ctl));
        a.off_act           = up(a.off_w + M * a.w_stride * 4);
        a.off_slot          = up(a.off_act + act_elems * 4);
        a.off_res           = up(a.off_slot + slot_view::bytes(a.S));
        const size_t smem   = (size_t)up(a.off_res + (a.res_in_smem ? M * L * 12 : 0) + 16);
        if (smem > (size_t)max_smem) return ll_fail(ctx, LL_EINVAL, "fast episode kernel: geometry needs " + std::to_string(smem) + " B of shared memory");

        const bool prof = env_int("LL_FAST_PROF", 0) != 0;
        auto kern       = prof ? ll_episode_fast_kernel<E, true> : ll_episode_fast_kernel<E, false>;
        if (prof)
        {
            LL_CUDA(ctx, cudaMalloc(&a.prof, 64 * 8));
            LL_CUDA(ctx, cudaMemsetAsync(a.prof, 0, 64 * 8, ctx->stream));
        }
        LL_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;
        LL_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem));
        if (per_sm < 1) return ll_fail(ctx, LL_EINVAL, "fast episode kernel does not fit on an SM");
        const unsigned long long blocks_needed = (h.n_members + M - 1) / M;
        unsigned int grid = (unsigned int)std::min<unsigned long long>((unsigned long long)per_sm * ctx->sm_count, blocks_needed);
        if (grid < 1) grid = 1;
        ll_env<float> env;
        ll_make_env<float>(ctx->env_cfg, 1, &env);
        LL_CUDA(ctx, cudaMemsetAsync(h.queue, 0, sizeof(unsigned int), ctx->stream));
        kern<<<grid, threads, smem, ctx->stream>>>(env, a);
        LL_CUDA(ctx, cudaGetLastError());
        if (prof)
        {
            unsigned long long pc[64];
            LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            LL_CUDA(ctx, cudaMemcpy(pc, a.prof, sizeof(pc), cudaMemcpyDeviceToHost));
            cudaFree(a.prof);
            double tot = 0;
            for (int i = 0; i < 4 + tp.n_layers; ++i)
                if (i != 2 && i != 3) tot += (double)pc[i];
            fprintf(stderr, "[ll_fast_prof] grid %u x %d thr, M %d G %d E %d smem %zu B, CTA-steps %llu, mean live groups %.2f, cycles/CTA-step %.0f |", grid,
                    threads, M, G, E, smem, pc[2], pc[2] ? (double)pc[3] / pc[2] : 0.0, pc[2] ? tot / pc[2] : 0.0);
            fprintf(stderr, " env %.1f%% manage %.1f%%", 100.0 * pc[0] / tot, 100.0 * pc[1] / tot);
            for (int l = 0; l < tp.n_layers; ++l) fprintf(stderr, " L%d(ks%d) %.1f%%", l, a.layer[l].ks, 100.0 * pc[4 + l] / tot);
            fprintf(stderr, "\n");
        }
        return LL_OK;
    }
} // namespace

// The pool kernel needs weight reuse across layouts (>= 4 of them) and the shape the light phase is written for: 5 inputs,
// at least one hidden layer, at most 8 outputs (the reference's networks are 5-h...-4, src/nn_engine_trainer.cpp:45-48).
bool ll_fast_applicable(const ll_ctx* ctx, int n_layouts)
{
    const ll_topo& tp = ctx->topo;
    return n_layouts >= 4 && tp.n_layers >= 2 && tp.in[0] == 5 && tp.out[tp.n_layers - 1] <= FO;
}

// Fast-mode launcher (evaluations with >= 4 layouts; single-layout evaluations and traces stay on the lane-per-episode
// kernel of ll_episode.cu, see ll_launch_episode).
int ll_launch_episode_fast(ll_ctx* ctx, const ll_episode_args& a)
{
    switch (pick_E(a.n_layouts))
    {
    case 8: return launch_fast_t<8>(ctx, a);
    case 5: return launch_fast_t<5>(ctx, a);
    case 4: return launch_fast_t<4>(ctx, a);
    case 2: return launch_fast_t<2>(ctx, a);
    default: return launch_fast_t<1>(ctx, a);
    }
}
