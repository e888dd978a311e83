// ll_episode_fast.cu — K1 fast (fp32, LL_MODE_FAST): the throughput variant of the fused episode kernel.
//
// Replaces the same reference code as ll_episode.cu (train_generation / sim_task src/nn_engine_trainer.cpp:443-539,
// GameLogicCpp<T> src/game_logic.cpp:124-388, the ma-libs feedforward call :501-502), with free summation order,
// fused multiply-add and tanh.approx.f32 — judged on outcome rates, not on bit patterns (tests/test_gpu_parity.py).
//
// Why a different mapping: a lane-per-episode GEMV delivers one shared-memory word per FMA and
// shared memory returns 32 words/clk/SM against 128 FMA lanes/clk/SM, i.e. <= 25 % of the FP32 pipe.  All L layouts of
// a member share its weights, so one forward pass for a member is a small GEMM  D[S x out] = A[S x in] * W^T[in x out]
// (S = concurrent episode slots).  A TEAM of threads owns one member at a time; each thread computes a register tile of
// E episodes x 8 outputs (E*8 FMAs per E+8 words fetched), layers wide in K are split over k-slices and reduced
// through shared memory.  Per step: env phase (one thread per slot: argmax -> lander update -> penalties -> next
// state) then one phase per layer, separated by CTA barriers.  A member's row is staged once by a TMA bulk copy
// (cp.async.bulk + mbarrier); teams pull members from a global queue, slots pull the member's layouts from a
// shared-memory counter, so episode lengths of 9..max_steps steps balance dynamically.
#include "ll_env.cuh"

namespace
{
    constexpr int FO          = 8;   // outputs per thread tile
    constexpr int MAX_THREADS = 256; // per CTA
    constexpr int MAX_RES     = 256; // layouts whose per-episode results are kept in shared memory

    struct fast_layer
    {
        int in, out, out_pad, w_off, b_off; // topology (element offsets inside a member row)
        int nog, tiles, ks, klen;           // output groups, (episode group x output group) tiles, k-slices, k per slice
        int in_off, out_off;                // element offsets of the activation buffers inside the team's region
        int last;                           // output layer: its activations go to the Z buffer of the env phase
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
        int n_layers, act, p_dev;
        int TS, NT, Ge, S;                                     // threads per team, teams per CTA, episode groups, slots = E * Ge
        int off_w, off_act, off_slot, off_res, team_bytes;     // byte offsets inside a team's shared-memory region
        int res_in_smem;
        fast_layer layer[LL_MAX_LAYERS];
    };

    // per-team control words (ints at byte 16 of the team region)
    enum
    {
        C_NEXT_LAYOUT = 0,
        C_N_ACTIVE    = 1,
        C_HAS_MEMBER  = 2,
        C_MEMBER_LO   = 3,
        C_MEMBER_HI   = 4,
        C_DIRTY       = 5
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

    // One (episode group, output group, k-slice) register tile: acc[e][j] = sum_k in[k][ge*E + e] * W[k][col(j)], where the
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

    // One layer for one team.  Slots are kept compacted (active slots are [0, n_act)) and a thread's E slots are contiguous
    // (slot = ge*E + e), so tiles of episode groups past n_act are skipped.  KS k-slices of a tile (rows k = KS*i + ksi)
    // live in one warp and are reduced by shuffles; the owners add the bias, apply the activation and store
    // out[o][slot].  The output layer stores into the Z buffer the env phase takes its argmax from.
    template <int E, int KS>
    __device__ __forceinline__ void layer_phase(const fast_args& a, const fast_layer& ly, const float* __restrict__ W, float* __restrict__ ACT,
                                                int tt, int n_act)
    {
        constexpr int TPW = 32 / KS;
        const int S = a.S;
        const float* in = ACT + ly.in_off;
        float* outb     = ACT + ly.out_off;
        int tile, stride, ksi = 0;
        unsigned mask = 0xffffffffu;
        if constexpr (KS == 1) tile = tt, stride = a.TS;
        else
        {
            const int lane = tt & 31, q = lane & (TPW - 1);
            ksi    = lane / TPW;
            tile   = (tt >> 5) * TPW + q;
            stride = (a.TS >> 5) * TPW;
            mask   = 0;
#pragma unroll
            for (int i = 0; i < KS; ++i) mask |= 1u << (i * TPW + q);
        }
        for (; tile < ly.tiles; tile += stride)
        {
            const int ge = tile / ly.nog, go = tile - ge * ly.nog; // episode group slowest: active tiles fill whole warps
            if (ge * E >= n_act) break;
            const int kl   = (ly.in - ksi + KS - 1) / KS;
            const int half = ly.out_pad >> 1;
            float acc[E][FO];
            gemm_tile<E>(W + ly.w_off + ksi * ly.out_pad + go * 4, in + ksi * S + ge * E, kl, KS * ly.out_pad, KS * S, half, acc);
            float res[E][FO / KS];
            const int j0   = reduce_scatter<E, KS>(acc, ksi, mask, res);
            const int o0   = go * 4 + (j0 & 3) + ((j0 & 4) ? half : 0); // first owned column (the owned ones are consecutive)
            const float* b = W + ly.b_off + o0;
            float* out     = outb + o0 * S + ge * E;
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
                        for (int e = 0; e < E; ++e) outb[o * S + ge * E + e] = fast_act(res[e][j] + bj, a.act);
                    }
                }
            }
            else
            {
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
    __device__ __forceinline__ void layer_dispatch(const fast_args& a, const fast_layer& ly, const float* __restrict__ W, float* __restrict__ ACT,
                                                   int tt, int n_act)
    {
        if (ly.ks == 4) layer_phase<E, 4>(a, ly, W, ACT, tt, n_act);
        else if (ly.ks == 2) layer_phase<E, 2>(a, ly, W, ACT, tt, n_act);
        else layer_phase<E, 1>(a, ly, W, ACT, tt, n_act);
    }

    // begin an episode of layout l in slot s (reset(spad_x1, lpad_x1), src/game_logic.cpp:143-154) and publish its state
    __device__ __forceinline__ void slot_begin(const fast_args& a, const ll_env<float>& env, slot_view& sv, float* A0, int s, int l)
    {
        const ll_layout<float> lay = a.layouts[l];
        sv.layout[s] = l | (lay.dir << 30);
        sv.cx[s] = lay.center_x, sv.lpad[s] = lay.lpad_x1, sv.lpad_r[s] = lay.lpad_right;
        ll_lander<float> p{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
        sv.x[s] = p.x, sv.y[s] = p.y, sv.vx[s] = p.vx, sv.vy[s] = p.vy, sv.fuel[s] = p.fuel;
        sv.steps[s] = 0, sv.acc[s] = 0.0, sv.active[s] = 1;
        ll_get_state<float, LL_MODE_FAST>(env, p, lay.center_x, A0 + s, a.S);
    }

    // env phase of one active slot: output layer reduce -> argmax -> update -> penalties -> next state / next layout
    __device__ __forceinline__ void slot_step(const fast_args& a, const ll_env<float>& env, slot_view& sv, const float* __restrict__ W,
                                              float* __restrict__ ACT, int* ctl, double* res_fit, int* res_meta, unsigned long long member, int s)
    {
        const fast_layer& ly = a.layer[a.n_layers - 1];
        const float* z       = ACT + ly.out_off + s;
        int action  = 0;
        float vbest = z[0];
        for (int o = 1; o < ly.out; ++o)
        {
            const float v = z[o * a.S];
            if (v > vbest) vbest = v, action = o;
        }
        ll_lander<float> p{sv.x[s], sv.y[s], sv.vx[s], sv.vy[s], sv.fuel[s]};
        const float cx = sv.cx[s];
        int steps      = sv.steps[s];
        bool landed, success;
        float dist;
        const float sp = ll_env_step<float, LL_MODE_FAST>(env, p, action, steps == 0, cx, sv.lpad[s], sv.lpad_r[s], &landed, &success, &dist);
        const double acc = sv.acc[s] + (double)sp;
        ++steps;
        float* A0 = ACT + a.layer[0].in_off;
        if (landed || steps >= a.max_steps)
        {
            const float term = ll_terminal_penalty<float, LL_MODE_FAST>(env, p, steps, landed, success, dist);
            const double fit = (double)term + acc;
            const int lw     = sv.layout[s];
            const int layout = lw & 0x3fffffff;
            const int meta   = steps | (landed ? (1 << 28) : 0) | ((landed && !success) ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
            if (a.res_in_smem) res_fit[layout] = fit, res_meta[layout] = meta;
            if (a.ep_steps)
            {
                a.ep_steps[member * (size_t)a.n_layouts + layout] = meta;
                a.ep_fit[member * (size_t)a.n_layouts + layout]   = fit;
            }
            const int nl = atomicAdd(&ctl[C_NEXT_LAYOUT], 1);
            if (nl < a.n_layouts) slot_begin(a, env, sv, A0, s, nl);
            else
            {
                sv.active[s] = 0;
                ctl[C_DIRTY] = 1; // compacted by the team leader before the next forward pass
            }
            return;
        }
        sv.x[s] = p.x, sv.y[s] = p.y, sv.vx[s] = p.vx, sv.vy[s] = p.vy, sv.fuel[s] = p.fuel;
        sv.steps[s] = steps, sv.acc[s] = acc;
        ll_get_state<float, LL_MODE_FAST>(env, p, cx, A0 + s, a.S);
    }

    // Keep the active slots packed in [0, n_act): holes left by finished episodes are filled from the top.  Rare (at most
    // once per finished episode), done serially by the team leader; the moved slot's input column is rebuilt.
    __device__ void team_compact(const fast_args& a, const ll_env<float>& env, slot_view& sv, float* A0, int* ctl)
    {
        int i = 0, j = ctl[C_N_ACTIVE] - 1;
        for (;;)
        {
            while (i <= j && sv.active[i]) ++i;
            while (j >= i && !sv.active[j]) --j;
            if (i >= j) break;
            sv.acc[i] = sv.acc[j];
            sv.x[i] = sv.x[j], sv.y[i] = sv.y[j], sv.vx[i] = sv.vx[j], sv.vy[i] = sv.vy[j], sv.fuel[i] = sv.fuel[j];
            sv.cx[i] = sv.cx[j], sv.lpad[i] = sv.lpad[j], sv.lpad_r[i] = sv.lpad_r[j];
            sv.steps[i] = sv.steps[j], sv.layout[i] = sv.layout[j];
            sv.active[i] = 1, sv.active[j] = 0;
            for (int q = 0; q < a.layer[0].in; ++q) A0[q * a.S + i] = A0[q * a.S + j];
        }
        ctl[C_N_ACTIVE] = j + 1; // j is the last active slot (or -1)
        ctl[C_DIRTY]    = 0;
    }

    // per-member totals in layout order (src/nn_engine_trainer.cpp:518-530), by the team leader
    __device__ void member_finalize(const fast_args& a, const ll_env<float>& env, const double* res_fit, const int* res_meta,
                                    unsigned long long member)
    {
        double total = 0.0, total_steps = 0.0;
        float bonus0 = 0.f, bonus1 = 0.f;
        for (int l = 0; l < a.n_layouts; ++l)
        {
            const int meta   = a.res_in_smem ? res_meta[l] : a.ep_steps[member * (size_t)a.n_layouts + l];
            const double fit = a.res_in_smem ? res_fit[l] : a.ep_fit[member * (size_t)a.n_layouts + l];
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

    template <int E>
    __global__ void __launch_bounds__(MAX_THREADS) ll_episode_fast_kernel(const __grid_constant__ ll_env<float> env, const __grid_constant__ fast_args a)
    {
        extern __shared__ __align__(128) unsigned char smem_raw[];
        __shared__ fast_layer s_layer[LL_MAX_LAYERS]; // the layer table in shared memory: indexed reads of kernel parameters are slow
        const int tid = threadIdx.x, team = tid / a.TS, tt = tid - team * a.TS;
        unsigned char* tb = smem_raw + (size_t)team * a.team_bytes;
        uint64_t* bar     = reinterpret_cast<uint64_t*>(tb);
        int* ctl          = reinterpret_cast<int*>(tb + 16);
        float* W          = reinterpret_cast<float*>(tb + a.off_w);
        float* ACT        = reinterpret_cast<float*>(tb + a.off_act);
        slot_view sv(tb + a.off_slot, a.S);
        double* res_fit = reinterpret_cast<double*>(tb + a.off_res);
        int* res_meta   = reinterpret_cast<int*>(res_fit + (a.res_in_smem ? a.n_layouts : 0));
        float* A0       = ACT + a.layer[0].in_off;
        const uint32_t row_bytes = (uint32_t)a.p_dev * sizeof(float);
        const int n_layers       = a.n_layers;

        for (int i = tid; i < n_layers; i += blockDim.x) s_layer[i] = a.layer[i];
        if (tt == 0)
        {
            ll_mbar_init(bar, 1);
            ctl[C_N_ACTIVE] = 0, ctl[C_HAS_MEMBER] = 0, ctl[C_NEXT_LAYOUT] = 0, ctl[C_DIRTY] = 0;
        }
        __syncthreads();
        uint32_t phase            = 0;
        bool running              = false; // the team holds a member whose episodes are in flight
        bool finished             = false; // the queue is exhausted for this team
        unsigned long long member = 0;
        int n_act                 = 0;

        for (;;)
        {
            // ---- env phase: one thread per active slot
            if (running)
                for (int s = tt; s < n_act; s += a.TS) slot_step(a, env, sv, W, ACT, ctl, res_fit, res_meta, member, s);
            __syncthreads();
            const bool dirty = running && ctl[C_DIRTY] != 0;
            if (__syncthreads_or(dirty || (!finished && !running)))
            {
                // ---- rare path: compaction after finished episodes; member hand-over when none is left
                if (dirty && tt == 0) team_compact(a, env, sv, A0, ctl);
                __syncthreads();
                if (running) n_act = ctl[C_N_ACTIVE];
                const bool need = !finished && (!running || n_act == 0);
                if (need && tt == 0)
                {
                    if (running) member_finalize(a, env, res_fit, res_meta, member);
                    const unsigned long long m = (unsigned long long)atomicAdd(a.queue, 1u);
                    if (m < a.n_members)
                    {
                        ctl[C_HAS_MEMBER] = 1, ctl[C_MEMBER_LO] = (int)(unsigned int)m, ctl[C_MEMBER_HI] = (int)(unsigned int)(m >> 32);
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        ll_mbar_expect_tx(bar, row_bytes);
                        ll_bulk_g2s(W, a.params + m * (size_t)a.p_dev, row_bytes, bar);
                    }
                    else ctl[C_HAS_MEMBER] = 0;
                }
                __syncthreads();
                if (need)
                {
                    if (ctl[C_HAS_MEMBER])
                    {
                        member  = (unsigned long long)(unsigned int)ctl[C_MEMBER_LO] | ((unsigned long long)(unsigned int)ctl[C_MEMBER_HI] << 32);
                        running = true;
                        ll_mbar_wait(bar, phase);
                        phase ^= 1u;
                    }
                    else running = false, finished = true, n_act = 0;
                }
                __syncthreads(); // every thread of the team has read the control words before the leader rewrites them
                if (need && running)
                {
                    n_act = min(a.S, a.n_layouts);
                    for (int s = tt; s < a.S; s += a.TS)
                    {
                        if (s < n_act) slot_begin(a, env, sv, A0, s, s);
                        else sv.active[s] = 0;
                    }
                    if (tt == 0) ctl[C_NEXT_LAYOUT] = n_act, ctl[C_N_ACTIVE] = n_act, ctl[C_DIRTY] = 0;
                }
                if (__syncthreads_and(finished)) break;
            }
            // ---- forward pass: one phase per layer
            for (int l = 0; l < n_layers; ++l)
            {
                const fast_layer ly = s_layer[l];
                if (running) layer_dispatch<E>(a, ly, W, ACT, tt, n_act);
                __syncthreads();
            }
        }
    }

    int pick_E(int L)
    { // episode slots per thread: the candidate with the best slot utilisation, larger tiles on ties
        const int cand[5] = {8, 5, 4, 2, 1};
        int best = 1;
        double best_u = -1.0;
        for (int E : cand)
        {
            int Ge = (L + E - 1) / E;
            if (E * Ge > 64) Ge = 64 / E;
            const int S = E * Ge, rounds = (L + S - 1) / S;
            const double u = (double)L / ((double)rounds * S);
            if (u > best_u + 1e-9) best_u = u, best = E;
        }
        return best;
    }

    template <int E> int launch_fast_t(ll_ctx* ctx, const ll_episode_args& h)
    {
        const ll_topo& tp = ctx->topo;
        fast_args a{};
        a.params = static_cast<const float*>(h.params), a.layouts = static_cast<const ll_layout<float>*>(h.layouts);
        a.n_layouts = h.n_layouts, a.max_steps = h.max_steps, a.n_members = h.n_members;
        a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps, a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit, a.queue = h.queue;
        a.n_layers = tp.n_layers, a.act = tp.activation, a.p_dev = tp.p_dev;
        const int L = h.n_layouts;
        int Ge      = (L + E - 1) / E;
        if (E * Ge > 64) Ge = 64 / E;
        a.Ge = Ge, a.S = E * Ge;
        a.res_in_smem = L <= MAX_RES ? 1 : 0;
        if (!a.res_in_smem && !h.ep_steps) return ll_fail(ctx, LL_EINVAL, "fast episode kernel: more than 256 layouts need the per-episode buffers");

        int nog_max = 1;
        for (int l = 0; l < tp.n_layers; ++l) nog_max = std::max(nog_max, tp.out_pad[l] / FO);
        int TS = Ge * nog_max;
        if (TS > MAX_THREADS) TS = MAX_THREADS;
        TS   = TS >= 32 ? ((TS + 31) & ~31) : (TS <= 8 ? 8 : (TS <= 16 ? 16 : 32));
        if (TS > MAX_THREADS) TS = MAX_THREADS;
        a.TS = TS;

        // activation buffers: even / odd ping-pong sized by the widest activation of each parity, then the output layer's Z
        int even = 0, odd = 0;
        for (int l = 0; l < tp.n_layers; ++l) ((l & 1) ? odd : even) = std::max((l & 1) ? odd : even, tp.in[l] * a.S);
        for (int l = 0; l < tp.n_layers; ++l)
        {
            fast_layer& ly = a.layer[l];
            ly.in = tp.in[l], ly.out = tp.out[l], ly.out_pad = tp.out_pad[l], ly.w_off = tp.w_off[l], ly.b_off = tp.b_off[l];
            ly.last  = l == tp.n_layers - 1;
            ly.nog   = tp.out_pad[l] / FO;
            ly.tiles = Ge * ly.nog;
            int ks   = 1; // k-slices per tile: fill the team's threads, keep >= 8 rows per slice, slices of a tile in one warp
            while (ks < 4 && (TS & 31) == 0 && ly.tiles * ks * 2 <= TS && ly.in / (ks * 2) >= 8) ks *= 2;
            ly.ks      = ks;
            ly.klen    = (ly.in + ks - 1) / ks;
            ly.in_off  = (l & 1) ? even : 0;
            ly.out_off = ly.last ? even + odd : ((l & 1) ? 0 : even);
        }
        const int act_elems = even + odd + tp.out[tp.n_layers - 1] * a.S;
        auto up             = [](int v) { return (v + 127) & ~127; };
        a.off_w             = 128;
        a.off_act           = up(a.off_w + tp.p_dev * 4);
        a.off_slot          = up(a.off_act + act_elems * 4);
        a.off_res           = up(a.off_slot + slot_view::bytes(a.S));
        a.team_bytes        = up(a.off_res + (a.res_in_smem ? L * 12 : 0) + 16);

        int max_smem = 0;
        LL_CUDA(ctx, cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device));
        if (a.team_bytes > max_smem) return ll_fail(ctx, LL_EINVAL, "fast episode kernel: a member team needs " + std::to_string(a.team_bytes) + " B of shared memory");
        // teams per CTA: fill up to 256 threads, but keep at least two CTAs per SM when the topology allows it
        int NT = std::max(1, MAX_THREADS / TS);
        while (NT > 1 && (size_t)NT * a.team_bytes > (size_t)max_smem / 2) --NT;
        const unsigned long long teams_needed = h.n_members;
        if ((unsigned long long)NT > teams_needed) NT = (int)teams_needed;
        a.NT              = NT;
        const int threads = NT * TS;
        const size_t smem = (size_t)NT * a.team_bytes;
        auto kern         = ll_episode_fast_kernel<E>;
        LL_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;
        LL_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem));
        if (per_sm < 1) return ll_fail(ctx, LL_EINVAL, "fast episode kernel does not fit on an SM");
        unsigned long long blocks_needed = (h.n_members + NT - 1) / NT;
        unsigned int grid                = (unsigned int)std::min<unsigned long long>((unsigned long long)per_sm * ctx->sm_count, blocks_needed);
        if (grid < 1) grid = 1;
        ll_env<float> env;
        ll_make_env<float>(ctx->env_cfg, 1, &env);
        LL_CUDA(ctx, cudaMemsetAsync(h.queue, 0, sizeof(unsigned int), ctx->stream));
        kern<<<grid, threads, smem, ctx->stream>>>(env, a);
        LL_CUDA(ctx, cudaGetLastError());
        return LL_OK;
    }
} // namespace

// Fast-mode launcher; returns LL_OK after enqueueing, or a negative value when the team kernel does not apply
// (single-layout evaluations and traces stay on the lane-per-episode kernel of ll_episode.cu).
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
