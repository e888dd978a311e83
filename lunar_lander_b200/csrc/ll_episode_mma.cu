// ll_episode_mma.cu — K1 fast, tensor-core variant (fp32 via 3xTF32 mma.sync, LL_MODE_FAST, >= 9 layouts).
//
// Replaces the same reference code as ll_episode.cu / ll_episode_fast.cu: train_generation / sim_task
// src/nn_engine_trainer.cpp:443-539, GameLogicCpp<T> src/game_logic.cpp:124-388 and the ma-libs feedforward call
// :501-502.  BASELINE.json north_star (a): "tensor cores (mma) are used only for the multi-start case, where each
// individual's forward pass becomes a real small dense GEMM".
//
// Mapping.  All layouts of a member share its weights, so one forward pass over S concurrent episode
// slots is D[S x out] = A[S x in] * W^T[in x out] per layer.  ONE WARP owns one member at a time and S = 16*MT slots:
//   * the slots are the M dimension of mma.sync.m16n8k8 (tf32 operands, fp32 accumulate); the activations never leave
//     the register file: the C fragment of layer l (rows g / g+8, columns 2t, 2t+1) IS the A fragment of layer l+1 once
//     the k index of that layer is enumerated as (t -> 2t, t+4 -> 2t+1), which costs nothing because the B fragments
//     (weights) are fetched as consecutive-k float2 pairs from shared memory in exactly that order;
//   * the member's weights are staged once per member into shared memory as W[out][in] with an XOR swizzle of the 8-float
//     k blocks that makes every LDS.64 fragment fetch bank-conflict free (no padding: 35.6 KB for 32_128_32, 6 members
//     resident per SM);
//   * fp32 accuracy on the tf32 tensor pipe: x = hi + lo with hi = x & 0xffffe000 (the 10-bit-mantissa part, exact) and
//     lo = x - hi (exact), D += A_lo*B_hi + A_hi*B_lo + A_hi*B_hi — the dropped lo*lo term is < 2^-20 relative;
//   * the lander of slot s lives in the registers of lane s; the argmax over the 4 outputs is two compares + one shuffle
//     in the quad that holds them, the action and the next state travel between the fragment layout and the slot lanes
//     by shuffles.  No barrier, no atomics and no shared-memory traffic other than the weight fetches in the step loop;
//   * a slot whose episode ends takes the member's next layout (ballot + prefix count); the warp moves on to the next
//     member of the global queue when all layouts are done.
// Layer sizes are template parameters (fragments are register arrays): the sample_runs topologies are instantiated,
// anything else stays on the FFMA kernels.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "ll_env.cuh"

namespace
{
    template <int... S> struct net_t
    {
        static constexpr int NS = sizeof...(S);
        static constexpr int NL = NS - 1;
        static __host__ __device__ constexpr int sz(int i)
        {
            constexpr int a[NS] = {S...};
            return a[i];
        }
        static __host__ __device__ constexpr int in8(int l) { return (sz(l) + 7) & ~7; }
        static __host__ __device__ constexpr int out8(int l) { return (sz(l + 1) + 7) & ~7; }
        // shared-memory (fragment-friendly) layout: per layer W[out8][in8] (swizzled) then b[out8]
        static __host__ __device__ constexpr int w_off(int l)
        {
            int o = 0;
            for (int i = 0; i < l; ++i) o += in8(i) * out8(i) + out8(i);
            return o;
        }
        static __host__ __device__ constexpr int b_off(int l) { return w_off(l) + in8(l) * out8(l); }
        static constexpr int total = w_off(NL);
        static __host__ __device__ constexpr bool ok()
        {
            for (int l = 0; l < NL; ++l)
            {
                const int kb = in8(l) / 8;
                if (!(kb == 1 || kb == 2 || kb % 4 == 0)) return false;
            }
            return sz(0) == 5 && sz(NL) <= 8;
        }
    };

    // XOR applied to the 8-float block index of row `o` so that rows o..o+3 land in four different 8-bank groups
    template <int KB> __device__ __forceinline__ int swz(int o) { return KB >= 4 ? (o & 3) : (KB == 2 ? ((o >> 1) & 1) : 0); }

    struct mma_args
    {
        const float* params;
        const ll_layout<float>* layouts;
        int n_layouts, max_steps;
        unsigned long long n_members;
        double *fitness, *lr, *steps;
        int* ep_steps;
        double* ep_fit;
        unsigned int* queue;
        int p_dev;
        int w_off[LL_MAX_LAYERS], b_off[LL_MAX_LAYERS], out_pad[LL_MAX_LAYERS]; // the HBM row layout (ll_topo)
    };

    __device__ __forceinline__ void mma_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1)
    {
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    __device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo)
    {
        hi = __float_as_uint(x) & 0xffffe000u;
        lo = __float_as_uint(x - __uint_as_float(hi));
    }

    // One layer for the warp's MT slot tiles.  a[mt][kt][4]: activations in C-fragment order ([0]=(g,2t) [1]=(g,2t+1)
    // [2]=(g+8,2t) [3]=(g+8,2t+1) of k block kt); c[mt][nt][4]: act(bias + sum), same order, for output block nt.
    //
    // Accumulators.  A dependent HMMA (same accumulator) cannot issue until its predecessor has left the tensor pipe, and
    // ptxas schedules the three HMMAs of one (kt, nt) tile next to each other (they share the freshly split B fragment).
    // The three terms therefore accumulate into three separate register tiles — the dependent ones are then a whole k block
    // apart — which are summed once per layer, small terms first.  Output blocks are processed in panels of PN <= 8 so that
    // the 3 * PN * MT accumulator tiles fit the register file; the 4-output layer (one block) also alternates by k parity.
    template <class N, int L, int MT, int EXP = 0>
    __device__ __forceinline__ void layer_mma(const float* __restrict__ sw, int g, int t, const float (&a)[MT][N::in8(L) / 8][4],
                                              float (&c)[MT][N::out8(L) / 8][4])
    {
        constexpr int IN8 = N::in8(L), KT = IN8 / 8, NT = N::out8(L) / 8;
        constexpr int PN = NT >= 8 ? 8 : NT;
        constexpr int KP = (PN * MT == 1 && KT >= 2) ? 2 : 1; // k-parity replicas
        static_assert(NT % PN == 0, "output blocks must fill whole panels");
        const float* W = sw + N::w_off(L) + g * IN8 + 2 * t;
        const float* B = sw + N::b_off(L) + 2 * t;
        int vr[4];
        const int sx = swz<KT>(g);
#pragma unroll
        for (int r = 0; r < 4; ++r) vr[r] = (r ^ sx) << 3;
#pragma unroll
        for (int np = 0; np < NT; np += PN)
        {
            float acc[KP][3][MT][PN][4]; // [.][0] A_lo*B_hi, [.][1] A_hi*B_lo, [.][2] bias + A_hi*B_hi
#pragma unroll
            for (int j = 0; j < PN; ++j)
            {
                const float2 b = *reinterpret_cast<const float2*>(B + (np + j) * 8);
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                    for (int p = 0; p < KP; ++p)
#pragma unroll
                        for (int r = 0; r < 3; ++r)
                        {
                            const bool main = p == 0 && r == 2;
                            acc[p][r][mt][j][0] = main ? b.x : 0.f, acc[p][r][mt][j][1] = main ? b.y : 0.f;
                            acc[p][r][mt][j][2] = main ? b.x : 0.f, acc[p][r][mt][j][3] = main ? b.y : 0.f;
                        }
            }
#pragma unroll
            for (int kt = 0; kt < KT; ++kt)
            {
                uint32_t ah[MT][4], al[MT][4];
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (EXP == 4 || EXP == 5) ah[mt][i] = __float_as_uint(a[mt][kt][i]), al[mt][i] = ah[mt][i] ^ 0x8000u;
                        else split_tf32(a[mt][kt][i], ah[mt][i], al[mt][i]);
                const int p = kt % KP;
#pragma unroll
                for (int j = 0; j < PN; ++j)
                {
                    float2 w;
                    if (EXP == 2) w = make_float2(__int_as_float(vr[kt & 3] + j), __int_as_float(vr[(kt + 1) & 3] + np));
                    else w = *reinterpret_cast<const float2*>(W + (np + j) * 8 * IN8 + (kt & ~3) * 8 + vr[kt & 3]);
                    uint32_t bh0, bl0, bh1, bl1;
                    if (EXP == 1 || EXP == 5) bh0 = __float_as_uint(w.x), bl0 = bh0 ^ 0x8000u, bh1 = __float_as_uint(w.y), bl1 = bh1 ^ 0x8000u;
                    else split_tf32(w.x, bh0, bl0), split_tf32(w.y, bh1, bl1);
                    // (Measured: dropping A_lo*B_hi halves the share of episodes whose step count equals the exact fp32
                    // path's, 0.98 -> 0.50, so all three terms stay.)
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt)
                    {
                        mma_tf32(acc[p][0][mt][j], al[mt][0], al[mt][2], al[mt][1], al[mt][3], bh0, bh1);
                        mma_tf32(acc[p][1][mt][j], ah[mt][0], ah[mt][2], ah[mt][1], ah[mt][3], bl0, bl1);
                        mma_tf32(acc[p][2][mt][j], ah[mt][0], ah[mt][2], ah[mt][1], ah[mt][3], bh0, bh1);
                    }
                }
            }
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int j = 0; j < PN; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                    {
                        float small = acc[0][0][mt][j][i] + acc[0][1][mt][j][i], big = acc[0][2][mt][j][i];
                        if (KP == 2) small += acc[1][0][mt][j][i] + acc[1][1][mt][j][i], big += acc[1][2][mt][j][i];
                        c[mt][np + j][i] = (EXP == 3 || EXP == 5) ? (big + small) * 0.01f : ll_tanh_approx(big + small);
                    }
        }
    }

    // layers L..NL-1; returns the output layer's fragments in `z`
    template <class N, int L, int MT, int EXP = 0> struct chain
    {
        static __device__ __forceinline__ void run(const float* __restrict__ sw, int g, int t, const float (&a)[MT][N::in8(L) / 8][4],
                                                   float (&z)[MT][4])
        {
            float c[MT][N::out8(L) / 8][4];
            layer_mma<N, L, MT, EXP>(sw, g, t, a, c);
            if constexpr (L == N::NL - 1)
            {
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                    for (int i = 0; i < 4; ++i) z[mt][i] = c[mt][0][i];
            }
            else chain<N, L + 1, MT, EXP>::run(sw, g, t, c, z);
        }
    };

    // Stage one member row (HBM layout of ll_topo: per layer Wt[k][out_pad] then b[out_pad]) into the swizzled W[out][in]
    // layout.  Coalesced 128-bit global reads, scattered 32-bit shared stores — once per member, i.e. once per
    // ~10^5 forward passes, so a TMA copy plus a second shared buffer for the transposition would only cost residency.
    template <class N, int L> __device__ __forceinline__ void stage_layer(const mma_args& a, const float* __restrict__ row, float* __restrict__ sw, int lane)
    {
        constexpr int IN = N::sz(L), IN8 = N::in8(L), OUT = N::sz(L + 1), OUT8 = N::out8(L), KB = IN8 / 8;
        const int op = a.out_pad[L]; // == OUT8
        const float4* src = reinterpret_cast<const float4*>(row + a.w_off[L]);
        for (int i = lane; i < IN * OUT8 / 4; i += 32)
        {
            const float4 v = __ldg(src + i);
            const int k = (i * 4) / op, o = (i * 4) - k * op;
            float* d       = sw + N::w_off(L) + (k & 7);
            const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) d[(o + j) * IN8 + (((k >> 3) ^ swz<KB>(o + j)) << 3)] = (o + j < OUT) ? e[j] : 0.f;
        }
        for (int o = lane; o < OUT8; o += 32) sw[N::b_off(L) + o] = (o < OUT) ? __ldg(row + a.b_off[L] + o) : 0.f;
        if constexpr (L + 1 < N::NL) stage_layer<N, L + 1>(a, row, sw, lane);
    }

    // warps per CTA (one CTA per SM, so that the warps spread evenly over the four SM sub-partitions): one member buffer
    // each within 227 KB of shared memory, at most 12 (MT = 1) / 8 (MT = 2) so that every thread can keep its fragments in
    // registers (65536 / (32 * warps) registers per thread)
    template <class N, int MT> constexpr int max_warps()
    {
        const int by_smem = (227 * 1024) / (N::total * 4), cap = MT == 1 ? 12 : 8;
        return by_smem < cap ? (by_smem < 1 ? 1 : by_smem) : cap;
    }

    template <class N, int MT, int EXP = 0>
    __global__ void __launch_bounds__(32 * max_warps<N, MT>(), 1) ll_episode_mma_kernel(const __grid_constant__ ll_env<float> env, const __grid_constant__ mma_args a)
    {
        static_assert(N::ok(), "topology not supported by the mma episode kernel");
        extern __shared__ __align__(16) float smem_all[];
        float* sw = smem_all + (threadIdx.x >> 5) * N::total; // this warp's weight buffer
        constexpr int S   = 16 * MT;
        const int lane    = threadIdx.x & 31;
        const int g       = lane >> 2, t = lane & 3;
        const int slot    = lane & (S - 1); // MT == 1: lanes 16..31 mirror lanes 0..15
        const bool owner  = lane < S;
        const unsigned sm = S == 32 ? 0xffffffffu : 0x0000ffffu;
        const int L       = a.n_layouts;

        for (int i = lane; i < N::total; i += 32) sw[i] = 0.f; // k / out padding stays zero for the whole kernel
        __syncwarp();

        for (;;)
        {
            unsigned long long member = 0;
            if (lane == 0) member = (unsigned long long)atomicAdd(a.queue, 1u);
            member = __shfl_sync(0xffffffffu, member, 0);
            if (member >= a.n_members) break;
            __syncwarp();
            stage_layer<N, 0>(a, a.params + member * (size_t)a.p_dev, sw, lane);
            __syncwarp();

            // ---- slot state (lane = slot)
            ll_lander<float> p{0.f, 0.f, 0.f, 0.f, 0.f};
            float cx = 0.f, lpad = 0.f, lpad_r = 0.f;
            int steps = 0, layout = -1, dir = 0;
            double acc = 0.0;
            bool active = false;
            double fit_sum = 0.0;
            unsigned int steps_sum = 0, succ0 = 0, succ1 = 0;
            int next_layout = L < S ? L : S;
            if (slot < L)
            {
                const ll_layout<float> lay = a.layouts[slot];
                layout = slot, dir = lay.dir;
                cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
                p      = ll_lander<float>{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
                active = true;
            }

            while (__ballot_sync(0xffffffffu, active) != 0u)
            {
                // ---- state -> A fragments of layer 0 (k block 0: positions (t, t+4) hold inputs (2t, 2t+1))
                float st[5];
                ll_get_state<float, LL_MODE_FAST>(env, p, cx, st, 1);
                float a0[MT][1][4];
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                    for (int h = 0; h < 2; ++h)
                    {
                        const int src = 16 * mt + 8 * h + g;
                        const float v0 = __shfl_sync(0xffffffffu, st[0], src), v1 = __shfl_sync(0xffffffffu, st[1], src);
                        const float v2 = __shfl_sync(0xffffffffu, st[2], src), v3 = __shfl_sync(0xffffffffu, st[3], src);
                        const float v4 = __shfl_sync(0xffffffffu, st[4], src);
                        a0[mt][0][2 * h + 0] = t == 0 ? v0 : (t == 1 ? v2 : (t == 2 ? v4 : 0.f));
                        a0[mt][0][2 * h + 1] = t == 0 ? v1 : (t == 1 ? v3 : 0.f);
                    }
                // ---- forward pass, activations chained through registers
                float z[MT][4];
                chain<N, 0, MT, EXP>::run(sw, g, t, a0, z);
                // ---- argmax over outputs 0..3 (first maximum wins): lanes t=0 hold outputs 0,1 and t=1 outputs 2,3
                unsigned packed = 0;
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                    for (int h = 0; h < 2; ++h)
                    {
                        float best = z[mt][2 * h];
                        int idx    = 2 * t;
                        if (z[mt][2 * h + 1] > best) best = z[mt][2 * h + 1], idx = 2 * t + 1;
                        const float ob = __shfl_xor_sync(0xffffffffu, best, 1);
                        const int oi   = __shfl_xor_sync(0xffffffffu, idx, 1);
                        if (ob > best) idx = oi; // meaningful in lanes t == 0 (partner t == 1 holds outputs 2,3)
                        packed |= (unsigned)idx << (8 * mt + 4 * h);
                    }
                const unsigned pk = __shfl_sync(0xffffffffu, packed, 4 * (slot & 7));
                const int action  = (pk >> (8 * (slot >> 4) + 4 * ((slot >> 3) & 1))) & 15;

                // ---- lander update + penalties (src/game_logic.cpp:295-320 through ll_env.cuh)
                bool done = false, landed = false, success = false;
                float dist = 0.f;
                if (active)
                {
                    const float sp = ll_env_step<float, LL_MODE_FAST>(env, p, action, steps == 0, cx, lpad, lpad_r, &landed, &success, &dist);
                    acc += (double)sp;
                    ++steps;
                    done = landed || steps >= a.max_steps;
                    if (EXP != 0) done = steps >= a.max_steps; // timing experiments: fixed-length episodes
                }
                const unsigned fin = __ballot_sync(0xffffffffu, done && owner) & sm;
                if (fin)
                { // some episode ended: book it, hand the slot the member's next layout
                    if (done)
                    {
                        const float term = ll_terminal_penalty<float, LL_MODE_FAST>(env, p, steps, landed, success, dist);
                        const double fit = (double)term + acc;
                        fit_sum += fit, steps_sum += (unsigned)steps;
                        if (success) (dir == 0 ? succ0 : succ1) += 1u;
                        if (a.ep_steps && owner)
                        {
                            const int meta = steps | (landed ? (1 << 28) : 0) | ((landed && !success) ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
                            a.ep_steps[member * (size_t)L + layout] = meta;
                            a.ep_fit[member * (size_t)L + layout]   = fit;
                        }
                        const int nl = next_layout + __popc(fin & ((1u << slot) - 1u));
                        if (nl < L)
                        { // reset(spad_x1, lpad_x1): src/game_logic.cpp:143-154
                            const ll_layout<float> lay = a.layouts[nl];
                            layout = nl, dir = lay.dir;
                            cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
                            p     = ll_lander<float>{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
                            steps = 0, acc = 0.0;
                        }
                        else active = false;
                    }
                    next_layout += __popc(fin);
                }
            }

            // ---- member totals (src/nn_engine_trainer.cpp:518-530; fast mode: free summation order)
            if (!owner) fit_sum = 0.0, steps_sum = 0, succ0 = 0, succ1 = 0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
            {
                fit_sum += __shfl_xor_sync(0xffffffffu, fit_sum, o);
                steps_sum += __shfl_xor_sync(0xffffffffu, steps_sum, o);
                succ0 += __shfl_xor_sync(0xffffffffu, succ0, o);
                succ1 += __shfl_xor_sync(0xffffffffu, succ1, o);
            }
            if (lane == 0)
            {
                float bonus0 = 0.f, bonus1 = 0.f;
                for (unsigned i = 0; i < succ0; ++i) bonus0 += env.tp_landed_lr_bonus;
                for (unsigned i = 0; i < succ1; ++i) bonus1 += env.tp_landed_lr_bonus;
                const double lr   = (double)(bonus1 < bonus0 ? bonus1 : bonus0);
                a.fitness[member] = fit_sum - lr;
                a.lr[member]      = lr;
                a.steps[member]   = (double)steps_sum;
            }
        }
    }

    int env_int(const char* name, int fallback)
    {
        const char* v = getenv(name);
        return v && *v ? atoi(v) : fallback;
    }

    template <class N, int MT, int EXP = 0> int launch_mma_t(ll_ctx* ctx, const ll_episode_args& h)
    {
        const ll_topo& tp = ctx->topo;
        mma_args a{};
        a.params = static_cast<const float*>(h.params), a.layouts = static_cast<const ll_layout<float>*>(h.layouts);
        a.n_layouts = h.n_layouts, a.max_steps = h.max_steps, a.n_members = h.n_members;
        a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps, a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit, a.queue = h.queue;
        a.p_dev = tp.p_dev;
        for (int l = 0; l < tp.n_layers; ++l) a.w_off[l] = tp.w_off[l], a.b_off[l] = tp.b_off[l], a.out_pad[l] = tp.out_pad[l];
        auto kern = ll_episode_mma_kernel<N, MT, EXP>;
        int max_smem = 0;
        LL_CUDA(ctx, cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device));
        int warps = std::min<int>(max_warps<N, MT>(), max_smem / (N::total * 4));
        warps     = std::min(warps, env_int("LL_MMA_WARPS_PER_SM", warps));
        if (warps < 1) return ll_fail(ctx, LL_EINVAL, "mma episode kernel: the member row does not fit in shared memory");
        const size_t smem = (size_t)warps * N::total * 4;
        LL_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const unsigned long long ctas_needed = (h.n_members + warps - 1) / warps;
        const unsigned int grid = (unsigned int)std::max<unsigned long long>(1, std::min<unsigned long long>((unsigned long long)ctx->sm_count, ctas_needed));
        ll_env<float> env;
        ll_make_env<float>(ctx->env_cfg, 1, &env);
        LL_CUDA(ctx, cudaMemsetAsync(h.queue, 0, sizeof(unsigned int), ctx->stream));
        kern<<<grid, 32 * warps, smem, ctx->stream>>>(env, a);
        LL_CUDA(ctx, cudaGetLastError());
        return LL_OK;
    }

    template <class N> int launch_mma_n(ll_ctx* ctx, const ll_episode_args& h)
    {
        return env_int("LL_MMA_MT", 1) == 2 ? launch_mma_t<N, 2>(ctx, h) : launch_mma_t<N, 1>(ctx, h);
    }

    using net_16_16        = net_t<5, 16, 16, 4>;
    using net_16_32        = net_t<5, 16, 32, 4>;
    using net_8_16_8       = net_t<5, 8, 16, 8, 4>;
    using net_16_64_32     = net_t<5, 16, 64, 32, 4>;
    using net_32_128_32    = net_t<5, 32, 128, 32, 4>;
    using net_8_16_32_16_8 = net_t<5, 8, 16, 32, 16, 8, 4>;

    template <class N> bool same_net(const ll_topo& tp)
    {
        if (tp.n_layers != N::NL || tp.in[0] != N::sz(0)) return false;
        for (int l = 0; l < N::NL; ++l)
            if (tp.out[l] != N::sz(l + 1)) return false;
        return true;
    }
} // namespace

// The tensor-core kernel covers the sample_runs topologies (src/nn_engine_trainer.cpp:45-48: 5-h...-4) when one member
// has enough layouts to fill a 16-row tile; LL_FAST_KERNEL=ffma forces the FFMA kernels, =mma lowers the layout threshold.
bool ll_mma_applicable(const ll_ctx* ctx, int n_layouts)
{
    const char* force = getenv("LL_FAST_KERNEL");
    if (force && !strcmp(force, "ffma")) return false;
    const ll_topo& tp = ctx->topo;
    if (tp.activation != LL_ACT_TANH) return false; // the sigmoid nets stay on the FFMA kernels
    const bool known  = same_net<net_16_16>(tp) || same_net<net_16_32>(tp) || same_net<net_8_16_8>(tp) || same_net<net_16_64_32>(tp) ||
                       same_net<net_32_128_32>(tp) || same_net<net_8_16_32_16_8>(tp);
    if (!known) return false;
    return n_layouts >= ((force && !strcmp(force, "mma")) ? 1 : 9);
}

int ll_launch_episode_mma(ll_ctx* ctx, const ll_episode_args& a)
{
    const ll_topo& tp = ctx->topo;
#ifdef LL_MMA_EXPERIMENTS // timing-only variants with parts of the step removed (wrong results): where does the time go?
    if (same_net<net_32_128_32>(tp))
        switch (env_int("LL_MMA_EXP", 0))
        {
        case 1: return launch_mma_t<net_32_128_32, 1, 1>(ctx, a);
        case 2: return launch_mma_t<net_32_128_32, 1, 2>(ctx, a);
        case 3: return launch_mma_t<net_32_128_32, 1, 3>(ctx, a);
        case 4: return launch_mma_t<net_32_128_32, 1, 4>(ctx, a);
        case 5: return launch_mma_t<net_32_128_32, 1, 5>(ctx, a);
        case 9: return launch_mma_t<net_32_128_32, 1, 9>(ctx, a);
        default: break;
        }
#endif
    if (same_net<net_32_128_32>(tp)) return launch_mma_n<net_32_128_32>(ctx, a);
    if (same_net<net_16_64_32>(tp)) return launch_mma_n<net_16_64_32>(ctx, a);
    if (same_net<net_8_16_32_16_8>(tp)) return launch_mma_n<net_8_16_32_16_8>(ctx, a);
    if (same_net<net_8_16_8>(tp)) return launch_mma_n<net_8_16_8>(ctx, a);
    if (same_net<net_16_32>(tp)) return launch_mma_n<net_16_32>(ctx, a);
    if (same_net<net_16_16>(tp)) return launch_mma_n<net_16_16>(ctx, a);
    return ll_fail(ctx, LL_EINVAL, "mma episode kernel: topology not instantiated");
}
