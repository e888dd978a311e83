// ll_episode_mma.cu — K1 fast, tensor-core variant (fp32 via 3 x fp16 split products on mma.sync, LL_MODE_FAST).
//
// Replaces the same reference code as ll_episode.cu / ll_episode_fast.cu: train_generation / sim_task
// src/nn_engine_trainer.cpp:443-539, GameLogicCpp<T> src/game_logic.cpp:124-388 and the ma-libs feedforward call
// :501-502.  BASELINE.json north_star (a): "tensor cores (mma) are used only for the multi-start case, where each
// individual's forward pass becomes a real small dense GEMM".
//
// Mapping.  All layouts of a member share its weights, so one forward pass over 16 concurrent episode
// slots is D[16 x out] = A[16 x in] * W^T[in x out] per layer.  ONE WARP owns one member at a time:
//   * the 16 slots are the M dimension of mma.sync.m16n8k16 (f16 operands, fp32 accumulate).  The activations never leave
//     the register file: the C fragment of layer l (rows g / g+8, columns 8nt+2t, 8nt+2t+1) packed to half2 IS register
//     (nt & 1) * 2 + {0, 1} of the A fragment of k block nt / 2 of layer l+1;
//   * fp32 accuracy from 11-bit operands: x = hi + lo with hi = the leading 11 significant bits and lo = x - hi, both fp16 and
//     unscaled (|lo| < 2^-11 |x|; where lo is an fp16 subnormal its absolute error is <= 2^-25: activations are <= 5, weights that
//     small contribute below fp32 resolution of a pre-activation).  D = bias + A_hi*B_hi + A_lo*B_hi + A_hi*B_lo in one fp32
//     accumulator tile; the dropped lo*lo term is < 2^-21 relative (measured: the share of episodes whose step count equals the
//     exact fp32 path's is the same 0.978 as for the FFMA kernel; dropping A_lo*B_hi as well lowers it to 0.50);
//   * the member's weights are split ONCE per member, while they are staged into shared memory, and stored as ready-made
//     B fragments {b0_hi, b1_hi, b0_lo, b1_lo} per (output block, k block, lane): one conflict-free LDS.128 per three
//     HMMAs, 4 bytes per weight (35.3 KB for 32_128_32);
//   * the lander of slot s lives in the registers of lane s (lanes 16..31 mirror 0..15); the argmax over the 4 outputs is
//     two compares + one shuffle in the quad that holds them, the action and the next state travel between the
//     fragment layout and the slot lanes by shuffles.  No barrier, no atomics and no shared-memory traffic other than
//     the weight fetches in the step loop;
//   * a slot whose episode ends takes the member's next layout (ballot + prefix count); the warp moves on to the next
//     member of the global queue when all layouts are done.
// Why these choices (measured on B200, ll_microbench_fma codes 19x/20xx): a legacy HMMA occupies its SM sub-partition's
// issue port for its whole 8.56 cycles — other instructions do not issue in its shadow, from the same warp or another —
// so the cost of a step is 8.56 * #HMMA + ~1.1 * #other instructions per sub-partition, and one warp per sub-partition
// saturates it.  m16n8k16.f16 takes the same 8.56 cycles as m16n8k8.tf32 for twice the MACs, and pre-split weights
// remove four ALU instructions per fragment: hence f16 splits instead of 3xTF32, and 4 or 8 warps per SM.
// Layer sizes are template parameters (fragments are register arrays): the sample_runs topologies are instantiated,
// anything else stays on the FFMA kernels.  fp16 range: weights beyond +-65504 saturate (such a pre-activation has
// saturated tanh long before).
#include <cuda_fp16.h>

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
        static __host__ __device__ constexpr int in16(int l) { return (sz(l) + 15) & ~15; }
        static __host__ __device__ constexpr int out8(int l) { return (sz(l + 1) + 7) & ~7; }
        // shared-memory layout in 32-bit words: per layer the B fragments [out8/8][in16/16][32 lanes][4 words], then the
        // bias as accumulator fragments [out8/8][4 (t)][4] = {b[8nt+2t], b[8nt+2t+1], b[8nt+2t], b[8nt+2t+1]}
        static __host__ __device__ constexpr int w_off(int l)
        {
            int o = 0;
            for (int i = 0; i < l; ++i) o += in16(i) * out8(i) + 2 * out8(i);
            return o;
        }
        static __host__ __device__ constexpr int b_off(int l) { return w_off(l) + in16(l) * out8(l); }
        static constexpr int total = w_off(NL);
        static __host__ __device__ constexpr bool ok() { return sz(0) <= 8 && sz(NL) <= 8; }
    };

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
        // small populations: a member's layouts are dealt to `chunks` warps (work item = member * chunks + chunk; chunk c flies the
        // layouts at positions c, c + chunks, ... of the start order); partial sums meet in `part`, the last warp to arrive adds them
        // in chunk order
        int chunks;
        double* part;             // [n_members][chunks][4] = fitness sum, steps, successes direction 0 / 1
        unsigned int* part_count; // [n_members], zero before the launch
    };

    __device__ __forceinline__ void mma_f16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
    {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }

    // Activations (x0, x1) -> half2 of the leading parts and half2 of the remainders.  hi = x & 0xffffe000 keeps 11
    // significant bits (exact in fp16 over its normal range), lo = x - hi is exact in fp32 and below 2^-10 |x|.  The
    // activations are bounded (tanh outputs, clamped state inputs <= 5), so lo needs no scaling: where it falls into the
    // fp16 subnormal range its absolute error is <= 2^-25, the rounding noise of one fp32 product.
    __device__ __forceinline__ void split_pair(float x0, float x1, uint32_t& hi, uint32_t& lo)
    {
        const float h0 = __uint_as_float(__float_as_uint(x0) & 0xffffe000u), h1 = __uint_as_float(__float_as_uint(x1) & 0xffffe000u);
        const __half2 ph = __floats2half2_rn(h0, h1);
        const __half2 pl = __floats2half2_rn(x0 - h0, x1 - h1);
        hi = *reinterpret_cast<const uint32_t*>(&ph);
        lo = *reinterpret_cast<const uint32_t*>(&pl);
    }

    // One layer for the warp's 16 slots.  ah/al[kt][4]: A fragments (hi / lo) of k block kt (16 inputs).  Output:
    // the next layer's A fragments nh/nl (tanh applied, split, packed), or for the output layer the raw fragment z[4]
    // (rows g / g+8, outputs 2t, 2t+1) with tanh applied.  Output blocks go in panels of PN <= 8 so that the PN
    // accumulator tiles stay in registers.
    template <class N, int L>
    __device__ __forceinline__ void layer_mma(const float* __restrict__ sw, int lane, int t, const uint32_t (&ah)[N::in16(L) / 16][4],
                                              const uint32_t (&al)[N::in16(L) / 16][4],
                                              uint32_t (&nh)[(N::out8(L) + 15) / 16][4], uint32_t (&nl)[(N::out8(L) + 15) / 16][4], float (&z)[4])
    {
        constexpr int KT = N::in16(L) / 16, NT = N::out8(L) / 8;
        constexpr int PN = NT >= 8 ? 8 : NT;
        constexpr bool LAST = L == N::NL - 1;
        static_assert(NT % PN == 0, "output blocks must fill whole panels");
        const uint4* W = reinterpret_cast<const uint4*>(sw + N::w_off(L)) + lane;
        const float4* B = reinterpret_cast<const float4*>(sw + N::b_off(L)) + t;
        if constexpr (!LAST && (NT & 1))
        { // odd number of output blocks: the upper half of the last k block of the next layer is padding
#pragma unroll
            for (int i = 2; i < 4; ++i) nh[NT / 2][i] = 0u, nl[NT / 2][i] = 0u;
        }
#pragma unroll
        for (int np = 0; np < NT; np += PN)
        {
            // bias + A_hi*B_hi + A_lo*B_hi + A_hi*B_lo: the three products of the split operands (A_lo*B_lo is below fp32
            // resolution) go into ONE fp32 accumulator tile per output block when the panel is wide: per k block the products
            // are issued pass by pass, so the HMMAs of one accumulator are PN instructions apart (a dependent HMMA waits 21
            // cycles for its predecessor).  Narrow panels keep one tile per product and add them at the end.
            constexpr int NA = PN >= 4 ? 1 : 3;
            float acc[NA][PN][4];
#pragma unroll
            for (int j = 0; j < PN; ++j)
            {
                const float4 b = B[(np + j) * 4]; // one LDS.128 straight into the accumulator tile
                acc[0][j][0] = b.x, acc[0][j][1] = b.y, acc[0][j][2] = b.z, acc[0][j][3] = b.w;
#pragma unroll
                for (int q = 1; q < NA; ++q)
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[q][j][i] = 0.f;
            }
#pragma unroll
            for (int kt = 0; kt < KT; ++kt)
            {
                uint4 w[PN]; // {b0_hi, b1_hi, b0_lo, b1_lo}
#pragma unroll
                for (int j = 0; j < PN; ++j) w[j] = W[((np + j) * KT + kt) * 32];
#pragma unroll
                for (int j = 0; j < PN; ++j) mma_f16(acc[1 % NA][j], al[kt], w[j].x, w[j].y);
#pragma unroll
                for (int j = 0; j < PN; ++j) mma_f16(acc[2 % NA][j], ah[kt], w[j].z, w[j].w);
#pragma unroll
                for (int j = 0; j < PN; ++j) mma_f16(acc[0][j], ah[kt], w[j].x, w[j].y);
            }
#pragma unroll
            for (int j = 0; j < PN; ++j)
            {
                float v[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = ll_tanh_approx(NA == 1 ? acc[0][j][i] : acc[0][j][i] + (acc[1 % NA][j][i] + acc[2 % NA][j][i]));
                if constexpr (LAST)
                {
#pragma unroll
                    for (int i = 0; i < 4; ++i) z[i] = v[i];
                }
                else
                {
                    const int nt = np + j;
                    split_pair(v[0], v[1], nh[nt / 2][(nt & 1) * 2 + 0], nl[nt / 2][(nt & 1) * 2 + 0]); // row g
                    split_pair(v[2], v[3], nh[nt / 2][(nt & 1) * 2 + 1], nl[nt / 2][(nt & 1) * 2 + 1]); // row g + 8
                }
            }
        }
    }

    // layers L..NL-1; the output layer's fragment ends up in z
    template <class N, int L> struct chain
    {
        static __device__ __forceinline__ void run(const float* __restrict__ sw, int lane, int t, const uint32_t (&ah)[N::in16(L) / 16][4],
                                                   const uint32_t (&al)[N::in16(L) / 16][4], float (&z)[4])
        {
            uint32_t nh[(N::out8(L) + 15) / 16][4], nl[(N::out8(L) + 15) / 16][4];
            layer_mma<N, L>(sw, lane, t, ah, al, nh, nl, z);
            if constexpr (L + 1 < N::NL) chain<N, L + 1>::run(sw, lane, t, nh, nl, z);
        }
    };

    // Stage one member row (HBM layout of ll_topo: per layer Wt[k][out_pad] then b[out_pad]) into the B-fragment layout,
    // splitting every weight into its fp16 parts on the way.  Coalesced 128-bit global reads, scattered 16-bit shared
    // stores — once per member, i.e. once per ~10^5 forward passes, so a TMA copy plus a second shared buffer for the
    // re-layout would only cost residency.  Element (k, o) of a layer lives in lane (o & 7) * 4 + ((k & 7) >> 1) of the
    // fragment of (output block o >> 3, k block k >> 4), word (k >> 3) & 1 (hi) / 2 + ((k >> 3) & 1) (lo), half k & 1.
    template <class N, int L> __device__ __forceinline__ void stage_layer(const mma_args& a, const float* __restrict__ row, float* __restrict__ sw, int lane)
    {
        constexpr int IN = N::sz(L), OUT = N::sz(L + 1), OUT8 = N::out8(L), KT = N::in16(L) / 16;
        const int op      = a.out_pad[L]; // == OUT8
        const float4* src = reinterpret_cast<const float4*>(row + a.w_off[L]);
        __half* dst       = reinterpret_cast<__half*>(sw + N::w_off(L));
        for (int i = lane; i < IN * OUT8 / 4; i += 32)
        {
            const float4 v = __ldg(src + i);
            const int k = (i * 4) / op, o0 = (i * 4) - k * op;
            const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                const int o = o0 + j;
                if (o < OUT)
                {
                    const __half hi = __float2half_rn(fminf(fmaxf(e[j], -65504.f), 65504.f));
                    const __half lo = __float2half_rn(fminf(fmaxf(e[j] - __half2float(hi), -65504.f), 65504.f)); // unscaled: exact to 2^-25 absolute
                    const int word  = ((((o >> 3) * KT + (k >> 4)) * 32 + (o & 7) * 4 + ((k & 7) >> 1)) << 2) + ((k >> 3) & 1);
                    dst[word * 2 + (k & 1)]       = hi;
                    dst[(word + 2) * 2 + (k & 1)] = lo;
                }
            }
        }
        for (int i = lane; i < 2 * OUT8; i += 32)
        { // word i of the bias fragments: block i >> 4, t = (i >> 2) & 3, element i & 3 -> output 8nt + 2t + (i & 1)
            const int o           = (i >> 4) * 8 + ((i >> 2) & 3) * 2 + (i & 1);
            sw[N::b_off(L) + i] = (o < OUT) ? __ldg(row + a.b_off[L] + o) : 0.f;
        }
        if constexpr (L + 1 < N::NL) stage_layer<N, L + 1>(a, row, sw, lane);
    }

    // warps per CTA (one CTA per SM): one member buffer each within 227 KB of shared memory - as many as fit up to 8, then 12 or 16
    // for the small networks.  Measured on configs[3] (32_128_32, 37.4 KB per member): 6 warps reach 97.9 % of the FP32 FMA roofline
    // against 84.4 % with 4 - the second warp of a sub-partition fills the issue slots the first leaves while it waits on its own
    // HMMA -> epilogue -> HMMA chain, which outweighs the uneven 2-2-1-1 split; configs[2]: 12 warps 54.4 % against 47.1 % with 8.
    template <class N> constexpr int max_warps()
    {
        const int by_smem = (227 * 1024) / (N::total * 4);
        return by_smem >= 16 ? 16 : (by_smem >= 12 ? 12 : (by_smem < 1 ? 1 : by_smem));
    }

    template <class N, int EXP = 0>
    __global__ void __launch_bounds__(32 * max_warps<N>(), 1) ll_episode_mma_kernel(const __grid_constant__ ll_env<float> env, const __grid_constant__ mma_args a)
    {
        static_assert(N::ok(), "topology not supported by the mma episode kernel");
        extern __shared__ __align__(16) float smem_all[];
        float* sw = smem_all + (threadIdx.x >> 5) * N::total; // this warp's weight buffer
        const int lane   = threadIdx.x & 31;
        const int g      = lane >> 2, t = lane & 3;
        const int slot   = lane & 15; // lanes 16..31 mirror lanes 0..15
        const bool owner = lane < 16;
        const int LT     = a.n_layouts; // row length of the per-episode outputs

        for (int i = lane; i < N::total; i += 32) sw[i] = 0.f; // k / out padding stays zero for the whole kernel
        __syncwarp();

        for (;;)
        {
            unsigned long long member = 0;
            if (lane == 0) member = (unsigned long long)atomicAdd(a.queue, 1u);
            member = __shfl_sync(0xffffffffu, member, 0);
            const int C = a.chunks;
            if (member >= a.n_members * (unsigned long long)C) break;
            const int chunk = (int)(member % (unsigned long long)C);
            member /= (unsigned long long)C;
            const int L = (a.n_layouts - chunk + C - 1) / C; // layouts of this chunk: start-order positions chunk + i * C
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
            unsigned long long steps_sum = 0; // 4096 layouts x 2^28 steps do not fit 32 bits
            unsigned int succ0 = 0, succ1 = 0;
            bool publish = true; // false for a chunk that is not the last of its member to arrive
            int next_layout = L < 16 ? L : 16;
            if (slot < L)
            {
                layout                     = a.layouts[chunk + slot * C].next; // start order: longest expected flight first
                const ll_layout<float> lay = a.layouts[layout];
                dir                        = lay.dir;
                cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
                p      = ll_lander<float>{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
                active = true;
            }

            while (__ballot_sync(0xffffffffu, active) != 0u)
            {
                // ---- state -> A fragment of layer 0: registers 0 / 1 hold inputs (2t, 2t+1) of rows g / g+8, k >= 8 is padding
                float st[5];
                ll_get_state<float, LL_MODE_FAST>(env, p, cx, st, 1);
                uint32_t a0h[1][4], a0l[1][4];
#pragma unroll
                for (int h = 0; h < 2; ++h)
                {
                    const int src = 8 * h + g;
                    const float v0 = __shfl_sync(0xffffffffu, st[0], src), v1 = __shfl_sync(0xffffffffu, st[1], src);
                    const float v2 = __shfl_sync(0xffffffffu, st[2], src), v3 = __shfl_sync(0xffffffffu, st[3], src);
                    const float v4 = __shfl_sync(0xffffffffu, st[4], src);
                    const float e0 = t == 0 ? v0 : (t == 1 ? v2 : (t == 2 ? v4 : 0.f));
                    const float e1 = t == 0 ? v1 : (t == 1 ? v3 : 0.f);
                    split_pair(e0, e1, a0h[0][h], a0l[0][h]);
                    a0h[0][2 + h] = 0u, a0l[0][2 + h] = 0u;
                }
                // ---- forward pass, activations chained through registers
                float z[4];
                chain<N, 0>::run(sw, lane, t, a0h, a0l, z);
                // ---- argmax over outputs 0..3 (first maximum wins): lanes t=0 hold outputs 0,1 and t=1 outputs 2,3
                unsigned packed = 0;
#pragma unroll
                for (int h = 0; h < 2; ++h)
                {
                    float best = z[2 * h];
                    int idx    = 2 * t;
                    if (z[2 * h + 1] > best) best = z[2 * h + 1], idx = 2 * t + 1;
                    const float ob = __shfl_xor_sync(0xffffffffu, best, 1);
                    const int oi   = __shfl_xor_sync(0xffffffffu, idx, 1);
                    if (ob > best) idx = oi; // meaningful in lanes t == 0 (partner t == 1 holds outputs 2,3)
                    packed |= (unsigned)idx << (4 * h);
                }
                const unsigned pk = __shfl_sync(0xffffffffu, packed, 4 * (slot & 7));
                const int action  = (pk >> (4 * (slot >> 3))) & 15;

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
                const unsigned fin = __ballot_sync(0xffffffffu, done && owner) & 0xffffu;
                if (fin)
                { // some episode ended: book it, hand the slot the member's next layout
                    if (done)
                    {
                        const float term = ll_terminal_penalty<float, LL_MODE_FAST>(env, p, steps, landed, success, dist);
                        const double fit = (double)term + acc;
                        fit_sum += fit, steps_sum += (unsigned long long)steps;
                        if (success) (dir == 0 ? succ0 : succ1) += 1u;
                        if (a.ep_steps && owner)
                        {
                            const int meta = steps | (landed ? (1 << 28) : 0) | ((landed && !success) ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
                            a.ep_steps[member * (size_t)LT + layout] = meta;
                            a.ep_fit[member * (size_t)LT + layout]   = fit;
                        }
                        const int nl = next_layout + __popc(fin & ((1u << slot) - 1u));
                        if (nl < L)
                        { // reset(spad_x1, lpad_x1): src/game_logic.cpp:143-154
                            layout                     = a.layouts[chunk + nl * C].next;
                            const ll_layout<float> lay = a.layouts[layout];
                            dir                        = lay.dir;
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
            if (lane == 0 && C > 1)
            { // hand the partial sums over; the last chunk of the member to arrive adds them in chunk order (deterministic)
                double* mine = a.part + (member * (size_t)C + chunk) * 4;
                mine[0] = fit_sum, mine[1] = (double)steps_sum, mine[2] = (double)succ0, mine[3] = (double)succ1;
                __threadfence();
                const unsigned int arrived = atomicAdd(a.part_count + member, 1u);
                if (arrived == (unsigned int)C - 1u)
                {
                    __threadfence();
                    const volatile double* all = a.part + member * (size_t)C * 4;
                    double f = 0.0, st = 0.0, s0 = 0.0, s1 = 0.0;
                    for (int c = 0; c < C; ++c) f += all[c * 4 + 0], st += all[c * 4 + 1], s0 += all[c * 4 + 2], s1 += all[c * 4 + 3];
                    if (a.ep_fit)
                    { // the chunk count depends on how many members this GPU holds: sum the per-episode values in LAYOUT order instead, so
                      // that a small population's fitness is bit-identical for any GPU count (the exact kernel's order, :518)
                        const volatile double* ef = a.ep_fit + member * (size_t)LT;
                        f = 0.0;
                        for (int l = 0; l < a.n_layouts; ++l) f += ef[l];
                    }
                    fit_sum = f, steps_sum = (unsigned long long)st, succ0 = (unsigned int)s0, succ1 = (unsigned int)s1;
                    a.part_count[member] = 0u;
                }
                else publish = false;
            }
            if (lane == 0 && publish)
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

    // ---- straggler pass behind the tcgen05 tile kernel (ll_episode_tc05.cu): a member's last <= 16 running episodes, handed over with
    // their lander state, step count and accumulated step penalty, finish here in ONE 16-row tile (no refill: nothing is left to
    // start).  The warp then closes the member: every layout's result, the tile kernel's and ours, is summed in layout order from the
    // per-episode arrays (src/nn_engine_trainer.cpp:518-530).
    struct resume_args
    {
        mma_args m;
        ll_resume_args r;
    };

    template <class N>
    __global__ void __launch_bounds__(32 * max_warps<N>(), 1) ll_episode_mma_resume_kernel(const __grid_constant__ ll_env<float> env, const __grid_constant__ resume_args ra)
    {
        const mma_args& a = ra.m;
        extern __shared__ __align__(16) float smem_all[];
        float* sw = smem_all + (threadIdx.x >> 5) * N::total;
        const int lane = threadIdx.x & 31;
        const int g = lane >> 2, t = lane & 3;
        const int slot   = lane & 15;
        const bool owner = lane < 16;
        const int LT     = a.n_layouts;
        for (int i = lane; i < N::total; i += 32) sw[i] = 0.f;
        __syncwarp();
        const unsigned int n_items = *ra.r.n;
        for (;;)
        {
            unsigned int item = 0;
            if (lane == 0) item = atomicAdd(a.queue, 1u);
            item = __shfl_sync(0xffffffffu, item, 0);
            if (item >= n_items) break;
            const unsigned long long member = ra.r.list[item];
            const int cnt                   = (int)ra.r.count[member];
            __syncwarp();
            stage_layer<N, 0>(a, a.params + member * (size_t)a.p_dev, sw, lane);
            __syncwarp();
            ll_lander<float> p{0.f, 0.f, 0.f, 0.f, 0.f};
            float cx = 0.f, lpad = 0.f, lpad_r = 0.f;
            int steps = 0, layout = 0;
            double acc  = 0.0;
            bool active = false;
            const ll_resume_rec* recs = ra.r.rec + member * (size_t)ra.r.stride;
            int next_rec              = cnt < 16 ? cnt : 16; // records beyond the 16 rows wait their turn (hand-over thresholds above 16)
            if (slot < cnt)
            {
                const ll_resume_rec r = recs[slot];
                layout                = r.layout;
                const ll_layout<float> lay = a.layouts[layout];
                cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
                p      = ll_lander<float>{r.x, r.y, r.vx, r.vy, r.fuel};
                steps = r.steps, acc = r.acc, active = true;
            }
            while (__ballot_sync(0xffffffffu, active) != 0u)
            {
                float st[5];
                ll_get_state<float, LL_MODE_FAST>(env, p, cx, st, 1);
                uint32_t a0h[1][4], a0l[1][4];
#pragma unroll
                for (int h = 0; h < 2; ++h)
                {
                    const int src = 8 * h + g;
                    const float v0 = __shfl_sync(0xffffffffu, st[0], src), v1 = __shfl_sync(0xffffffffu, st[1], src);
                    const float v2 = __shfl_sync(0xffffffffu, st[2], src), v3 = __shfl_sync(0xffffffffu, st[3], src);
                    const float v4 = __shfl_sync(0xffffffffu, st[4], src);
                    const float e0 = t == 0 ? v0 : (t == 1 ? v2 : (t == 2 ? v4 : 0.f));
                    const float e1 = t == 0 ? v1 : (t == 1 ? v3 : 0.f);
                    split_pair(e0, e1, a0h[0][h], a0l[0][h]);
                    a0h[0][2 + h] = 0u, a0l[0][2 + h] = 0u;
                }
                float z[4];
                chain<N, 0>::run(sw, lane, t, a0h, a0l, z);
                unsigned packed = 0;
#pragma unroll
                for (int h = 0; h < 2; ++h)
                {
                    float best = z[2 * h];
                    int idx    = 2 * t;
                    if (z[2 * h + 1] > best) best = z[2 * h + 1], idx = 2 * t + 1;
                    const float ob = __shfl_xor_sync(0xffffffffu, best, 1);
                    const int oi   = __shfl_xor_sync(0xffffffffu, idx, 1);
                    if (ob > best) idx = oi;
                    packed |= (unsigned)idx << (4 * h);
                }
                const unsigned pk = __shfl_sync(0xffffffffu, packed, 4 * (slot & 7));
                const int action  = (pk >> (4 * (slot >> 3))) & 15;
                bool done = false, landed = false, success = false;
                float dist = 0.f;
                if (active)
                {
                    const float sp = ll_env_step<float, LL_MODE_FAST>(env, p, action, steps == 0, cx, lpad, lpad_r, &landed, &success, &dist);
                    acc += (double)sp;
                    ++steps;
                    done = landed || steps >= a.max_steps;
                }
                const unsigned fin = __ballot_sync(0xffffffffu, done && owner) & 0xffffu;
                if (fin)
                {
                    if (done)
                    {
                        const float term = ll_terminal_penalty<float, LL_MODE_FAST>(env, p, steps, landed, success, dist);
                        if (owner)
                        {
                            a.ep_steps[member * (size_t)LT + layout] = steps | (landed ? (1 << 28) : 0) | ((landed && !success) ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
                            a.ep_fit[member * (size_t)LT + layout]   = (double)term + acc;
                        }
                        const int nr = next_rec + __popc(fin & ((1u << slot) - 1u));
                        if (nr < cnt)
                        { // the row takes the member's next waiting episode
                            const ll_resume_rec r = recs[nr];
                            layout                = r.layout;
                            const ll_layout<float> lay = a.layouts[layout];
                            cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
                            p     = ll_lander<float>{r.x, r.y, r.vx, r.vy, r.fuel};
                            steps = r.steps, acc = r.acc;
                        }
                        else active = false;
                    }
                    next_rec += __popc(fin);
                }
            }
            __syncwarp();
            if (lane == 0)
            { // close the member: all layouts in layout order (the tile kernel wrote the others before this kernel started)
                double f = 0.0, stp = 0.0;
                float bonus0 = 0.f, bonus1 = 0.f;
                for (int l = 0; l < LT; ++l)
                {
                    const int meta = a.ep_steps[member * (size_t)LT + l];
                    f += a.ep_fit[member * (size_t)LT + l], stp += (double)(meta & 0x0fffffff);
                    if (meta & (1 << 30)) (a.layouts[l].dir == 0 ? bonus0 : bonus1) += env.tp_landed_lr_bonus;
                }
                const double lr   = (double)(bonus1 < bonus0 ? bonus1 : bonus0);
                a.fitness[member] = f - lr, a.lr[member] = lr, a.steps[member] = stp;
            }
        }
    }

    int env_int(const char* name, int fallback)
    {
        const char* v = getenv(name);
        return v && *v ? atoi(v) : fallback;
    }

    template <class N, int EXP = 0> int launch_mma_t(ll_ctx* ctx, const ll_episode_args& h)
    {
        const ll_topo& tp = ctx->topo;
        mma_args a{};
        a.params = static_cast<const float*>(h.params), a.layouts = static_cast<const ll_layout<float>*>(h.layouts);
        a.n_layouts = h.n_layouts, a.max_steps = h.max_steps, a.n_members = h.n_members;
        a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps, a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit, a.queue = h.queue;
        a.p_dev = tp.p_dev;
        for (int l = 0; l < tp.n_layers; ++l) a.w_off[l] = tp.w_off[l], a.b_off[l] = tp.b_off[l], a.out_pad[l] = tp.out_pad[l];
        auto kern = ll_episode_mma_kernel<N, EXP>;
        int max_smem = 0;
        LL_CUDA(ctx, cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device));
        int warps = std::min<int>(max_warps<N>(), max_smem / (N::total * 4));
        warps     = std::min(warps, env_int("LL_MMA_WARPS_PER_SM", warps));
        if (warps < 1) return ll_fail(ctx, LL_EINVAL, "mma episode kernel: the member row does not fit in shared memory");
        const size_t smem = (size_t)warps * N::total * 4;
        LL_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        // a population smaller than the GPU's warp slots: deal each member's layouts to several warps so that they fly
        // concurrently instead of in rounds of 16 (a generation then lasts one longest episode, not ceil(L / 16) of them)
        const unsigned long long slots_w = (unsigned long long)ctx->sm_count * warps;
        int chunks = 1;
        if (h.n_members < slots_w) chunks = (int)std::max<unsigned long long>(1, std::min<unsigned long long>(slots_w / h.n_members, (unsigned long long)(h.n_layouts / 2)));
        chunks = std::max(1, std::min(env_int("LL_MMA_CHUNKS", chunks), h.n_layouts));
        a.chunks = chunks, a.part = nullptr, a.part_count = nullptr;
        if (chunks > 1)
        {
            const size_t part_bytes = (size_t)h.n_members * chunks * 4 * sizeof(double), count_bytes = (size_t)h.n_members * sizeof(unsigned int);
            if (ctx->part_cap < part_bytes + count_bytes)
            {
                if (ctx->d_part) cudaFree(ctx->d_part);
                ctx->d_part = nullptr, ctx->part_cap = 0;
                LL_CUDA(ctx, cudaMalloc(&ctx->d_part, part_bytes + count_bytes));
                ctx->part_cap = part_bytes + count_bytes;
            }
            a.part       = static_cast<double*>(ctx->d_part);
            a.part_count = reinterpret_cast<unsigned int*>(static_cast<char*>(ctx->d_part) + part_bytes);
            LL_CUDA(ctx, cudaMemsetAsync(a.part_count, 0, count_bytes, ctx->stream));
        }
        const unsigned long long ctas_needed = (h.n_members * chunks + warps - 1) / warps;
        const unsigned int grid = (unsigned int)std::max<unsigned long long>(1, std::min<unsigned long long>((unsigned long long)ctx->sm_count, ctas_needed));
        ll_env<float> env;
        ll_make_env<float>(ctx->env_cfg, 1, &env);
        LL_CUDA(ctx, cudaMemsetAsync(h.queue, 0, sizeof(unsigned int), ctx->stream));
        kern<<<grid, 32 * warps, smem, ctx->stream>>>(env, a);
        LL_CUDA(ctx, cudaGetLastError());
        return LL_OK;
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

// The tensor-core kernel covers the sample_runs topologies (src/nn_engine_trainer.cpp:45-48: 5-h...-4) with tanh at any
// layout count: even a single layout per member (1 live row of the 16-row tile) runs 2.7x faster than the lane-per-episode
// FFMA kernel (configs[1]: 5.9e8 vs 2.1e8 lander-steps/s), 8 layouts 8x faster than the slot-pool FFMA kernel
// (configs[2]).  LL_FAST_KERNEL=ffma forces the FFMA kernels.
bool ll_mma_applicable(const ll_ctx* ctx, int n_layouts)
{
    const char* force = getenv("LL_FAST_KERNEL");
    if (force && !strcmp(force, "ffma")) return false;
    const ll_topo& tp = ctx->topo;
    if (tp.activation != LL_ACT_TANH) return false; // the sigmoid nets stay on the FFMA kernels
    const bool known  = same_net<net_16_16>(tp) || same_net<net_16_32>(tp) || same_net<net_8_16_8>(tp) || same_net<net_16_64_32>(tp) ||
                       same_net<net_32_128_32>(tp) || same_net<net_8_16_32_16_8>(tp);
    return known && n_layouts >= 1;
}

int ll_launch_episode_mma(ll_ctx* ctx, const ll_episode_args& a)
{
    const ll_topo& tp = ctx->topo;
#ifdef LL_MMA_EXPERIMENTS // timing-only variant with fixed-length episodes
    if (same_net<net_32_128_32>(tp) && env_int("LL_MMA_EXP", 0) == 9) return launch_mma_t<net_32_128_32, 9>(ctx, a);
#endif
    if (same_net<net_32_128_32>(tp)) return launch_mma_t<net_32_128_32>(ctx, a);
    if (same_net<net_16_64_32>(tp)) return launch_mma_t<net_16_64_32>(ctx, a);
    if (same_net<net_8_16_32_16_8>(tp)) return launch_mma_t<net_8_16_32_16_8>(ctx, a);
    if (same_net<net_8_16_8>(tp)) return launch_mma_t<net_8_16_8>(ctx, a);
    if (same_net<net_16_32>(tp)) return launch_mma_t<net_16_32>(ctx, a);
    if (same_net<net_16_16>(tp)) return launch_mma_t<net_16_16>(ctx, a);
    return ll_fail(ctx, LL_EINVAL, "mma episode kernel: topology not instantiated");
}

namespace
{
    template <class N> int launch_resume_t(ll_ctx* ctx, const ll_episode_args& h, const ll_resume_args& r)
    {
        const ll_topo& tp = ctx->topo;
        resume_args ra{};
        mma_args& a = ra.m;
        a.params = static_cast<const float*>(h.params), a.layouts = static_cast<const ll_layout<float>*>(h.layouts);
        a.n_layouts = h.n_layouts, a.max_steps = h.max_steps, a.n_members = h.n_members;
        a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps, a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit;
        a.queue = h.queue + 2; // the tile kernel uses words 0 (work counter) and 1 (error flag)
        a.p_dev = tp.p_dev, a.chunks = 1;
        for (int l = 0; l < tp.n_layers; ++l) a.w_off[l] = tp.w_off[l], a.b_off[l] = tp.b_off[l], a.out_pad[l] = tp.out_pad[l];
        ra.r      = r;
        auto kern = ll_episode_mma_resume_kernel<N>;
        int max_smem = 0;
        LL_CUDA(ctx, cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device));
        const int warps = std::max(1, std::min<int>(max_warps<N>(), max_smem / (N::total * 4)));
        const size_t smem = (size_t)warps * N::total * 4;
        LL_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ll_env<float> env;
        ll_make_env<float>(ctx->env_cfg, 1, &env);
        LL_CUDA(ctx, cudaMemsetAsync(a.queue, 0, sizeof(unsigned int), ctx->stream));
        const unsigned long long ctas_needed = (h.n_members + warps - 1) / warps;
        const unsigned int grid = (unsigned int)std::max<unsigned long long>(1, std::min<unsigned long long>((unsigned long long)ctx->sm_count, ctas_needed));
        kern<<<grid, 32 * warps, smem, ctx->stream>>>(env, ra);
        LL_CUDA(ctx, cudaGetLastError());
        return LL_OK;
    }
} // namespace

int ll_launch_episode_mma_resume(ll_ctx* ctx, const ll_episode_args& a, const ll_resume_args& r)
{
    const ll_topo& tp = ctx->topo;
    if (same_net<net_32_128_32>(tp)) return launch_resume_t<net_32_128_32>(ctx, a, r);
    if (same_net<net_16_64_32>(tp)) return launch_resume_t<net_16_64_32>(ctx, a, r);
    return ll_fail(ctx, LL_EINVAL, "mma resume kernel: topology not instantiated");
}
