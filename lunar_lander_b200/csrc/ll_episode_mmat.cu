// ll_episode_mmat.cu — K1 fast, tensor-core variant with the episode slots on the N side of the MMA (LL_MODE_FAST).
//
// Replaces the same reference code as the other episode kernels: train_generation / sim_task src/nn_engine_trainer.cpp:443-539,
// GameLogicCpp<T> src/game_logic.cpp:124-388 and the ma-libs feedforward call :501-502.
//
// ll_episode_mma.cu puts the 16 episode slots of a member on the M side of mma.sync.m16n8k16 (D[slots x out] = A[slots x in] *
// W^T).  A member with 8 layouts (BASELINE configs[2]) then fills half of every tile, one with a single layout a sixteenth,
// and a 50-layout member drains its last episodes in tiles that are mostly empty.  Here the product is transposed:
//     D[out x slots] = W[out x in] * X[in x slots],   M = 16 output neurons, N = 8 slots, K = 16 inputs per HMMA
//   * an N tile is 8 slots: 8 layouts fill it completely, and a warp that owns 16 slots (two N tiles) skips the HMMAs, the
//     epilogue and the environment of a tile whose 8 episodes have all ended (a warp-uniform switch to a one-tile pass);
//   * the weights are the A operand: split once per member into fp16 hi / lo and stored as ready-made A fragments,
//     one LDS.128 each per (output tile, k tile, lane), shared by the warp's N tiles; the bias is a float2 per (tile, g);
//   * activations stay in registers: the D fragment of layer l (rows = outputs g / g+8, columns = slots 2t, 2t+1), after
//     tanh and the hi/lo split, becomes the B fragment of k tile = output tile of layer l+1 (k = 2t, 2t+1 / 2t+8, 2t+9, column =
//     slot g) through one movmatrix.m8n8.trans per 8x8 block — the only extra cost of the transposed form;
//   * the lander of slot g lives, replicated, in the four lanes of quad g: each lane builds its own pair of layer-0 inputs
//     without a shuffle; the 4 outputs of a slot sit in lanes g = 0..3 of one column and are gathered with three shuffles.
// fp32 accuracy from fp16 operands exactly as in ll_episode_mma.cu: D = bias + W_hi X_hi + W_hi X_lo + W_lo X_hi, one accumulator tile.
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "ll_env.cuh"

#ifndef LL_MMAT_ONE_ACC
#define LL_MMAT_ONE_ACC 2 /* layers with at least this many output tiles keep ONE accumulator tile per output tile */
#endif
#ifndef LL_MMAT_PANEL
#define LL_MMAT_PANEL 8 /* output tiles of a layer processed together (their accumulator tiles and weight fragments live in registers) */
#endif


namespace
{
    template <int... S> struct tnet_t
    {
        static constexpr int NS = sizeof...(S);
        static constexpr int NL = NS - 1;
        static __host__ __device__ constexpr int sz(int i)
        {
            constexpr int a[NS] = {S...};
            return a[i];
        }
        static __host__ __device__ constexpr int k16(int l) { return (sz(l) + 15) & ~15; }
        static __host__ __device__ constexpr int m16(int l) { return (sz(l + 1) + 15) & ~15; }
        // shared-memory layout in 32-bit words: per layer the A fragments [m16/16][k16/16][hi, lo][32 lanes][4 words] (one word per
        // weight), then the bias [m16/16][8 (g)][2] = {b[16mt + g], b[16mt + g + 8]}.  The first layer has at most 8 inputs: its
        // fragments are those of mma.m16n8k8 (2 words per lane; k 8..15 would be padding).
        static __host__ __device__ constexpr bool half_k(int l) { return l == 0 && sz(0) <= 8; }
        static __host__ __device__ constexpr int w_words(int l) { return half_k(l) ? m16(l) * k16(l) / 2 : m16(l) * k16(l); }
        static __host__ __device__ constexpr int w_off(int l)
        {
            int o = 0;
            for (int i = 0; i < l; ++i) o += w_words(i) + m16(i);
            return o;
        }
        static __host__ __device__ constexpr int b_off(int l) { return w_off(l) + w_words(l); }
        static constexpr int total = w_off(NL);
        static __host__ __device__ constexpr bool ok() { return sz(0) <= 8 && sz(NL) <= 8; }
    };

    struct mmat_args
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

    
    __device__ __forceinline__ void mma_w(float (&c)[4], const uint4& w, uint32_t b0, uint32_t b1)
    {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(w.x), "r"(w.y), "r"(w.z), "r"(w.w), "r"(b0), "r"(b1));
    }

    __device__ __forceinline__ uint32_t movm_t(uint32_t v)
    {
        uint32_t r;
        asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(r) : "r"(v));
        return r;
    }

    // (x0, x1) -> half2 of the leading 11 significant bits and half2 of the remainders (see ll_episode_mma.cu: split_pair)
    __device__ __forceinline__ void split2(float x0, float x1, uint32_t& hi, uint32_t& lo)
    {
        const float h0 = __uint_as_float(__float_as_uint(x0) & 0xffffe000u), h1 = __uint_as_float(__float_as_uint(x1) & 0xffffe000u);
        const __half2 ph = __floats2half2_rn(h0, h1);
        const __half2 pl = __floats2half2_rn(x0 - h0, x1 - h1);
        hi = *reinterpret_cast<const uint32_t*>(&ph);
        lo = *reinterpret_cast<const uint32_t*>(&pl);
    }

    // m16n8k8 for the first layer (inputs 0..7 are the whole k extent): the fragment is a uint2
    __device__ __forceinline__ void mma_w(float (&c)[4], const uint2& w, uint32_t b0, uint32_t)
    {
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(w.x), "r"(w.y), "r"(b0));
    }
    template <bool HALF> struct wfrag { using type = uint4; };
    template <> struct wfrag<true> { using type = uint2; };

    // A fragment (hi: hl = 0, lo: hl = 1) of tile mt * KT + kt of layer L from a member's staged row
    template <class N, int L> __device__ __forceinline__ typename wfrag<N::half_k(L)>::type load_w(const float* __restrict__ row, int lane, int tile, int hl)
    {
        using F = typename wfrag<N::half_k(L)>::type;
        return reinterpret_cast<const F*>(row + N::w_off(L))[(tile * 2 + hl) * 32 + lane];
    }

    // One layer for the warp's NT slot tiles.  xh/xl[j][kt][2]: B fragments (hi / lo) of slot tile j, k tile kt.  Output: the next
    // layer's B fragments yh/yl[j][mt][2] (tanh applied, split, transposed), or for the output layer z[j][2] = tanh'd outputs
    // (row g, slots 2t and 2t+1).  Output tiles go in panels of PM so that the 3 * PM * NT accumulator tiles stay in registers.
    template <class N, int L, int NT>
    __device__ __forceinline__ void layer_t(const float* __restrict__ sw, int lane, int g, const uint32_t (&xh)[NT][N::k16(L) / 16][2],
                                            const uint32_t (&xl)[NT][N::k16(L) / 16][2], uint32_t (&yh)[NT][N::m16(L) / 16][2],
                                            uint32_t (&yl)[NT][N::m16(L) / 16][2], float (&z)[NT][2])
    {
        constexpr int KT = N::k16(L) / 16, MT = N::m16(L) / 16;
        constexpr int PMAX = LL_MMAT_PANEL / NT;
        constexpr int PM   = MT >= PMAX ? PMAX : MT;
        constexpr bool LAST = L == N::NL - 1;
        static_assert(MT % PM == 0, "output tiles must fill whole panels");
        const float2* B = reinterpret_cast<const float2*>(sw + N::b_off(L)) + g;
#pragma unroll
        for (int mp = 0; mp < MT; mp += PM)
        {
            // bias + W_hi*X_hi + W_hi*X_lo + W_lo*X_hi in ONE fp32 accumulator tile per (output tile, slot tile) when the layer has
            // at least LL_MMAT_ONE_ACC output tiles (per k tile the products go pass by pass, so the HMMAs of one accumulator are
            // PM * NT instructions apart); single-tile layers keep one tile per product and add them at the end.  The rule looks
            // at the layer only, so every kernel variant rounds alike.
            constexpr int NA = MT >= LL_MMAT_ONE_ACC ? 1 : 3;
            float acc[NA][PM][NT][4];
#pragma unroll
            for (int i = 0; i < PM; ++i)
            {
                const float2 b = B[(mp + i) * 8];
#pragma unroll
                for (int j = 0; j < NT; ++j)
                {
                    acc[0][i][j][0] = b.x, acc[0][i][j][1] = b.x, acc[0][i][j][2] = b.y, acc[0][i][j][3] = b.y;
#pragma unroll
                    for (int q = 1; q < NA; ++q)
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc[q][i][j][e] = 0.f;
                }
            }
#pragma unroll
            for (int kt = 0; kt < KT; ++kt)
            {
                typename wfrag<N::half_k(L)>::type wh[PM], wl[PM];
#pragma unroll
                for (int i = 0; i < PM; ++i) wh[i] = load_w<N, L>(sw, lane, (mp + i) * KT + kt, 0), wl[i] = load_w<N, L>(sw, lane, (mp + i) * KT + kt, 1);
#pragma unroll
                for (int i = 0; i < PM; ++i)
#pragma unroll
                    for (int j = 0; j < NT; ++j) mma_w(acc[1 % NA][i][j], wh[i], xl[j][kt][0], xl[j][kt][1]);
#pragma unroll
                for (int i = 0; i < PM; ++i)
#pragma unroll
                    for (int j = 0; j < NT; ++j) mma_w(acc[2 % NA][i][j], wl[i], xh[j][kt][0], xh[j][kt][1]);
#pragma unroll
                for (int i = 0; i < PM; ++i)
#pragma unroll
                    for (int j = 0; j < NT; ++j) mma_w(acc[0][i][j], wh[i], xh[j][kt][0], xh[j][kt][1]);
            }
#pragma unroll
            for (int i = 0; i < PM; ++i)
#pragma unroll
                for (int j = 0; j < NT; ++j)
                    {
                        float v[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            v[e] = ll_tanh_approx(NA == 1 ? acc[0][i][j][e] : acc[0][i][j][e] + (acc[1 % NA][i][j][e] + acc[2 % NA][i][j][e]));
                        if constexpr (LAST)
                        {
                            if (mp + i == 0) z[j][0] = v[0], z[j][1] = v[1];
                        }
                        else
                        {
                            uint32_t h01, l01, h23, l23;
                            split2(v[0], v[1], h01, l01); // row g     (output 16mt + g),     slots 2t, 2t+1
                            split2(v[2], v[3], h23, l23); // row g + 8 (output 16mt + g + 8), slots 2t, 2t+1
                            yh[j][mp + i][0] = movm_t(h01), yh[j][mp + i][1] = movm_t(h23);
                            yl[j][mp + i][0] = movm_t(l01), yl[j][mp + i][1] = movm_t(l23);
                        }
                    }
        }
    }

    template <class N, int L, int NT> struct tchain
    {
        static __device__ __forceinline__ void run(const float* __restrict__ sw, int lane, int g,
                                                   const uint32_t (&xh)[NT][N::k16(L) / 16][2], const uint32_t (&xl)[NT][N::k16(L) / 16][2],
                                                   float (&z)[NT][2])
        {
            uint32_t yh[NT][N::m16(L) / 16][2], yl[NT][N::m16(L) / 16][2];
            layer_t<N, L, NT>(sw, lane, g, xh, xl, yh, yl, z);
            if constexpr (L + 1 < N::NL) tchain<N, L + 1, NT>::run(sw, lane, g, yh, yl, z);
        }
    };

    // Stage one member row (HBM layout of ll_topo: per layer Wt[k][out_pad] then b[out_pad]) into the A-fragment layout, splitting
    // every weight into its fp16 parts on the way.  Element (k, o): tile (o >> 4, k >> 4), row r = o & 15, column c = k & 15 ->
    // lane (r & 7) * 4 + ((c & 7) >> 1), word (r >> 3) + 2 * (c >> 3), half c & 1.
    // lane / nthreads: index and number of the threads that share the staging (a warp; the straggler pass: the two warps of a pair)
    template <class N, int L>
    __device__ __forceinline__ void stage_t(const mmat_args& a, const float* __restrict__ row, float* __restrict__ sw, int lane, int nthreads = 32)
    {
        constexpr int IN = N::sz(L), OUT = N::sz(L + 1), KT = N::k16(L) / 16, M16 = N::m16(L);
        const int op      = a.out_pad[L];
        const float4* src = reinterpret_cast<const float4*>(row + a.w_off[L]);
        __half* dst       = reinterpret_cast<__half*>(sw + N::w_off(L));
        for (int i = lane; i < IN * op / 4; i += nthreads)
        {
            const float4 v = __ldg(src + i);
            const int k = (i * 4) / op, o0 = (i * 4) - k * op;
            const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int q = 0; q < 4; ++q)
            {
                const int o = o0 + q;
                if (o < OUT)
                {
                    const __half hi = __float2half_rn(fminf(fmaxf(e[q], -65504.f), 65504.f));
                    const __half lo = __float2half_rn(fminf(fmaxf(e[q] - __half2float(hi), -65504.f), 65504.f)); // unscaled: exact to 2^-25 absolute
                    const int r = o & 15, c = k & 15;
                    const int tile = (o >> 4) * KT + (k >> 4);
                    const int ln   = (r & 7) * 4 + ((c & 7) >> 1);
                    constexpr int FW = N::half_k(L) ? 2 : 4; // words per lane in a fragment
                    const int word = N::half_k(L) ? (r >> 3) : (r >> 3) + 2 * (c >> 3);
                    const int base = ((tile * 2) * 32 + ln) * FW + word; // hi fragment; the lo fragment is 32 * FW words further
                    dst[base * 2 + (c & 1)]             = hi;
                    dst[(base + 32 * FW) * 2 + (c & 1)] = lo;
                }
            }
        }
        for (int i = lane; i < M16; i += nthreads)
        { // word i of the bias block: tile i >> 4, g = (i >> 1) & 7, half i & 1 -> output 16mt + g + 8h
            const int o         = (i >> 4) * 16 + ((i >> 1) & 7) + 8 * (i & 1);
            sw[N::b_off(L) + i] = (o < OUT) ? __ldg(row + a.b_off[L] + o) : 0.f;
        }
        if constexpr (L + 1 < N::NL) stage_t<N, L + 1>(a, row, sw, lane, nthreads);
    }

    // warps per CTA (one CTA per SM): as many member rows as fit in 227 KB of shared memory, at most 16.  With one slot tile per warp
    // the register count (<= 160) never limits this; with two tiles (up to 255 registers) the count stays at 12 / 8 / what fits.
    template <class N, int NT> constexpr int t_max_warps()
    {
        const int by_smem = (227 * 1024) / (N::total * 4);
        if (NT == 1) return by_smem >= 28 ? 28 : (by_smem >= 16 ? 16 : (by_smem < 1 ? 1 : by_smem));
        return by_smem >= 12 ? 12 : (by_smem >= 8 ? 8 : (by_smem < 1 ? 1 : by_smem)); // two tiles per warp need > 128 registers
    }

    template <class N, int NT>
    __global__ void __launch_bounds__(32 * t_max_warps<N, NT>(), 1) ll_episode_mmat_kernel(const __grid_constant__ ll_env<float> env, const __grid_constant__ mmat_args a)
    {
        static_assert(N::ok(), "topology not supported by the mma episode kernel");
        extern __shared__ __align__(16) float smem_all[];
        float* sw = smem_all + (threadIdx.x >> 5) * N::total; // this warp's weight buffer
        const int lane   = threadIdx.x & 31;
        const int g      = lane >> 2, t = lane & 3; // quad g holds slot g of every tile, replicated in its four lanes
        const bool owner = t == 0;
        const int L      = a.n_layouts;
        constexpr int SLOTS = 8 * NT;

        for (int i = lane; i < N::total; i += 32) sw[i] = 0.f; // k / output padding stays zero for the whole kernel
        __syncwarp();

        for (;;)
        {
            unsigned long long member = 0;
            if (lane == 0) member = (unsigned long long)atomicAdd(a.queue, 1u);
            member = __shfl_sync(0xffffffffu, member, 0);
            if (member >= a.n_members) break;
            __syncwarp();
            stage_t<N, 0>(a, a.params + member * (size_t)a.p_dev, sw, lane);
            __syncwarp();

            ll_lander<float> p[NT];
            float cx[NT], lpad[NT], lpad_r[NT];
            int steps[NT], layout[NT], dir[NT];
            double acc[NT];
            bool active[NT];
            double fit_sum = 0.0;
            unsigned int steps_sum = 0, succ0 = 0, succ1 = 0;
            int next_layout = L < SLOTS ? L : SLOTS;
#pragma unroll
            for (int j = 0; j < NT; ++j)
            {
                const int slot = j * 8 + g;
                p[j]  = ll_lander<float>{0.f, 0.f, 0.f, 0.f, 0.f};
                cx[j] = 0.f, lpad[j] = 0.f, lpad_r[j] = 0.f, steps[j] = 0, layout[j] = -1, dir[j] = 0, acc[j] = 0.0, active[j] = false;
                if (slot < L)
                {
                    layout[j]                  = a.layouts[slot].next; // start order: longest expected flight first
                    const ll_layout<float> lay = a.layouts[layout[j]];
                    dir[j]                     = lay.dir;
                    cx[j] = lay.center_x, lpad[j] = lay.lpad_x1, lpad_r[j] = lay.lpad_right;
                    p[j]      = ll_lander<float>{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
                    active[j] = true;
                }
            }

            for (;;)
            {
                unsigned mask = 0;
#pragma unroll
                for (int j = 0; j < NT; ++j)
                    if (__ballot_sync(0xffffffffu, active[j]) != 0u) mask |= 1u << j;
                if (mask == 0u) break;

                // ---- state -> B fragment of layer 0: word 0 holds inputs (2t, 2t+1) of slot g, k >= 8 is padding
                uint32_t xh[NT][1][2], xl[NT][1][2];
#pragma unroll
                for (int j = 0; j < NT; ++j)
                {
                    xh[j][0][0] = 0u, xl[j][0][0] = 0u, xh[j][0][1] = 0u, xl[j][0][1] = 0u;
                    if ((mask >> j) & 1u)
                    {
                        float st[5];
                        ll_get_state<float, LL_MODE_FAST>(env, p[j], cx[j], st, 1);
                        const float e0 = t == 0 ? st[0] : (t == 1 ? st[2] : (t == 2 ? st[4] : 0.f));
                        const float e1 = t == 0 ? st[1] : (t == 1 ? st[3] : 0.f);
                        split2(e0, e1, xh[j][0][0], xl[j][0][0]);
                    }
                }
                // ---- forward pass, activations chained through registers
                float z[NT][2];
#pragma unroll
                for (int j = 0; j < NT; ++j) z[j][0] = 0.f, z[j][1] = 0.f;
                if (NT == 1 || mask == (1u << NT) - 1u) tchain<N, 0, NT>::run(sw, lane, g, xh, xl, z);
                else
                { // NT == 2 and the 8 episodes of one tile have all ended: a single-tile pass over the live one (warp-uniform)
                    const bool first = (mask & 1u) != 0u;
                    uint32_t sh[1][1][2], sl[1][1][2];
                    float z1[1][2] = {{0.f, 0.f}};
#pragma unroll
                    for (int w = 0; w < 2; ++w) sh[0][0][w] = first ? xh[0][0][w] : xh[NT - 1][0][w], sl[0][0][w] = first ? xl[0][0][w] : xl[NT - 1][0][w];
                    tchain<N, 0, 1>::run(sw, lane, g, sh, sl, z1);
                    if (first) z[0][0] = z1[0][0], z[0][1] = z1[0][1];
                    else z[NT - 1][0] = z1[0][0], z[NT - 1][1] = z1[0][1];
                }

                unsigned fin[NT];
                bool done[NT], landed[NT], success[NT];
                float dist[NT];
                unsigned any_fin = 0;
#pragma unroll
                for (int j = 0; j < NT; ++j)
                {
                    fin[j] = 0u, done[j] = false, landed[j] = false, success[j] = false, dist[j] = 0.f;
                    if ((mask >> j) & 1u)
                    {
                        // ---- argmax over outputs 0..3 of a slot (first maximum wins): lane (g', t') holds output g' of slots 2t', 2t'+1;
                        //      lanes g' = 0 gather outputs 1..3 from lanes +4, +8, +12 and are the only ones read below
                        unsigned packed = 0;
#pragma unroll
                        for (int e = 0; e < 2; ++e)
                        {
                            const float o0 = z[j][e];
                            const float o1 = __shfl_xor_sync(0xffffffffu, o0, 4), o2 = __shfl_xor_sync(0xffffffffu, o0, 8);
                            const float o3 = __shfl_xor_sync(0xffffffffu, o0, 12);
                            float best = o0;
                            unsigned idx = 0;
                            if (o1 > best) best = o1, idx = 1;
                            if (o2 > best) best = o2, idx = 2;
                            if (o3 > best) best = o3, idx = 3;
                            packed |= idx << (4 * e);
                        }
                        const unsigned pk = __shfl_sync(0xffffffffu, packed, g >> 1); // lane (0, g >> 1) holds slots g & ~1, g | 1
                        const int action  = (pk >> (4 * (g & 1))) & 15;

                        // ---- lander update + penalties (src/game_logic.cpp:295-320 through ll_env.cuh), identical in the quad's lanes
                        if (active[j])
                        {
                            const float sp = ll_env_step<float, LL_MODE_FAST>(env, p[j], action, steps[j] == 0, cx[j], lpad[j], lpad_r[j], &landed[j],
                                                                              &success[j], &dist[j]);
                            acc[j] += (double)sp;
                            ++steps[j];
                            done[j] = landed[j] || steps[j] >= a.max_steps;
                        }
                        fin[j] = __ballot_sync(0xffffffffu, done[j] && owner); // bit 4g' = slot g' of this tile ended
                        any_fin |= fin[j];
                    }
                }
                if (any_fin)
                { // some episode ended: book it, hand the slot the member's next layout (slots served in (tile, g) order)
                    int before = 0;
#pragma unroll
                    for (int j = 0; j < NT; ++j)
                    {
                        if (done[j])
                        {
                            const float term = ll_terminal_penalty<float, LL_MODE_FAST>(env, p[j], steps[j], landed[j], success[j], dist[j]);
                            const double fit = (double)term + acc[j];
                            if (owner)
                            {
                                fit_sum += fit, steps_sum += (unsigned)steps[j];
                                if (success[j]) (dir[j] == 0 ? succ0 : succ1) += 1u;
                                if (a.ep_steps)
                                {
                                    const int meta = steps[j] | (landed[j] ? (1 << 28) : 0) | ((landed[j] && !success[j]) ? (1 << 29) : 0) |
                                                     (success[j] ? (1 << 30) : 0);
                                    a.ep_steps[member * (size_t)L + layout[j]] = meta;
                                    a.ep_fit[member * (size_t)L + layout[j]]   = fit;
                                }
                            }
                            const int nl = next_layout + before + __popc(fin[j] & ((1u << (4 * g)) - 1u));
                            if (nl < L)
                            { // reset(spad_x1, lpad_x1): src/game_logic.cpp:143-154
                                layout[j]                  = a.layouts[nl].next;
                                const ll_layout<float> lay = a.layouts[layout[j]];
                                dir[j]                     = lay.dir;
                                cx[j] = lay.center_x, lpad[j] = lay.lpad_x1, lpad_r[j] = lay.lpad_right;
                                p[j]     = ll_lander<float>{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
                                steps[j] = 0, acc[j] = 0.0;
                            }
                            else active[j] = false;
                        }
                        before += __popc(fin[j]);
                    }
                    next_layout += before;
                }
            }

            // ---- member totals (src/nn_engine_trainer.cpp:518-530; fast mode: free summation order); only t == 0 lanes carry sums
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

    // ---- straggler pass behind the tcgen05 tile kernel (ll_episode_tc05.cu), slots on the N side: a member's handed-over episodes
    // (lander state, step count, accumulated step penalty: ll_resume_rec) finish here.  TWO warps share one staged member: each flies an
    // 8-column tile and both refill their columns from the member's waiting records (a shared-memory counter) as episodes end — the
    // pass is a handful of long dependent chains per SM (shared memory holds 6 members of 32_128_32), so a second chain per member is
    // worth more than the 45 % fewer HMMAs per step an 8-column tile costs against the 16-row M-side tile (ll_episode_mma.cu).  The
    // pair then closes the member: every layout's result, the tile kernel's and ours, summed in layout order from the per-episode
    // arrays (src/nn_engine_trainer.cpp:518-530).
    struct mmat_resume_args
    {
        mmat_args m;
        ll_resume_args r;
    };
    struct pair_ctl
    {
        unsigned int item;
        int next_rec;
    };

    // member buffers per CTA of the straggler pass: what fits in shared memory, at most 6 (12 warps x <= 168 registers, barrier ids 1..6)
    template <class N> __host__ __device__ constexpr int resume_pairs()
    {
        const int by_smem = (227 * 1024 - 256) / (N::total * 4);
        return by_smem >= 6 ? 6 : (by_smem < 1 ? 1 : by_smem);
    }
    // WPM warps share one member
    template <int WPM> __device__ __forceinline__ void pair_sync(int pair) { asm volatile("bar.sync %0, %1;" : : "r"(1 + pair), "n"(32 * WPM) : "memory"); }

    template <class N, int WPM>
    __global__ void __launch_bounds__(32 * WPM * resume_pairs<N>(), 1) ll_episode_mmat_resume_kernel(const __grid_constant__ ll_env<float> env,
                                                                                                const __grid_constant__ mmat_resume_args ra)
    {
        const mmat_args& a = ra.m;
        extern __shared__ __align__(16) float smem_all[];
        const int warp = threadIdx.x >> 5, pair = warp / WPM, half = warp % WPM; // half: which of the member's WPM tiles
        float* sw      = smem_all + pair * N::total;
        pair_ctl* ctl  = reinterpret_cast<pair_ctl*>(smem_all + resume_pairs<N>() * N::total) + pair;
        const int lane   = threadIdx.x & 31;
        const int g      = lane >> 2, t = lane & 3; // quad g holds column (slot) g, replicated in its four lanes
        const bool owner = t == 0;
        const int LT     = a.n_layouts;
        for (int i = lane + 32 * half; i < N::total; i += 32 * WPM) sw[i] = 0.f;
        const unsigned int n_items = *ra.r.n;
        for (;;)
        {
            if (half == 0 && lane == 0) ctl->item = atomicAdd(a.queue, 1u);
            pair_sync<WPM>(pair);
            const unsigned int item = ctl->item;
            if (item >= n_items) break;
            const unsigned long long member = ra.r.list[item];
            const int cnt                   = (int)ra.r.count[member];
            const ll_resume_rec* recs       = ra.r.rec + member * (size_t)ra.r.stride;
            stage_t<N, 0>(a, a.params + member * (size_t)a.p_dev, sw, lane + 32 * half, 32 * WPM);
            if (half == 0 && lane == 0) ctl->next_rec = cnt < 8 * WPM ? cnt : 8 * WPM;
            pair_sync<WPM>(pair);
            ll_lander<float> p{0.f, 0.f, 0.f, 0.f, 0.f};
            float cx = 0.f, lpad = 0.f, lpad_r = 0.f;
            int steps = 0, layout = 0;
            double acc  = 0.0;
            bool active = false;
            auto take = [&](int i) {
                const ll_resume_rec r      = recs[i];
                layout                     = r.layout;
                const ll_layout<float> lay = a.layouts[layout];
                cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
                p     = ll_lander<float>{r.x, r.y, r.vx, r.vy, r.fuel};
                steps = r.steps, acc = r.acc;
            };
            if (8 * half + g < cnt)
            {
                take(8 * half + g);
                active = true;
            }
            while (__ballot_sync(0xffffffffu, active) != 0u)
            {
                uint32_t xh[1][1][2], xl[1][1][2];
                xh[0][0][1] = 0u, xl[0][0][1] = 0u;
                {
                    float st[5];
                    ll_get_state<float, LL_MODE_FAST>(env, p, cx, st, 1);
                    const float e0 = t == 0 ? st[0] : (t == 1 ? st[2] : (t == 2 ? st[4] : 0.f));
                    const float e1 = t == 0 ? st[1] : (t == 1 ? st[3] : 0.f);
                    split2(e0, e1, xh[0][0][0], xl[0][0][0]);
                }
                float z[1][2] = {{0.f, 0.f}};
                tchain<N, 0, 1>::run(sw, lane, g, xh, xl, z);
                unsigned packed = 0;
#pragma unroll
                for (int e = 0; e < 2; ++e)
                { // argmax over outputs 0..3 of a slot (first maximum wins), as in the kernel above
                    const float o0 = z[0][e];
                    const float o1 = __shfl_xor_sync(0xffffffffu, o0, 4), o2 = __shfl_xor_sync(0xffffffffu, o0, 8);
                    const float o3 = __shfl_xor_sync(0xffffffffu, o0, 12);
                    float best = o0;
                    unsigned idx = 0;
                    if (o1 > best) best = o1, idx = 1;
                    if (o2 > best) best = o2, idx = 2;
                    if (o3 > best) best = o3, idx = 3;
                    packed |= idx << (4 * e);
                }
                const unsigned pk = __shfl_sync(0xffffffffu, packed, g >> 1);
                const int action  = (pk >> (4 * (g & 1))) & 15;
                bool done = false, landed = false, success = false;
                float dist = 0.f;
                if (active)
                {
                    const float sp = ll_env_step<float, LL_MODE_FAST>(env, p, action, steps == 0, cx, lpad, lpad_r, &landed, &success, &dist);
                    acc += (double)sp;
                    ++steps;
                    done = landed || steps >= a.max_steps;
                }
                const unsigned fin = __ballot_sync(0xffffffffu, done && owner); // bit 4g' = column g' ended
                if (fin)
                {
                    int base = 0; // the pair's two tiles draw the waiting records from one counter
                    if (lane == 0) base = atomicAdd(&ctl->next_rec, __popc(fin));
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (done)
                    {
                        const float term = ll_terminal_penalty<float, LL_MODE_FAST>(env, p, steps, landed, success, dist);
                        if (owner)
                        {
                            a.ep_steps[member * (size_t)LT + layout] = steps | (landed ? (1 << 28) : 0) | ((landed && !success) ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
                            a.ep_fit[member * (size_t)LT + layout]   = (double)term + acc;
                        }
                        const int nr = base + __popc(fin & ((1u << (4 * g)) - 1u));
                        if (nr < cnt) take(nr); // the column takes the member's next waiting episode
                        else active = false;
                    }
                }
            }
            pair_sync<WPM>(pair); // both tiles are done, their results visible to the closing lane
            if (half == 0 && lane == 0)
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

    int t_env_int(const char* name, int fallback)
    {
        const char* v = getenv(name);
        return v && *v ? atoi(v) : fallback;
    }

    template <class N, int NT> int launch_mmat_t(ll_ctx* ctx, const ll_episode_args& h)
    {
        const ll_topo& tp = ctx->topo;
        mmat_args a{};
        a.params = static_cast<const float*>(h.params), a.layouts = static_cast<const ll_layout<float>*>(h.layouts);
        a.n_layouts = h.n_layouts, a.max_steps = h.max_steps, a.n_members = h.n_members;
        a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps, a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit, a.queue = h.queue;
        a.p_dev = tp.p_dev;
        for (int l = 0; l < tp.n_layers; ++l) a.w_off[l] = tp.w_off[l], a.b_off[l] = tp.b_off[l], a.out_pad[l] = tp.out_pad[l];
        auto kern = ll_episode_mmat_kernel<N, NT>;
        int max_smem = 0;
        LL_CUDA(ctx, cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device));
        int warps = std::min<int>(t_max_warps<N, NT>(), max_smem / (N::total * 4));
        warps     = std::min(warps, t_env_int("LL_MMA_WARPS_PER_SM", warps));
        if (warps < 1) return ll_fail(ctx, LL_EINVAL, "mmat episode kernel: the member row does not fit in shared memory");
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

    // one member per warp with one 8-slot tile (up to 8 layouts) or two
    template <class N> int launch_mmat(ll_ctx* ctx, const ll_episode_args& h)
    {
        const int nt = t_env_int("LL_MMAT_TILES", h.n_layouts <= 8 ? 1 : 2);
        return nt == 1 ? launch_mmat_t<N, 1>(ctx, h) : launch_mmat_t<N, 2>(ctx, h);
    }

    using tnet_16_16        = tnet_t<5, 16, 16, 4>;
    using tnet_16_32        = tnet_t<5, 16, 32, 4>;
    using tnet_8_16_8       = tnet_t<5, 8, 16, 8, 4>;
    using tnet_16_64_32     = tnet_t<5, 16, 64, 32, 4>;
    using tnet_32_128_32    = tnet_t<5, 32, 128, 32, 4>;
    using tnet_8_16_32_16_8 = tnet_t<5, 8, 16, 32, 16, 8, 4>;

    template <class N> bool t_same_net(const ll_topo& tp)
    {
        if (tp.n_layers != N::NL || tp.in[0] != N::sz(0)) return false;
        for (int l = 0; l < N::NL; ++l)
            if (tp.out[l] != N::sz(l + 1)) return false;
        return true;
    }
} // namespace

namespace
{
    template <class N, int WPM> int launch_mmat_resume_t(ll_ctx* ctx, const ll_episode_args& h, const ll_resume_args& r)
    {
        const ll_topo& tp = ctx->topo;
        mmat_resume_args ra{};
        mmat_args& a = ra.m;
        a.params = static_cast<const float*>(h.params), a.layouts = static_cast<const ll_layout<float>*>(h.layouts);
        a.n_layouts = h.n_layouts, a.max_steps = h.max_steps, a.n_members = h.n_members;
        a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps, a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit;
        a.queue = h.queue + 2; // the tile kernel uses words 0 (work counter) and 1 (error flag)
        a.p_dev = tp.p_dev;
        for (int l = 0; l < tp.n_layers; ++l) a.w_off[l] = tp.w_off[l], a.b_off[l] = tp.b_off[l], a.out_pad[l] = tp.out_pad[l];
        ra.r      = r;
        auto kern = ll_episode_mmat_resume_kernel<N, WPM>;
        int max_smem = 0;
        LL_CUDA(ctx, cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device));
        constexpr int pairs = resume_pairs<N>();
        const size_t smem   = (size_t)pairs * N::total * 4 + pairs * sizeof(pair_ctl);
        if ((size_t)max_smem < smem) return ll_fail(ctx, LL_EINVAL, "mmat resume kernel: the member row does not fit in shared memory");
        LL_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ll_env<float> env;
        ll_make_env<float>(ctx->env_cfg, 1, &env);
        LL_CUDA(ctx, cudaMemsetAsync(a.queue, 0, sizeof(unsigned int), ctx->stream));
        const unsigned long long ctas_needed = (h.n_members + pairs - 1) / pairs;
        const unsigned int grid = (unsigned int)std::max<unsigned long long>(1, std::min<unsigned long long>((unsigned long long)ctx->sm_count, ctas_needed));
        kern<<<grid, 32 * WPM * pairs, smem, ctx->stream>>>(env, ra);
        LL_CUDA(ctx, cudaGetLastError());
        return LL_OK;
    }
} // namespace

int ll_launch_episode_mmat_resume(ll_ctx* ctx, const ll_episode_args& a, const ll_resume_args& r)
{
    const ll_topo& tp = ctx->topo;
    // two warps per staged member (one: 204.0 ms on the 32 768-member sample, two: 201.4, three: 202.4 at 92 registers)
    if (t_same_net<tnet_32_128_32>(tp)) return launch_mmat_resume_t<tnet_32_128_32, 2>(ctx, a, r);
    if (t_same_net<tnet_16_64_32>(tp)) return launch_mmat_resume_t<tnet_16_64_32, 2>(ctx, a, r);
    return ll_fail(ctx, LL_EINVAL, "mmat resume kernel: topology not instantiated");
}

// Same coverage as ll_mma_applicable (the sample_runs topologies with tanh).  Which of the two tensor-core kernels runs is
// decided in ll_episode.cu (ll_launch_episode): LL_FAST_KERNEL = mmat | mma | ffma forces one.
bool ll_mmat_applicable(const ll_ctx* ctx, int n_layouts)
{
    const ll_topo& tp = ctx->topo;
    if (tp.activation != LL_ACT_TANH) return false;
    const bool known = t_same_net<tnet_16_16>(tp) || t_same_net<tnet_16_32>(tp) || t_same_net<tnet_8_16_8>(tp) || t_same_net<tnet_16_64_32>(tp) ||
                       t_same_net<tnet_32_128_32>(tp) || t_same_net<tnet_8_16_32_16_8>(tp);
    return known && n_layouts >= 1;
}

int ll_launch_episode_mmat(ll_ctx* ctx, const ll_episode_args& a)
{
    const ll_topo& tp = ctx->topo;
    if (t_same_net<tnet_32_128_32>(tp)) return launch_mmat<tnet_32_128_32>(ctx, a);
    if (t_same_net<tnet_16_64_32>(tp)) return launch_mmat<tnet_16_64_32>(ctx, a);
    if (t_same_net<tnet_8_16_32_16_8>(tp)) return launch_mmat<tnet_8_16_32_16_8>(ctx, a);
    if (t_same_net<tnet_8_16_8>(tp)) return launch_mmat<tnet_8_16_8>(ctx, a);
    if (t_same_net<tnet_16_32>(tp)) return launch_mmat<tnet_16_32>(ctx, a);
    if (t_same_net<tnet_16_16>(tp)) return launch_mmat<tnet_16_16>(ctx, a);
    return ll_fail(ctx, LL_EINVAL, "mmat episode kernel: topology not instantiated");
}
