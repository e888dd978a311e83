// ll_episode_tc05.cu — K1 fast on the Blackwell tensor path: tcgen05.mma with TMEM accumulators AND TMEM activations (LL_MODE_FAST).
//
// Replaces the same reference code as the other episode kernels: train_generation / sim_task src/nn_engine_trainer.cpp:443-539,
// GameLogicCpp<T> src/game_logic.cpp:124-388 and the ma-libs feedforward call :501-502.
//
// Mapping (measured building blocks in profiles/r02_tc05_probe_*.jsonl):
//   * all layouts of a member share its weights, so one forward pass over its S = 64 episode slots is D[S x out] = X[S x in] * W^T per
//     layer.  The SLOTS are the M dimension of tcgen05.mma (M = 64), the layer's output neurons the N dimension, the weights the
//     B operand in shared memory (K-major, no swizzle, staged and split into f16 hi / lo once per member), and the activations the
//     A operand read FROM TENSOR MEMORY (a shared-memory A operand costs 4 KB of reads per instruction: 37-41 cycles against 19-26);
//   * one THREAD owns one slot for the whole episode: TMEM lane = slot.  It reads its row of D with tcgen05.ld, adds the bias,
//     applies tanh.approx, splits the activation into f16 hi + lo and writes them back into its own TMEM lane with tcgen05.st as
//     the next layer's A operand, in place over the columns it has just read — activations never touch shared memory or HBM,
//     and no thread ever needs another thread's data (no shuffles, no shared-memory exchange).  The lander state lives in the
//     thread's registers; argmax, physics, penalties and the next state follow the last layer in the same thread;
//   * an M = 64 tile occupies 16 of the 32 lanes of every TMEM sub-partition (measured: rows 0-15 -> lanes 0-15, 16-31 -> 32-47, ...),
//     and a second tile addressed with lane offset 16 fills the other half.  A QUAD of 4 warps (128 threads) therefore flies TWO
//     members at once in the same TMEM columns: lanes 0-15 of each warp are slots of member A, lanes 16-31 of member B;
//   * fp32 accuracy from 11-bit operands as in the mma.sync kernels: x = hi + lo, D = X_hi*W_hi + X_lo*W_hi + X_hi*W_lo (+ bias in the
//     epilogue), fp32 accumulation in TMEM.  The first layer (5 inputs) packs hi, the constant 1 (bias) and lo into ONE K = 16 block:
//     two instructions instead of three;
//   * 3 quads per CTA (6 members in flight per SM, 168 TMEM columns per quad, 36.7 KB of shared memory per member) hide the
//     mma -> commit -> mbarrier -> tcgen05.ld hop (~400 cycles, four per step) behind the other quads' epilogues; the tensor pipe is
//     asynchronous, so unlike the legacy HMMA (which blocks its sub-partition's issue port) the epilogue warps keep issuing;
//   * every wait is bounded: a descriptor or barrier mistake ends the kernel with an error flag instead of hanging the GPU.
// Supported: tanh networks whose hidden widths are multiples of 16 (<= 256), <= 5 inputs, <= 8 outputs, 17..64 layouts per member
// (one slot per layout, all layouts of a member in flight at once).
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "ll_env.cuh"
#include "ll_tc05.cuh"

#ifndef LL_TC05_CUT_DEFAULT
#define LL_TC05_CUT_DEFAULT 24
#endif
#ifndef LL_TC05_RESUME_MMAT_DEFAULT
#define LL_TC05_RESUME_MMAT_DEFAULT true
#endif

namespace
{
    template <int... S> struct tnet
    {
        static constexpr int NS = sizeof...(S);
        static constexpr int NL = NS - 1;
        static __host__ __device__ constexpr int sz(int i)
        {
            constexpr int a[NS] = {S...};
            return a[i];
        }
        // K of layer l (multiple of 16) and N of layer l (multiple of 16 for hidden layers: it is the next layer's K; 8 for the output)
        static __host__ __device__ constexpr int kpad(int l) { return l == 0 ? 16 : (sz(l) + 15) & ~15; }
        static __host__ __device__ constexpr int npad(int l) { return l == NL - 1 ? 8 : (sz(l + 1) + 15) & ~15; }
        // TMEM columns of a quad: even layers' D at column 0, odd layers' D behind them, the packed state (X0, 8 columns) last.
        // The next layer's A operand overwrites D in place, so two regions + X0 are enough.
        static __host__ __device__ constexpr int cols_even()
        {
            int m = 0;
            for (int l = 0; l < NL; l += 2) m = npad(l) > m ? npad(l) : m;
            return m;
        }
        static __host__ __device__ constexpr int cols_odd()
        {
            int m = 0;
            for (int l = 1; l < NL; l += 2) m = npad(l) > m ? npad(l) : m;
            return m;
        }
        static __host__ __device__ constexpr int dcol(int l) { return (l & 1) ? cols_even() : 0; }
        static constexpr int x0col    = cols_even() + cols_odd();
        static constexpr int cols     = x0col + 8;
        // shared memory per member, bytes: per layer W_hi then W_lo ([npad x kpad] f16, canonical K-major core-matrix layout), then the
        // biases of layers >= 1 as fp32 [npad]
        static __host__ __device__ constexpr int w_bytes(int l) { return npad(l) * kpad(l) * 2; }
        static __host__ __device__ constexpr int w_off(int l)
        {
            int o = 0;
            for (int i = 0; i < l; ++i) o += 2 * w_bytes(i);
            return o;
        }
        static __host__ __device__ constexpr int b_off(int l)
        {
            int o = w_off(NL);
            for (int i = 1; i < l; ++i) o += npad(i) * 4;
            return o;
        }
        static constexpr int member_bytes = (b_off(NL) + 127) & ~127;
        static __host__ __device__ constexpr bool ok()
        {
            if (sz(0) > 5 || sz(NL) > 8 || NL < 2) return false;
            for (int l = 0; l < NL; ++l)
                if (npad(l) > 256) return false;
            return true;
        }
    };

    struct tc_args
    {
        const float* params;
        const ll_layout<float>* layouts;
        int n_layouts, max_steps;
        unsigned long long n_members;
        double *fitness, *lr, *steps;
        int* ep_steps;
        double* ep_fit;
        unsigned int* queue;
        int* error; // set to 1 when a bounded wait timed out
        ll_resume_args resume; // hand-over of a member's last episodes to the 16-row mma.sync kernel (rec == nullptr: off)
        int cut;               // a member leaves the tile when at most this many of its episodes are still running
        int p_dev;
        int w_off[LL_MAX_LAYERS], b_off[LL_MAX_LAYERS], out_pad[LL_MAX_LAYERS]; // the HBM row layout (ll_topo)
    };

    constexpr int QUAD_THREADS = 128;
    constexpr int SLOTS        = 64; // per member: one M = 64 tile

    // per-quad control block in shared memory
    struct quad_ctl
    {
        unsigned long long mbar;          // tcgen05.commit target
        unsigned long long member[2];     // member flown by half 0 / 1 (~0ull: none)
        unsigned int live[4];             // per warp: ballot of slots whose episode is running (low 16 bits half 0, high 16 half 1)
        int exhausted[2];                 // the queue is empty for this half
        int pad_[2];
        double ep_fit[2][SLOTS];          // per-slot episode results of the member in flight
        int ep_meta[2][SLOTS];            // steps | flags << 28
    };

    // byte offset of element (n, k) of a K-major no-swizzle operand with `rows` rows: 8 x 16-byte core matrices, 8-row groups contiguous
    __device__ __forceinline__ int op_off(int n, int k, int rows) { return (n >> 3) * 128 + (k >> 3) * (rows * 16) + (n & 7) * 16 + (k & 7) * 2; }

    // packed fp32 pairs (FADD2: one issue slot for two IEEE round-to-nearest adds — same results as two FADDs)
#ifndef LL_TC05_PACKED
#define LL_TC05_PACKED 1
#endif
    __device__ __forceinline__ void add2(float& a0, float& a1, float b0, float b1, bool negate_b = false)
    {
#if LL_TC05_PACKED
        unsigned long long a, b, c;
        asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
        asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(b0), "f"(b1));
        if (negate_b) asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(c) : "l"(a), "l"(b));
        else asm("add.rn.f32x2 %0, %1, %2;" : "=l"(c) : "l"(a), "l"(b));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(a0), "=f"(a1) : "l"(c));
#else
        a0 = negate_b ? a0 - b0 : a0 + b0, a1 = negate_b ? a1 - b1 : a1 + b1;
#endif
    }

    __device__ __forceinline__ void split_pair(float x0, float x1, uint32_t& hi, uint32_t& lo)
    { // hi = leading 11 significant bits (exact in f16), lo = x - hi (exact in fp32, |lo| < 2^-11 |x|): see ll_episode_mma.cu
        const float h0 = __uint_as_float(__float_as_uint(x0) & 0xffffe000u), h1 = __uint_as_float(__float_as_uint(x1) & 0xffffe000u);
        const __half2 ph = __floats2half2_rn(h0, h1);
        add2(x0, x1, h0, h1, true);
        const __half2 pl = __floats2half2_rn(x0, x1);
        hi = *reinterpret_cast<const uint32_t*>(&ph);
        lo = *reinterpret_cast<const uint32_t*>(&pl);
    }

    // Stage one member row (HBM layout of ll_topo: per layer Wt[k][out_pad] then b[out_pad]) into the B-operand layout of its half,
    // splitting every weight into f16 hi / lo.  Run by the 64 threads of the half (t = 0..63), once per member (~10^5 forward passes).
    template <class N, int L> __device__ __forceinline__ void stage_layer(const tc_args& a, const float* __restrict__ row, unsigned char* __restrict__ sm, int t)
    {
        constexpr int IN = N::sz(L), OUT = N::sz(L + 1), NP = N::npad(L);
        const int op = a.out_pad[L];
        __half* whi  = reinterpret_cast<__half*>(sm + N::w_off(L));
        __half* wlo  = reinterpret_cast<__half*>(sm + N::w_off(L) + N::w_bytes(L));
        for (int i = t; i < IN * OUT; i += SLOTS)
        {
            const int k = i / OUT, o = i - k * OUT;
            const float v   = fminf(fmaxf(__ldg(row + a.w_off[L] + k * op + o), -65504.f), 65504.f);
            const __half hi = __float2half_rn(v);
            const __half lo = __float2half_rn(v - __half2float(hi));
            whi[op_off(o, k, NP) >> 1] = hi;
            wlo[op_off(o, k, NP) >> 1] = lo;
            if (L == 0) whi[op_off(o, k + 6, NP) >> 1] = hi; // multiplies the lo part of the state, packed at k = 6..10
        }
        if (L == 0)
        { // the bias rides in the K block: input k = 5 is the constant 1
            for (int o = t; o < OUT; o += SLOTS)
            {
                const float v   = fminf(fmaxf(__ldg(row + a.b_off[0] + o), -65504.f), 65504.f);
                const __half hi = __float2half_rn(v);
                whi[op_off(o, 5, NP) >> 1] = hi;
                wlo[op_off(o, 5, NP) >> 1] = __float2half_rn(v - __half2float(hi));
            }
        }
        else
        {
            float* sb = reinterpret_cast<float*>(sm + N::b_off(L));
            for (int o = t; o < NP; o += SLOTS) sb[o] = o < OUT ? __ldg(row + a.b_off[L] + o) : 0.f;
        }
        if constexpr (L + 1 < N::NL) stage_layer<N, L + 1>(a, row, sm, t);
    }

    template <class N> __host__ __device__ constexpr int chunk_of(int l) { return (N::npad(l) % 32 == 0) ? 32 : (N::npad(l) % 16 == 0 ? 16 : 8); }

    // tcgen05.mma chain of layer L for one half: D[64 x npad] (+)= X[64 x kpad] * W^T, three split products per K = 16 block.
    // tq = TMEM address of the quad's column 0 at the half's lane offset (0 or 16); sm = shared address of the member's region.
    template <class N, int L> __device__ __forceinline__ void issue_layer(uint32_t tq, uint32_t sm)
    {
        constexpr int KP = N::kpad(L), NP = N::npad(L);
        constexpr uint32_t idesc = tc05::idesc_f16(64, NP, 0, 0);
        constexpr uint32_t LBO = NP * 16, SBO = 128;
        const uint32_t d   = tq + N::dcol(L);
        const uint32_t whi = sm + N::w_off(L), wlo = whi + N::w_bytes(L);
        if constexpr (L == 0)
        { // one K block holds hi(state), 1, lo(state): W_hi covers X_hi*W_hi + bias_hi + X_lo*W_hi, W_lo adds X_hi*W_lo + bias_lo
            const uint32_t x0 = tq + N::x0col;
            tc05::mma_f16_ts(d, x0, tc05::smem_desc(whi, LBO, SBO), idesc, 0u);
            tc05::mma_f16_ts(d, x0, tc05::smem_desc(wlo, LBO, SBO), idesc, 1u);
        }
        else
        {
            constexpr int CH  = chunk_of<N>(L - 1);
            const uint32_t xin = tq + N::dcol(L - 1); // the activations overwrote D of the previous layer in place, chunk by chunk
#pragma unroll
            for (int kc = 0; kc < KP / 16; ++kc)
            {
                const int k0 = 16 * kc;
                const uint32_t a_hi = xin + (k0 / CH) * CH + (k0 % CH) / 2, a_lo = a_hi + CH / 2;
                const uint64_t bhi = tc05::smem_desc(whi + 2 * kc * LBO, LBO, SBO), blo = tc05::smem_desc(wlo + 2 * kc * LBO, LBO, SBO);
                tc05::mma_f16_ts(d, a_lo, bhi, idesc, kc > 0 ? 1u : 0u);
                tc05::mma_f16_ts(d, a_hi, blo, idesc, 1u);
                tc05::mma_f16_ts(d, a_hi, bhi, idesc, 1u);
            }
        }
    }

    // issue layers L.. of both running halves, then commit; called by ONE elected lane of the quad's first warp
    template <class N, int L> __device__ __forceinline__ void issue_both(uint32_t tq, uint32_t sm0, bool run0, bool run1, uint32_t mbar)
    { // one half after the other (alternating the halves K block by K block measured 3 % slower: 169 vs 163 ms)
        if (run0) issue_layer<N, L>(tq, sm0);
        if (run1) issue_layer<N, L>(tq + (16u << 16), sm0 + N::member_bytes);
        tc05::commit(mbar);
    }

    template <int CH> __device__ __forceinline__ void ld_chunk(uint32_t taddr, uint32_t (&v)[CH])
    {
        if constexpr (CH == 32) tc05::ld_x32(taddr, v);
        else if constexpr (CH == 16) tc05::ld_x16(taddr, v);
        else tc05::ld_x8(taddr, v);
    }
    template <int W> __device__ __forceinline__ void st_words(uint32_t taddr, const uint32_t (&v)[W])
    {
        if constexpr (W == 16) tc05::st_x16(taddr, v);
        else if constexpr (W == 8) tc05::st_x8(taddr, v);
        else tc05::st_x4(taddr, v);
    }

    // Hidden layer L, one thread = one slot: D row -> + bias -> tanh.approx -> f16 hi / lo -> back into the same TMEM columns as the
    // next layer's A operand (chunk c: hi words at column c*CH, lo words at c*CH + CH/2).
    template <class N, int L> __device__ __forceinline__ void epilogue_hidden(uint32_t tt, const unsigned char* __restrict__ sm)
    {
        constexpr int NP = N::npad(L), CH = chunk_of<N>(L);
        const float4* sb = reinterpret_cast<const float4*>(sm + N::b_off(L));
#pragma unroll
        for (int c = 0; c < NP / CH; ++c)
        {
            uint32_t v[CH];
            ld_chunk<CH>(tt + N::dcol(L) + c * CH, v);
            tc05::wait_ld();
            uint32_t hi[CH / 2], lo[CH / 2];
#pragma unroll
            for (int j = 0; j < CH / 4; ++j)
            {
                float z0 = __uint_as_float(v[4 * j + 0]), z1 = __uint_as_float(v[4 * j + 1]), z2 = __uint_as_float(v[4 * j + 2]), z3 = __uint_as_float(v[4 * j + 3]);
                if constexpr (L > 0)
                { // (layer 0's bias rides in its K block)
                    const float4 b = sb[(c * CH) / 4 + j];
                    add2(z0, z1, b.x, b.y);
                    add2(z2, z3, b.z, b.w);
                }
                const float x0 = ll_tanh_approx(z0), x1 = ll_tanh_approx(z1), x2 = ll_tanh_approx(z2), x3 = ll_tanh_approx(z3);
                split_pair(x0, x1, hi[2 * j], lo[2 * j]);
                split_pair(x2, x3, hi[2 * j + 1], lo[2 * j + 1]);
            }
            st_words<CH / 2>(tt + N::dcol(L) + c * CH, hi);
            st_words<CH / 2>(tt + N::dcol(L) + c * CH + CH / 2, lo);
        }
    }

    // the packed first-layer input of a slot: k 0..4 = hi(state), k 5 = 1 (bias), k 6..10 = lo(state), k 11..15 = 0
    __device__ __forceinline__ void store_state(uint32_t taddr, const float (&st)[5])
    {
        uint32_t w[8], h01, l01, h23, l23, h4, l4;
        split_pair(st[0], st[1], h01, l01);
        split_pair(st[2], st[3], h23, l23);
        split_pair(st[4], 1.0f, h4, l4); // hi(1) = 1, lo(1) = 0
        w[0] = h01, w[1] = h23, w[2] = h4;                    // k 0..5
        w[3] = l01, w[4] = l23, w[5] = l4 & 0xffffu;          // k 6..10, k 11 = 0
        w[6] = 0u, w[7] = 0u;
        tc05::st_x8(taddr, w);
    }

    template <class N> __host__ __device__ constexpr int quads_per_cta()
    {
        int q = 512 / N::cols;
        const int by_smem = (227 * 1024 - 64) / (2 * N::member_bytes + (int)sizeof(quad_ctl));
        q = q < by_smem ? q : by_smem;
        return q > 3 ? 3 : (q < 1 ? 1 : q);
    }

    template <class N> struct step_loop;

    template <class N>
    __global__ void __launch_bounds__(QUAD_THREADS* quads_per_cta<N>(), 1) ll_episode_tc05_kernel(const __grid_constant__ ll_env<float> env, const __grid_constant__ tc_args a)
    {
        static_assert(N::ok(), "topology not supported by the tcgen05 episode kernel");
        constexpr int NQ = quads_per_cta<N>();
        extern __shared__ __align__(1024) unsigned char smem[];
        __shared__ uint32_t tmem_base_s;
        quad_ctl* ctls = reinterpret_cast<quad_ctl*>(smem + 2 * NQ * N::member_bytes);
        const int tid = threadIdx.x, q = tid / QUAD_THREADS, wq = (tid >> 5) & 3, lane = tid & 31;
        const int h = lane >> 4, slot = 16 * wq + (lane & 15);
        quad_ctl& ctl      = ctls[q];
        unsigned char* smq = smem + (size_t)(2 * q) * N::member_bytes; // the quad's two member regions
        unsigned char* smh = smq + h * N::member_bytes;                // this thread's half
        const uint32_t mbar = tc05::smem_u32(&ctl.mbar);

        for (int i = tid; i < 2 * NQ * N::member_bytes / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u; // K / N padding stays zero
        if ((tid & (QUAD_THREADS - 1)) == 0)
        {
            tc05::mbar_init(mbar, 1);
            ctl.member[0] = ctl.member[1] = ~0ull;
            ctl.live[0] = ctl.live[1] = ctl.live[2] = ctl.live[3] = 0u;
            ctl.exhausted[0] = ctl.exhausted[1] = 0;
            tc05::fence_mbar_init();
        }
        if (tid < 32) tc05::tmem_alloc(tc05::smem_u32(&tmem_base_s), 512);
        tc05::fence_before_sync();
        __syncthreads();
        tc05::fence_after_sync();
        const uint32_t tbase = tmem_base_s;
        const uint32_t tq    = tbase + (uint32_t)(q * N::cols);        // MMA addresses: lane offset 0 (half 0) / 16 (half 1) added at issue
        const uint32_t tt    = tq + ((uint32_t)(32 * wq) << 16);       // this warp's sub-partition: thread i <-> lane 32 wq + i
        const uint32_t smq_a = tc05::smem_u32(smq);

        // ---- slot state (thread = slot)
        ll_lander<float> p{0.f, 0.f, 0.f, 0.f, 0.f};
        float cx = 0.f, lpad = 0.f, lpad_r = 0.f;
        int steps = 0;
        double acc = 0.0;
        bool active = false, failed = false;
        unsigned long long cur_member = ~0ull; // the member this thread's half is flying (kept in a register: the leader replaces ctl.member)
        uint32_t phase = 0;
        const int L    = a.n_layouts; // <= SLOTS: slot s flies layout s

        for (;;)
        {
            // ---- who is still flying?  (this barrier is also the one in front of the first layer's MMAs)
            const unsigned bal = __ballot_sync(0xffffffffu, active);
            if (lane == 0) ctl.live[wq] = bal;
            tc05::wait_st();
            tc05::fence_before_sync();
            tc05::named_bar_sync(1 + q, QUAD_THREADS);
            tc05::fence_after_sync();
            const unsigned lv = ctl.live[0] | ctl.live[1] | ctl.live[2] | ctl.live[3];
            (void)lv;
            // A tile runs until its member's LONGEST episode ends; measured on the benchmark population only 40-54 % of the 64 rows are
            // live on average, and the last 16 episodes of a member hold a tenth of its steps but a third of its tile time
            // (scripts/liveness.py).  So when at most `cut` episodes are left the half hands them to the 16-row mma.sync kernel
            // (one warp = one member x 16 slots, launched behind this kernel) and takes the next member.
            const int c0 = __popc(ctl.live[0] & 0xffffu) + __popc(ctl.live[1] & 0xffffu) + __popc(ctl.live[2] & 0xffffu) + __popc(ctl.live[3] & 0xffffu);
            const int c1 = __popc(ctl.live[0] >> 16) + __popc(ctl.live[1] >> 16) + __popc(ctl.live[2] >> 16) + __popc(ctl.live[3] >> 16);
            const bool leave0 = !ctl.exhausted[0] && c0 <= a.cut, leave1 = !ctl.exhausted[1] && c1 <= a.cut;
            const bool mine_live = !(h ? leave1 : leave0);
            if (leave0 || leave1)
            { // ---- rare path (once per member): a half has finished (or hands over) its member -> book it, take the next one from the queue
                const int mine_left = h ? c1 : c0;
                if (!mine_live && active)
                { // my episode moves: rank among the half's running slots, in slot order
                    const unsigned long long member = cur_member;
                    int rank = 0;
                    for (int w = 0; w < wq; ++w) rank += __popc(h ? (ctl.live[w] >> 16) : (ctl.live[w] & 0xffffu));
                    rank += __popc((h ? (ctl.live[wq] >> 16) : (ctl.live[wq] & 0xffffu)) & ((1u << (lane & 15)) - 1u));
                    ll_resume_rec r;
                    r.x = p.x, r.y = p.y, r.vx = p.vx, r.vy = p.vy, r.fuel = p.fuel, r.layout = slot, r.steps = steps, r.pad_ = 0, r.acc = acc;
                    a.resume.rec[member * (size_t)a.resume.stride + rank] = r;
                    active = false;
                }
                if (slot == 0 && !mine_live)
                {
                    const unsigned long long done = cur_member;
                    if (done != ~0ull && mine_left > 0)
                    { // the straggler pass finishes the member (it sums every layout's result, ours included, from the per-episode arrays)
                        a.resume.count[done] = (unsigned int)mine_left;
                        a.resume.list[atomicAdd(a.resume.n, 1u)] = (unsigned int)done;
                    }
                    else if (done != ~0ull)
                    { // member totals, summed in layout order (src/nn_engine_trainer.cpp:518-530; deterministic, shard-independent)
                        double f = 0.0, st = 0.0;
                        float bonus0 = 0.f, bonus1 = 0.f;
                        for (int s = 0; s < L; ++s)
                        {
                            const int meta = ctl.ep_meta[h][s];
                            f += ctl.ep_fit[h][s], st += (double)(meta & 0x0fffffff);
                            if (meta & (1 << 30)) (a.layouts[s].dir == 0 ? bonus0 : bonus1) += env.tp_landed_lr_bonus;
                        }
                        const double lr = (double)(bonus1 < bonus0 ? bonus1 : bonus0);
                        a.fitness[done] = f - lr, a.lr[done] = lr, a.steps[done] = st;
                    }
                    const unsigned long long next = (unsigned long long)atomicAdd(a.queue, 1u);
                    if (next < a.n_members) ctl.member[h] = next;
                    else ctl.member[h] = ~0ull, ctl.exhausted[h] = 1;
                }
                tc05::named_bar_sync(1 + q, QUAD_THREADS);
                if (!mine_live && ctl.member[h] != ~0ull && !ctl.exhausted[h])
                { // the 64 threads of the half: stage the new member's weights, start one episode per layout
                    const unsigned long long member = ctl.member[h];
                    cur_member                      = member;
                    stage_layer<N, 0>(a, a.params + member * (size_t)a.p_dev, smh, slot);
                    steps = 0, acc = 0.0, active = slot < L;
                    const ll_layout<float> lay = a.layouts[active ? slot : 0];
                    cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
                    p = ll_lander<float>{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
                }
                { // every thread (re)writes its first-layer input: tcgen05.st is warp-collective, so it sits outside the divergent branch
                    float st[5];
                    ll_get_state<float, LL_MODE_FAST>(env, p, cx, st, 1);
                    store_state(tt + N::x0col, st);
                }
                tc05::wait_st();
                tc05::fence_proxy_async(); // the staged weights (generic proxy) -> visible to the tensor core's operand reads
                tc05::fence_before_sync();
                tc05::named_bar_sync(1 + q, QUAD_THREADS);
                tc05::fence_after_sync();
            }
            const bool run0 = ctl.member[0] != ~0ull, run1 = ctl.member[1] != ~0ull;
            if (!run0 && !run1) break; // both halves drained
            if (step_loop<N>::run(env, a, ctl, tq, tt, smq_a, smh, mbar, q, wq, h, slot, run0, run1, phase, p, cx, lpad, lpad_r, steps, acc, active, cur_member))
            {
                failed = true;
                break;
            }
        }
        if (failed && (tid & (QUAD_THREADS - 1)) == 0) *a.error = 1;
        tc05::fence_before_sync();
        __syncthreads();
        if (tid < 32) tc05::tmem_dealloc(tbase, 512);
    }

    // One lander step of the quad: the layers in order, each = barrier -> tcgen05.mma chain (one elected lane) -> commit -> wait -> epilogue.
    // Returns true when a bounded wait timed out.
    // What was tried to overlap a quad's instructions with its own epilogue, measured on B200 and removed (all bit-identical to this
    // loop): issuing a 128-wide layer as two N halves with separate commits (-0.7 %: 2 x N = 64 costs 84 pipe cycles against 74);
    // additionally issuing the next layer's first K half between the two epilogue halves behind one more quad barrier (-3.3 %);
    // issuing the next layer chunk by chunk from the epilogue (-7 %: a quad-wide bar.sync waits for the slowest of four warps that
    // share their MUFU pipes with two other quads); alternating the two members K block by K block (-3 %).
    template <class N> struct step_loop
    {
        template <int L>
        static __device__ __forceinline__ bool layers(const ll_env<float>& env, const tc_args& a, quad_ctl& ctl, uint32_t tq, uint32_t tt, uint32_t smq_a,
                                                      const unsigned char* smh, uint32_t mbar, int q, int wq, int h, int slot, bool run0, bool run1,
                                                      uint32_t& phase, ll_lander<float>& p, float cx, float lpad, float lpad_r, int& steps, double& acc,
                                                      bool& active, unsigned long long member)
        {
            if constexpr (L > 0)
            { // the activations this layer reads were written by every thread of the quad: order them before the issue
                tc05::wait_st();
                tc05::fence_before_sync();
                tc05::named_bar_sync(1 + q, QUAD_THREADS);
                tc05::fence_after_sync();
            }
            if (wq == 0)
            {
                if (tc05::elect_one()) issue_both<N, L>(tq, smq_a, run0, run1, mbar);
                __syncwarp();
            }
            if (!tc05::mbar_wait(mbar, phase)) return true;
            phase ^= 1u;
            tc05::fence_after_sync();
            if constexpr (L + 1 < N::NL)
            {
                epilogue_hidden<N, L>(tt, smh);
                return layers<L + 1>(env, a, ctl, tq, tt, smq_a, smh, mbar, q, wq, h, slot, run0, run1, phase, p, cx, lpad, lpad_r, steps, acc, active, member);
            }
            else
            { // ---- output layer: tanh, argmax (first maximum wins), lander update + penalties, next state
                constexpr int NO = N::sz(N::NL);
                uint32_t v[8];
                tc05::ld_x8(tt + N::dcol(L), v);
                tc05::wait_ld();
                const float* sb = reinterpret_cast<const float*>(smh + N::b_off(L));
                float best = 0.f;
                int action = 0;
#pragma unroll
                for (int j = 0; j < NO; ++j)
                {
                    const float z = ll_tanh_approx(__uint_as_float(v[j]) + sb[j]);
                    if (j == 0 || z > best) best = z, action = j;
                }
                if (active)
                {
                    bool landed = false, success = false;
                    float dist  = 0.f;
                    const float sp = ll_env_step<float, LL_MODE_FAST>(env, p, action, steps == 0, cx, lpad, lpad_r, &landed, &success, &dist);
                    acc += (double)sp;
                    ++steps;
                    if (landed || steps >= a.max_steps)
                    { // the episode ends: book it (src/nn_engine_trainer.cpp:514-519); the slot idles until the member's last one ends
                        const float term = ll_terminal_penalty<float, LL_MODE_FAST>(env, p, steps, landed, success, dist);
                        const double fit = (double)term + acc;
                        const int meta   = steps | (landed ? (1 << 28) : 0) | ((landed && !success) ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
                        ctl.ep_fit[h][slot] = fit, ctl.ep_meta[h][slot] = meta;
                        if (a.ep_steps)
                        {
                            a.ep_steps[member * (size_t)a.n_layouts + slot] = meta;
                            a.ep_fit[member * (size_t)a.n_layouts + slot]   = fit;
                        }
                        active = false;
                    }
                }
                float st[5];
                ll_get_state<float, LL_MODE_FAST>(env, p, cx, st, 1);
                store_state(tt + N::x0col, st);
                return false;
            }
        }

        static __device__ __forceinline__ bool run(const ll_env<float>& env, const tc_args& a, quad_ctl& ctl, uint32_t tq, uint32_t tt, uint32_t smq_a,
                                                   const unsigned char* smh, uint32_t mbar, int q, int wq, int h, int slot, bool run0, bool run1,
                                                   uint32_t& phase, ll_lander<float>& p, float cx, float lpad, float lpad_r, int& steps, double& acc,
                                                   bool& active, unsigned long long member)
        {
            return layers<0>(env, a, ctl, tq, tt, smq_a, smh, mbar, q, wq, h, slot, run0, run1, phase, p, cx, lpad, lpad_r, steps, acc, active, member);
        }
    };

    // LL_TC05_CUT: a member leaves its 64-row tile for the straggler pass when at most this many episodes are still running
    // (0 = never; default 24: three 8-column tiles' worth for the two warps that share the member in ll_episode_mmat_resume_kernel)
    int env_cut()
    {
        const char* v = getenv("LL_TC05_CUT");
        return v && *v ? atoi(v) : LL_TC05_CUT_DEFAULT;
    }

    template <class N> int launch_tc05_t(ll_ctx* ctx, const ll_episode_args& h)
    {
        const ll_topo& tp = ctx->topo;
        constexpr int NQ  = quads_per_cta<N>();
        tc_args a{};
        a.params = static_cast<const float*>(h.params), a.layouts = static_cast<const ll_layout<float>*>(h.layouts);
        a.n_layouts = h.n_layouts, a.max_steps = h.max_steps, a.n_members = h.n_members;
        a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps, a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit, a.queue = h.queue;
        a.p_dev = tp.p_dev;
        for (int l = 0; l < tp.n_layers; ++l) a.w_off[l] = tp.w_off[l], a.b_off[l] = tp.b_off[l], a.out_pad[l] = tp.out_pad[l];
        a.error = reinterpret_cast<int*>(h.queue + 1); // the work counter's neighbour: zeroed with it, read back by ll_read_results
        // hand-over to the 16-row mma.sync kernel (needs the per-episode arrays: that is where the two kernels' results meet)
        const int cut = h.ep_steps && h.ep_fit ? std::max(0, std::min(env_cut(), LL_RESUME_MAX)) : 0;
        a.cut         = cut;
        if (cut > 0)
        {
            a.resume.stride = cut;
            const size_t rec_bytes = (size_t)h.n_members * cut * sizeof(ll_resume_rec), cnt_bytes = (size_t)h.n_members * sizeof(unsigned int);
            const size_t need = rec_bytes + 2 * cnt_bytes + 64;
            if (ctx->resume_cap < need)
            {
                if (ctx->d_resume) cudaFree(ctx->d_resume);
                ctx->d_resume = nullptr, ctx->resume_cap = 0;
                LL_CUDA(ctx, cudaMalloc(&ctx->d_resume, need));
                ctx->resume_cap = need;
            }
            char* base      = static_cast<char*>(ctx->d_resume);
            a.resume.rec    = reinterpret_cast<ll_resume_rec*>(base);
            a.resume.count  = reinterpret_cast<unsigned int*>(base + rec_bytes);
            a.resume.list   = reinterpret_cast<unsigned int*>(base + rec_bytes + cnt_bytes);
            a.resume.n      = reinterpret_cast<unsigned int*>(base + rec_bytes + 2 * cnt_bytes);
            LL_CUDA(ctx, cudaMemsetAsync(a.resume.n, 0, sizeof(unsigned int), ctx->stream));
        }
        auto kern         = ll_episode_tc05_kernel<N>;
        const size_t smem = (size_t)2 * NQ * N::member_bytes + (size_t)NQ * sizeof(quad_ctl);
        LL_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const unsigned long long ctas_needed = (h.n_members + 2 * NQ - 1) / (2 * NQ);
        const unsigned int grid = (unsigned int)std::max<unsigned long long>(1, std::min<unsigned long long>((unsigned long long)ctx->sm_count, ctas_needed));
        ll_env<float> env;
        ll_make_env<float>(ctx->env_cfg, 1, &env);
        LL_CUDA(ctx, cudaMemsetAsync(h.queue, 0, 2 * sizeof(unsigned int), ctx->stream));
        kern<<<grid, QUAD_THREADS * NQ, smem, ctx->stream>>>(env, a);
        LL_CUDA(ctx, cudaGetLastError());
        if (cut > 0 && !getenv("LL_TC05_DEBUG_NORESUME")) // debug switch: time the tile kernel alone (results incomplete)
        { // LL_TC05_RESUME = mma (16-row tile, slots on M) | mmat (8-column tile, slots on N)
            const char* which = getenv("LL_TC05_RESUME");
            const bool n_side = which && *which ? !strcmp(which, "mmat") : LL_TC05_RESUME_MMAT_DEFAULT;
            return n_side ? ll_launch_episode_mmat_resume(ctx, h, a.resume) : ll_launch_episode_mma_resume(ctx, h, a.resume);
        }
        return LL_OK;
    }

    using tnet_32_128_32 = tnet<5, 32, 128, 32, 4>;
    using tnet_16_64_32  = tnet<5, 16, 64, 32, 4>;

    template <class N> bool same_net(const ll_topo& tp)
    {
        if (tp.n_layers != N::NL || tp.in[0] != N::sz(0)) return false;
        for (int l = 0; l < N::NL; ++l)
            if (tp.out[l] != N::sz(l + 1)) return false;
        return true;
    }
} // namespace

// The tcgen05 kernel flies one slot per layout with all of a member's layouts in flight at once, so it wants many layouts per
// member (17..64) and a population that fills 6 members per SM.  LL_FAST_KERNEL=tc05 forces it where it applies, =mma / mmat / ffma
// force the others; the default is decided in ll_episode.cu from measurements.
bool ll_tc05_applicable(const ll_ctx* ctx, int n_layouts)
{
    const ll_topo& tp = ctx->topo;
    if (tp.activation != LL_ACT_TANH || ctx->precision != 32 || ctx->mode != LL_MODE_FAST) return false;
    if (n_layouts < 17 || n_layouts > SLOTS) return false;
    return same_net<tnet_32_128_32>(tp) || same_net<tnet_16_64_32>(tp);
}

int ll_launch_episode_tc05(ll_ctx* ctx, const ll_episode_args& a)
{
    const ll_topo& tp = ctx->topo;
    if (same_net<tnet_32_128_32>(tp)) return launch_tc05_t<tnet_32_128_32>(ctx, a);
    if (same_net<tnet_16_64_32>(tp)) return launch_tc05_t<tnet_16_64_32>(ctx, a);
    return ll_fail(ctx, LL_EINVAL, "tcgen05 episode kernel: topology not instantiated");
}
