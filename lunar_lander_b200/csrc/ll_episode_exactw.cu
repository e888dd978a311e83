// ll_episode_exactw.cu — K1 EXACT for SMALL populations (BASELINE.json configs[0]: the reference's own scale, 16_16, 100 members x 50
// start positions, fp64): ONE WARP = ONE EPISODE, the neurons of a layer across the lanes, the topology compiled in.
//
// Replaces the same reference code as ll_episode.cu: train_generation / sim_task src/nn_engine_trainer.cpp:443-539,
// GameLogicCpp<T> src/game_logic.cpp:124-388 and the ma-libs feedforward call :501-502 — in LL_MODE_EXACT arithmetic, i.e. bit for
// bit the oracle's results (every multiply and add individually rounded, ascending-k dot products from 0, bias last, ll_tanh).
//
// Why another exact kernel.  With 5 000 episodes for 148 SMs a generation lasts as long as ONE episode's chain of dependent steps
// (the longest flies max_steps = 5 000 of them), so what counts is the latency of a step, not throughput.  The general kernel
// (ll_episode.cu) takes its topology at run time: ncu on configs[0] shows ~1 200 warp instructions per step of which 24 % are fp64
// arithmetic — the rest is loop control and address arithmetic around the run-time sizes (IMAD 18 %, UMOV 8 %, ISETP + BRA 12 %) — at
// one instruction per ~5 cycles of a lone warp (profiles/r02_ncu_exact_f64_c1_100x50_q16.txt).  Here the layer sizes are template
// parameters: the k loops are unrolled with immediate offsets, lane o computes neuron o (o + 32, ... for wider layers), the weights
// come straight from the member's row in global memory (read-only path, L1-resident: 3.5 KB for 16_16 fp64; all of a step's loads
// are independent of the arithmetic chain), the activations of the previous layer are read as shared-memory broadcasts, and the
// lander is replicated in every lane (SIMT: the redundant lanes cost nothing).  A neuron's own arithmetic is that of
// ll_episode.cu::layer_tile, so the results are bit-identical (tests/test_gpu_parity.py::test_exact_warp_per_episode_kernel_bit_exact).
// Work items (member, layout) come from a global queue, longest expected flight first (ll_layout::next); each episode's result goes
// to the per-episode arrays and the LAST episode of a member to finish sums all of them in layout order, the reference's order
// (src/nn_engine_trainer.cpp:518-530).  Chosen when all work items fit on the GPU at once (about two waves at most); larger
// populations use the tiled kernel, whose lanes each fly their own episode.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "ll_env.cuh"

namespace
{
    template <int... S> struct xnet
    {
        static constexpr int NS = sizeof...(S);
        static constexpr int NL = NS - 1;
        static __host__ __device__ constexpr int sz(int i)
        {
            constexpr int a[NS] = {S...};
            return a[i];
        }
        // the device row layout of ll_build_topo (ll_host.cpp): per layer Wt[k][out_pad] then b[out_pad], out_pad = out rounded up to 8
        static __host__ __device__ constexpr int out_pad(int l) { return (sz(l + 1) + 7) & ~7; }
        static __host__ __device__ constexpr int w_off(int l)
        {
            int o = 0;
            for (int i = 0; i < l; ++i) o += (sz(i) + 1) * out_pad(i);
            return o;
        }
        static __host__ __device__ constexpr int b_off(int l) { return w_off(l) + sz(l) * out_pad(l); }
        static constexpr int p_dev = w_off(NL);
        // two activation buffers, alternating layer by layer: buffer 0 holds the inputs of the even layers, buffer 1 of the odd ones
        static __host__ __device__ constexpr int width(int parity)
        {
            int m = 8;
            for (int i = parity; i < NS; i += 2) m = sz(i) > m ? sz(i) : m;
            return m;
        }
    };

    struct xw_args
    {
        const void* params;
        const void* layouts;
        int n_layouts, max_steps;
        unsigned long long n_members;
        double *fitness, *lr, *steps;
        int* ep_steps;
        double* ep_fit;
        unsigned int* queue;
        unsigned int* part_count; // [n_members] finished episodes of the member, zero before the launch
        int activation;
    };

    constexpr int XW_WARPS = 4;

    // layers L..NL-1: ain -> aout -> ain ...; returns the buffer that holds the outputs
    template <typename T, class N, int L> __device__ __forceinline__ const T* xw_layers(const T* __restrict__ row, int act, int lane, T* ain, T* aout)
    {
        using O = ll_ops<T, LL_MODE_EXACT>;
        constexpr int IN = N::sz(L), OUT = N::sz(L + 1), OP = N::out_pad(L);
        const T* w = row + N::w_off(L);
        const T* b = row + N::b_off(L);
#pragma unroll
        for (int o0 = 0; o0 < OUT; o0 += 32)
        {
            const int o = o0 + lane;
            if (OUT - o0 >= 32 || o < OUT)
            {
                T acc = T(0);
#pragma unroll
                for (int k = 0; k < IN; ++k) acc = O::madd(acc, __ldg(w + k * OP + o), ain[k]);
                T v[1] = {O::add(acc, __ldg(b + o))};
                O::template act_n<1>(v, act);
                aout[o] = v[0];
            }
        }
        __syncwarp();
        if constexpr (L + 1 < N::NL) return xw_layers<T, N, L + 1>(row, act, lane, aout, ain);
        else return aout;
    }

    template <typename T, class N>
    __global__ void __launch_bounds__(32 * XW_WARPS) ll_episode_exactw_kernel(const __grid_constant__ ll_env<T> env, const __grid_constant__ xw_args a)
    {
        using O = ll_ops<T, LL_MODE_EXACT>;
        __shared__ T s_act[XW_WARPS][N::width(0) + N::width(1)];
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        T* buf0 = s_act[warp];
        T* buf1 = buf0 + N::width(0);
        const ll_layout<T>* layouts = static_cast<const ll_layout<T>*>(a.layouts);
        const int L                 = a.n_layouts;
        const unsigned long long n_items = a.n_members * (unsigned long long)L;
        for (;;)
        {
            unsigned long long item = 0;
            if (lane == 0) item = (unsigned long long)atomicAdd(a.queue, 1u);
            item = __shfl_sync(0xffffffffu, item, 0);
            if (item >= n_items) break;
            // layout-major, longest expected flight first: every member's long episodes start before anybody's short ones
            const unsigned long long member = item % a.n_members;
            const int layout                = layouts[(int)(item / a.n_members)].next;
            const ll_layout<T> lay          = layouts[layout];
            const T* row                    = static_cast<const T*>(a.params) + member * (size_t)N::p_dev;
            ll_lander<T> p{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
            const T cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
            int steps  = 0;
            double acc = 0.0;
            for (;;)
            {
                ll_get_state<T, LL_MODE_EXACT>(env, p, cx, buf0, 1); // every lane writes the same five values
                __syncwarp();
                const T* out = xw_layers<T, N, 0>(row, a.activation, lane, buf0, buf1);
                int action   = 0;
                T vbest      = out[0];
#pragma unroll
                for (int j = 1; j < N::sz(N::NL); ++j)
                {
                    const T v = out[j];
                    if (v > vbest) vbest = v, action = j; // first maximum wins
                }
                __syncwarp(); // the outputs are read before the next step's state overwrites the buffers
                bool landed = false, success = false;
                T dist      = T(0);
                const T sp  = ll_env_step<T, LL_MODE_EXACT>(env, p, action, steps == 0, cx, lpad, lpad_r, &landed, &success, &dist);
                acc         = __dadd_rn(acc, (double)sp);
                ++steps;
                if (landed || steps >= a.max_steps)
                {
                    const T term = ll_terminal_penalty<T, LL_MODE_EXACT>(env, p, steps, landed, success, dist);
                    if (lane == 0)
                    {
                        const bool crashed = landed && !success;
                        a.ep_steps[member * (size_t)L + layout] = steps | (landed ? (1 << 28) : 0) | (crashed ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
                        a.ep_fit[member * (size_t)L + layout]   = __dadd_rn((double)term, acc);
                    }
                    break;
                }
            }
            // ---- the member's last episode to finish books the member: all layouts in layout order (src/nn_engine_trainer.cpp:518-530)
            if (lane == 0)
            {
                __threadfence(); // this episode's result is visible device-wide before the arrival is counted
                if (atomicAdd(a.part_count + member, 1u) == (unsigned int)L - 1u)
                {
                    __threadfence();
                    a.part_count[member]         = 0u;
                    const volatile int* g_meta   = a.ep_steps + member * (size_t)L;
                    const volatile double* g_fit = a.ep_fit + member * (size_t)L;
                    double total = 0.0, total_steps = 0.0;
                    T bonus0 = T(0), bonus1 = T(0);
                    for (int l = 0; l < L; ++l)
                    {
                        const int meta = g_meta[l];
                        total          = __dadd_rn(total, g_fit[l]);
                        total_steps    = __dadd_rn(total_steps, (double)(meta & 0x0fffffff));
                        if (meta & (1 << 30))
                        {
                            if (layouts[l].dir == 0) bonus0 = O::add(bonus0, env.tp_landed_lr_bonus);
                            else bonus1 = O::add(bonus1, env.tp_landed_lr_bonus);
                        }
                    }
                    const double lr   = (double)((bonus1 < bonus0) ? bonus1 : bonus0);
                    a.fitness[member] = __dsub_rn(total, lr);
                    a.lr[member]      = lr;
                    a.steps[member]   = total_steps;
                }
            }
            __syncwarp();
        }
    }

    template <class N> bool same_net(const ll_topo& tp)
    {
        if (tp.n_layers != N::NL || tp.in[0] != N::sz(0) || tp.p_dev != N::p_dev) return false;
        for (int l = 0; l < N::NL; ++l)
            if (tp.out[l] != N::sz(l + 1) || tp.w_off[l] != N::w_off(l) || tp.b_off[l] != N::b_off(l)) return false;
        return true;
    }

    using xnet_16_16        = xnet<5, 16, 16, 4>;
    using xnet_16_32        = xnet<5, 16, 32, 4>;
    using xnet_8_16_8       = xnet<5, 8, 16, 8, 4>;
    using xnet_16_64_32     = xnet<5, 16, 64, 32, 4>;
    using xnet_32_128_32    = xnet<5, 32, 128, 32, 4>;
    using xnet_8_16_32_16_8 = xnet<5, 8, 16, 32, 16, 8, 4>;

    template <typename T, class N> int resident_warps(const ll_ctx* ctx)
    {
        int per_sm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ll_episode_exactw_kernel<T, N>, 32 * XW_WARPS, 0) != cudaSuccess) return 0;
        return per_sm * XW_WARPS * ctx->sm_count;
    }
    template <typename T> int resident_warps_t(const ll_ctx* ctx)
    {
        const ll_topo& tp = ctx->topo;
        if (same_net<xnet_16_16>(tp)) return resident_warps<T, xnet_16_16>(ctx);
        if (same_net<xnet_16_32>(tp)) return resident_warps<T, xnet_16_32>(ctx);
        if (same_net<xnet_8_16_8>(tp)) return resident_warps<T, xnet_8_16_8>(ctx);
        if (same_net<xnet_16_64_32>(tp)) return resident_warps<T, xnet_16_64_32>(ctx);
        if (same_net<xnet_32_128_32>(tp)) return resident_warps<T, xnet_32_128_32>(ctx);
        if (same_net<xnet_8_16_32_16_8>(tp)) return resident_warps<T, xnet_8_16_32_16_8>(ctx);
        return 0;
    }

    template <typename T, class N> int launch_xw(ll_ctx* ctx, const ll_episode_args& h)
    {
        xw_args a{};
        a.params = h.params, a.layouts = h.layouts, a.n_layouts = h.n_layouts, a.max_steps = h.max_steps, a.n_members = h.n_members;
        a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps, a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit, a.queue = h.queue;
        a.activation             = ctx->topo.activation;
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
        LL_CUDA(ctx, cudaMemsetAsync(h.queue, 0, sizeof(unsigned int), ctx->stream));
        const unsigned long long items = h.n_members * (unsigned long long)h.n_layouts;
        const unsigned long long want  = (items + XW_WARPS - 1) / XW_WARPS;
        const unsigned long long fit   = (unsigned long long)std::max(1, resident_warps<T, N>(ctx) / XW_WARPS);
        const unsigned int grid        = (unsigned int)std::max<unsigned long long>(1, std::min(want, fit));
        ll_env<T> env;
        ll_make_env<T>(ctx->env_cfg, 0, &env);
        ll_episode_exactw_kernel<T, N><<<grid, 32 * XW_WARPS, 0, ctx->stream>>>(env, a);
        LL_CUDA(ctx, cudaGetLastError());
        return LL_OK;
    }

    template <typename T> int launch_xw_t(ll_ctx* ctx, const ll_episode_args& h)
    {
        const ll_topo& tp = ctx->topo;
        if (same_net<xnet_16_16>(tp)) return launch_xw<T, xnet_16_16>(ctx, h);
        if (same_net<xnet_16_32>(tp)) return launch_xw<T, xnet_16_32>(ctx, h);
        if (same_net<xnet_8_16_8>(tp)) return launch_xw<T, xnet_8_16_8>(ctx, h);
        if (same_net<xnet_16_64_32>(tp)) return launch_xw<T, xnet_16_64_32>(ctx, h);
        if (same_net<xnet_32_128_32>(tp)) return launch_xw<T, xnet_32_128_32>(ctx, h);
        if (same_net<xnet_8_16_32_16_8>(tp)) return launch_xw<T, xnet_8_16_32_16_8>(ctx, h);
        return ll_fail(ctx, LL_EINVAL, "exact warp-per-episode kernel: topology not instantiated");
    }
} // namespace

// LL_EXACT_WARP = 1 / 0 forces / forbids the kernel where it applies (exact mode, no trace, per-episode arrays, an instantiated
// topology, layouts that carry their start order); default: when all (member, layout) work items fit on the GPU within two waves.
bool ll_exactw_applicable(const ll_ctx* ctx, const ll_episode_args& a)
{
    if (ctx->mode != LL_MODE_EXACT || a.trace != nullptr || !a.ep_steps || !a.ep_fit || a.n_layouts < 1) return false;
    const int warps = ctx->precision == 64 ? resident_warps_t<double>(ctx) : resident_warps_t<float>(ctx);
    if (warps <= 0) return false;
    const char* force = getenv("LL_EXACT_WARP");
    if (force && *force) return atoi(force) != 0;
    if (getenv("LL_EXACT_Q") || getenv("LL_EXACT_CHUNKS")) return false; // switches of the tiled kernel: the caller wants that one
    return a.n_members * (unsigned long long)a.n_layouts <= 2ull * (unsigned long long)warps;
}

int ll_launch_episode_exactw(ll_ctx* ctx, const ll_episode_args& a)
{
    return ctx->precision == 64 ? launch_xw_t<double>(ctx, a) : launch_xw_t<float>(ctx, a);
}
