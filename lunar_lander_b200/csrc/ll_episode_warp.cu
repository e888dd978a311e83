// ll_episode_warp.cu — K1 fast for FEW layouts per member (BASELINE.json configs[1]: 8_16_8, population 4096, ONE start position).
//
// Replaces the same reference code as the other episode kernels: train_generation / sim_task src/nn_engine_trainer.cpp:443-539,
// GameLogicCpp<T> src/game_logic.cpp:124-388 and the ma-libs feedforward call :501-502.
//
// With a single layout there is no weight reuse to turn the forward pass into a GEMM: the tensor-core kernels then pay a full 8- or
// 16-slot tile for one live slot (configs[1]: 1 of 8 columns, ~930 cycles per lander step per warp).  This kernel is the mapping
// BASELINE.json north_star (a) describes: ONE WARP = ONE MEMBER, the neurons of a layer across the lanes, the member's weights in
// REGISTERS (lane j keeps row j of every layer: sum of the layer widths = 37 floats for 8_16_8), activations exchanged with warp
// shuffles, the argmax taken in-warp, the lander state replicated in every lane (so the first layer needs no exchange and the
// environment step no broadcast).  fp32 FFMA on unsplit weights, tanh.approx.f32: LL_MODE_FAST arithmetic, judged on outcomes.
// A generation of configs[1] is as long as its longest episode (5000 dependent steps), so what counts is the latency of one step:
// ~70 FFMA / SHFL pairs + 36 MUFU for 8_16_8 instead of 45 HMMA + 520 other instructions.
// Supported: tanh networks whose layers are at most 32 wide (the sample_runs topologies 8_16_8, 16_16, 16_32, 8_16_32_16_8).
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "ll_env.cuh"

bool ll_warp_per_episode(const ll_ctx* ctx, unsigned long long n_members, int n_layouts, bool have_episode_arrays);

namespace
{
    template <int... S> struct wnet
    {
        static constexpr int NS = sizeof...(S);
        static constexpr int NL = NS - 1;
        static __host__ __device__ constexpr int sz(int i)
        {
            constexpr int a[NS] = {S...};
            return a[i];
        }
        static __host__ __device__ constexpr int w_at(int l) // first register of layer l's weight row
        {
            int o = 0;
            for (int i = 0; i < l; ++i) o += sz(i);
            return o;
        }
        static constexpr int n_w = w_at(NL);
        static __host__ __device__ constexpr bool ok()
        {
            for (int i = 0; i < NS; ++i)
                if (sz(i) > 32) return false;
            return sz(0) <= 5 && sz(NL) <= 4;
        }
    };

    struct warp_args
    {
        const float* params;
        const ll_layout<float>* layouts;
        int n_layouts, max_steps;
        unsigned long long n_members;
        double *fitness, *lr, *steps;
        int* ep_steps;
        double* ep_fit;
        unsigned int* queue;
        unsigned int* part_count; // [n_members] finished episodes of the member (warp-per-episode kernel), zero before the launch
        int p_dev;
        int w_off[LL_MAX_LAYERS], b_off[LL_MAX_LAYERS], out_pad[LL_MAX_LAYERS]; // the HBM row layout (ll_topo)
    };

    // How a layer's activations reach every lane.  1: through a per-warp shared-memory buffer (one STS, then K / 4 broadcast LDS.128);
    // 0: one __shfl_sync per input.  A lone warp gets a shuffle result only every ~7.5 cycles (ncu source view), and
    // a step of 8_16_32_16_8 needs 84 of them against 21 LDS.128 + 6 STS.
#ifndef LL_WARP_SMEM_EXCHANGE
#define LL_WARP_SMEM_EXCHANGE 1
#endif

    // layers L..NL-1 of the forward pass; x = this lane's activation of layer L's input (layer 0: the state, replicated in st);
    // xb: the warp's exchange buffer, 2 x 32 floats used alternately layer by layer
    template <class N, int L> __device__ __forceinline__ float forward(const float (&w)[N::n_w], const float (&b)[N::NL], const float (&st)[5], float x, float* xb)
    {
        constexpr int IN = N::sz(L);
        float acc0 = b[L], acc1 = 0.f; // two chains: the dependent FFMA latency is what a step is made of
#if LL_WARP_SMEM_EXCHANGE
        float* buf = xb + 32 * (L & 1);
        if constexpr (L > 0)
        {
            buf[threadIdx.x & 31] = x;
            __syncwarp();
        }
#pragma unroll
        for (int k4 = 0; k4 < (IN + 3) / 4; ++k4)
        {
            float4 a4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if constexpr (L > 0) a4 = reinterpret_cast<const float4*>(buf)[k4];
            const float av[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                const int k = 4 * k4 + j;
                if (k < IN)
                {
                    const float a = L == 0 ? st[k] : av[j];
                    if (k & 1) acc1 = fmaf(w[N::w_at(L) + k], a, acc1);
                    else acc0 = fmaf(w[N::w_at(L) + k], a, acc0);
                }
            }
        }
#else
#pragma unroll
        for (int k = 0; k < IN; ++k)
        {
            const float a = L == 0 ? st[k] : __shfl_sync(0xffffffffu, x, k);
            if (k & 1) acc1 = fmaf(w[N::w_at(L) + k], a, acc1);
            else acc0 = fmaf(w[N::w_at(L) + k], a, acc0);
        }
#endif
        const float y = ll_tanh_approx(acc0 + acc1);
        if constexpr (L + 1 < N::NL) return forward<N, L + 1>(w, b, st, y, xb);
        else return y;
    }

    // first maximum of the network's outputs (held by lanes 0 .. n_out - 1), computed by every lane
    template <class N> __device__ __forceinline__ int argmax_outputs(float z, float* xb)
    {
        constexpr int NO = N::sz(N::NL);
#if LL_WARP_SMEM_EXCHANGE
        static_assert(NO <= 4, "one LDS.128 holds the outputs");
        float* buf = xb + 32 * (N::NL & 1);
        buf[threadIdx.x & 31] = z;
        __syncwarp();
        const float4 o4  = reinterpret_cast<const float4*>(buf)[0];
        __syncwarp(); // every lane has the outputs before the next step's first exchange may reuse this buffer
        const float o[4] = {o4.x, o4.y, o4.z, o4.w};
        float best = o[0];
        int action = 0;
#pragma unroll
        for (int j = 1; j < NO; ++j)
            if (o[j] > best) best = o[j], action = j;
        return action;
#else
        float best = __shfl_sync(0xffffffffu, z, 0);
        int action = 0;
#pragma unroll
        for (int j = 1; j < NO; ++j)
        {
            const float zj = __shfl_sync(0xffffffffu, z, j);
            if (zj > best) best = zj, action = j;
        }
        return action;
#endif
    }

    template <class N> __global__ void __launch_bounds__(256) ll_episode_warp_kernel(const __grid_constant__ ll_env<float> env, const __grid_constant__ warp_args a)
    {
        static_assert(N::ok(), "topology not supported by the warp-per-member episode kernel");
        __shared__ __align__(16) float s_xb[8][64];
        float* xb      = s_xb[threadIdx.x >> 5];
        const int lane = threadIdx.x & 31;
        for (;;)
        {
            unsigned long long member = 0;
            if (lane == 0) member = (unsigned long long)atomicAdd(a.queue, 1u);
            member = __shfl_sync(0xffffffffu, member, 0);
            if (member >= a.n_members) break;
            // ---- the member's weights: lane j <- row j of every layer (coalesced: the HBM row is k-major, Wt[k][out_pad])
            const float* row = a.params + member * (size_t)a.p_dev;
            float w[N::n_w], b[N::NL];
#pragma unroll
            for (int l = 0; l < N::NL; ++l)
            {
                const bool mine = lane < N::sz(l + 1);
#pragma unroll
                for (int k = 0; k < N::sz(l); ++k) w[N::w_at(l) + k] = mine ? __ldg(row + a.w_off[l] + k * a.out_pad[l] + lane) : 0.f;
                b[l] = mine ? __ldg(row + a.b_off[l] + lane) : 0.f;
            }
            double fit_sum = 0.0, steps_sum = 0.0;
            float bonus0 = 0.f, bonus1 = 0.f;
            for (int layout = 0; layout < a.n_layouts; ++layout)
            { // the layouts one after the other, like sim_task (:486): this kernel is for one or two of them
                const ll_layout<float> lay = a.layouts[layout];
                ll_lander<float> p{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
                const float cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
                int steps = 0;
                double acc = 0.0;
                bool landed = false, success = false;
                float dist = 0.f;
                while (steps < a.max_steps)
                {
                    float st[5];
                    ll_get_state<float, LL_MODE_FAST>(env, p, cx, st, 1);
                    const float z    = forward<N, 0>(w, b, st, 0.f, xb);
                    const int action = argmax_outputs<N>(z, xb); // first maximum wins, every lane computes it
                    const float sp = ll_env_step<float, LL_MODE_FAST>(env, p, action, steps == 0, cx, lpad, lpad_r, &landed, &success, &dist);
                    acc += (double)sp;
                    ++steps;
                    if (landed) break;
                }
                const float term = ll_terminal_penalty<float, LL_MODE_FAST>(env, p, steps, landed, success, dist);
                const double fit = (double)term + acc;
                fit_sum += fit, steps_sum += (double)steps;
                if (success) (lay.dir == 0 ? bonus0 : bonus1) += env.tp_landed_lr_bonus;
                if (a.ep_steps && lane == 0)
                {
                    a.ep_steps[member * (size_t)a.n_layouts + layout] = steps | (landed ? (1 << 28) : 0) | ((landed && !success) ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
                    a.ep_fit[member * (size_t)a.n_layouts + layout]   = fit;
                }
            }
            if (lane == 0)
            {
                const double lr   = (double)(bonus1 < bonus0 ? bonus1 : bonus0);
                a.fitness[member] = fit_sum - lr, a.lr[member] = lr, a.steps[member] = steps_sum;
            }
        }
    }

    // ---- small populations with SEVERAL layouts per member (BASELINE.json configs[4]: 8_16_32_16_8, 100 members x 50 start positions,
    // main_train_f): one warp = one EPISODE.  All members x layouts work items fly at once (longest expected flight first), so a
    // generation lasts as long as its longest episode's chain of steps instead of a member's episodes one after the other; each
    // episode's result goes to the per-episode arrays and the member's last episode to finish sums them in layout order.
    template <class N> __global__ void __launch_bounds__(256) ll_episode_warp_ep_kernel(const __grid_constant__ ll_env<float> env, const __grid_constant__ warp_args a)
    {
        static_assert(N::ok(), "topology not supported by the warp-per-episode kernel");
        __shared__ __align__(16) float s_xb[8][64];
        float* xb      = s_xb[threadIdx.x >> 5];
        const int lane = threadIdx.x & 31;
        const int L    = a.n_layouts;
        const unsigned long long n_items = a.n_members * (unsigned long long)L;
        for (;;)
        {
            unsigned long long item = 0;
            if (lane == 0) item = (unsigned long long)atomicAdd(a.queue, 1u);
            item = __shfl_sync(0xffffffffu, item, 0);
            if (item >= n_items) break;
            const unsigned long long member = item % a.n_members;
            const int layout                = a.layouts[(int)(item / a.n_members)].next;
            const float* row                = a.params + member * (size_t)a.p_dev;
            float w[N::n_w], b[N::NL];
#pragma unroll
            for (int l = 0; l < N::NL; ++l)
            {
                const bool mine = lane < N::sz(l + 1);
#pragma unroll
                for (int k = 0; k < N::sz(l); ++k) w[N::w_at(l) + k] = mine ? __ldg(row + a.w_off[l] + k * a.out_pad[l] + lane) : 0.f;
                b[l] = mine ? __ldg(row + a.b_off[l] + lane) : 0.f;
            }
            const ll_layout<float> lay = a.layouts[layout];
            ll_lander<float> p{lay.init_x, env.y0, env.vx0, env.vy0, env.max_fuel};
            const float cx = lay.center_x, lpad = lay.lpad_x1, lpad_r = lay.lpad_right;
            int steps = 0;
            double acc = 0.0;
            bool landed = false, success = false;
            float dist = 0.f;
            while (steps < a.max_steps)
            {
                float st[5];
                ll_get_state<float, LL_MODE_FAST>(env, p, cx, st, 1);
                const float z    = forward<N, 0>(w, b, st, 0.f, xb);
                const int action = argmax_outputs<N>(z, xb); // first maximum wins
                const float sp = ll_env_step<float, LL_MODE_FAST>(env, p, action, steps == 0, cx, lpad, lpad_r, &landed, &success, &dist);
                acc += (double)sp;
                ++steps;
                if (landed) break;
            }
            if (lane == 0)
            {
                const float term = ll_terminal_penalty<float, LL_MODE_FAST>(env, p, steps, landed, success, dist);
                a.ep_steps[member * (size_t)L + layout] = steps | (landed ? (1 << 28) : 0) | ((landed && !success) ? (1 << 29) : 0) | (success ? (1 << 30) : 0);
                a.ep_fit[member * (size_t)L + layout]   = (double)term + acc;
                __threadfence(); // the result is visible device-wide before the arrival is counted
                if (atomicAdd(a.part_count + member, 1u) == (unsigned int)L - 1u)
                { // the member's last episode: all layouts in layout order (src/nn_engine_trainer.cpp:518-530)
                    __threadfence();
                    a.part_count[member]         = 0u;
                    const volatile int* g_meta   = a.ep_steps + member * (size_t)L;
                    const volatile double* g_fit = a.ep_fit + member * (size_t)L;
                    double f = 0.0, stp = 0.0;
                    float bonus0 = 0.f, bonus1 = 0.f;
                    for (int l = 0; l < L; ++l)
                    {
                        const int meta = g_meta[l];
                        f += g_fit[l], stp += (double)(meta & 0x0fffffff);
                        if (meta & (1 << 30)) (a.layouts[l].dir == 0 ? bonus0 : bonus1) += env.tp_landed_lr_bonus;
                    }
                    const double lr   = (double)(bonus1 < bonus0 ? bonus1 : bonus0);
                    a.fitness[member] = f - lr, a.lr[member] = lr, a.steps[member] = stp;
                }
            }
            __syncwarp();
        }
    }

    template <class N> int launch_warp_t(ll_ctx* ctx, const ll_episode_args& h)
    {
        const ll_topo& tp = ctx->topo;
        warp_args a{};
        a.params = static_cast<const float*>(h.params), a.layouts = static_cast<const ll_layout<float>*>(h.layouts);
        a.n_layouts = h.n_layouts, a.max_steps = h.max_steps, a.n_members = h.n_members;
        a.fitness = h.fitness, a.lr = h.lr, a.steps = h.steps, a.ep_steps = h.ep_steps, a.ep_fit = h.ep_fit, a.queue = h.queue;
        a.p_dev = tp.p_dev;
        for (int l = 0; l < tp.n_layers; ++l) a.w_off[l] = tp.w_off[l], a.b_off[l] = tp.b_off[l], a.out_pad[l] = tp.out_pad[l];
        // spread the members over the SMs first (a generation lasts as long as its longest episode: fewer warps per scheduler =
        // shorter steps), 8 warps per block, persistent warps drawing members from the queue
        // several layouts per member and room for every episode at once (two waves at most): one warp per EPISODE
        const unsigned long long items = h.n_members * (unsigned long long)h.n_layouts;
        if (ll_warp_per_episode(ctx, h.n_members, h.n_layouts, h.ep_steps && h.ep_fit))
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
            int per_sm_e = 0;
            LL_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_e, ll_episode_warp_ep_kernel<N>, 256, 0));
            const unsigned long long resident_e = (unsigned long long)ctx->sm_count * std::max(1, per_sm_e);
            const unsigned int grid_e = (unsigned int)std::max<unsigned long long>(1, std::min(resident_e, (items + 7) / 8));
            ll_env<float> env_e;
            ll_make_env<float>(ctx->env_cfg, 1, &env_e);
            LL_CUDA(ctx, cudaMemsetAsync(h.queue, 0, sizeof(unsigned int), ctx->stream));
            ll_episode_warp_ep_kernel<N><<<grid_e, 256, 0, ctx->stream>>>(env_e, a);
            LL_CUDA(ctx, cudaGetLastError());
            return LL_OK;
        }
        int per_sm = 0;
        LL_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ll_episode_warp_kernel<N>, 256, 0));
        const unsigned long long resident = (unsigned long long)ctx->sm_count * std::max(1, per_sm);
        const unsigned long long wanted   = (h.n_members + 7) / 8;
        const unsigned int grid = (unsigned int)std::max<unsigned long long>(1, std::min(resident, wanted));
        ll_env<float> env;
        ll_make_env<float>(ctx->env_cfg, 1, &env);
        LL_CUDA(ctx, cudaMemsetAsync(h.queue, 0, sizeof(unsigned int), ctx->stream));
        ll_episode_warp_kernel<N><<<grid, 256, 0, ctx->stream>>>(env, a);
        LL_CUDA(ctx, cudaGetLastError());
        return LL_OK;
    }

    using wnet_16_16        = wnet<5, 16, 16, 4>;
    using wnet_16_32        = wnet<5, 16, 32, 4>;
    using wnet_8_16_8       = wnet<5, 8, 16, 8, 4>;
    using wnet_8_16_32_16_8 = wnet<5, 8, 16, 32, 16, 8, 4>;

    template <class N> bool same_net(const ll_topo& tp)
    {
        if (tp.n_layers != N::NL || tp.in[0] != N::sz(0)) return false;
        for (int l = 0; l < N::NL; ++l)
            if (tp.out[l] != N::sz(l + 1)) return false;
        return true;
    }
} // namespace

// One warp per EPISODE instead of per member: several layouts per member, the per-episode arrays at hand, and room for every
// (member, layout) work item within about two waves of resident warps (24 per SM).  LL_WARP_PER_EPISODE = 1 / 0 forces / forbids it.
bool ll_warp_per_episode(const ll_ctx* ctx, unsigned long long n_members, int n_layouts, bool have_episode_arrays)
{
    if (n_layouts <= 1 || !have_episode_arrays) return false;
    const char* force = getenv("LL_WARP_PER_EPISODE");
    if (force && *force) return atoi(force) != 0;
    return n_members * (unsigned long long)n_layouts <= 2ull * (unsigned long long)ctx->sm_count * 24ull;
}

bool ll_warp_applicable(const ll_ctx* ctx, int n_layouts)
{
    const ll_topo& tp = ctx->topo;
    if (tp.activation != LL_ACT_TANH || ctx->precision != 32 || ctx->mode != LL_MODE_FAST || n_layouts < 1) return false;
    return same_net<wnet_16_16>(tp) || same_net<wnet_16_32>(tp) || same_net<wnet_8_16_8>(tp) || same_net<wnet_8_16_32_16_8>(tp);
}

int ll_launch_episode_warp(ll_ctx* ctx, const ll_episode_args& a)
{
    const ll_topo& tp = ctx->topo;
    if (same_net<wnet_8_16_8>(tp)) return launch_warp_t<wnet_8_16_8>(ctx, a);
    if (same_net<wnet_16_16>(tp)) return launch_warp_t<wnet_16_16>(ctx, a);
    if (same_net<wnet_16_32>(tp)) return launch_warp_t<wnet_16_32>(ctx, a);
    if (same_net<wnet_8_16_32_16_8>(tp)) return launch_warp_t<wnet_8_16_32_16_8>(ctx, a);
    return ll_fail(ctx, LL_EINVAL, "warp-per-member episode kernel: topology not instantiated");
}
