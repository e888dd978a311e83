// ll_sim.cu — one device-resident lander stepped from the host: the GameLogicCpp<T> object of nn_play / play mode
// (src/game_logic.h:15-93; driven by main.py:127-238 through the lunar_lander_cpp bindings, src/lunar_lander_cpp.cpp:21-106).
// The state lives in HBM and every call is one single-thread kernel built from the same device functions the episode
// kernel uses (ll_env.cuh, LL_MODE_EXACT arithmetic), so a lander flown step by step through this interface goes through
// exactly the states K1 computes for the same actions.  A frame of nn_play costs two launches (policy + update): this is an
// interactive, 60-calls-per-second consumer, not a throughput path.
#include <cstring>

#include "ll_env.cuh"
#include "ll_internal.h"

template <typename T> struct ll_sim_dev
{
    ll_lander<T> p;
    T init_x, lpad_x1, lpad_right, center_x;
    int landed, crashed, success, time_step, last_action, pad;
};

namespace
{
    template <typename T> __device__ void sim_view(const ll_env<T>& e, const ll_sim_dev<T>& s, ll_sim_view* v)
    {
        T st[5];
        ll_get_state<T, LL_MODE_EXACT>(e, s.p, s.center_x, st, 1);
        v->x = (double)s.p.x, v->y = (double)s.p.y, v->vx = (double)s.p.vx, v->vy = (double)s.p.vy, v->fuel = (double)s.p.fuel;
        v->landing_pad_center_x = (double)s.center_x, v->landing_pad_y = (double)e.pad_y;
        v->landed = s.landed, v->crashed = s.crashed, v->landed_successfully = s.success;
        v->time_step = s.time_step, v->last_action = s.last_action;
        v->done = (s.landed || s.crashed) ? 1 : 0; // is_done(), src/game_logic.cpp:253-256
        for (int i = 0; i < 5; ++i) v->state[i] = (double)st[i];
    }

    // reset(): src/game_logic.cpp:124-140
    template <typename T> __global__ void sim_reset_kernel(ll_env<T> e, ll_sim_dev<T>* s, T init_x, T lpad_x1, T lpad_right, T center_x, ll_sim_view* v)
    {
        s->init_x = init_x, s->lpad_x1 = lpad_x1, s->lpad_right = lpad_right, s->center_x = center_x;
        s->p.x = init_x, s->p.y = e.y0, s->p.vx = e.vx0, s->p.vy = e.vy0, s->p.fuel = e.max_fuel;
        s->landed = s->crashed = s->success = 0, s->time_step = 0, s->last_action = 0;
        sim_view(e, *s, v);
    }

    // update(action, state_out, n): src/game_logic.cpp:295-310
    template <typename T> __global__ void sim_update_kernel(ll_env<T> e, ll_sim_dev<T>* s, int action, ll_sim_view* v)
    {
        if (!(s->landed || s->crashed))
        {
            bool landed, success;
            T dist;
            s->last_action = action; // apply_action stores it first, src/game_logic.cpp:158
            (void)ll_env_step<T, LL_MODE_EXACT>(e, s->p, action, s->time_step == 0, s->center_x, s->lpad_x1, s->lpad_right, &landed, &success, &dist);
            if (landed) s->landed = 1, s->success = success ? 1 : 0, s->crashed = success ? 0 : 1;
            s->time_step += 1;
        }
        sim_view(e, *s, v);
    }

    template <typename T> __global__ void sim_set_kernel(ll_env<T> e, ll_sim_dev<T>* s, ll_sim_view in, ll_sim_view* v)
    {
        s->p.x = (T)in.x, s->p.y = (T)in.y, s->p.vx = (T)in.vx, s->p.vy = (T)in.vy, s->p.fuel = (T)in.fuel;
        s->center_x = (T)in.landing_pad_center_x;
        s->landed = in.landed != 0, s->crashed = in.crashed != 0, s->success = in.landed_successfully != 0;
        s->last_action = in.last_action;
        sim_view(e, *s, v);
    }

    // out[0] = calculate_step_penalty(action) (src/game_logic.cpp:312-320); out[1] = calculate_terminal_penalty(steps)
    // (:322-388, both overloads compute the same value); out[2] = the L/R bonus the 3-argument overload books
    template <typename T> __global__ void sim_penalty_kernel(ll_env<T> e, const ll_sim_dev<T>* s, int action, int steps, double* out)
    {
        using O      = ll_ops<T, LL_MODE_EXACT>;
        const T dist = O::add(ll_abs(O::sub(s->p.x, s->center_x)), ll_abs(O::sub(s->p.y, e.pad_y)));
        T sp         = O::add(T(0), O::mul(dist, e.sp_dist));
        if (action > 0) sp = O::add(sp, e.sp_action);
        out[0] = (double)sp;
        out[1] = (double)ll_terminal_penalty<T, LL_MODE_EXACT>(e, s->p, steps, s->landed != 0, s->success != 0, dist);
        out[2] = s->success ? (double)e.tp_landed_lr_bonus : 0.0;
    }

    template <typename T> int sim_prepare(ll_ctx* ctx, ll_env<T>* env)
    {
        cudaSetDevice(ctx->device);
        ll_make_env<T>(ctx->env_cfg, 0, env);
        if (!ctx->d_sim)
        {
            LL_CUDA(ctx, cudaMalloc(&ctx->d_sim, 1024));
            LL_CUDA(ctx, cudaMemsetAsync(ctx->d_sim, 0, 1024, ctx->stream));
        }
        return LL_OK;
    }
    template <typename T> ll_sim_dev<T>* sim_of(ll_ctx* c) { return static_cast<ll_sim_dev<T>*>(c->d_sim); }
    ll_sim_view* view_of(ll_ctx* c) { return reinterpret_cast<ll_sim_view*>(static_cast<char*>(c->d_sim) + 256); }
    double* pen_of(ll_ctx* c) { return reinterpret_cast<double*>(static_cast<char*>(c->d_sim) + 768); }

    int sim_finish(ll_ctx* ctx, ll_sim_view* out)
    {
        LL_CUDA(ctx, cudaGetLastError());
        ctx->times.other_launches += 1;
        ll_sim_view host;
        LL_CUDA(ctx, cudaMemcpyAsync(&host, view_of(ctx), sizeof(host), cudaMemcpyDeviceToHost, ctx->stream));
        LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (out) *out = host;
        return LL_OK;
    }

    // reset(T spad_x1, T lpad_x1) (src/game_logic.cpp:143-154) or reset() with the configured pad and GameCfg.x0
    template <typename T> int sim_reset(ll_ctx* ctx, int use_pads, double spad_x1, double lpad_x1, ll_sim_view* out)
    {
        ll_env<T> env;
        if (int rc = sim_prepare<T>(ctx, &env)) return rc;
        const ll_env_cfg& c = ctx->env_cfg;
        const T lpad_w = static_cast<T>(c.lpad_width), width = static_cast<T>(c.display_width), height = static_cast<T>(c.display_height);
        T lpad, init_x;
        if (use_pads)
        {
            lpad   = static_cast<T>(lpad_x1);
            init_x = static_cast<T>(spad_x1) + (static_cast<T>(c.spad_width) / T(2.0)) - (static_cast<T>(c.lander_width) / T(2.0));
            ctx->sim_lpad_x1 = lpad_x1, ctx->sim_init_x = (double)init_x, ctx->sim_custom = true;
        }
        else if (ctx->sim_custom) lpad = static_cast<T>(ctx->sim_lpad_x1), init_x = static_cast<T>(ctx->sim_init_x); // reset() keeps the last pads
        else lpad = static_cast<T>(c.lpad_x1), init_x = static_cast<T>(c.x0);
        const T center = (width > T(0) && height > T(0)) ? lpad + lpad_w / T(2.0) : T(0);
        sim_reset_kernel<T><<<1, 1, 0, ctx->stream>>>(env, sim_of<T>(ctx), init_x, lpad, lpad + lpad_w, center, view_of(ctx));
        ctx->sim_ready = true;
        return sim_finish(ctx, out);
    }
} // namespace

#define LL_SIM_DISPATCH(ctx, expr64, expr32) ((ctx)->precision == 64 ? (expr64) : (expr32))

extern "C" int ll_sim_reset(ll_ctx* ctx, int32_t use_pads, double spad_x1, double lpad_x1, ll_sim_view* out)
{
    if (!ctx) return ll_fail(nullptr, LL_EINVAL, "null context");
    return LL_SIM_DISPATCH(ctx, sim_reset<double>(ctx, use_pads, spad_x1, lpad_x1, out), sim_reset<float>(ctx, use_pads, spad_x1, lpad_x1, out));
}

template <typename T> static int sim_update(ll_ctx* ctx, int action, ll_sim_view* out)
{
    ll_env<T> env;
    if (int rc = sim_prepare<T>(ctx, &env)) return rc;
    sim_update_kernel<T><<<1, 1, 0, ctx->stream>>>(env, sim_of<T>(ctx), action, view_of(ctx));
    return sim_finish(ctx, out);
}

extern "C" int ll_sim_update(ll_ctx* ctx, int32_t action, ll_sim_view* out)
{
    if (!ctx) return ll_fail(nullptr, LL_EINVAL, "null context");
    if (!ctx->sim_ready) LL_SIM_DISPATCH(ctx, sim_reset<double>(ctx, 0, 0, 0, nullptr), sim_reset<float>(ctx, 0, 0, 0, nullptr));
    return LL_SIM_DISPATCH(ctx, sim_update<double>(ctx, action, out), sim_update<float>(ctx, action, out));
}

template <typename T> static int sim_set(ll_ctx* ctx, const ll_sim_view& in, ll_sim_view* out)
{
    ll_env<T> env;
    if (int rc = sim_prepare<T>(ctx, &env)) return rc;
    sim_set_kernel<T><<<1, 1, 0, ctx->stream>>>(env, sim_of<T>(ctx), in, view_of(ctx));
    return sim_finish(ctx, out);
}

extern "C" int ll_sim_set(ll_ctx* ctx, const ll_sim_view* in, ll_sim_view* out)
{
    if (!ctx || !in) return ll_fail(ctx, LL_EINVAL, "ll_sim_set: null argument");
    if (!ctx->sim_ready) LL_SIM_DISPATCH(ctx, sim_reset<double>(ctx, 0, 0, 0, nullptr), sim_reset<float>(ctx, 0, 0, 0, nullptr));
    return LL_SIM_DISPATCH(ctx, sim_set<double>(ctx, *in, out), sim_set<float>(ctx, *in, out));
}

extern "C" int ll_sim_get(ll_ctx* ctx, ll_sim_view* out)
{
    if (!ctx || !out) return ll_fail(ctx, LL_EINVAL, "ll_sim_get: null argument");
    if (!ctx->sim_ready) return LL_SIM_DISPATCH(ctx, sim_reset<double>(ctx, 0, 0, 0, out), sim_reset<float>(ctx, 0, 0, 0, out));
    cudaSetDevice(ctx->device);
    LL_CUDA(ctx, cudaMemcpyAsync(out, view_of(ctx), sizeof(*out), cudaMemcpyDeviceToHost, ctx->stream));
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LL_OK;
}

template <typename T> static int sim_penalties(ll_ctx* ctx, int action, int steps, double* host3)
{
    ll_env<T> env;
    if (int rc = sim_prepare<T>(ctx, &env)) return rc;
    sim_penalty_kernel<T><<<1, 1, 0, ctx->stream>>>(env, sim_of<T>(ctx), action, steps, pen_of(ctx));
    LL_CUDA(ctx, cudaGetLastError());
    ctx->times.other_launches += 1;
    LL_CUDA(ctx, cudaMemcpyAsync(host3, pen_of(ctx), 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    LL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LL_OK;
}

extern "C" int ll_sim_step_penalty(ll_ctx* ctx, int32_t action, double* penalty)
{
    if (!ctx || !penalty) return ll_fail(ctx, LL_EINVAL, "ll_sim_step_penalty: null argument");
    if (!ctx->sim_ready) LL_SIM_DISPATCH(ctx, sim_reset<double>(ctx, 0, 0, 0, nullptr), sim_reset<float>(ctx, 0, 0, 0, nullptr));
    double r[3];
    const int rc = LL_SIM_DISPATCH(ctx, sim_penalties<double>(ctx, action, 0, r), sim_penalties<float>(ctx, action, 0, r));
    if (rc == LL_OK) *penalty = r[0];
    return rc;
}

extern "C" int ll_sim_terminal_penalty(ll_ctx* ctx, int32_t steps_taken, int32_t direction, double* penalty, double* bonus_lr)
{
    if (!ctx || !penalty) return ll_fail(ctx, LL_EINVAL, "ll_sim_terminal_penalty: null argument");
    if (direction > 1) return ll_fail(ctx, LL_EINVAL, "ll_sim_terminal_penalty: direction must be 0 or 1"); // assert, src/game_logic.cpp:356
    if (!ctx->sim_ready) LL_SIM_DISPATCH(ctx, sim_reset<double>(ctx, 0, 0, 0, nullptr), sim_reset<float>(ctx, 0, 0, 0, nullptr));
    double r[3];
    const int rc = LL_SIM_DISPATCH(ctx, sim_penalties<double>(ctx, 0, steps_taken, r), sim_penalties<float>(ctx, 0, steps_taken, r));
    if (rc != LL_OK) return rc;
    *penalty = r[1];
    if (direction >= 0 && bonus_lr)
    {
        // landing_bonus_lr[direction] += tp_landed_lr_bonus, accumulated in T like the reference's std::array<T,2>
        if (ctx->precision == 64) bonus_lr[direction] = bonus_lr[direction] + r[2];
        else bonus_lr[direction] = (double)((float)bonus_lr[direction] + (float)r[2]);
    }
    return LL_OK;
}
