// ll_env.cuh — the lander environment as device functions: one GameLogicCpp<T>::update + calculate_step_penalty,
// and calculate_terminal_penalty, written once for every arithmetic policy (ll_ops.cuh).
//
// Follows (reference file:line, azimonti/lunar-lander):
//   apply_action            src/game_logic.cpp:156-183      update_physics   :185-195
//   check_landing_crash     src/game_logic.cpp:197-239      get_state        :275-293
//   calculate_step_penalty  src/game_logic.cpp:312-320      calculate_terminal_penalty (3-arg)  :352-388
#ifndef LL_ENV_CUH_
#define LL_ENV_CUH_

#include "ll_ops.cuh"

template <typename T> struct ll_lander
{
    T x, y, vx, vy, fuel;
};

template <typename T> __device__ __forceinline__ T ll_clamp(T v, T lo, T hi) { return v < lo ? lo : (hi < v ? hi : v); }
__device__ __forceinline__ float ll_abs(float v) { return fabsf(v); }
__device__ __forceinline__ double ll_abs(double v) { return fabs(v); }

// get_state(): 5 normalised inputs, element i written to s[i * stride]
template <typename T, int MODE>
__device__ __forceinline__ void ll_get_state(const ll_env<T>& e, const ll_lander<T>& p, T cx, T* s, int stride)
{
    using O = ll_ops<T, MODE>;
    const T dx = O::sub(p.x, cx), dy = O::sub(p.y, e.pad_y);
    T s0, s1, s2, s3, s4;
    if (O::fast)
    {
        s0 = O::mul(p.vx, e.inv_max_vx), s1 = O::mul(p.vy, e.inv_max_vy), s2 = O::mul(dx, e.inv_width);
        s3 = O::mul(dy, e.inv_height), s4 = O::mul(p.fuel, e.inv_max_fuel);
    }
    else
    {
        s0 = O::div(p.vx, e.max_vx), s1 = O::div(p.vy, e.max_vy), s2 = O::div(dx, e.width);
        s3 = O::div(dy, e.height), s4 = O::div(p.fuel, e.max_fuel);
    }
    s[0 * stride] = ll_clamp(s0, T(-2), T(2));
    s[1 * stride] = ll_clamp(s1, T(-2), T(2));
    s[2 * stride] = ll_clamp(s2, T(-5), T(5));
    s[3 * stride] = ll_clamp(s3, T(-5), T(5));
    s[4 * stride] = s4;
}

// One update(action) on a lander that is not done, followed by calculate_step_penalty(action) on the post-update
// state.  `first` = (time_step == 0): no landing test on the very first step.  Returns the step penalty (T);
// *landed / *success report a touchdown in this step (velocity is zeroed on touchdown, before any penalty).
template <typename T, int MODE>
__device__ __forceinline__ T ll_env_step(const ll_env<T>& e, ll_lander<T>& p, int action, bool first, T cx, T lpad, T lpad_r, bool* landed,
                                         bool* success, T* dist_out)
{
    using O = ll_ops<T, MODE>;
    if (p.fuel > T(0))
    {
        if (action == 1) p.vy = O::sub(p.vy, T(0.3)), p.fuel = O::sub(p.fuel, T(1.0));
        else if (action == 2) p.vx = O::sub(p.vx, T(0.05)), p.fuel = O::sub(p.fuel, T(0.5));
        else if (action == 3) p.vx = O::add(p.vx, T(0.05)), p.fuel = O::sub(p.fuel, T(0.5));
        p.fuel = (T(0) < p.fuel) ? p.fuel : T(0);
    }
    p.vy = O::add(p.vy, e.ay);
    p.vx = O::mul(p.vx, e.one_minus_mu_x);
    p.vy = O::mul(p.vy, e.one_minus_mu_y);
    p.x  = O::add(p.x, p.vx);
    p.y  = O::add(p.y, p.vy);
    *landed = false, *success = false;
    if (!first && p.y >= e.ground)
    {
        p.y     = e.ground;
        *landed = true;
        const bool on_pad = (p.x >= lpad) && (O::add(p.x, e.lander_w) <= lpad_r);
        const bool safe   = (ll_abs(p.vx) < e.max_vx) && (ll_abs(p.vy) < e.max_vy);
        *success          = on_pad && safe;
        p.vx = T(0), p.vy = T(0);
    }
    const T dist = O::add(ll_abs(O::sub(p.x, cx)), ll_abs(O::sub(p.y, e.pad_y)));
    T sp         = O::add(T(0), O::mul(dist, e.sp_dist));
    if (action > 0) sp = O::add(sp, e.sp_action);
    *dist_out = dist;
    return sp;
}

// calculate_terminal_penalty(steps, direction, bonus) without the bonus side effect (the caller books it)
template <typename T, int MODE>
__device__ __forceinline__ T ll_terminal_penalty(const ll_env<T>& e, const ll_lander<T>& p, int steps, bool landed, bool success, T dist)
{
    using O = ll_ops<T, MODE>;
    const bool crashed = landed && !success;
    T term             = O::add(T(0), O::mul((T)steps, e.tp_steps));
    term               = O::add(term, O::mul(dist, e.tp_dist));
    const T vmag       = O::sqrt(O::add(O::mul(p.vx, p.vx), O::mul(p.vy, p.vy)));
    if (success)
    {
        term = O::sub(term, e.tp_landed_bonus);
        term = O::sub(term, O::mul(p.fuel, e.tp_fuel_bonus));
    }
    else if (crashed)
    {
        term = O::add(term, e.tp_crashed);
        term = O::add(term, O::mul(vmag, e.tp_crash_v_mag));
    }
    else if (p.fuel <= T(0) && !landed) { term = O::add(term, e.tp_no_fuel); }
    if (!success) term = O::add(term, O::mul(vmag, T(1.0)));
    return term;
}

#endif
