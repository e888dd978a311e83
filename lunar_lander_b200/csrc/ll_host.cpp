// ll_host.cpp — host-only pieces of libll_gpu.so: topology / constant preparation and the two layout helpers
// that feed the hot path (they run 50 times per training run and stay on the CPU).
#include <algorithm>
#include <cmath>
#include <random>

#include "ll_internal.h"

// ---- topology ---------------------------------------------------------------------------------------------------
// canonical order of a member (ABI, weight file): per layer W[out][in] row-major, then b[out].
void ll_build_topo(const std::vector<uint64_t>& sizes, int activation, ll_topo* t, std::vector<int>* dev2canon)
{
    *t            = ll_topo{};
    t->n_layers   = (int)sizes.size() - 1;
    t->activation = activation;
    t->max_width  = (int)sizes[0];
    int canon = 0, dev = 0;
    for (int l = 0; l < t->n_layers; ++l)
    {
        const int in = (int)sizes[l], out = (int)sizes[l + 1], op = (out + 7) & ~7;
        t->in[l] = in, t->out[l] = out, t->out_pad[l] = op;
        t->w_off[l] = dev;
        t->b_off[l] = dev + in * op;
        dev += (in + 1) * op;
        canon += in * out + out;
        t->max_width = std::max(t->max_width, op);
    }
    t->p_canon = canon, t->p_dev = dev;
    if (dev2canon)
    {
        dev2canon->assign((size_t)dev, -1);
        int c0 = 0;
        for (int l = 0; l < t->n_layers; ++l)
        {
            const int in = t->in[l], out = t->out[l], op = t->out_pad[l];
            for (int o = 0; o < out; ++o)
                for (int k = 0; k < in; ++k) (*dev2canon)[(size_t)t->w_off[l] + (size_t)k * op + o] = c0 + o * in + k;
            for (int o = 0; o < out; ++o) (*dev2canon)[(size_t)t->b_off[l] + o] = c0 + in * out + o;
            c0 += in * out + out;
        }
    }
}

// ---- environment constants: the conversions and derived values of GameLogicCpp<T>'s constructor, reset() and
//      recalculate_derived_values() (src/game_logic.cpp:48-140), evaluated once on the host in T ------------------
template <typename T> void ll_make_env(const ll_env_cfg& c, int fast, ll_env<T>* e)
{
    auto cv = [](double v) { return static_cast<T>(v); };
    const T pad_y1 = cv(c.pad_y1), terrain_y = cv(c.terrain_y);
    T max_vx = std::abs(cv(c.max_vx)), max_vy = std::abs(cv(c.max_vy));
    if (max_vx == T(0)) max_vx = T(1e-6);
    if (max_vy == T(0)) max_vy = T(1e-6);
    const T width = cv(c.display_width), height = cv(c.display_height), lander_h = cv(c.lander_height);
    T max_fuel = static_cast<T>(static_cast<int>(c.max_fuel)); // read with getInt in the reference (:64)
    if (max_fuel <= T(0)) max_fuel = T(1e-6);
    const bool world_ok = width > T(0) && height > T(0);
    e->pad_y            = pad_y1;
    e->ground           = world_ok ? (height - lander_h) - terrain_y : T(0);
    e->max_vx = max_vx, e->max_vy = max_vy;
    e->ay             = cv(c.ay0) + cv(c.g);
    e->one_minus_mu_x = T(1.0) - cv(c.mu_x);
    e->one_minus_mu_y = T(1.0) - cv(c.mu_y);
    e->lander_w = cv(c.lander_width), e->lpad_w = cv(c.lpad_width);
    e->width = width, e->height = height, e->max_fuel = max_fuel;
    e->y0 = cv(c.y0), e->vx0 = cv(c.vx0), e->vy0 = cv(c.vy0);
    e->sp_dist = cv(c.sp_dist_factor), e->sp_action = cv(c.sp_action_factor);
    e->tp_steps = cv(c.tp_steps_factor), e->tp_dist = cv(c.tp_dist_factor);
    e->tp_landed_bonus = cv(c.tp_landed_bonus), e->tp_landed_lr_bonus = cv(c.tp_landed_lr_bonus);
    e->tp_fuel_bonus = cv(c.tp_fuel_bonus_factor), e->tp_crashed = cv(c.tp_crashed_penalty);
    e->tp_crash_v_mag = cv(c.tp_crash_v_mag_factor), e->tp_no_fuel = cv(c.tp_no_fuel_penalty);
    (void)fast;
    e->inv_max_vx = T(1) / max_vx, e->inv_max_vy = T(1) / max_vy;
    e->inv_width = T(1) / width, e->inv_height = T(1) / height, e->inv_max_fuel = T(1) / max_fuel;
}
template void ll_make_env<float>(const ll_env_cfg&, int, ll_env<float>*);
template void ll_make_env<double>(const ll_env_cfg&, int, ll_env<double>*);

// reset(spad_x1, lpad_x1): src/game_logic.cpp:143-154 (+ :101-122)
template <typename T> void ll_make_layout(const ll_env_cfg& c, int spad_x1, int lpad_x1, int dir, ll_layout<T>* out)
{
    const T spad_w = static_cast<T>(c.spad_width), lander_w = static_cast<T>(c.lander_width), lpad_w = static_cast<T>(c.lpad_width);
    const T width = static_cast<T>(c.display_width), height = static_cast<T>(c.display_height);
    const T lpad  = static_cast<T>(lpad_x1);
    out->init_x     = (static_cast<T>(spad_x1) + spad_w / T(2.0)) - lander_w / T(2.0);
    out->lpad_x1    = lpad;
    out->lpad_right = lpad + lpad_w;
    out->center_x   = (width > T(0) && height > T(0)) ? lpad + lpad_w / T(2.0) : T(0);
    out->dir        = dir;
    out->next       = 0;
}
template void ll_make_layout<float>(const ll_env_cfg&, int, int, int, ll_layout<float>*);
template void ll_make_layout<double>(const ll_env_cfg&, int, int, int, ll_layout<double>*);

// ---- layout helpers -----------------------------------------------------------------------------------------------
extern "C" int ll_generate_random_pad_positions(double screen_width, int32_t spad_width, int32_t lpad_width, int32_t has_seed,
                                                int32_t seed, int32_t force_left_to_right, int32_t force_right_to_left,
                                                int32_t* spad_x1, int32_t* lpad_x1)
{
    // Semantics of src/game_utils.cpp:12-74: std::mt19937 + uniform_int_distribution<int>(10, width); reject pairs that
    // leave the right 10-px border or whose centres are closer than half-widths + 10; optional mirroring.
    if (!spad_x1 || !lpad_x1) return LL_EINVAL;
    const int border = 10, min_gap = 10;
    const int w = static_cast<int>(screen_width);
    if (w < border || spad_width + lpad_width + 2 * border + min_gap > w) return LL_EINVAL; // would never terminate
    std::mt19937 rng;
    if (has_seed) rng.seed(static_cast<unsigned int>(seed));
    else
    {
        std::random_device rd;
        rng.seed(rd());
    }
    std::uniform_int_distribution<int> pick(border, w);
    const double half_s = static_cast<double>(spad_width) / 2.0, half_l = static_cast<double>(lpad_width) / 2.0;
    for (;;)
    {
        int s = pick(rng);
        int l = pick(rng);
        if (s + spad_width > w - border || l + lpad_width > w - border) continue;
        const double cs = static_cast<double>(s) + half_s, cl = static_cast<double>(l) + half_l;
        if (std::abs(cs - cl) < half_s + half_l + min_gap) continue;
        const bool mirror = (force_left_to_right && cs >= cl) || (force_right_to_left && cs <= cl);
        if (mirror)
        {
            const double mid = static_cast<double>(border) + static_cast<double>(w - 2 * border) / 2.0;
            s = static_cast<int>((2.0 * mid - cs) - half_s);
            l = static_cast<int>((2.0 * mid - cl) - half_l);
        }
        *spad_x1 = s, *lpad_x1 = l;
        return LL_OK;
    }
}

extern "C" int ll_balance_layouts(int32_t* spad_x1, int32_t* lpad_x1, double* spad_center, double* lpad_center, uint64_t* directions,
                                  int32_t n, double target_ltr_ratio, double screen_width, double spad_width, double lpad_width)
{
    // Semantics of src/nn_engine_trainer.cpp:196-260, including the early return that leaves `directions` untouched
    // when the left-to-right count already matches.
    if (n < 0 || (n > 0 && (!spad_x1 || !lpad_x1 || !spad_center || !lpad_center || !directions))) return LL_EINVAL;
    if (target_ltr_ratio < 0.0 || target_ltr_ratio > 1.0) target_ltr_ratio = 0.5;
    int ltr = 0;
    for (int i = 0; i < n; ++i) ltr += spad_center[i] < lpad_center[i] ? 1 : 0;
    const int target = static_cast<int>(std::round(static_cast<double>(n) * target_ltr_ratio));
    int to_flip      = std::abs(ltr - target);
    if (to_flip == 0) return LL_OK;
    const bool make_rtl = ltr > target;
    const double axis   = screen_width / 2.0;
    for (int i = 0; i < n && to_flip > 0; ++i)
    {
        const bool is_ltr = spad_center[i] < lpad_center[i];
        if (is_ltr != make_rtl) continue;
        const double cs = 2.0 * axis - spad_center[i], cl = 2.0 * axis - lpad_center[i];
        const double sx = std::max(0.0, std::min(cs - spad_width / 2.0, screen_width - spad_width));
        const double lx = std::max(0.0, std::min(cl - lpad_width / 2.0, screen_width - lpad_width));
        spad_x1[i] = static_cast<int>(std::round(sx)), lpad_x1[i] = static_cast<int>(std::round(lx));
        spad_center[i] = cs, lpad_center[i] = cl;
        --to_flip;
    }
    for (int i = 0; i < n; ++i) directions[i] = spad_center[i] < lpad_center[i] ? 0u : 1u;
    return LL_OK;
}
