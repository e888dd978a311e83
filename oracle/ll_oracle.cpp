// ORACLE / TEST INFRASTRUCTURE — not part of the product.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load this library; the product (libll_gpu.so) never does.
//
// Plain CPU restatement ("port") of the hot path of azimonti/lunar-lander, independent of the reference sources so
// that it can be rebuilt on the GPU box (where /root/reference does not exist).  Each function cites the reference
// lines it follows.  Pinning status:
//   * lander physics / penalties / per-member accumulation: PINNED — checked bit-for-bit against the reference's own
//     sources compiled unchanged (oracle/_ref/libll_ref.so, see oracle/Makefile) in tests/test_oracle.py, and against
//     the constant-action known answers of SURVEY.md App. D (tests/golden/env_known_answers.json).
//   * policy network forward pass and GA: PARITY UNPINNED — they live in the un-vendored submodule
//     azimonti/ma-libs @ dd81ee426fe6b8c83dc427082c62f2a8f2578cd0 (absent; the reference ships no test or golden
//     vector for it).  Restated from the call sites and the 49 sample weight files; decisions recorded in DESIGN.md.
// Build: oracle/Makefile, g++ -O2 -ffp-contract=off -fno-fast-math (every operation individually rounded).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <thread>
#include <type_traits>
#include <vector>

#include "../include/ll_gpu.h"
#include "ll_math.h"
#include "ll_philox.h"
#include "ll_devlayout.h"

namespace
{
    // ---- constants: GameLogicCpp<T>::GameLogicCpp + recalculate_derived_values + reset  (src/game_logic.cpp:33-140)
    template <typename T> struct Env
    {
        T pad_y1, terrain_y, max_safe_vx, max_safe_vy, g, mu_x, mu_y, lander_w, lander_h, max_fuel, lpad_w, width, height, spad_w;
        T y0, vx0, vy0, ay0;
        T sp_dist, sp_action, tp_steps, tp_dist, tp_landed_bonus, tp_landed_lr_bonus, tp_fuel_bonus, tp_crashed, tp_crash_v_mag, tp_no_fuel;
        explicit Env(const ll_env_cfg& c)
        {
            pad_y1 = static_cast<T>(c.pad_y1), terrain_y = static_cast<T>(c.terrain_y);
            max_safe_vx = std::abs(static_cast<T>(c.max_vx)), max_safe_vy = std::abs(static_cast<T>(c.max_vy)); // :51-56
            if (max_safe_vx == T(0.0)) max_safe_vx = T(1e-6);
            if (max_safe_vy == T(0.0)) max_safe_vy = T(1e-6);
            g = static_cast<T>(c.g), mu_x = static_cast<T>(c.mu_x), mu_y = static_cast<T>(c.mu_y);
            lander_w = static_cast<T>(c.lander_width), lander_h = static_cast<T>(c.lander_height);
            max_fuel = static_cast<T>(static_cast<int>(c.max_fuel)); // getInt, :64
            if (max_fuel <= T(0.0)) max_fuel = T(1e-6);              // :121
            lpad_w = static_cast<T>(c.lpad_width), width = static_cast<T>(c.display_width), height = static_cast<T>(c.display_height);
            spad_w = static_cast<T>(c.spad_width);
            y0 = static_cast<T>(c.y0), vx0 = static_cast<T>(c.vx0), vy0 = static_cast<T>(c.vy0), ay0 = static_cast<T>(c.ay0);
            sp_dist = static_cast<T>(c.sp_dist_factor), sp_action = static_cast<T>(c.sp_action_factor);
            tp_steps = static_cast<T>(c.tp_steps_factor), tp_dist = static_cast<T>(c.tp_dist_factor);
            tp_landed_bonus = static_cast<T>(c.tp_landed_bonus), tp_landed_lr_bonus = static_cast<T>(c.tp_landed_lr_bonus);
            tp_fuel_bonus = static_cast<T>(c.tp_fuel_bonus_factor), tp_crashed = static_cast<T>(c.tp_crashed_penalty);
            tp_crash_v_mag = static_cast<T>(c.tp_crash_v_mag_factor), tp_no_fuel = static_cast<T>(c.tp_no_fuel_penalty);
        }
    };

    // ---- one lander: the state machine of GameLogicCpp<T> (src/game_logic.cpp:124-320, 352-388)
    template <typename T> struct Lander
    {
        const Env<T>& e;
        T x, y, vx, vy, fuel, ay, lpad_x1, center_x, ground;
        bool landed, crashed, success;
        int t;
        explicit Lander(const Env<T>& env) : e(env) {}

        void reset(T spad_x1, T lpad) // :143-154 -> :101-122 -> :124-140
        {
            lpad_x1       = lpad;
            const T init_x = spad_x1 + (e.spad_w / T(2.0)) - (e.lander_w / T(2.0));
            if (e.width > T(0.0) && e.height > T(0.0))
            {
                center_x = lpad_x1 + e.lpad_w / T(2.0);
                ground   = e.height - e.lander_h - e.terrain_y;
            }
            else { center_x = T(0.0), ground = T(0.0); }
            x = init_x, y = e.y0, vx = e.vx0, vy = e.vy0;
            ay   = e.ay0 + e.g;
            fuel = e.max_fuel;
            landed = crashed = success = false;
            t = 0;
        }
        bool done() const { return landed || crashed; } // :253-256

        void state(T* s) const // :275-293
        {
            const T dx = x - center_x, dy = y - e.pad_y1;
            s[0] = vx / e.max_safe_vx, s[1] = vy / e.max_safe_vy, s[2] = dx / e.width, s[3] = dy / e.height, s[4] = fuel / e.max_fuel;
            s[0] = std::clamp(s[0], T(-2.0), T(2.0)), s[1] = std::clamp(s[1], T(-2.0), T(2.0));
            s[2] = std::clamp(s[2], T(-5.0), T(5.0)), s[3] = std::clamp(s[3], T(-5.0), T(5.0));
        }

        bool update(int action, T* s) // :295-310
        {
            if (!done())
            {
                if (!(fuel <= T(0.0))) // apply_action :156-183
                {
                    if (action == 1) { vy -= T(0.3), fuel -= T(1.0); }
                    else if (action == 2) { vx -= T(0.05), fuel -= T(0.5); }
                    else if (action == 3) { vx += T(0.05), fuel -= T(0.5); }
                    fuel = std::max(T(0.0), fuel);
                }
                vy += ay; // update_physics :185-195
                vx *= (T(1.0) - e.mu_x);
                vy *= (T(1.0) - e.mu_y);
                x += vx;
                y += vy;
                if (t != 0 && y >= ground) // check_landing_crash :197-239
                {
                    y      = ground;
                    landed = true;
                    const bool on_pad = (x >= lpad_x1 && (x + e.lander_w) <= (lpad_x1 + e.lpad_w));
                    const bool safe   = (std::abs(vx) < e.max_safe_vx && std::abs(vy) < e.max_safe_vy);
                    success           = on_pad && safe;
                    crashed           = !success;
                    vx = T(0.0), vy = T(0.0);
                }
                ++t;
            }
            state(s);
            return done();
        }

        T step_penalty(int action) const // :312-320
        {
            T p = T(0.0);
            p += (std::abs(x - center_x) + std::abs(y - e.pad_y1)) * e.sp_dist;
            if (action > 0) p += e.sp_action;
            return p;
        }

        T terminal_penalty(int steps, size_t dir, T* bonus_lr) const // :352-388
        {
            T p = T(0.0);
            p += static_cast<T>(steps) * e.tp_steps;
            p += (std::abs(x - center_x) + std::abs(y - e.pad_y1)) * e.tp_dist;
            if (success)
            {
                p -= e.tp_landed_bonus;
                p -= fuel * e.tp_fuel_bonus;
                bonus_lr[dir] += e.tp_landed_lr_bonus;
            }
            else if (crashed)
            {
                p += e.tp_crashed;
                p += std::sqrt(vx * vx + vy * vy) * e.tp_crash_v_mag;
            }
            else if (fuel <= T(0.0) && !landed) { p += e.tp_no_fuel; }
            if (!success) p += std::sqrt(vx * vx + vy * vy) * T(1.0);
            return p;
        }
    };

    // ---- policy: restated nn::ANN_MLP_GA<T>::feedforward (ma-libs; call site src/nn_engine_trainer.cpp:501-502)
    // a_{l+1} = act(W_l a_l + b_l) on every layer, W_l [out][in] row-major (canonical member layout), ascending-k
    // accumulation from 0 with separate multiply and add, bias added last, activation = ll_math.h evaluated in double,
    // action = first maximum.  PARITY UNPINNED (see header).
    int g_libm_activation = 0; // orc_set_activation_impl

    template <typename T> struct Net
    {
        std::vector<uint64_t> sizes;
        int act;
        size_t P = 0;
        std::vector<size_t> w_at, b_at;
        Net(const uint64_t* s, int n, int activation) : sizes(s, s + n), act(activation)
        {
            for (size_t l = 0; l + 1 < sizes.size(); ++l)
            {
                w_at.push_back(P);
                b_at.push_back(P + sizes[l] * sizes[l + 1]);
                P += sizes[l] * sizes[l + 1] + sizes[l + 1];
            }
        }
        T activate(T z) const
        {
            // SURVEY.md H2 cross-check: libm's tanh / exp evaluated in T (what a C++ template library most plausibly calls), independent
            // of the product's csrc/ll_math.h — only used to measure how far the deterministic ll_tanh moves the action sequences
            if (g_libm_activation) return act == LL_ACT_TANH ? std::tanh(z) : T(1) / (T(1) + std::exp(-z));
            const double zd = static_cast<double>(z);
            return static_cast<T>(act == LL_ACT_TANH ? ll_tanh(zd) : ll_sigmoid(zd));
        }
        int forward(const T* p, const T* in) const
        {
            T a[1024], b[1024];
            T *cur = a, *nxt = b;
            for (uint64_t i = 0; i < sizes[0]; ++i) cur[i] = in[i];
            for (size_t l = 0; l + 1 < sizes.size(); ++l)
            {
                const size_t ni = sizes[l], no = sizes[l + 1];
                const T* w = p + w_at[l];
                const T* bias = p + b_at[l];
                for (size_t o = 0; o < no; ++o)
                {
                    T acc = T(0);
                    for (size_t k = 0; k < ni; ++k) acc = acc + w[o * ni + k] * cur[k];
                    nxt[o] = activate(acc + bias[o]);
                }
                std::swap(cur, nxt);
            }
            int best = 0;
            for (uint64_t j = 1; j < sizes.back(); ++j)
                if (cur[j] > cur[best]) best = (int)j;
            return best;
        }
    };

    struct EpisodeOut
    {
        int steps, flags;
        double acc, term;
    };

    // one (member, layout) episode: body of sim_task, src/nn_engine_trainer.cpp:486-519
    template <typename T>
    EpisodeOut run_episode(const Env<T>& env, const Net<T>& net, const T* params, int spad, int lpad, size_t dir, int max_steps, T* bonus_lr,
                           int fixed_action, double* rec, int cap)
    {
        Lander<T> sim(env);
        T s[5];
        sim.reset(static_cast<T>(spad), static_cast<T>(lpad));
        sim.state(s);
        double acc = 0.0;
        int steps  = 0;
        while (steps < max_steps)
        {
            const int a     = fixed_action >= 0 ? fixed_action : net.forward(params, s);
            const bool done = sim.update(a, s);
            acc += static_cast<double>(sim.step_penalty(a));
            if (rec && steps < cap)
            {
                double* r = rec + (size_t)steps * 11;
                r[0]      = a;
                for (int i = 0; i < 5; ++i) r[1 + i] = static_cast<double>(s[i]);
                r[6] = sim.x, r[7] = sim.y, r[8] = sim.vx, r[9] = sim.vy, r[10] = sim.fuel;
            }
            ++steps;
            if (done) break;
        }
        EpisodeOut o;
        o.term  = static_cast<double>(sim.terminal_penalty(steps, dir, bonus_lr));
        o.acc   = acc;
        o.steps = steps;
        o.flags = (sim.landed ? 1 : 0) | (sim.crashed ? 2 : 0) | (sim.success ? 4 : 0);
        return o;
    }

    // train_generation: src/nn_engine_trainer.cpp:443-539 (one task per member; `threads` workers instead of tp::thread_pool)
    template <typename T>
    int evaluate(const ll_env_cfg* cfg, const uint64_t* sizes, int n_sizes, int act, uint64_t pop, const T* params, const int32_t* spad,
                 const int32_t* lpad, const uint64_t* dirs, int L, int max_steps, int threads, double* fitness, double* lr, double* steps,
                 int32_t* ep_meta, double* ep_fit, double* seconds)
    {
        const Env<T> env(*cfg);
        const Net<T> net(sizes, n_sizes, act);
        std::atomic<uint64_t> next{0};
        auto worker = [&]() {
            for (;;)
            {
                const uint64_t m = next.fetch_add(1);
                if (m >= pop) return;
                T bonus[2]   = {T(0.0), T(0.0)};
                double total = 0.0, total_steps = 0.0;
                for (int l = 0; l < L; ++l)
                {
                    const EpisodeOut o = run_episode<T>(env, net, params + m * net.P, spad[l], lpad[l], (size_t)dirs[l], max_steps, bonus, -1, nullptr, 0);
                    double f           = o.term;
                    f += o.acc;
                    total += f;
                    total_steps += static_cast<double>(o.steps);
                    if (ep_meta) ep_meta[m * L + l] = o.steps | (o.flags << 28);
                    if (ep_fit) ep_fit[m * L + l] = f;
                }
                const double b = static_cast<double>(std::min(bonus[0], bonus[1])); // :521-525
                total -= b;
                fitness[m] = total, lr[m] = b, steps[m] = total_steps;
            }
        };
        const auto t0 = std::chrono::steady_clock::now();
        if (threads <= 1) worker();
        else
        {
            std::vector<std::thread> th;
            for (int i = 0; i < threads; ++i) th.emplace_back(worker);
            for (auto& t : th) t.join();
        }
        if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        return 0;
    }

    // GA (restated design shared with the product: Philox-keyed, see ll_philox.h and oracle/shim/ann_mlp_ga_v1.h)
    template <typename T>
    void create_population(const uint64_t* sizes, int n_sizes, uint32_t seed, uint64_t pop, uint64_t top, uint32_t gen, int strategy, double ratio,
                           int elitism, int has_elites, double sigma_init, double sigma_mut, T* params)
    {
        const Net<T> net(sizes, n_sizes, 1);
        const size_t P   = net.P;
        const uint64_t first = has_elites ? (elitism ? top : 0) : 0;
        uint64_t n_inject    = pop - first;
        if (has_elites)
        {
            n_inject = 0;
            if (strategy == LL_STRATEGY_MIXED_WITH_RANDOM_INJECTION) n_inject = (uint64_t)std::floor(ratio * (double)(pop - first) + 0.5);
            n_inject = std::min<uint64_t>(n_inject, pop - first);
        }
        std::vector<T> elites;
        if (has_elites) elites.assign(params, params + top * P);
        const T s_init = static_cast<T>(sigma_init), s_mut = static_cast<T>(sigma_mut);
        const std::vector<int> d2c = orc_dev2canon(sizes, n_sizes);
        for (uint64_t i = first; i < pop; ++i)
        {
            const bool inject = ll_is_injected(i - first, pop - first, n_inject);
            const T* parent   = inject ? nullptr : &elites[(size_t)ll_parent_pick(seed, gen, (uint32_t)i, (uint32_t)top) * P];
            T* dst            = params + i * P;
            if (std::is_same<T, float>::value)
            { // float populations: four Box-Muller variates per DEVICE-order quad (ll_normal4_f32), padding positions discarded
                for (size_t qd = 0; qd * 4 < d2c.size(); ++qd)
                {
                    float g[4];
                    ll_normal4_f32(seed, gen, (uint32_t)i, (uint32_t)qd, g);
                    for (int e = 0; e < 4; ++e)
                    {
                        const int c = d2c[qd * 4 + e];
                        if (c >= 0) dst[c] = inject ? s_init * (T)g[e] : parent[c] + s_mut * (T)g[e];
                    }
                }
                continue;
            }
            for (size_t j = 0; j < P; j += 2)
            {
                double n0, n1;
                ll_normal_pair(seed, gen, (uint32_t)i, (uint32_t)(j / 2), &n0, &n1);
                const T g0 = static_cast<T>(n0), g1 = static_cast<T>(n1);
                dst[j] = inject ? s_init * g0 : parent[j] + s_mut * g0;
                if (j + 1 < P) dst[j + 1] = inject ? s_init * g1 : parent[j + 1] + s_mut * g1;
            }
        }
    }
} // namespace

extern "C"
{
    // 0: ll_tanh / ll_sigmoid of csrc/ll_math.h in double (default, what the exact kernels use); 1: libm std::tanh / std::exp in T
    void orc_set_activation_impl(int libm) { g_libm_activation = libm ? 1 : 0; }

    unsigned orc_hardware_threads() { return std::max(1u, std::thread::hardware_concurrency()); }

    int orc_evaluate_f64(const ll_env_cfg* cfg, const uint64_t* sizes, int n_sizes, int act, uint64_t pop, const double* params,
                         const int32_t* spad, const int32_t* lpad, const uint64_t* dirs, int L, int max_steps, int threads, double* fitness,
                         double* lr, double* steps, int32_t* ep_meta, double* ep_fit, double* seconds)
    {
        return evaluate<double>(cfg, sizes, n_sizes, act, pop, params, spad, lpad, dirs, L, max_steps, threads, fitness, lr, steps, ep_meta, ep_fit, seconds);
    }
    int orc_evaluate_f32(const ll_env_cfg* cfg, const uint64_t* sizes, int n_sizes, int act, uint64_t pop, const float* params,
                         const int32_t* spad, const int32_t* lpad, const uint64_t* dirs, int L, int max_steps, int threads, double* fitness,
                         double* lr, double* steps, int32_t* ep_meta, double* ep_fit, double* seconds)
    {
        return evaluate<float>(cfg, sizes, n_sizes, act, pop, params, spad, lpad, dirs, L, max_steps, threads, fitness, lr, steps, ep_meta, ep_fit, seconds);
    }

    // rec: 11 doubles per step (action, state[5] after the update, x, y, vx, vy, fuel); params may be NULL with fixed_action >= 0
    int orc_trace_f64(const ll_env_cfg* cfg, const uint64_t* sizes, int n_sizes, int act, const double* params, int spad, int lpad, uint64_t dir,
                      int max_steps, int fixed_action, double* rec, int cap, int* n_steps, int* flags, double* acc, double* term, double* bonus)
    {
        const Env<double> env(*cfg);
        const Net<double> net(sizes, n_sizes, act);
        double b[2] = {0.0, 0.0};
        const EpisodeOut o = run_episode<double>(env, net, params, spad, lpad, (size_t)dir, max_steps, b, fixed_action, rec, cap);
        *n_steps = o.steps, *flags = o.flags, *acc = o.acc, *term = o.term, bonus[0] = b[0], bonus[1] = b[1];
        return 0;
    }
    int orc_trace_f32(const ll_env_cfg* cfg, const uint64_t* sizes, int n_sizes, int act, const float* params, int spad, int lpad, uint64_t dir,
                      int max_steps, int fixed_action, double* rec, int cap, int* n_steps, int* flags, double* acc, double* term, double* bonus)
    {
        const Env<float> env(*cfg);
        const Net<float> net(sizes, n_sizes, act);
        float b[2] = {0.f, 0.f};
        const EpisodeOut o = run_episode<float>(env, net, params, spad, lpad, (size_t)dir, max_steps, b, fixed_action, rec, cap);
        *n_steps = o.steps, *flags = o.flags, *acc = o.acc, *term = o.term, bonus[0] = b[0], bonus[1] = b[1];
        return 0;
    }

    int orc_forward_f64(const uint64_t* sizes, int n_sizes, int act, const double* params, const double* state)
    {
        return Net<double>(sizes, n_sizes, act).forward(params, state);
    }
    int orc_forward_f32(const uint64_t* sizes, int n_sizes, int act, const float* params, const float* state)
    {
        return Net<float>(sizes, n_sizes, act).forward(params, state);
    }

    void orc_create_population_f64(const uint64_t* sizes, int n_sizes, uint32_t seed, uint64_t pop, uint64_t top, uint32_t gen, int strategy,
                                   double ratio, int elitism, int has_elites, double sigma_init, double sigma_mut, double* params)
    {
        create_population<double>(sizes, n_sizes, seed, pop, top, gen, strategy, ratio, elitism, has_elites, sigma_init, sigma_mut, params);
    }
    void orc_create_population_f32(const uint64_t* sizes, int n_sizes, uint32_t seed, uint64_t pop, uint64_t top, uint32_t gen, int strategy,
                                   double ratio, int elitism, int has_elites, double sigma_init, double sigma_mut, float* params)
    {
        create_population<float>(sizes, n_sizes, seed, pop, top, gen, strategy, ratio, elitism, has_elites, sigma_init, sigma_mut, params);
    }

    // selection order the product defines: ascending (fitness, index); -0.0 == +0.0  (std::sort at
    // src/nn_engine_trainer.cpp:376-379 is unstable: any order inside a tie group is a valid reference answer)
    void orc_topk(const double* fitness, uint64_t n, uint64_t k, uint64_t* idx_out)
    {
        std::vector<uint64_t> idx(n);
        std::iota(idx.begin(), idx.end(), 0);
        std::stable_sort(idx.begin(), idx.end(), [&](uint64_t a, uint64_t b) { return fitness[a] < fitness[b]; });
        for (uint64_t i = 0; i < k && i < n; ++i) idx_out[i] = idx[i];
    }

    double orc_tanh(double x) { return ll_tanh(x); }
    // ll_tanh_n (the exact kernel's tile-wide tanh, csrc/ll_math.h) on n values, 8 at a time: must equal orc_tanh bit for bit
    void orc_tanh_n(const double* x, double* y, uint64_t n)
    {
        for (uint64_t o = 0; o < n; o += 8)
        {
            double xi[8], yi[8];
            for (int i = 0; i < 8; ++i) xi[i] = o + i < n ? x[o + i] : 0.0;
            ll_tanh_n<8>(xi, yi);
            for (int i = 0; i < 8 && o + i < n; ++i) y[o + i] = yi[i];
        }
    }
    double orc_sigmoid(double x) { return ll_sigmoid(x); }
    void orc_normal_pair(uint32_t seed, uint32_t gen, uint32_t member, uint32_t pair, double* n0, double* n1) { ll_normal_pair(seed, gen, member, pair, n0, n1); }
    uint32_t orc_parent_pick(uint32_t seed, uint32_t gen, uint32_t member, uint32_t top) { return ll_parent_pick(seed, gen, member, top); }
    void orc_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out)
    {
        const ll_u32x4 r = ll_philox4x32_10(c0, c1, c2, c3, k0, k1);
        for (int i = 0; i < 4; ++i) out[i] = r.v[i];
    }
}
