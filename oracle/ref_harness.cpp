// ORACLE / TEST INFRASTRUCTURE — not part of the product.
//
// C entry points around the reference's OWN, UNMODIFIED sources (compiled from where they lie under
// /root/reference by oracle/Makefile into oracle/_ref/libll_ref*.so):
//     src/game_logic.{h,cpp}   src/game_utils.{h,cpp}   src/nn_engine_trainer.{h,cpp}
// linked against the ma-libs stand-ins in oracle/shim/.  Everything the reference owns on the hot path —
// GameLogicCpp<T> (physics, landing test, penalties) and NNEngineTrainer<T>::train_generation (the
// population x layout x step loop, src/nn_engine_trainer.cpp:443-539) — therefore runs as shipped; only
// the policy network / GA come from the shim (parity unpinned there, see shim/ann_mlp_ga_v1.h).
//
// train_generation and balance_layouts are private members; this translation unit widens the access
// specifiers while including the reference header (class layout and mangled names are unaffected).
#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <optional>
#include <random>
#include <sstream>
#include <string>
#include <vector>

#include "ann_mlp_ga_v1.h"
#include "config_loader.h"
#include "thread/thread_pool.hpp"

#define private public
#include "game_logic.h"
#include "game_utils.h"
#include "nn_engine_trainer.h"
#undef private

namespace
{
    thread_local std::string g_err;

    void apply_overrides(const char* config_path, const char* overrides)
    {
        Config::clear();
        if (!Config::loadConfiguration(config_path)) throw std::runtime_error(std::string("cannot load ") + config_path);
        // overrides: "key=value;key=value"
        if (overrides)
        {
            std::stringstream ss(overrides);
            std::string item;
            while (std::getline(ss, item, ';'))
            {
                size_t eq = item.find('=');
                if (eq != std::string::npos) Config::set(Config::strip(item.substr(0, eq)), Config::strip(item.substr(eq + 1)));
            }
        }
        Config::set("NNConfig.save_nn", "false");
    }

    std::vector<size_t> sizes_from_config()
    {
        std::vector<size_t> s{5};
        for (int h : Config::getVectorInt("NNConfig.hlayers")) s.push_back(static_cast<size_t>(h));
        s.push_back(4);
        return s;
    }

    template <typename T>
    int evaluate(const char* config_path, const char* overrides, uint64_t pop, const T* params, const int32_t* spad,
                 const int32_t* lpad, const uint64_t* dirs, int n_layouts, double* fitness, double* lr, double* steps,
                 uint64_t* ff_calls, double* seconds)
    {
        try
        {
            apply_overrides(config_path, overrides);
            Config::set("NNConfig.population_size", std::to_string(pop));
            Training::NNEngineTrainer<T> trainer; // reads the Config store (src/nn_engine_trainer.cpp:26-68)
            const auto sizes = sizes_from_config();
            trainer.net_     = std::make_unique<nn::ANN_MLP_GA<T>>(sizes, Config::getInt("NNConfig.seed"), pop,
                                                               std::min<size_t>(pop, Config::getInt("NNConfig.top_individuals")),
                                                               static_cast<size_t>(Config::getInt("NNConfig.activation_id")));
            const size_t P   = trainer.net_->ParamCount();
            for (uint64_t m = 0; m < pop; ++m) trainer.net_->SetMemberParams(m, params + m * P);

            std::vector<Training::LayoutInfo> layouts(static_cast<size_t>(n_layouts));
            std::vector<size_t> directions(static_cast<size_t>(n_layouts));
            for (int i = 0; i < n_layouts; ++i)
            {
                layouts[i].spad_x1     = spad[i];
                layouts[i].lpad_x1     = lpad[i];
                layouts[i].spad_center = spad[i] + Config::getDouble("GameCfg.spad_width") / 2.0;
                layouts[i].lpad_center = lpad[i] + Config::getDouble("GameCfg.lpad_width") / 2.0;
                directions[i]          = static_cast<size_t>(dirs[i]);
            }
            std::vector<double> f(pop), l(pop), s(pop);
            std::vector<GameLogicCpp<T>> sims; // as in train(): src/nn_engine_trainer.cpp:325-330
            sims.reserve(pop);
            for (uint64_t i = 0; i < pop; ++i) sims.emplace_back(true);

            auto t0 = std::chrono::steady_clock::now();
            trainer.train_generation(layouts, directions, pop, f, l, s, sims);
            auto t1 = std::chrono::steady_clock::now();

            std::memcpy(fitness, f.data(), pop * sizeof(double));
            std::memcpy(lr, l.data(), pop * sizeof(double));
            std::memcpy(steps, s.data(), pop * sizeof(double));
            if (ff_calls) *ff_calls = trainer.net_->FeedforwardCalls();
            if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
            return 0;
        } catch (const std::exception& e)
        {
            g_err = e.what();
            return 1;
        }
    }

    // One (member, layout) episode stepped exactly as the body of sim_task does
    // (src/nn_engine_trainer.cpp:491-517), recording every step.  rec layout per step (11 doubles):
    // action, state[0..4] (after update), x, y, vx, vy, fuel.
    template <typename T>
    int trace(const char* config_path, const char* overrides, const T* member_params, int spad, int lpad, uint64_t dir,
              int fixed_action, double* rec, int cap, int* n_steps, int* flags, double* step_penalty_sum,
              double* terminal_penalty, double* bonus_lr)
    {
        try
        {
            apply_overrides(config_path, overrides);
            const auto sizes = sizes_from_config();
            nn::ANN_MLP_GA<T> net(sizes, 0, 1, 1, static_cast<size_t>(Config::getInt("NNConfig.activation_id")));
            if (member_params) net.SetMemberParams(0, member_params);
            const int max_steps = Config::getInt("GameCfg.max_steps");
            GameLogicCpp<T> sim(true);
            std::vector<T> state(5);
            std::array<T, 2> bonus{T(0), T(0)};
            sim.reset(static_cast<T>(spad), static_cast<T>(lpad));
            sim.get_state(state.data(), state.size());
            double acc = 0.0;
            int steps  = 0;
            while (steps < max_steps)
            {
                size_t a  = fixed_action >= 0 ? static_cast<size_t>(fixed_action) : net.feedforward(state.data(), 5, 0);
                bool done = sim.update(static_cast<int>(a), state.data(), state.size());
                acc += static_cast<double>(sim.calculate_step_penalty(static_cast<int>(a)));
                if (rec && steps < cap)
                {
                    double* r = rec + static_cast<size_t>(steps) * 11;
                    r[0]      = static_cast<double>(a);
                    for (int i = 0; i < 5; ++i) r[1 + i] = static_cast<double>(state[i]);
                    r[6] = sim.x, r[7] = sim.y, r[8] = sim.vx, r[9] = sim.vy, r[10] = sim.fuel;
                }
                steps++;
                if (done) break;
            }
            double term = static_cast<double>(sim.calculate_terminal_penalty(steps, static_cast<size_t>(dir), bonus));
            *n_steps    = steps;
            *flags      = (sim.landed ? 1 : 0) | (sim.crashed ? 2 : 0) | (sim.landed_successfully ? 4 : 0);
            *step_penalty_sum = acc;
            *terminal_penalty = term;
            bonus_lr[0] = bonus[0], bonus_lr[1] = bonus[1];
            return 0;
        } catch (const std::exception& e)
        {
            g_err = e.what();
            return 1;
        }
    }
} // namespace

extern "C"
{
    const char* ref_last_error() { return g_err.c_str(); }
    unsigned ref_threads() { return tp::thread_pool::default_threads(); }

    int ref_evaluate_f64(const char* cfg, const char* ov, uint64_t pop, const double* params, const int32_t* spad,
                         const int32_t* lpad, const uint64_t* dirs, int n_layouts, double* fitness, double* lr,
                         double* steps, uint64_t* ff_calls, double* seconds)
    {
        return evaluate<double>(cfg, ov, pop, params, spad, lpad, dirs, n_layouts, fitness, lr, steps, ff_calls, seconds);
    }
    int ref_evaluate_f32(const char* cfg, const char* ov, uint64_t pop, const float* params, const int32_t* spad,
                         const int32_t* lpad, const uint64_t* dirs, int n_layouts, double* fitness, double* lr,
                         double* steps, uint64_t* ff_calls, double* seconds)
    {
        return evaluate<float>(cfg, ov, pop, params, spad, lpad, dirs, n_layouts, fitness, lr, steps, ff_calls, seconds);
    }
    int ref_trace_f64(const char* cfg, const char* ov, const double* p, int spad, int lpad, uint64_t dir, int fixed_action,
                      double* rec, int cap, int* n, int* flags, double* sp, double* tp_, double* bonus)
    {
        return trace<double>(cfg, ov, p, spad, lpad, dir, fixed_action, rec, cap, n, flags, sp, tp_, bonus);
    }
    int ref_trace_f32(const char* cfg, const char* ov, const float* p, int spad, int lpad, uint64_t dir, int fixed_action,
                      double* rec, int cap, int* n, int* flags, double* sp, double* tp_, double* bonus)
    {
        return trace<float>(cfg, ov, p, spad, lpad, dir, fixed_action, rec, cap, n, flags, sp, tp_, bonus);
    }

    // src/game_utils.cpp:12-74 verbatim
    void ref_generate_random_pad_positions(double width, int spad_w, int lpad_w, int has_seed, int seed, int force_ltr,
                                           int force_rtl, int* spad_x1, int* lpad_x1)
    {
        auto p   = generate_random_pad_positions(width, spad_w, lpad_w, has_seed ? std::optional<int>(seed) : std::nullopt,
                                                 force_ltr != 0, force_rtl != 0);
        *spad_x1 = p.first;
        *lpad_x1 = p.second;
    }

    // Layout set exactly as train() builds it (src/nn_engine_trainer.cpp:283-309): seeds base+i, then
    // balance_layouts (:196-260).  `directions` is in/out (the early return at :214 leaves it untouched).
    int ref_make_layouts(const char* cfg, const char* ov, unsigned base_seed, int n, int32_t* spad, int32_t* lpad,
                         uint64_t* directions)
    {
        try
        {
            apply_overrides(cfg, ov);
            Training::NNEngineTrainer<double> trainer;
            std::vector<Training::LayoutInfo> v;
            for (int i = 0; i < n; ++i)
            {
                auto p = generate_random_pad_positions(trainer.cfg_width_, static_cast<int>(trainer.game_cfg_spad_width_),
                                                       static_cast<int>(trainer.game_cfg_lpad_width_),
                                                       static_cast<int>(base_seed + static_cast<unsigned>(i)));
                Training::LayoutInfo li;
                li.spad_x1     = p.first;
                li.lpad_x1     = p.second;
                li.spad_center = static_cast<double>(p.first) + trainer.game_cfg_spad_width_ / 2.0;
                li.lpad_center = static_cast<double>(p.second) + trainer.game_cfg_lpad_width_ / 2.0;
                v.push_back(li);
            }
            std::vector<size_t> d(static_cast<size_t>(n));
            for (int i = 0; i < n; ++i) d[i] = static_cast<size_t>(directions[i]);
            trainer.balance_layouts(v, d, n, trainer.left_right_ratio_);
            for (int i = 0; i < n; ++i) spad[i] = v[i].spad_x1, lpad[i] = v[i].lpad_x1, directions[i] = d[i];
            return 0;
        } catch (const std::exception& e)
        {
            g_err = e.what();
            return 1;
        }
    }

    // main.py --mode=nn_play without the window (main.py:127-238): the reference's GameLogicCpp<double> and
    // generate_random_pad_positions, the shim network loaded from `weight_file` (mod_nn_play.py:25-59), the UI loop's start
    // gating with time = frame / fps, penalty accumulated in a Python float (double).  Test-only counterpart of
    // lunar_lander_b200/nn_play.py.  rec (may be null): per update 6 doubles = action, state[0..4].
    int ref_nn_play(const char* cfg, const char* ov, const char* weight_file, int has_seed, int seed, int force_ltr,
                    int force_rtl, double fps, int max_frames, int* out_int /*spad,lpad,frames,steps,start_frame,flags*/,
                    double* out_f64 /*penalty,x,y,fuel*/, double* rec, int cap)
    {
        try
        {
            apply_overrides(cfg, ov);
            GameLogicCpp<double> logic(true);
            auto pads = generate_random_pad_positions(Config::getDouble("DisplayCfg.width"),
                                                      static_cast<int>(Config::getDouble("GameCfg.spad_width")),
                                                      static_cast<int>(Config::getDouble("GameCfg.lpad_width")),
                                                      has_seed ? std::optional<int>(seed) : std::nullopt, force_ltr != 0,
                                                      force_rtl != 0);
            logic.reset(static_cast<double>(pads.first), static_cast<double>(pads.second));
            nn::ANN_MLP_GA<double> net;
            net.Deserialize(weight_file);
            double state[5] = {0, 0, 0, 0, 0};
            bool started = false, over = false;
            double penalty = 0.0;
            int steps = 0, frame = 0, start_frame = -1;
            while (frame < max_frames && !over)
            {
                int action = 0;
                logic.get_state(state, 5);
                const int nn_action = static_cast<int>(net.feedforward(state, 5, 0));
                if (!started)
                {
                    const double elapsed_ms = frame * 1000.0 / fps;
                    if (elapsed_ms > 2000) action = 1, started = true;
                    else if (nn_action != 0) action = 1, started = true;
                    if (started) start_frame = frame;
                }
                else action = nn_action;
                if (started)
                {
                    const bool done = logic.update(action, state, 5);
                    penalty += logic.calculate_step_penalty(action);
                    if (rec && steps < cap)
                    {
                        rec[steps * 6] = action;
                        for (int i = 0; i < 5; ++i) rec[steps * 6 + 1 + i] = state[i];
                    }
                    ++steps;
                    if (done) over = true;
                }
                if (over) penalty += logic.calculate_terminal_penalty(steps);
                ++frame;
            }
            out_int[0] = pads.first, out_int[1] = pads.second, out_int[2] = frame, out_int[3] = steps, out_int[4] = start_frame;
            out_int[5] = (logic.landed ? 1 : 0) | (logic.crashed ? 2 : 0) | (logic.landed_successfully ? 4 : 0) | (over ? 8 : 0);
            out_f64[0] = penalty, out_f64[1] = logic.x, out_f64[2] = logic.y, out_f64[3] = logic.fuel;
            return 0;
        } catch (const std::exception& e)
        {
            g_err = e.what();
            return 1;
        }
    }

    // GA of the shim, exposed so the product's GA kernels can be checked against it.
    // params: [pop][P] canonical, in/out.  has_elites: 0 -> initial population, 1 -> children of slots [0,top).
    int ref_create_population_f64(const uint64_t* sizes, int n_sizes, int seed, uint64_t pop, uint64_t top, uint64_t epochs,
                                  double ratio, int elitism, int has_elites, double* params)
    {
        try
        {
            nn::ANN_MLP_GA<double> net(std::vector<size_t>(sizes, sizes + n_sizes), seed, pop, top, 1);
            net.SetPopulationStrategy(nn::PopulationStrategy::MIXED_WITH_RANDOM_INJECTION, ratio);
            net.UpdateEpochs(epochs);
            const size_t P = net.ParamCount();
            if (has_elites)
            {
                for (uint64_t m = 0; m < top; ++m) net.SetMemberParams(m, params + m * P);
                net.MarkElites();
            }
            net.CreatePopulation(elitism != 0);
            for (uint64_t m = 0; m < pop; ++m) std::memcpy(params + m * P, net.MemberParams(m), P * sizeof(double));
            return 0;
        } catch (const std::exception& e)
        {
            g_err = e.what();
            return 1;
        }
    }
    int ref_create_population_f32(const uint64_t* sizes, int n_sizes, int seed, uint64_t pop, uint64_t top, uint64_t epochs,
                                  double ratio, int elitism, int has_elites, float* params)
    {
        try
        {
            nn::ANN_MLP_GA<float> net(std::vector<size_t>(sizes, sizes + n_sizes), seed, pop, top, 1);
            net.SetPopulationStrategy(nn::PopulationStrategy::MIXED_WITH_RANDOM_INJECTION, ratio);
            net.UpdateEpochs(epochs);
            const size_t P = net.ParamCount();
            if (has_elites)
            {
                for (uint64_t m = 0; m < top; ++m) net.SetMemberParams(m, params + m * P);
                net.MarkElites();
            }
            net.CreatePopulation(elitism != 0);
            for (uint64_t m = 0; m < pop; ++m) std::memcpy(params + m * P, net.MemberParams(m), P * sizeof(float));
            return 0;
        } catch (const std::exception& e)
        {
            g_err = e.what();
            return 1;
        }
    }

    // Weight-file round trip through the shim reader (loads every sample_runs file).
    int ref_load_weight_file(const char* path, uint64_t* sizes, int* n_sizes, uint64_t* pop, uint64_t* top,
                             uint64_t* epochs, double* elite_params, uint64_t cap)
    {
        try
        {
            nn::ANN_MLP_GA<double> net;
            net.Deserialize(path);
            auto s   = net.GetNetworkSize();
            *n_sizes = static_cast<int>(s.size());
            for (size_t i = 0; i < s.size(); ++i) sizes[i] = s[i];
            *pop = net.GetPopSize(), *top = net.GetTopPerformersSize(), *epochs = net.GetEpochs();
            const size_t P = net.ParamCount();
            if (elite_params)
            {
                if (cap < *top * P) throw std::runtime_error("buffer too small");
                for (uint64_t m = 0; m < *top; ++m) std::memcpy(elite_params + m * P, net.MemberParams(m), P * sizeof(double));
            }
            return 0;
        } catch (const std::exception& e)
        {
            g_err = e.what();
            return 1;
        }
    }
}
