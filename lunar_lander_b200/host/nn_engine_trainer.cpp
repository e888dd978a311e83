#include "nn_engine_trainer.h"

#include <algorithm>
#include <cmath>
#include <filesystem>
#include <iomanip>
#include <iostream>
#include <limits>
#include <numeric>
#include <sstream>
#include <stdexcept>

#include "config_loader.h"

namespace fs = std::filesystem;

namespace Training
{
    int LayoutSet::left_to_right() const
    {
        int n = 0;
        for (size_t i = 0; i < size(); ++i) n += spad_center[i] < lpad_center[i] ? 1 : 0;
        return n;
    }

    // Everything the reference's constructor caches (src/nn_engine_trainer.cpp:29-65), same keys.
    template <typename T> NNEngineTrainer<T>::NNEngineTrainer()
    {
        name_            = Config::getString("NNConfig.name");
        save_dir_        = Config::getString("NNConfig.save_path_nn");
        save_enabled_    = Config::getBool("NNConfig.save_nn");
        overwrite_       = Config::getBool("NNConfig.overwrite");
        save_interval_   = Config::getInt("NNConfig.save_interval");
        epochs_          = Config::getInt("NNConfig.epochs");
        seed_            = Config::getInt("NNConfig.seed");
        population_      = Config::getInt("NNConfig.population_size");
        top_             = Config::getInt("NNConfig.top_individuals");
        activation_      = Config::getInt("NNConfig.activation_id");
        elitism_         = Config::getBool("NNConfig.elitism");
        injection_ratio_ = Config::getDouble("NNConfig.random_injection_ratio");

        sizes_.push_back(5); // the five state variables of get_state (src/game_logic.cpp:275-293)
        for (int h : Config::getVectorInt("NNConfig.hlayers")) sizes_.push_back(static_cast<size_t>(h));
        sizes_.push_back(4); // noop / up / left / right

        layout_nb_     = Config::getInt("NNTrainingCfg.layout_nb");
        ltr_ratio_     = Config::getDouble("NNTrainingCfg.left_right_ratio");
        reset_period_  = Config::getInt("NNTrainingCfg.reset_period");
        (void)Config::getBool("NNTrainingCfg.multithread"); // read like the reference; the GPU grid replaces the pool
        lr_bonus_unit_ = static_cast<int>(static_cast<T>(Config::getDouble("NNTrainingCfg.tp_landed_lr_bonus")));

        width_      = Config::getDouble("DisplayCfg.width");
        height_     = Config::getDouble("DisplayCfg.height");
        spad_width_ = Config::getDouble("GameCfg.spad_width");
        lpad_width_ = Config::getDouble("GameCfg.lpad_width");
        max_steps_  = Config::getInt("GameCfg.max_steps");

        stagnated_   = 0;
        last_best_   = std::numeric_limits<double>::max();
        select_mode_ = Config::getStringOr("GpuCfg.select", "auto");
    }

    template <typename T> bool NNEngineTrainer<T>::init()
    {
        if (save_enabled_ && !save_dir_.empty())
        {
            std::error_code ec;
            if (!fs::exists(save_dir_, ec)) fs::create_directories(save_dir_, ec);
            if (ec)
            {
                std::cerr << "Error creating save directory " << save_dir_ << ": " << ec.message() << std::endl;
                return false;
            }
        }
        try
        {
            net_ = std::make_unique<nn::ANN_MLP_GA<T>>(sizes_, seed_, static_cast<size_t>(population_), static_cast<size_t>(top_),
                                                       static_cast<size_t>(activation_));
            net_->SetName(name_.c_str());
            net_->SetPopulationStrategy(nn::PopulationStrategy::MIXED_WITH_RANDOM_INJECTION, injection_ratio_);
            net_->CreatePopulation(elitism_);
        } catch (const std::exception& e)
        {
            std::cerr << "Error initializing the GPU network: " << e.what() << std::endl;
            net_.reset();
            return false;
        }
        generation_ = 0;
        t0_         = std::chrono::steady_clock::now();
        return true;
    }

    // "last" -> <name>_last.txt, anything else -> <name>_<step>.txt (src/nn_engine_trainer.cpp:107-114)
    template <typename T> bool NNEngineTrainer<T>::load(const std::string& step)
    {
        const fs::path file = fs::path(save_dir_) / (name_ + "_" + step + ".txt");
        if (!fs::exists(file))
        {
            std::cerr << "Error loading network: File not found at " << file.string() << std::endl;
            net_.reset();
            return false;
        }
        try
        {
            net_ = std::make_unique<nn::ANN_MLP_GA<T>>();
            net_->SetName(name_.c_str());
            net_->Deserialize(file.string());
            generation_ = static_cast<int>(net_->GetEpochs());
            sizes_      = net_->GetNetworkSize();
            t0_         = std::chrono::steady_clock::now();
            return true;
        } catch (const std::exception& e)
        {
            std::cerr << "Error loading network from " << file.string() << ": " << e.what() << std::endl;
            net_.reset();
            return false;
        }
    }

    // <name>_<s0>_<s1>_..._s_<generation, 4 digits>.txt (or <name>_latest_overwrite.txt), then a copy as <name>_last.txt
    // (src/nn_engine_trainer.cpp:140-183)
    template <typename T> bool NNEngineTrainer<T>::save()
    {
        if (!net_)
        {
            std::cerr << "Save error: Network not initialized." << std::endl;
            return false;
        }
        std::ostringstream fname;
        fname << name_;
        if (overwrite_) fname << "_latest_overwrite.txt";
        else
        {
            for (size_t s : net_->GetNetworkSize()) fname << "_" << s;
            fname << "_s_" << std::setw(4) << std::setfill('0') << generation_ << ".txt";
        }
        const fs::path numbered = fs::path(save_dir_) / fname.str();
        const fs::path last     = fs::path(save_dir_) / (name_ + "_last.txt");
        try
        {
            fs::remove(numbered);
            net_->Serialize(numbered.string());
            fs::remove(last);
            fs::copy_file(numbered, last, fs::copy_options::overwrite_existing);
            return true;
        } catch (const std::exception& e)
        {
            std::cerr << "Error during network serialization/copy: " << e.what() << std::endl;
            return false;
        }
    }

    template <typename T> std::string NNEngineTrainer<T>::mmss(long long s)
    {
        if (s < 0) return "00:00";
        char buf[48];
        std::snprintf(buf, sizeof buf, "%02lld:%02lld", s / 60, s % 60);
        return buf;
    }

    // layout i of a set is drawn from seed_base + i, then the set is balanced to the configured left/right ratio
    // (src/nn_engine_trainer.cpp:283-309 and :338-365; generator src/game_utils.cpp:12-74, balance :196-260)
    template <typename T> LayoutSet NNEngineTrainer<T>::draw_layouts(unsigned int seed_base)
    {
        const size_t n = static_cast<size_t>(std::max(layout_nb_, 0));
        LayoutSet ls;
        ls.spad_x1.resize(n), ls.lpad_x1.resize(n), ls.spad_center.resize(n), ls.lpad_center.resize(n);
        if (directions_.size() != n) directions_.assign(n, 0);
        for (size_t i = 0; i < n; ++i)
        {
            const int32_t seed = static_cast<int32_t>(seed_base + static_cast<unsigned int>(i));
            if (ll_generate_random_pad_positions(width_, static_cast<int32_t>(spad_width_), static_cast<int32_t>(lpad_width_), 1, seed,
                                                 0, 0, &ls.spad_x1[i], &ls.lpad_x1[i]) != LL_OK)
                throw std::runtime_error(std::string("generate_random_pad_positions: ") + ll_last_error(nullptr));
            ls.spad_center[i] = static_cast<double>(ls.spad_x1[i]) + spad_width_ / 2.0;
            ls.lpad_center[i] = static_cast<double>(ls.lpad_x1[i]) + lpad_width_ / 2.0;
        }
        if (n && ll_balance_layouts(ls.spad_x1.data(), ls.lpad_x1.data(), ls.spad_center.data(), ls.lpad_center.data(),
                                    directions_.data(), static_cast<int32_t>(n), ltr_ratio_, width_, spad_width_, lpad_width_) != LL_OK)
            throw std::runtime_error(std::string("balance_layouts: ") + ll_last_error(nullptr));
        ls.direction = directions_;
        return ls;
    }

    template <typename T> int NNEngineTrainer<T>::landings_both_directions(double lr_bonus) const
    {
        if (lr_bonus_unit_ == 0) return 0; // the reference divides by zero here
        return static_cast<int>(std::round(lr_bonus / lr_bonus_unit_)); // src/nn_engine_trainer.cpp:542-545
    }

    // host: fitness comes back and is argsorted with std::sort exactly like src/nn_engine_trainer.cpp:376-379 (identical tie
    // order, since it is the same algorithm on the same values); device: radix top-k ordered by (fitness, index), only
    // the statistics cross PCIe.  auto = host up to 65 536 members.
    template <typename T> bool NNEngineTrainer<T>::select_on_device() const
    {
        if (select_mode_ == "device") return true;
        if (select_mode_ == "host") return false;
        return net_->GetPopSize() > 65536;
    }

    // One generation: evaluate (the fused episode kernel), rank, stagnation bookkeeping, install the elites, breed,
    // epochs += 1  (src/nn_engine_trainer.cpp:370-407).
    template <typename T> GenerationReport NNEngineTrainer<T>::run_generation(const LayoutSet& layouts)
    {
        GenerationReport rep;
        const size_t pop = net_->GetPopSize();
        if (select_on_device())
        {
            ll_gen_stats st{};
            net_->EvaluateGeneration(layouts.spad_x1, layouts.lpad_x1, layouts.direction, max_steps_, nullptr, nullptr, nullptr, &st);
            const std::vector<size_t> elites = net_->SelectOnDevice();
            rep.best_member  = elites.empty() ? static_cast<size_t>(st.best_index) : elites[0];
            rep.best_fitness = st.best_fitness;
            rep.best_lr      = st.best_lr;
            rep.sum_fitness  = st.sum_fitness;
            rep.sum_steps    = st.sum_steps;
        }
        else
        {
            fitness_.resize(pop), lr_.resize(pop), steps_.resize(pop);
            net_->EvaluateGeneration(layouts.spad_x1, layouts.lpad_x1, layouts.direction, max_steps_, fitness_.data(), lr_.data(),
                                     steps_.data(), nullptr);
            std::vector<size_t> order(pop);
            std::iota(order.begin(), order.end(), size_t(0));
            std::sort(order.begin(), order.end(), [this](size_t a, size_t b) { return fitness_[a] < fitness_[b]; });
            rep.best_member  = order[0];
            rep.best_fitness = fitness_[order[0]];
            rep.best_lr      = lr_[order[0]];
            for (double f : fitness_) rep.sum_fitness += f; // sequential, like the log's mean (:417-418)
            for (double s : steps_) rep.sum_steps += s;
            net_->UpdateWeightsAndBiases(order);
        }

        // stagnation: the incumbent (slot 0 = last generation's best) is still the best and did not improve (:382-402)
        const double epsilon = 1e-5;
        if (rep.best_member == 0 && rep.best_fitness >= last_best_ - epsilon) ++stagnated_;
        else stagnated_ = 0;
        last_best_ = rep.best_fitness;

        net_->CreatePopulation(elitism_);
        net_->UpdateEpochs(1);
        generation_    = static_cast<int>(net_->GetEpochs());
        rep.generation = generation_;
        return rep;
    }

    template <typename T> void NNEngineTrainer<T>::train()
    {
        if (!net_)
        {
            std::cerr << "Train error: Network not initialized or loaded." << std::endl;
            return;
        }
        std::ostream& out = std::cout;
        out << "Starting C++ training for " << epochs_ << " generations..." << std::endl;
        out << "Population size: " << net_->GetPopSize() << ", Top performers: " << net_->GetTopPerformersSize()
            << ", Elitism: " << elitism_ << std::endl;
        out << "Network structure: ";
        for (size_t i = 0; i < sizes_.size(); ++i) out << (i ? "-" : "") << sizes_[i];
        out << std::endl;

        try
        {
            LayoutSet layouts = draw_layouts(static_cast<unsigned int>(seed_));
            const int ltr     = layouts.left_to_right();
            out << "Layouts after balancing: " << ltr << " left-to-right, " << (layout_nb_ - ltr) << " right-to-left." << std::endl;

            const double per_layout = static_cast<double>(layout_nb_);
            const double per_member = static_cast<double>(net_->GetPopSize());
            for (int k = 0; k < epochs_; ++k)
            {
                const auto g0 = std::chrono::steady_clock::now();
                const int shown = generation_ + 1;
                bool reset      = false;
                if (stagnated_ >= reset_period_)
                {
                    out << "--- Stagnation detected. Resetting training layouts after " << stagnated_ << " generations. --- "
                        << std::endl;
                    // a fresh seed base per reset (:344-345)
                    layouts    = draw_layouts(static_cast<unsigned int>(seed_ + generation_ + k));
                    stagnated_ = 0;
                    last_best_ = std::numeric_limits<double>::max();
                    reset      = true;
                }
                out << "\n--- Generation " << shown << " ---" << std::endl;

                GenerationReport rep = run_generation(layouts);
                rep.layouts_reset    = reset;

                const auto g1       = std::chrono::steady_clock::now();
                rep.seconds         = std::chrono::duration<double>(g1 - g0).count();
                const long long gs  = std::chrono::duration_cast<std::chrono::seconds>(g1 - g0).count();
                const long long tot = std::chrono::duration_cast<std::chrono::seconds>(g1 - t0_).count();
                const bool any      = layout_nb_ > 0;
                out << "Generation " << generation_ << " complete. (Took: " << mmss(gs) << ", Total: " << mmss(tot) << ")" << std::endl;
                out << std::fixed << std::setprecision(4);
                out << "  Best Avg Fitness: " << rep.best_fitness / per_layout << " (Member " << rep.best_member << ")" << std::endl;
                out << "  Min Number of Landing in Both Directions: " << landings_both_directions(rep.best_lr) << std::endl;
                out << "  Avg Fitness:      " << (any ? rep.sum_fitness / per_member / per_layout : 0.0) << std::endl;
                out << std::fixed << std::setprecision(1);
                out << "  Avg Steps/Layout: " << (any ? rep.sum_steps / per_member / per_layout : 0.0) << std::endl;
                history_.push_back(rep);

                if (save_enabled_ && (generation_ % save_interval_ == 0 || k == epochs_ - 1)) save();
            }
        } catch (const std::exception& e)
        {
            std::cerr << "Train error: " << e.what() << std::endl;
            return;
        }
        out << "\nC++ Training finished." << std::endl;
    }

    template class NNEngineTrainer<float>;
    template class NNEngineTrainer<double>;
} // namespace Training
