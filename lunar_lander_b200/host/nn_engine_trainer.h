// nn_engine_trainer.h — Training::NNEngineTrainer<T> with the public interface of the reference's trainer
// (src/nn_engine_trainer.h:31-40: NNEngineTrainer(); bool init(); bool load(step = "last"); bool save(); void train();)
// driving the B200 library instead of a CPU thread pool.  Console output, file names, the stagnation / layout-reset
// schedule and the checkpoint files follow the reference so that a run can be diffed against `main_train_d`
// (tests/test_gpu_host.py does exactly that).  Error convention as there: bool + std::cerr.
#ifndef LL_HOST_NN_ENGINE_TRAINER_H_
#define LL_HOST_NN_ENGINE_TRAINER_H_

#include <chrono>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "ann_mlp_ga_gpu.h"

namespace Training
{
    // One training layout = (start pad, landing pad, flight direction); struct-of-arrays because that is what the
    // kernel consumes (LayoutInfo + layout_directions of src/nn_engine_trainer.h:23-29 / :283-309).
    struct LayoutSet
    {
        std::vector<int32_t> spad_x1, lpad_x1;
        std::vector<double> spad_center, lpad_center;
        std::vector<uint64_t> direction; // 0 left-to-right, 1 right-to-left
        size_t size() const { return spad_x1.size(); }
        int left_to_right() const;
    };

    // What one generation reports (the numbers of the log block, src/nn_engine_trainer.cpp:413-435)
    struct GenerationReport
    {
        int generation        = 0;
        size_t best_member    = 0;
        double best_fitness   = 0.0; // sum over layouts
        double best_lr        = 0.0;
        double sum_fitness    = 0.0;
        double sum_steps      = 0.0;
        double seconds        = 0.0;
        bool layouts_reset    = false;
    };

    template <typename T> class NNEngineTrainer
    {
      public:
        NNEngineTrainer();
        ~NNEngineTrainer() = default;

        bool init();
        bool load(const std::string& step = "last");
        bool save();
        void train();

        // additions for embedding and tests
        const std::vector<GenerationReport>& history() const { return history_; }
        nn::ANN_MLP_GA<T>* net() { return net_.get(); }

      private:
        static std::string mmss(long long total_seconds);
        LayoutSet draw_layouts(unsigned int seed_base);
        GenerationReport run_generation(const LayoutSet& layouts);
        int landings_both_directions(double lr_bonus) const;
        bool select_on_device() const;

        std::unique_ptr<nn::ANN_MLP_GA<T>> net_;
        int generation_ = 0; // generations completed (== net epochs)
        std::chrono::steady_clock::time_point t0_;
        std::vector<GenerationReport> history_;

        // NNConfig.*
        std::string name_, save_dir_;
        bool save_enabled_ = false, overwrite_ = false, elitism_ = true;
        int save_interval_ = 1, epochs_ = 0, seed_ = 0, population_ = 0, top_ = 0, activation_ = 1;
        double injection_ratio_ = 0.0;
        std::vector<size_t> sizes_;
        // NNTrainingCfg.* / GameCfg.* / DisplayCfg.*
        int layout_nb_ = 0, reset_period_ = 0, max_steps_ = 0;
        double ltr_ratio_ = 0.5, width_ = 0.0, height_ = 0.0, spad_width_ = 0.0, lpad_width_ = 0.0;
        int lr_bonus_unit_ = 1; // an int in the reference too (src/nn_engine_trainer.h:98): only the printed landing count uses it
        // stagnation tracking (src/nn_engine_trainer.cpp:382-402)
        int stagnated_       = 0;
        double last_best_    = 0.0;
        // GpuCfg.select: host (std::sort like the reference, identical tie order) | device (radix top-k) | auto
        std::string select_mode_;
        // layout_directions outlives a layout reset: balance_layouts leaves it untouched when nothing needs flipping
        // (src/nn_engine_trainer.cpp:214 vs :255-259), so a reset can run with the previous set's directions
        std::vector<uint64_t> directions_;
        // host-side result vectors, reused across generations
        std::vector<double> fitness_, lr_, steps_;
    };
} // namespace Training

#endif
