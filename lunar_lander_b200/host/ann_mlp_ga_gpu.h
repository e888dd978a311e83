// ann_mlp_ga_gpu.h — nn::ANN_MLP_GA<T> as the reference's trainer uses it (ma-libs cpp/libnn ann_mlp_ga_v1.h; call
// sites src/nn_engine_trainer.cpp:87-94, 104-128, 150-172, 271, 404-407, 501-502 and mod_nn_play.py:28-56), with the
// population living in the HBM of one B200 behind the C ABI of include/ll_gpu.h.  Same method names and argument
// meaning; failures throw std::runtime_error carrying ll_last_error() (the reference wraps Serialize/Deserialize in
// try/catch(std::exception), src/nn_engine_trainer.cpp:124-137, 165-182).
//
// One addition: EvaluateGeneration() — the whole of train_generation (src/nn_engine_trainer.cpp:443-539) as one call,
// because on the GPU the episode loop, the environment and the forward pass are one fused kernel; the per-step
// feedforward(state, n, member) stays available for nn_play-style single-lander use.
#ifndef LL_HOST_ANN_MLP_GA_GPU_H_
#define LL_HOST_ANN_MLP_GA_GPU_H_

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "ll_gpu.h"

namespace nn
{
    enum class PopulationStrategy : int
    {
        ALL_FROM_ELITES             = LL_STRATEGY_ALL_FROM_ELITES,
        MIXED_WITH_RANDOM_INJECTION = LL_STRATEGY_MIXED_WITH_RANDOM_INJECTION
    };

    template <typename T> class ANN_MLP_GA
    {
      public:
        ANN_MLP_GA(); // empty net: Deserialize() gives it a shape (src/nn_engine_trainer.cpp:104)
        ANN_MLP_GA(const std::vector<size_t>& sizes, int seed, size_t pop, size_t top, size_t act);
        ~ANN_MLP_GA();
        ANN_MLP_GA(const ANN_MLP_GA&)            = delete;
        ANN_MLP_GA& operator=(const ANN_MLP_GA&) = delete;

        void SetName(const char* name) { name_ = name ? name : ""; }
        void SetPopulationStrategy(PopulationStrategy strategy, double random_injection_ratio);
        void CreatePopulation(bool elitism);
        void UpdateWeightsAndBiases(const std::vector<size_t>& sorted_indices);
        void UpdateEpochs(size_t n);
        size_t GetEpochs() const;
        std::vector<size_t> GetNetworkSize() const { return sizes_; }
        size_t GetPopSize() const { return pop_; }
        size_t GetTopPerformersSize() const { return top_; }
        void Serialize(const std::string& path);
        void Deserialize(const std::string& path);
        size_t feedforward(const T* inputs, size_t n_inputs, size_t member);

        // ---- GPU-side additions -------------------------------------------------------------------------------------
        // train_generation in one call; any of fitness / last_gen_lr / steps may be null (pop doubles each otherwise)
        void EvaluateGeneration(const std::vector<int32_t>& spad_x1, const std::vector<int32_t>& lpad_x1,
                                const std::vector<uint64_t>& directions, int max_steps, double* fitness, double* last_gen_lr,
                                double* steps, ll_gen_stats* stats);
        // std::sort + UpdateWeightsAndBiases without the population-sized host round trip: device top-k ordered by
        // (fitness, index); returns the elite indices in rank order
        std::vector<size_t> SelectOnDevice();
        // evaluate -> device top-k -> gather -> children -> epochs += 1, nothing but the statistics crosses PCIe
        void StepGeneration(int max_steps, bool elitism, ll_gen_stats* stats);
        void SetLayouts(const std::vector<int32_t>& spad_x1, const std::vector<int32_t>& lpad_x1,
                        const std::vector<uint64_t>& directions);
        void GetMember(size_t member, T* params) const;
        void SetMember(size_t member, const T* params);
        size_t ParamCount() const;
        double TotalLanderSteps(bool reset);
        std::string KernelName() const;
        ll_ctx* Context() const { return ctx_; }
        // GpuCfg.gpus = N shards the population over N GPUs of this box: one context and one host thread per GPU, the exchange between
        // them (NCCL) is owned by the contexts (ll_comm_init) — the GPU-side counterpart of the reference's per-member thread pool
        // (src/nn_engine_trainer.cpp:460-538)
        int Gpus() const { return static_cast<int>(ctxs_.size()); }
        // LL_MODE_EXACT / LL_MODE_FAST for the next open (constructor with sizes or Deserialize); default: GpuCfg.mode
        void SetMode(int mode) { mode_override_ = mode; }

      private:
        void open(const std::vector<size_t>& sizes, uint64_t seed, size_t pop, size_t top, size_t act);
        void close();
        void require() const;
        [[noreturn]] void raise(const char* what) const;

        template <class F> void each(F f) const; // f(rank) on every shard, one host thread per GPU; rethrows the first failure
        size_t owner_of(size_t member) const;
        [[noreturn]] void raise_on(ll_ctx* ctx, const char* what) const;

        ll_ctx* ctx_ = nullptr;        // rank 0's context (the only one when GpuCfg.gpus = 1)
        std::vector<ll_ctx*> ctxs_;    // every rank's context
        std::vector<uint64_t> first_;  // first GLOBAL member of every rank's shard, plus pop_ at the end
        std::vector<size_t> sizes_;
        size_t pop_ = 0, top_ = 0, act_ = LL_ACT_TANH;
        uint64_t seed_ = 0;
        std::string name_ = "net";
        int strategy_     = LL_STRATEGY_MIXED_WITH_RANDOM_INJECTION;
        double ratio_     = 0.0;
        bool strategy_set_ = false;
        int mode_override_ = -1;
    };
} // namespace nn

#endif
