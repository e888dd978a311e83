#include "ann_mlp_ga_gpu.h"

#include <cstdio>
#include <stdexcept>
#include <type_traits>

#include "config_loader.h"

namespace nn
{
    template <typename T> ANN_MLP_GA<T>::ANN_MLP_GA() = default;

    template <typename T>
    ANN_MLP_GA<T>::ANN_MLP_GA(const std::vector<size_t>& sizes, int seed, size_t pop, size_t top, size_t act)
    {
        open(sizes, static_cast<uint64_t>(static_cast<uint32_t>(seed)), pop, top, act);
    }

    template <typename T> ANN_MLP_GA<T>::~ANN_MLP_GA() { close(); }

    template <typename T> void ANN_MLP_GA<T>::close()
    {
        if (ctx_) ll_destroy(ctx_);
        ctx_ = nullptr;
    }

    template <typename T> void ANN_MLP_GA<T>::raise(const char* what) const
    {
        const char* why = ll_last_error(ctx_);
        throw std::runtime_error(std::string(what) + ": " + (why && *why ? why : "unknown error"));
    }

    template <typename T> void ANN_MLP_GA<T>::require() const
    {
        if (!ctx_) throw std::runtime_error("ANN_MLP_GA: the network has no shape yet (construct with sizes or Deserialize first)");
    }

    // The environment constants travel with the context (one copy instead of the reference's `pop` GameLogicCpp
    // instances, src/nn_engine_trainer.cpp:325-330); they come from the same process-global Config.
    template <typename T>
    void ANN_MLP_GA<T>::open(const std::vector<size_t>& sizes, uint64_t seed, size_t pop, size_t top, size_t act)
    {
        close();
        sizes_ = sizes, pop_ = pop, top_ = top, act_ = act, seed_ = seed;
        std::vector<uint64_t> s(sizes.begin(), sizes.end());
        const bool is_float    = std::is_same<T, float>::value;
        const std::string mode = mode_override_ == LL_MODE_EXACT  ? "exact"
                                 : mode_override_ == LL_MODE_FAST ? "fast"
                                                                  : Config::getStringOr("GpuCfg.mode", is_float ? "fast" : "exact");
        if (mode != "exact" && mode != "fast") throw std::runtime_error("GpuCfg.mode must be exact or fast, got " + mode);
        if (mode == "fast" && !is_float) throw std::runtime_error("GpuCfg.mode = fast needs the float build (main_train_f)");
        ll_env_cfg env = Config::envCfg();
        ll_net_cfg net{};
        net.layer_sizes   = s.data();
        net.n_layers      = static_cast<int32_t>(s.size());
        net.activation_id = static_cast<int32_t>(act);
        net.population    = pop;
        net.top           = top;
        net.seed          = seed;
        net.precision     = is_float ? 32 : 64;
        net.mode          = mode == "fast" ? LL_MODE_FAST : LL_MODE_EXACT;
        net.rank          = 0;
        net.world         = 1;
        net.device        = Config::getIntOr("GpuCfg.device", 0);
        net.sigma_init    = Config::getDoubleOr("GpuCfg.sigma_init", 0.0);
        net.sigma_mut     = Config::getDoubleOr("GpuCfg.sigma_mut", 0.0);
        if (ll_create(&ctx_, &env, &net) != LL_OK)
        {
            const char* why = ll_last_error(nullptr);
            ctx_            = nullptr;
            throw std::runtime_error(std::string("ANN_MLP_GA: ll_create failed: ") + (why ? why : "?"));
        }
        if (strategy_set_ && ll_set_population_strategy(ctx_, strategy_, ratio_) != LL_OK) raise("SetPopulationStrategy");
    }

    template <typename T> void ANN_MLP_GA<T>::SetPopulationStrategy(PopulationStrategy strategy, double ratio)
    {
        strategy_ = static_cast<int>(strategy), ratio_ = ratio, strategy_set_ = true;
        if (ctx_ && ll_set_population_strategy(ctx_, strategy_, ratio_) != LL_OK) raise("SetPopulationStrategy");
    }

    template <typename T> void ANN_MLP_GA<T>::CreatePopulation(bool elitism)
    {
        require();
        if (ll_create_population(ctx_, elitism ? 1 : 0) != LL_OK) raise("CreatePopulation");
    }

    template <typename T> void ANN_MLP_GA<T>::UpdateWeightsAndBiases(const std::vector<size_t>& sorted_indices)
    {
        require();
        if (sorted_indices.size() < top_) throw std::runtime_error("UpdateWeightsAndBiases: fewer indices than top performers");
        // only the first `top` entries matter (the parents); the rest of the order is never looked at
        std::vector<uint64_t> head(sorted_indices.begin(), sorted_indices.begin() + static_cast<long>(top_));
        if (ll_select(ctx_, head.data(), head.size(), nullptr) != LL_OK) raise("UpdateWeightsAndBiases");
    }

    template <typename T> std::vector<size_t> ANN_MLP_GA<T>::SelectOnDevice()
    {
        require();
        std::vector<uint64_t> idx(top_);
        if (ll_select(ctx_, nullptr, 0, idx.data()) != LL_OK) raise("SelectOnDevice");
        return std::vector<size_t>(idx.begin(), idx.end());
    }

    template <typename T> void ANN_MLP_GA<T>::UpdateEpochs(size_t n)
    {
        require();
        if (ll_update_epochs(ctx_, n) != LL_OK) raise("UpdateEpochs");
    }

    template <typename T> size_t ANN_MLP_GA<T>::GetEpochs() const
    {
        if (!ctx_) return 0;
        uint64_t e = 0;
        if (ll_get_epochs(ctx_, &e) != LL_OK) raise("GetEpochs");
        return static_cast<size_t>(e);
    }

    template <typename T> void ANN_MLP_GA<T>::Serialize(const std::string& path)
    {
        require();
        if (ll_serialize(ctx_, path.c_str(), name_.c_str(), 6) != LL_OK) raise("Serialize");
    }

    // The file decides the shape (sizes, population, top, activation, strategy), as in the reference where load()
    // starts from an empty net and reads nn_size_config_ back from it (src/nn_engine_trainer.cpp:104, 126-128).
    template <typename T> void ANN_MLP_GA<T>::Deserialize(const std::string& path)
    {
        uint64_t sizes[16], pop = 0, top = 0, epochs = 0;
        int32_t n_sizes = 0, act = LL_ACT_TANH, strategy = LL_STRATEGY_MIXED_WITH_RANDOM_INJECTION;
        double ratio = 0.0;
        char stored_name[256];
        if (ll_peek_weight_file(path.c_str(), sizes, 16, &n_sizes, &pop, &top, &epochs, &act, &strategy, &ratio, stored_name,
                                static_cast<int32_t>(sizeof stored_name)) != LL_OK)
        {
            const char* why = ll_last_error(nullptr);
            throw std::runtime_error(std::string("Deserialize: ") + (why ? why : "cannot read ") + " (" + path + ")");
        }
        if (n_sizes < 2 || n_sizes > 16) throw std::runtime_error("Deserialize: unsupported number of layers in " + path);
        std::vector<size_t> s(sizes, sizes + n_sizes);
        strategy_ = strategy, ratio_ = ratio, strategy_set_ = true;
        open(s, seed_ ? seed_ : static_cast<uint64_t>(static_cast<uint32_t>(Config::getIntOr("NNConfig.seed", 0))), pop, top,
             static_cast<size_t>(act));
        if (ll_deserialize(ctx_, path.c_str()) != LL_OK) raise("Deserialize");
        if (*stored_name) name_ = stored_name;
    }

    template <typename T> size_t ANN_MLP_GA<T>::feedforward(const T* inputs, size_t n_inputs, size_t member)
    {
        require();
        if (!inputs || n_inputs < sizes_.front()) throw std::runtime_error("feedforward: input buffer shorter than the input layer");
        uint64_t action = 0;
        if (ll_feedforward_index(ctx_, inputs, member, &action) != LL_OK) raise("feedforward");
        return static_cast<size_t>(action);
    }

    template <typename T>
    void ANN_MLP_GA<T>::SetLayouts(const std::vector<int32_t>& spad_x1, const std::vector<int32_t>& lpad_x1,
                                   const std::vector<uint64_t>& directions)
    {
        require();
        if (spad_x1.size() != lpad_x1.size() || spad_x1.size() != directions.size())
            throw std::runtime_error("SetLayouts: the three layout arrays differ in length");
        if (ll_set_layouts(ctx_, spad_x1.data(), lpad_x1.data(), directions.data(), static_cast<int32_t>(spad_x1.size())) != LL_OK)
            raise("SetLayouts");
    }

    template <typename T>
    void ANN_MLP_GA<T>::EvaluateGeneration(const std::vector<int32_t>& spad_x1, const std::vector<int32_t>& lpad_x1,
                                           const std::vector<uint64_t>& directions, int max_steps, double* fitness,
                                           double* last_gen_lr, double* steps, ll_gen_stats* stats)
    {
        require();
        if (spad_x1.size() != lpad_x1.size() || spad_x1.size() != directions.size())
            throw std::runtime_error("EvaluateGeneration: the three layout arrays differ in length");
        if (ll_evaluate(ctx_, spad_x1.data(), lpad_x1.data(), directions.data(), static_cast<int32_t>(spad_x1.size()), max_steps,
                        fitness, last_gen_lr, steps, stats) != LL_OK)
            raise("EvaluateGeneration");
    }

    template <typename T> void ANN_MLP_GA<T>::StepGeneration(int max_steps, bool elitism, ll_gen_stats* stats)
    {
        require();
        if (ll_step_generation(ctx_, max_steps, elitism ? 1 : 0, stats) != LL_OK) raise("StepGeneration");
    }

    template <typename T> void ANN_MLP_GA<T>::GetMember(size_t member, T* params) const
    {
        require();
        if (ll_get_member(ctx_, member, params) != LL_OK) raise("GetMember");
    }
    template <typename T> void ANN_MLP_GA<T>::SetMember(size_t member, const T* params)
    {
        require();
        if (ll_set_member(ctx_, member, params) != LL_OK) raise("SetMember");
    }
    template <typename T> size_t ANN_MLP_GA<T>::ParamCount() const
    {
        require();
        ll_info info{};
        if (ll_get_info(ctx_, &info) != LL_OK) raise("ParamCount");
        return static_cast<size_t>(info.params);
    }
    template <typename T> double ANN_MLP_GA<T>::TotalLanderSteps(bool reset)
    {
        require();
        double total = 0.0;
        if (ll_step_counter(ctx_, &total, nullptr, reset ? 1 : 0) != LL_OK) raise("TotalLanderSteps");
        return total;
    }
    template <typename T> std::string ANN_MLP_GA<T>::KernelName() const
    {
        require();
        ll_info info{};
        if (ll_get_info(ctx_, &info) != LL_OK) raise("KernelName");
        return std::string(info.kernel);
    }

    template class ANN_MLP_GA<float>;
    template class ANN_MLP_GA<double>;
} // namespace nn
