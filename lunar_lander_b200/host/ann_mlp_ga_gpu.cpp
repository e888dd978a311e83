#include "ann_mlp_ga_gpu.h"

#include <cstdio>
#include <exception>
#include <stdexcept>
#include <thread>
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
        for (ll_ctx* c : ctxs_)
            if (c) ll_destroy(c);
        ctxs_.clear(), first_.clear();
        ctx_ = nullptr;
    }

    template <typename T> void ANN_MLP_GA<T>::raise_on(ll_ctx* ctx, const char* what) const
    {
        const char* why = ll_last_error(ctx);
        throw std::runtime_error(std::string(what) + ": " + (why && *why ? why : "unknown error"));
    }
    template <typename T> void ANN_MLP_GA<T>::raise(const char* what) const { raise_on(ctx_, what); }

    template <typename T> void ANN_MLP_GA<T>::require() const
    {
        if (!ctx_) throw std::runtime_error("ANN_MLP_GA: the network has no shape yet (construct with sizes or Deserialize first)");
    }

    // f(rank) for every shard.  One GPU: inline.  Several: one host thread per GPU for the duration of the call — the collectives
    // inside the library (NCCL on each context's stream) need every rank to be in the same call at the same time.
    template <typename T> template <class F> void ANN_MLP_GA<T>::each(F f) const
    {
        const size_t n = ctxs_.size();
        if (n <= 1)
        {
            f(size_t(0));
            return;
        }
        std::vector<std::exception_ptr> err(n);
        std::vector<std::thread> th;
        th.reserve(n);
        for (size_t r = 0; r < n; ++r)
            th.emplace_back([&, r] {
                try
                {
                    f(r);
                } catch (...)
                {
                    err[r] = std::current_exception();
                }
            });
        for (auto& t : th) t.join();
        for (auto& e : err)
            if (e) std::rethrow_exception(e);
    }

    template <typename T> size_t ANN_MLP_GA<T>::owner_of(size_t member) const
    {
        for (size_t r = 0; r + 1 < first_.size(); ++r)
            if (member >= first_[r] && member < first_[r + 1]) return r;
        throw std::runtime_error("ANN_MLP_GA: member index out of range");
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
        const int gpus = Config::getIntOr("GpuCfg.gpus", 1);
        if (gpus < 1 || gpus > 64) throw std::runtime_error("GpuCfg.gpus must be in 1..64");
        const int base    = Config::getIntOr("GpuCfg.device", 0);
        ll_env_cfg env    = Config::envCfg();
        ctxs_.assign(static_cast<size_t>(gpus), nullptr);
        for (int r = 0; r < gpus; ++r)
        {
            ll_net_cfg net{};
            net.layer_sizes   = s.data();
            net.n_layers      = static_cast<int32_t>(s.size());
            net.activation_id = static_cast<int32_t>(act);
            net.population    = pop;
            net.top           = top;
            net.seed          = seed;
            net.precision     = is_float ? 32 : 64;
            net.mode          = mode == "fast" ? LL_MODE_FAST : LL_MODE_EXACT;
            net.rank          = r;
            net.world         = gpus;
            net.device        = base + r;
            net.sigma_init    = Config::getDoubleOr("GpuCfg.sigma_init", 0.0);
            net.sigma_mut     = Config::getDoubleOr("GpuCfg.sigma_mut", 0.0);
            if (ll_create(&ctxs_[static_cast<size_t>(r)], &env, &net) != LL_OK)
            {
                const char* why       = ll_last_error(nullptr);
                const std::string msg = std::string("ANN_MLP_GA: ll_create failed: ") + (why ? why : "?");
                ctxs_[static_cast<size_t>(r)] = nullptr;
                close();
                throw std::runtime_error(msg);
            }
        }
        ctx_ = ctxs_[0];
        first_.clear();
        for (ll_ctx* c : ctxs_)
        {
            ll_info info{};
            if (ll_get_info(c, &info) != LL_OK) raise_on(c, "ll_get_info");
            first_.push_back(info.first_global);
        }
        first_.push_back(pop);
        if (gpus > 1)
        {
            unsigned char id[LL_COMM_ID_BYTES];
            if (ll_comm_get_unique_id(id) != LL_OK) raise_on(nullptr, "ll_comm_get_unique_id");
            each([&](size_t r) {
                if (ll_comm_init(ctxs_[r], id) != LL_OK) raise_on(ctxs_[r], "ll_comm_init");
            });
        }
        if (strategy_set_)
            for (ll_ctx* c : ctxs_)
                if (ll_set_population_strategy(c, strategy_, ratio_) != LL_OK) raise_on(c, "SetPopulationStrategy");
    }

    template <typename T> void ANN_MLP_GA<T>::SetPopulationStrategy(PopulationStrategy strategy, double ratio)
    {
        strategy_ = static_cast<int>(strategy), ratio_ = ratio, strategy_set_ = true;
        for (ll_ctx* c : ctxs_)
            if (ll_set_population_strategy(c, strategy_, ratio_) != LL_OK) raise_on(c, "SetPopulationStrategy");
    }

    template <typename T> void ANN_MLP_GA<T>::CreatePopulation(bool elitism)
    {
        require();
        each([&](size_t r) {
            if (ll_create_population(ctxs_[r], elitism ? 1 : 0) != LL_OK) raise_on(ctxs_[r], "CreatePopulation");
        });
    }

    template <typename T> void ANN_MLP_GA<T>::UpdateWeightsAndBiases(const std::vector<size_t>& sorted_indices)
    {
        require();
        if (sorted_indices.size() < top_) throw std::runtime_error("UpdateWeightsAndBiases: fewer indices than top performers");
        // only the first `top` entries matter (the parents); the rest of the order is never looked at
        std::vector<uint64_t> head(sorted_indices.begin(), sorted_indices.begin() + static_cast<long>(top_));
        each([&](size_t r) {
            if (ll_select(ctxs_[r], head.data(), head.size(), nullptr) != LL_OK) raise_on(ctxs_[r], "UpdateWeightsAndBiases");
        });
    }

    template <typename T> std::vector<size_t> ANN_MLP_GA<T>::SelectOnDevice()
    {
        require();
        std::vector<uint64_t> idx(top_);
        each([&](size_t r) {
            if (ll_select(ctxs_[r], nullptr, 0, r == 0 ? idx.data() : nullptr) != LL_OK) raise_on(ctxs_[r], "SelectOnDevice");
        });
        return std::vector<size_t>(idx.begin(), idx.end());
    }

    template <typename T> void ANN_MLP_GA<T>::UpdateEpochs(size_t n)
    {
        require();
        for (ll_ctx* c : ctxs_)
            if (ll_update_epochs(c, n) != LL_OK) raise_on(c, "UpdateEpochs");
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
        if (ll_serialize(ctx_, path.c_str(), name_.c_str(), 6) != LL_OK) raise("Serialize"); // every rank holds the parent set
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
        each([&](size_t r) { // every rank reads the parent set and rebuilds its own slots (Philox keyed by global member)
            if (ll_deserialize(ctxs_[r], path.c_str()) != LL_OK) raise_on(ctxs_[r], "Deserialize");
        });
        if (*stored_name) name_ = stored_name;
    }

    template <typename T> size_t ANN_MLP_GA<T>::feedforward(const T* inputs, size_t n_inputs, size_t member)
    {
        require();
        if (!inputs || n_inputs < sizes_.front()) throw std::runtime_error("feedforward: input buffer shorter than the input layer");
        uint64_t action = 0;
        ll_ctx* c       = ctxs_[owner_of(member)];
        if (ll_feedforward_index(c, inputs, member, &action) != LL_OK) raise_on(c, "feedforward");
        return static_cast<size_t>(action);
    }

    template <typename T>
    void ANN_MLP_GA<T>::SetLayouts(const std::vector<int32_t>& spad_x1, const std::vector<int32_t>& lpad_x1,
                                   const std::vector<uint64_t>& directions)
    {
        require();
        if (spad_x1.size() != lpad_x1.size() || spad_x1.size() != directions.size())
            throw std::runtime_error("SetLayouts: the three layout arrays differ in length");
        for (ll_ctx* c : ctxs_)
            if (ll_set_layouts(c, spad_x1.data(), lpad_x1.data(), directions.data(), static_cast<int32_t>(spad_x1.size())) != LL_OK)
                raise_on(c, "SetLayouts");
    }

    template <typename T>
    void ANN_MLP_GA<T>::EvaluateGeneration(const std::vector<int32_t>& spad_x1, const std::vector<int32_t>& lpad_x1,
                                           const std::vector<uint64_t>& directions, int max_steps, double* fitness,
                                           double* last_gen_lr, double* steps, ll_gen_stats* stats)
    {
        require();
        if (spad_x1.size() != lpad_x1.size() || spad_x1.size() != directions.size())
            throw std::runtime_error("EvaluateGeneration: the three layout arrays differ in length");
        std::vector<ll_gen_stats> st(ctxs_.size());
        each([&](size_t r) { // every shard writes its own slice of the caller's population-sized vectors
            const uint64_t f = first_[r];
            if (ll_evaluate(ctxs_[r], spad_x1.data(), lpad_x1.data(), directions.data(), static_cast<int32_t>(spad_x1.size()), max_steps,
                            fitness ? fitness + f : nullptr, last_gen_lr ? last_gen_lr + f : nullptr, steps ? steps + f : nullptr,
                            &st[r]) != LL_OK)
                raise_on(ctxs_[r], "EvaluateGeneration");
        });
        if (stats)
        { // whole-population statistics: best by (fitness, index), sums in rank order
            *stats = st[0];
            for (size_t r = 1; r < st.size(); ++r)
            {
                const bool better = st[r].best_fitness < stats->best_fitness ||
                                    (st[r].best_fitness == stats->best_fitness && st[r].best_index < stats->best_index);
                if (better) stats->best_fitness = st[r].best_fitness, stats->best_index = st[r].best_index, stats->best_lr = st[r].best_lr;
                stats->sum_fitness += st[r].sum_fitness, stats->sum_steps += st[r].sum_steps, stats->members += st[r].members;
            }
        }
    }

    template <typename T> void ANN_MLP_GA<T>::StepGeneration(int max_steps, bool elitism, ll_gen_stats* stats)
    {
        require();
        each([&](size_t r) { // with several GPUs the statistics a rank reads back already describe the whole population
            if (ll_step_generation(ctxs_[r], max_steps, elitism ? 1 : 0, r == 0 ? stats : nullptr) != LL_OK) raise_on(ctxs_[r], "StepGeneration");
        });
    }

    template <typename T> void ANN_MLP_GA<T>::GetMember(size_t member, T* params) const
    {
        require();
        ll_ctx* c = ctxs_[owner_of(member)];
        if (ll_get_member(c, member, params) != LL_OK) raise_on(c, "GetMember");
    }
    template <typename T> void ANN_MLP_GA<T>::SetMember(size_t member, const T* params)
    {
        require();
        ll_ctx* c = ctxs_[owner_of(member)];
        if (ll_set_member(c, member, params) != LL_OK) raise_on(c, "SetMember");
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
        for (ll_ctx* c : ctxs_)
        {
            double t = 0.0;
            if (ll_step_counter(c, &t, nullptr, reset ? 1 : 0) != LL_OK) raise_on(c, "TotalLanderSteps");
            total += t;
        }
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
