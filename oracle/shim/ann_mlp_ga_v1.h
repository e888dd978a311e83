// ORACLE / TEST INFRASTRUCTURE — not part of the product.   PARITY UNPINNED at this boundary.
//
// CPU restatement of ma-libs' nn::ANN_MLP_GA<T> (cpp/libnn/src/ann_mlp_ga_v1.h in azimonti/ma-libs
// @ dd81ee426fe6b8c83dc427082c62f2a8f2578cd0).  That submodule is NOT present under /root/reference
// (.gitmodules:1-3, externals/ma-libs is empty) and the reference ships no test or golden vector for
// it, so the arithmetic below is reconstructed from:
//   * the call sites  src/nn_engine_trainer.cpp:87-94 (ctor, SetName, SetPopulationStrategy,
//     CreatePopulation), :104-128 (Deserialize, GetEpochs, GetNetworkSize), :150-172 (Serialize),
//     :271 (GetPopSize, GetTopPerformersSize), :404-407 (UpdateWeightsAndBiases, CreatePopulation,
//     UpdateEpochs, GetEpochs), :501-502 (feedforward)  and  mod_nn_play.py:28-56;
//   * the 49 weight files in sample_runs/ (file format, [out x in] row-major weights, act id,
//     populationStrategy 2, mutation noise ~ N(0, 0.1^2) measured between parent/child elites);
//   * behaviour: with  a_{l+1} = act(W_l a_l + b_l)  on every layer and first-max argmax the saved
//     elites land on the pads; with transposed weights they crash at step 9.
// Decisions where the reference cannot be recovered (recorded in DESIGN.md §Oracle):
//   dot product = ascending-k, separate multiply and add, bias added last; activation on every layer
//   including the output; argmax tie -> lowest index; tanh/sigmoid = ll_math.h (deterministic);
//   GA = Philox-keyed design of ll_philox.h instead of the sequential std::mt19937 stream.
#ifndef ORACLE_SHIM_ANN_MLP_GA_V1_H_
#define ORACLE_SHIM_ANN_MLP_GA_V1_H_

#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <random>
#include <sstream>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "ll_math.h"
#include "ll_philox.h"
#include "../ll_devlayout.h"

namespace nn
{
    enum class PopulationStrategy : int
    {
        ALL_FROM_ELITES             = 1,
        MIXED_WITH_RANDOM_INJECTION = 2 // value stored in the sample files ("populationStrategy 2")
    };

    enum : size_t
    {
        ACT_SIGMOID = 0,
        ACT_TANH    = 1 // config.txt:40
    };

    template <typename T> class ANN_MLP_GA
    {
      public:
        ANN_MLP_GA() = default;
        ANN_MLP_GA(const std::vector<size_t>& sizes, int seed, size_t pop, size_t top, size_t act)
            : sizes_(sizes), seed_(static_cast<uint32_t>(seed)), pop_(pop), top_(top), act_(act)
        {
            layout();
            params_.assign(pop_ * P_, T(0));
            fwd_.assign(pop_ * P_, T(0));
        }

        void SetName(const char* name) { name_ = name; }
        void SetPopulationStrategy(PopulationStrategy s, double ratio)
        {
            strategy_ = s;
            ratio_    = ratio;
        }
        size_t GetEpochs() const { return epochs_; }
        void UpdateEpochs(size_t n) { epochs_ += n; }
        std::vector<size_t> GetNetworkSize() const { return sizes_; }
        size_t GetPopSize() const { return pop_; }
        size_t GetTopPerformersSize() const { return top_; }

        // ---- GA -------------------------------------------------------------------------------------
        // First call (no elites yet): every member ~ N(0, sigma_init^2), generation key = epochs.
        // Later calls: slots [first, pop) are regenerated from the elites in slots [0, top) with
        // generation key = epochs + 1 (the population being built is the one evaluated next).
        void CreatePopulation(bool elitism)
        {
            if (!has_elites_) { regenerate(static_cast<uint32_t>(epochs_), 0, /*from_elites=*/false); }
            else { regenerate(static_cast<uint32_t>(epochs_ + 1), elitism ? top_ : 0, true); }
        }

        // sorted_indices: population indices by ascending fitness (src/nn_engine_trainer.cpp:376-379, 404).
        void UpdateWeightsAndBiases(const std::vector<size_t>& sorted_indices)
        {
            std::vector<T> best(top_ * P_);
            for (size_t r = 0; r < top_; ++r)
                std::memcpy(&best[r * P_], &params_[sorted_indices[r] * P_], P_ * sizeof(T));
            std::memcpy(params_.data(), best.data(), best.size() * sizeof(T));
            for (size_t r = 0; r < top_; ++r) rebuild_forward(r);
            has_elites_ = true;
        }

        // ---- policy ----------------------------------------------------------------------------------
        // returns argmax_j act(W a + b)_j of the last layer, first maximum wins.
        size_t feedforward(const T* inputs, size_t n_inputs, size_t member)
        {
            (void)n_inputs;
            ++calls_[member];
            T bufA[kMaxWidth], bufB[kMaxWidth];
            T* cur = bufA;
            T* nxt = bufB;
            for (size_t i = 0; i < sizes_[0]; ++i) cur[i] = inputs[i];
            const T* p = &fwd_[member * P_];
            for (size_t l = 0; l + 1 < sizes_.size(); ++l)
            {
                const size_t in = sizes_[l], out = sizes_[l + 1];
                const T* wt = p;            // [in][out]  (k-major copy: same per-neuron k order, vectorisable over o)
                const T* b  = p + in * out; // [out]
                for (size_t o = 0; o < out; ++o) nxt[o] = T(0);
                for (size_t k = 0; k < in; ++k)
                {
                    const T a   = cur[k];
                    const T* wk = wt + k * out;
                    for (size_t o = 0; o < out; ++o) nxt[o] = nxt[o] + wk[o] * a;
                }
                for (size_t o = 0; o < out; ++o) nxt[o] = activate(nxt[o] + b[o]);
                p += in * out + out;
                T* t = cur;
                cur  = nxt;
                nxt  = t;
            }
            const size_t n_out = sizes_.back();
            size_t best        = 0;
            for (size_t j = 1; j < n_out; ++j)
                if (cur[j] > cur[best]) best = j;
            return best;
        }

        // ---- file format --------------------------------------------------------
        void Serialize(const std::string& path, int precision = 6) const
        {
            std::ofstream f(path);
            if (!f) throw std::runtime_error("Serialize: cannot open " + path);
            f << std::setprecision(precision);
            f << "sName " << name_ << "\nversion 1\nnEpochs " << epochs_ << "\nnPopSize " << pop_ << "\nnTop " << top_
              << "\nact " << act_ << "\npopulationStrategy " << static_cast<int>(strategy_)
              << "\nrandomInjectionRatio " << ratio_ << "\nnLayers " << sizes_.size() << "\nvSize ";
            for (size_t s : sizes_) f << s << " ";
            f << "\nvBiases_start\n";
            for (size_t m = 0; m < top_; ++m)
                for (size_t l = 0; l + 1 < sizes_.size(); ++l)
                {
                    f << "pop " << m << " layer " << l << "\ndims " << sizes_[l + 1] << " 1\n";
                    const T* b = &params_[m * P_ + off_[l] + sizes_[l] * sizes_[l + 1]];
                    for (size_t o = 0; o < sizes_[l + 1]; ++o) f << b[o] << " \n";
                }
            f << "vBiases_end\nvWeights_start\n";
            for (size_t m = 0; m < top_; ++m)
                for (size_t l = 0; l + 1 < sizes_.size(); ++l)
                {
                    f << "pop " << m << " layer " << l << "\ndims " << sizes_[l + 1] << " " << sizes_[l] << "\n";
                    const T* w = &params_[m * P_ + off_[l]];
                    for (size_t o = 0; o < sizes_[l + 1]; ++o)
                    {
                        for (size_t k = 0; k < sizes_[l]; ++k) f << w[o * sizes_[l] + k] << " ";
                        f << "\n";
                    }
                }
            f << "vWeights_end\ngenerator";
            // 624 state words, the libc++ stream form found in every sample file; the GA itself is
            // Philox-keyed, the words only keep the file shape.
            std::mt19937 g(seed_);
            std::ostringstream os;
            os << g;
            std::istringstream is(os.str());
            for (int i = 0; i < 624; ++i)
            {
                unsigned long w;
                is >> w;
                f << " " << w;
            }
            f << "\n";
        }

        void Deserialize(const std::string& path)
        {
            std::ifstream f(path);
            if (!f) throw std::runtime_error("Deserialize: cannot open " + path);
            std::string tok;
            size_t n_layers = 0;
            int strat       = 2;
            sizes_.clear();
            while (f >> tok)
            {
                if (tok == "sName") f >> name_;
                else if (tok == "version") { int v; f >> v; }
                else if (tok == "nEpochs") f >> epochs_;
                else if (tok == "nPopSize") f >> pop_;
                else if (tok == "nTop") f >> top_;
                else if (tok == "act") f >> act_;
                else if (tok == "populationStrategy") f >> strat;
                else if (tok == "randomInjectionRatio") f >> ratio_;
                else if (tok == "nLayers") f >> n_layers;
                else if (tok == "vSize")
                {
                    for (size_t i = 0; i < n_layers; ++i) { size_t s; f >> s; sizes_.push_back(s); }
                    break;
                }
                else throw std::runtime_error("Deserialize: unexpected token " + tok);
            }
            if (sizes_.size() < 2 || top_ == 0 || pop_ < top_) throw std::runtime_error("Deserialize: bad header");
            strategy_ = static_cast<PopulationStrategy>(strat);
            layout();
            params_.assign(pop_ * P_, T(0));
            fwd_.assign(pop_ * P_, T(0));
            read_section(f, "vBiases_start", "vBiases_end", true);
            read_section(f, "vWeights_start", "vWeights_end", false);
            // "generator": 624 (libc++) or 625 (libstdc++) integers.  A line that starts like a FRESHLY SEEDED mt19937 (word 1 follows
            // from word 0 by the seeding recurrence — what this stand-in and libll_gpu write) carries the Philox key of the run that
            // wrote the file in word 0, and the resumed run continues with it, the way the real Deserialize restores its mt19937.  Any
            // other line (a file written by the real ma-libs: an arbitrary engine state) leaves the configured seed alone.
            {
                std::string gtok;
                unsigned long w0 = 0, w1 = 0;
                if ((f >> gtok) && gtok == "generator" && (f >> w0) && (f >> w1))
                {
                    const uint32_t s0 = static_cast<uint32_t>(w0);
                    if (static_cast<uint32_t>(w1) == 1812433253u * (s0 ^ (s0 >> 30)) + 1u) seed_ = s0;
                }
            }
            for (size_t r = 0; r < top_; ++r) rebuild_forward(r);
            has_elites_ = true;
            // The file holds only the elites; the rest of the population is rebuilt exactly as the
            // CreatePopulation that preceded the save did (generation key = stored epoch count).
            regenerate(static_cast<uint32_t>(epochs_), top_, true);
        }

        // ---- extras used by the test / baseline harness only -----------------------------------------
        size_t ParamCount() const { return P_; }
        // canonical order: per layer, W[out][in] row-major then b[out]
        const T* MemberParams(size_t m) const { return &params_[m * P_]; }
        void SetMemberParams(size_t m, const T* p)
        {
            std::memcpy(&params_[m * P_], p, P_ * sizeof(T));
            rebuild_forward(m);
        }
        void MarkElites() { has_elites_ = true; }
        uint64_t FeedforwardCalls() const
        {
            uint64_t s = 0;
            for (uint64_t c : calls_) s += c;
            return s;
        }
        void ResetCounters() { calls_.assign(pop_, 0); }
        double sigma_init = 1.0;
        double sigma_mut  = 0.1;

      private:
        static constexpr size_t kMaxWidth = 1024;

        void layout()
        {
            off_.clear();
            P_ = 0;
            for (size_t l = 0; l + 1 < sizes_.size(); ++l)
            {
                if (sizes_[l] > kMaxWidth || sizes_[l + 1] > kMaxWidth) throw std::runtime_error("layer too wide");
                off_.push_back(P_);
                P_ += sizes_[l] * sizes_[l + 1] + sizes_[l + 1];
            }
            calls_.assign(pop_, 0);
        }

        T activate(T z) const
        {
            const double zd = static_cast<double>(z);
            return static_cast<T>(act_ == ACT_TANH ? ll_tanh(zd) : ll_sigmoid(zd));
        }

        void rebuild_forward(size_t m)
        {
            for (size_t l = 0; l + 1 < sizes_.size(); ++l)
            {
                const size_t in = sizes_[l], out = sizes_[l + 1];
                const T* w = &params_[m * P_ + off_[l]];
                T* wt      = &fwd_[m * P_ + off_[l]];
                for (size_t o = 0; o < out; ++o)
                    for (size_t k = 0; k < in; ++k) wt[k * out + o] = w[o * in + k];
                std::memcpy(wt + in * out, w + in * out, out * sizeof(T));
            }
        }

        // slots [first, pop): n_inject of them, spread evenly (ll_is_injected), are fresh N(0, sigma_init^2) individuals, the
        // others are  elite[pick] + sigma_mut * N(0,1)  with pick = ll_parent_pick(seed, gen, slot, top).
        void regenerate(uint32_t gen, size_t first, bool from_elites)
        {
            const size_t n_children = pop_ - first;
            size_t n_inject         = n_children;
            if (from_elites)
            {
                n_inject = 0;
                if (strategy_ == PopulationStrategy::MIXED_WITH_RANDOM_INJECTION)
                    n_inject = static_cast<size_t>(std::floor(ratio_ * static_cast<double>(n_children) + 0.5));
                if (n_inject > n_children) n_inject = n_children;
            }
            std::vector<T> elites;
            if (from_elites) elites.assign(params_.begin(), params_.begin() + static_cast<long>(top_ * P_));
            const T s_init = static_cast<T>(sigma_init), s_mut = static_cast<T>(sigma_mut);
            std::vector<uint64_t> sz64(sizes_.begin(), sizes_.end());
            const std::vector<int> d2c = orc_dev2canon(sz64.data(), static_cast<int>(sz64.size()));
            for (size_t i = first; i < pop_; ++i)
            {
                const bool inject = ll_is_injected(i - first, n_children, n_inject);
                const T* parent   = nullptr;
                if (!inject)
                    parent = &elites[ll_parent_pick(seed_, gen, static_cast<uint32_t>(i), static_cast<uint32_t>(top_)) * P_];
                T* dst = &params_[i * P_];
                if (std::is_same<T, float>::value)
                { // float populations: four Box-Muller variates per DEVICE-order quad (ll_normal4_f32), padding positions discarded
                    for (size_t qd = 0; qd * 4 < d2c.size(); ++qd)
                    {
                        float g[4];
                        ll_normal4_f32(seed_, gen, static_cast<uint32_t>(i), static_cast<uint32_t>(qd), g);
                        for (int e = 0; e < 4; ++e)
                        {
                            const int c = d2c[qd * 4 + e];
                            if (c >= 0) dst[c] = inject ? s_init * static_cast<T>(g[e]) : parent[c] + s_mut * static_cast<T>(g[e]);
                        }
                    }
                    rebuild_forward(i);
                    continue;
                }
                for (size_t j = 0; j < P_; j += 2)
                {
                    double n0, n1;
                    ll_normal_pair(seed_, gen, static_cast<uint32_t>(i), static_cast<uint32_t>(j / 2), &n0, &n1);
                    const T g0 = static_cast<T>(n0), g1 = static_cast<T>(n1);
                    dst[j] = inject ? s_init * g0 : parent[j] + s_mut * g0;
                    if (j + 1 < P_) dst[j + 1] = inject ? s_init * g1 : parent[j + 1] + s_mut * g1;
                }
                rebuild_forward(i);
            }
        }

        void read_section(std::ifstream& f, const char* start, const char* end, bool biases)
        {
            std::string tok;
            f >> tok;
            if (tok != start) throw std::runtime_error(std::string("Deserialize: expected ") + start);
            for (;;)
            {
                f >> tok;
                if (tok == end) return;
                if (tok != "pop") throw std::runtime_error("Deserialize: expected pop, got " + tok);
                size_t m, l, rows, cols;
                f >> m >> tok >> l >> tok >> rows >> cols;
                if (m >= top_ || l + 1 >= sizes_.size() || rows != sizes_[l + 1] || cols != (biases ? 1 : sizes_[l]))
                    throw std::runtime_error("Deserialize: dims do not match vSize");
                T* dst = &params_[m * P_ + off_[l] + (biases ? sizes_[l] * sizes_[l + 1] : 0)];
                for (size_t i = 0; i < rows * cols; ++i)
                {
                    double v;
                    if (!(f >> v)) throw std::runtime_error("Deserialize: truncated matrix");
                    dst[i] = static_cast<T>(v);
                }
            }
        }

        std::vector<size_t> sizes_;
        uint32_t seed_ = 0;
        size_t pop_ = 0, top_ = 0, act_ = ACT_TANH, epochs_ = 0, P_ = 0;
        std::string name_        = "net";
        PopulationStrategy strategy_ = PopulationStrategy::MIXED_WITH_RANDOM_INJECTION;
        double ratio_            = 0.0;
        bool has_elites_         = false;
        std::vector<size_t> off_;
        std::vector<T> params_; // canonical
        std::vector<T> fwd_;    // k-major copy used by feedforward
        std::vector<uint64_t> calls_;
    };
} // namespace nn

#endif
