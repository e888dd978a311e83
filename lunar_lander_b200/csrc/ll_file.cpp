// ll_file.cpp — the weight checkpoint text format of nn::ANN_MLP_GA<T>::Serialize / Deserialize
// (call sites src/nn_engine_trainer.cpp:126, 172; consumer mod_nn_play.py:33).  The ma-libs writer is not in the
// reference tree; the format below is the one found in all 49 files under sample_runs/ (SURVEY.md App. B):
//   header lines, "vBiases_start" ... "vBiases_end", "vWeights_start" ... "vWeights_end", "generator <624 words>".
// Only the top `nTop` members are stored; matrices are [rows x cols] row-major, one row per line, each value
// followed by one space, default ostream formatting (6 significant digits).
#include <fstream>
#include <iomanip>
#include <random>
#include <sstream>

#include "ll_internal.h"

int ll_file_write(ll_ctx* ctx, const char* path, const char* name, int digits, const std::vector<double>& elites_canon)
{
    const std::vector<uint64_t>& sz = ctx->sizes;
    const size_t pc                 = (size_t)ctx->topo.p_canon;
    std::ofstream f(path);
    if (!f) return ll_fail(ctx, LL_EIO, std::string("cannot open for writing: ") + path);
    f << std::setprecision(digits);
    f << "sName " << name << "\n"
      << "version 1\n"
      << "nEpochs " << ctx->epochs << "\n"
      << "nPopSize " << ctx->pop_global << "\n"
      << "nTop " << ctx->top << "\n"
      << "act " << ctx->topo.activation << "\n"
      << "populationStrategy " << ctx->strategy << "\n"
      << "randomInjectionRatio " << ctx->ratio << "\n"
      << "nLayers " << sz.size() << "\n"
      << "vSize ";
    for (uint64_t s : sz) f << s << " ";
    f << "\n";
    // canonical member: per layer W[out][in] then b[out]
    std::vector<size_t> w_at(sz.size() - 1), b_at(sz.size() - 1);
    size_t at = 0;
    for (size_t l = 0; l + 1 < sz.size(); ++l)
    {
        w_at[l] = at;
        b_at[l] = at + sz[l] * sz[l + 1];
        at      = b_at[l] + sz[l + 1];
    }
    f << "vBiases_start\n";
    for (uint64_t m = 0; m < ctx->top; ++m)
        for (size_t l = 0; l + 1 < sz.size(); ++l)
        {
            f << "pop " << m << " layer " << l << "\n"
              << "dims " << sz[l + 1] << " 1\n";
            for (uint64_t o = 0; o < sz[l + 1]; ++o) f << elites_canon[m * pc + b_at[l] + o] << " \n";
        }
    f << "vBiases_end\nvWeights_start\n";
    for (uint64_t m = 0; m < ctx->top; ++m)
        for (size_t l = 0; l + 1 < sz.size(); ++l)
        {
            f << "pop " << m << " layer " << l << "\n"
              << "dims " << sz[l + 1] << " " << sz[l] << "\n";
            for (uint64_t o = 0; o < sz[l + 1]; ++o)
            {
                for (uint64_t k = 0; k < sz[l]; ++k) f << elites_canon[m * pc + w_at[l] + o * sz[l] + k] << " ";
                f << "\n";
            }
        }
    f << "vWeights_end\n";
    // The GA here is Philox-keyed, so there is no sequential generator state to save; the line keeps the shape of the
    // sample files (624 words = the stream form of a std::mt19937 in libc++) with the state of a freshly seeded engine,
    // whose first word is the seed itself: ll_file_read takes the Philox key back from it, so a resumed run continues
    // with the generator of the run that wrote the file (as Deserialize restores the mt19937 in the reference).
    f << "generator";
    {
        std::mt19937 eng(static_cast<uint32_t>(ctx->seed));
        std::ostringstream os;
        os << eng;
        std::istringstream is(os.str());
        for (int i = 0; i < 624; ++i)
        {
            unsigned long w = 0;
            is >> w;
            f << " " << w;
        }
    }
    f << "\n";
    f.flush();
    if (!f) return ll_fail(ctx, LL_EIO, std::string("write failed: ") + path);
    return LL_OK;
}

int ll_file_read(const char* path, std::vector<uint64_t>* sizes, uint64_t* pop, uint64_t* top, uint64_t* epochs, int* act, int* strategy,
                 double* ratio, std::string* name, std::vector<double>* elites_canon, std::string* err, uint64_t* generator_word0)
{
    auto bad = [&](const std::string& m) {
        *err = std::string(path) + ": " + m;
        return LL_EIO;
    };
    std::ifstream f(path);
    if (!f) return bad("cannot open");
    std::string tok;
    uint64_t n_layers = 0;
    sizes->clear();
    bool have_sizes = false;
    while (!have_sizes && (f >> tok))
    {
        if (tok == "sName") f >> *name;
        else if (tok == "version")
        {
            int v;
            f >> v;
            if (v != 1) return bad("unsupported version");
        }
        else if (tok == "nEpochs") f >> *epochs;
        else if (tok == "nPopSize") f >> *pop;
        else if (tok == "nTop") f >> *top;
        else if (tok == "act") f >> *act;
        else if (tok == "populationStrategy") f >> *strategy;
        else if (tok == "randomInjectionRatio") f >> *ratio;
        else if (tok == "nLayers") f >> n_layers;
        else if (tok == "vSize")
        {
            if (n_layers < 2 || n_layers > LL_MAX_LAYERS + 1) return bad("bad nLayers");
            for (uint64_t i = 0; i < n_layers; ++i)
            {
                uint64_t s = 0;
                if (!(f >> s) || s == 0 || s > LL_MAX_WIDTH) return bad("bad vSize");
                sizes->push_back(s);
            }
            have_sizes = true;
        }
        else return bad("unexpected header token '" + tok + "'");
        if (!f) return bad("malformed header");
    }
    if (!have_sizes || *top == 0 || *pop < *top) return bad("incomplete header");
    if (!elites_canon) return LL_OK;

    const std::vector<uint64_t>& sz = *sizes;
    std::vector<size_t> w_at(sz.size() - 1), b_at(sz.size() - 1);
    size_t pc = 0;
    for (size_t l = 0; l + 1 < sz.size(); ++l)
    {
        w_at[l] = pc;
        b_at[l] = pc + sz[l] * sz[l + 1];
        pc      = b_at[l] + sz[l + 1];
    }
    elites_canon->assign(*top * pc, 0.0);
    std::vector<char> seen(*top * (sz.size() - 1) * 2, 0);
    for (int section = 0; section < 2; ++section)
    {
        const char* start = section == 0 ? "vBiases_start" : "vWeights_start";
        const char* end   = section == 0 ? "vBiases_end" : "vWeights_end";
        if (!(f >> tok) || tok != start) return bad(std::string("expected ") + start);
        for (;;)
        {
            if (!(f >> tok)) return bad(std::string("missing ") + end);
            if (tok == end) break;
            uint64_t m = 0, l = 0, rows = 0, cols = 0;
            std::string t1, t2;
            if (tok != "pop" || !(f >> m >> t1 >> l >> t2 >> rows >> cols) || t1 != "layer" || t2 != "dims") return bad("malformed block header");
            if (m >= *top || l + 1 >= sz.size() || rows != sz[l + 1] || cols != (section == 0 ? 1 : sz[l])) return bad("dims do not match vSize");
            double* dst = elites_canon->data() + m * pc + (section == 0 ? b_at[l] : w_at[l]);
            for (uint64_t i = 0; i < rows * cols; ++i)
                if (!(f >> dst[i])) return bad("truncated matrix");
            seen[(m * (sz.size() - 1) + l) * 2 + section] = 1;
        }
    }
    for (char s : seen)
        if (!s) return bad("missing pop/layer block");
    // "generator": 624 (libc++) or 625 (libstdc++) integers follow.  Word 0 is the Philox key of the run that wrote the file IF the line
    // starts like a freshly seeded std::mt19937 (word 1 = 1812433253 (w0 ^ w0 >> 30) + 1: what ll_file_write emits); a file written by
    // the real ma-libs holds an arbitrary engine state there, which is not a seed: it is left alone and the configured seed stays.
    if (generator_word0)
    {
        uint64_t w0 = 0, w1 = 0;
        if ((f >> tok) && tok == "generator" && (f >> w0) && (f >> w1))
        {
            const uint32_t s0 = static_cast<uint32_t>(w0);
            if (static_cast<uint32_t>(w1) == 1812433253u * (s0 ^ (s0 >> 30)) + 1u) *generator_word0 = s0;
        }
    }
    return LL_OK;
}
