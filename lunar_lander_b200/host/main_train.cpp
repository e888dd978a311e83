// main_train — command-line trainer with the reference's interface (src/main_train.cpp:14-106): reads ./config.txt, trains
// NNConfig.epochs generations, `--continue` resumes from <name>_last.txt, `--step N` from <name>_gen_<N, 4 digits>.txt.
// Built twice: main_train_d (double, bit-exact mode) and main_train_f (-DUSE_FLOAT; GpuCfg.mode = fast by default).
// Exit codes 0 / 1 as there.  stdout carries the reference's log lines; GPU facts go to stderr.
// Extensions: --config <file> (default config.txt) and --set Section.key=value (repeatable).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>
#include <type_traits>

#include "config_loader.h"
#include "nn_engine_trainer.h"

#if defined(USE_FLOAT)
using Real = float;
#else
using Real = double;
#endif

namespace
{
    bool parse_int(const char* text, int& value)
    {
        char* end = nullptr;
        const long v = std::strtol(text, &end, 10);
        if (end == text) return false;
        value = static_cast<int>(v);
        return true;
    }
} // namespace

int main(int argc, char** argv)
{
    std::string config_path = "config.txt", checkpoint = "last";
    bool resume = false;
    std::vector<std::pair<std::string, std::string>> overrides;
    for (int i = 1; i < argc; ++i)
    {
        const std::string arg = argv[i];
        if (arg == "--continue") resume = true;
        else if (arg == "--step")
        {
            if (i + 1 >= argc)
            {
                std::cerr << "Error: --step option requires an integer argument." << std::endl;
                return 1;
            }
            int n = 0;
            if (!parse_int(argv[++i], n))
            {
                std::cerr << "Error: Invalid argument for --step: " << argv[i] << " is not a valid integer." << std::endl;
                return 1;
            }
            char buf[32];
            std::snprintf(buf, sizeof buf, "gen_%04d", n);
            checkpoint = buf;
            resume     = true; // --step implies --continue
        }
        else if (arg == "--config" && i + 1 < argc) config_path = argv[++i];
        else if (arg == "--set" && i + 1 < argc)
        {
            const std::string kv = argv[++i];
            const size_t eq      = kv.find('=');
            if (eq == std::string::npos)
            {
                std::cerr << "Error: --set expects Section.key=value, got " << kv << std::endl;
                return 1;
            }
            overrides.emplace_back(kv.substr(0, eq), kv.substr(eq + 1));
        }
        else std::cerr << "Warning: Unknown argument: " << arg << std::endl;
    }

    if (!Config::loadConfiguration(config_path))
    {
        std::cerr << "FATAL: Could not load configuration from " << config_path << ". Exiting." << std::endl;
        return 1;
    }
    for (const auto& kv : overrides) Config::set(kv.first, kv.second);

    std::cout << "Using data type: " << (std::is_same<Real, float>::value ? "float" : "double") << std::endl;

    try
    {
        Training::NNEngineTrainer<Real> trainer;
        if (!trainer.init())
        {
            std::cerr << "Failed to initialize NNEngineTrainer." << std::endl;
            return 1;
        }
        if (resume)
        {
            std::cout << "Attempting to load network from checkpoint: " << checkpoint << std::endl;
            if (trainer.load(checkpoint))
                std::cout << "Successfully loaded network from checkpoint '" << checkpoint << "'." << std::endl;
            else
            {
                std::cerr << "Failed to load network from checkpoint '" << checkpoint
                          << "'. Training will start from scratch (using the newly initialized network)." << std::endl;
                if (!trainer.init()) return 1; // load() dropped the network; the reference keeps the fresh one
            }
        }
        else
            std::cout << "Starting training from scratch (no --continue specified, or --step without --continue)." << std::endl;

        const auto t0 = std::chrono::steady_clock::now();
        trainer.train();
        const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (trainer.net())
        {
            const double steps = trainer.net()->TotalLanderSteps(false);
            std::cerr << "[ll_gpu] kernel: " << trainer.net()->KernelName() << "; " << trainer.history().size() << " generations, "
                      << steps << " lander steps in " << wall << " s = " << (wall > 0 ? steps / wall : 0.0) << " lander-steps/s"
                      << std::endl;
        }
    } catch (const std::exception& e)
    {
        std::cerr << "FATAL: " << e.what() << std::endl;
        return 1;
    }
    std::cout << "Training process finished in C++." << std::endl;
    return 0;
}
