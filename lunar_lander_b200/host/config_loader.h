// config_loader.h — the `Config` namespace the reference's C++ reads its settings through
// (ma-libs cpp/libconfig_loader; call sites src/main_train.cpp:17, src/game_logic.cpp:42-95,
// src/nn_engine_trainer.cpp:29-65).  Same function names, argument meaning and failure behaviour:
// loadConfiguration returns false when the file cannot be read, a getter on a missing key throws
// std::runtime_error.  File syntax: `Section.key = value   # comment`, vectors comma separated,
// booleans true/false.  Unknown keys are kept (older sample configs carry extras).
#ifndef LL_HOST_CONFIG_LOADER_H_
#define LL_HOST_CONFIG_LOADER_H_

#include <string>
#include <vector>

#include "ll_gpu.h"

namespace Config
{
    bool loadConfiguration(const std::string& path);
    void clear();
    void set(const std::string& key, const std::string& value);
    bool has(const std::string& key);

    std::string getString(const std::string& key);
    double getDouble(const std::string& key);
    int getInt(const std::string& key);
    bool getBool(const std::string& key);
    std::vector<double> getVectorDouble(const std::string& key);
    std::vector<float> getVectorFloat(const std::string& key);
    std::vector<int> getVectorInt(const std::string& key);

    // optional keys (the GpuCfg.* extensions): value when present, fallback otherwise
    std::string getStringOr(const std::string& key, const std::string& fallback);
    int getIntOr(const std::string& key, int fallback);
    double getDoubleOr(const std::string& key, double fallback);

    // every value GameLogicCpp<T>'s constructor reads (src/game_logic.cpp:48-95), packed for ll_create
    ll_env_cfg envCfg();
} // namespace Config

#endif
