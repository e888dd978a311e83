// ORACLE / TEST INFRASTRUCTURE — not part of the product.
//
// Stand-in for ma-libs' cpp/libconfig_loader/config_loader.h (azimonti/ma-libs @ dd81ee42, an
// un-vendored submodule that is absent from /root/reference: .gitmodules:1-3).  It exists only so
// that the reference's own sources compile UNCHANGED into oracle/_ref (see oracle/Makefile).  The
// interface is reconstructed from the call sites: Config::loadConfiguration (src/main_train.cpp:17,
// src/game_logic.cpp:42), getDouble/getInt/getBool/getString/getVectorDouble/getVectorFloat/
// getVectorInt (src/game_logic.cpp:48-95, src/nn_engine_trainer.cpp:29-65).
// File syntax from config.txt: `Section.key = value   # comment`, vectors comma separated.
#ifndef ORACLE_SHIM_CONFIG_LOADER_H_
#define ORACLE_SHIM_CONFIG_LOADER_H_

#include <cstdlib>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace Config
{
    inline std::map<std::string, std::string>& store()
    {
        static std::map<std::string, std::string> kv;
        return kv;
    }

    inline std::string strip(const std::string& s)
    {
        size_t b = s.find_first_not_of(" \t\r\n");
        if (b == std::string::npos) return "";
        size_t e = s.find_last_not_of(" \t\r\n");
        return s.substr(b, e - b + 1);
    }

    // Process-global, like the reference's loader (SURVEY.md §5 "Config / flags").  Re-loading merges.
    inline bool loadConfiguration(const std::string& path)
    {
        std::ifstream in(path);
        if (!in) return false;
        std::string line;
        while (std::getline(in, line))
        {
            size_t hash = line.find('#');
            if (hash != std::string::npos) line.erase(hash);
            size_t eq = line.find('=');
            if (eq == std::string::npos) continue;
            std::string key = strip(line.substr(0, eq)), val = strip(line.substr(eq + 1));
            if (!key.empty()) store()[key] = val;
        }
        return true;
    }

    inline void clear() { store().clear(); }
    inline void set(const std::string& key, const std::string& value) { store()[key] = value; }

    inline const std::string& raw(const std::string& key)
    {
        auto it = store().find(key);
        if (it == store().end()) throw std::runtime_error("Config: missing key " + key);
        return it->second;
    }

    inline std::string getString(const std::string& key) { return raw(key); }
    inline double getDouble(const std::string& key) { return std::strtod(raw(key).c_str(), nullptr); }
    inline int getInt(const std::string& key) { return static_cast<int>(std::strtol(raw(key).c_str(), nullptr, 10)); }
    inline bool getBool(const std::string& key)
    {
        const std::string& v = raw(key);
        return v == "true" || v == "True" || v == "TRUE" || v == "1";
    }

    inline std::vector<std::string> split(const std::string& key)
    {
        std::vector<std::string> out;
        std::stringstream ss(raw(key));
        std::string item;
        while (std::getline(ss, item, ','))
        {
            item = strip(item);
            if (!item.empty()) out.push_back(item);
        }
        return out;
    }

    inline std::vector<double> getVectorDouble(const std::string& key)
    {
        std::vector<double> v;
        for (auto& s : split(key)) v.push_back(std::strtod(s.c_str(), nullptr));
        return v;
    }
    // value -> double -> float, the same conversion the scalar keys go through (static_cast<T>(getDouble))
    inline std::vector<float> getVectorFloat(const std::string& key)
    {
        std::vector<float> v;
        for (auto& s : split(key)) v.push_back(static_cast<float>(std::strtod(s.c_str(), nullptr)));
        return v;
    }
    inline std::vector<int> getVectorInt(const std::string& key)
    {
        std::vector<int> v;
        for (auto& s : split(key)) v.push_back(static_cast<int>(std::strtol(s.c_str(), nullptr, 10)));
        return v;
    }
} // namespace Config

#endif
