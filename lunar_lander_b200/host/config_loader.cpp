#include "config_loader.h"

#include <cstdlib>
#include <fstream>
#include <map>
#include <stdexcept>

namespace
{
    std::map<std::string, std::string>& table()
    {
        static std::map<std::string, std::string> t; // process-global like the reference's loader
        return t;
    }

    std::string trimmed(const std::string& s)
    {
        const char* ws = " \t\r\n";
        const size_t a = s.find_first_not_of(ws);
        if (a == std::string::npos) return std::string();
        return s.substr(a, s.find_last_not_of(ws) - a + 1);
    }

    const std::string& lookup(const std::string& key)
    {
        auto it = table().find(key);
        if (it == table().end()) throw std::runtime_error("Config: missing key " + key);
        return it->second;
    }

    std::vector<std::string> items(const std::string& key)
    {
        std::vector<std::string> out;
        const std::string& v = lookup(key);
        size_t pos = 0;
        while (pos <= v.size())
        {
            size_t comma = v.find(',', pos);
            if (comma == std::string::npos) comma = v.size();
            std::string piece = trimmed(v.substr(pos, comma - pos));
            if (!piece.empty()) out.push_back(piece);
            pos = comma + 1;
        }
        return out;
    }
} // namespace

bool Config::loadConfiguration(const std::string& path)
{
    std::ifstream in(path);
    if (!in) return false;
    for (std::string line; std::getline(in, line);)
    {
        line = line.substr(0, line.find('#'));
        const size_t eq = line.find('=');
        if (eq == std::string::npos) continue;
        const std::string key = trimmed(line.substr(0, eq));
        if (!key.empty()) table()[key] = trimmed(line.substr(eq + 1));
    }
    return true;
}

void Config::clear() { table().clear(); }
void Config::set(const std::string& key, const std::string& value) { table()[key] = value; }
bool Config::has(const std::string& key) { return table().count(key) != 0; }

std::string Config::getString(const std::string& key) { return lookup(key); }
double Config::getDouble(const std::string& key) { return std::strtod(lookup(key).c_str(), nullptr); }
int Config::getInt(const std::string& key) { return static_cast<int>(std::strtol(lookup(key).c_str(), nullptr, 10)); }
bool Config::getBool(const std::string& key)
{
    const std::string& v = lookup(key);
    return v == "true" || v == "True" || v == "TRUE" || v == "1";
}

std::vector<double> Config::getVectorDouble(const std::string& key)
{
    std::vector<double> out;
    for (const std::string& s : items(key)) out.push_back(std::strtod(s.c_str(), nullptr));
    return out;
}
std::vector<float> Config::getVectorFloat(const std::string& key)
{
    std::vector<float> out; // text -> double -> float, like static_cast<T>(getDouble(...))
    for (const std::string& s : items(key)) out.push_back(static_cast<float>(std::strtod(s.c_str(), nullptr)));
    return out;
}
std::vector<int> Config::getVectorInt(const std::string& key)
{
    std::vector<int> out;
    for (const std::string& s : items(key)) out.push_back(static_cast<int>(std::strtol(s.c_str(), nullptr, 10)));
    return out;
}

std::string Config::getStringOr(const std::string& key, const std::string& fallback) { return has(key) ? getString(key) : fallback; }
int Config::getIntOr(const std::string& key, int fallback) { return has(key) ? getInt(key) : fallback; }
double Config::getDoubleOr(const std::string& key, double fallback) { return has(key) ? getDouble(key) : fallback; }

ll_env_cfg Config::envCfg()
{
    auto pair_of = [](const char* key, double& a, double& b) {
        std::vector<double> v = Config::getVectorDouble(key);
        if (v.size() < 2) throw std::runtime_error(std::string("Config: ") + key + " needs two components");
        a = v[0], b = v[1];
    };
    ll_env_cfg e{};
    e.pad_y1         = getDouble("GameCfg.pad_y1");
    e.terrain_y      = getDouble("GameCfg.terrain_y");
    e.max_vx         = getDouble("GameCfg.max_vx");
    e.max_vy         = getDouble("GameCfg.max_vy");
    e.g              = getDouble("PlanetCfg.g");
    e.mu_x           = getDouble("PlanetCfg.mu_x");
    e.mu_y           = getDouble("PlanetCfg.mu_y");
    e.lander_width   = getDouble("LanderCfg.width");
    e.lander_height  = getDouble("LanderCfg.height");
    e.max_fuel       = static_cast<double>(getInt("LanderCfg.max_fuel")); // an int in the reference (src/game_logic.cpp:64)
    e.lpad_x1        = getDouble("GameCfg.lpad_x1");
    e.lpad_width     = getDouble("GameCfg.lpad_width");
    e.display_width  = getDouble("DisplayCfg.width");
    e.display_height = getDouble("DisplayCfg.height");
    e.spad_width     = getDouble("GameCfg.spad_width");
    pair_of("GameCfg.x0", e.x0, e.y0);
    pair_of("GameCfg.v0", e.vx0, e.vy0);
    pair_of("GameCfg.a0", e.ax0, e.ay0);
    e.sp_dist_factor        = getDouble("NNTrainingCfg.sp_dist_factor");
    e.sp_action_factor      = getDouble("NNTrainingCfg.sp_action_factor");
    e.tp_steps_factor       = getDouble("NNTrainingCfg.tp_steps_factor");
    e.tp_dist_factor        = getDouble("NNTrainingCfg.tp_dist_factor");
    e.tp_landed_bonus       = getDouble("NNTrainingCfg.tp_landed_bonus");
    e.tp_landed_lr_bonus    = getDouble("NNTrainingCfg.tp_landed_lr_bonus");
    e.tp_fuel_bonus_factor  = getDouble("NNTrainingCfg.tp_fuel_bonus_factor");
    e.tp_crashed_penalty    = getDouble("NNTrainingCfg.tp_crashed_penalty");
    e.tp_crash_v_mag_factor = getDouble("NNTrainingCfg.tp_crash_v_mag_factor");
    e.tp_no_fuel_penalty    = getDouble("NNTrainingCfg.tp_no_fuel_penalty");
    return e;
}
