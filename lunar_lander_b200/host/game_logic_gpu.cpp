#include "game_logic_gpu.h"

#include <iostream>
#include <stdexcept>
#include <type_traits>

#include "config_loader.h"

template <typename T> void GameLogicCpp<T>::raise(const char* what) const
{
    const char* why = ll_last_error(ctx_);
    throw std::runtime_error(std::string(what) + ": " + (why && *why ? why : "unknown error"));
}

template <typename T> GameLogicCpp<T>::GameLogicCpp(bool no_print_flag, const std::string& config_file) : no_print(no_print_flag)
{
    if (!config_file.empty() && !Config::loadConfiguration(config_file))
        throw std::runtime_error("FATAL: Could not load configuration from " + config_file + ". Exiting.");
    ll_env_cfg env = Config::envCfg(); // throws std::runtime_error on a missing key, like the reference's getters
    lander_w_ = env.lander_width, lpad_w_ = env.lpad_width;
    const uint64_t sizes[2] = {5, 4}; // a context needs a network shape; the stepped lander never evaluates it
    ll_net_cfg net{};
    net.layer_sizes = sizes, net.n_layers = 2, net.activation_id = LL_ACT_TANH, net.population = 1, net.top = 1, net.seed = 0;
    net.precision = std::is_same<T, float>::value ? 32 : 64, net.mode = LL_MODE_EXACT, net.rank = 0, net.world = 1;
    net.device = Config::getIntOr("GpuCfg.device", -1);
    if (ll_create(&ctx_, &env, &net) != LL_OK)
    {
        const char* why = ll_last_error(nullptr);
        ctx_            = nullptr;
        throw std::runtime_error(std::string("GameLogicCpp: ll_create failed: ") + (why ? why : "?"));
    }
    if (ll_sim_reset(ctx_, 0, 0.0, 0.0, &v_) != LL_OK) raise("GameLogicCpp");
}

template <typename T> GameLogicCpp<T>::~GameLogicCpp()
{
    if (ctx_) ll_destroy(ctx_);
}

template <typename T> void GameLogicCpp<T>::reset()
{
    if (ll_sim_reset(ctx_, 0, 0.0, 0.0, &v_) != LL_OK) raise("reset");
}

template <typename T> void GameLogicCpp<T>::reset(T spad_x1_new, T lpad_x1_new)
{
    if (ll_sim_reset(ctx_, 1, static_cast<double>(spad_x1_new), static_cast<double>(lpad_x1_new), &v_) != LL_OK) raise("reset");
}

template <typename T> void GameLogicCpp<T>::get_state(T* pOutputs, size_t outputsSize) const
{
    if (!pOutputs || outputsSize < 5) throw std::invalid_argument("get_state: the output buffer must hold at least 5 values"); // assert, src/game_logic.cpp:277
    for (int i = 0; i < 5; ++i) pOutputs[i] = static_cast<T>(v_.state[i]);
}

template <typename T> std::vector<T> GameLogicCpp<T>::get_state() const
{
    std::vector<T> s(5);
    get_state(s.data(), s.size());
    return s;
}

template <typename T> bool GameLogicCpp<T>::update(int action, T* pStateOutput, size_t stateOutputSize)
{
    const bool was_done = v_.done != 0;
    if (ll_sim_update(ctx_, action, &v_) != LL_OK) raise("update");
    get_state(pStateOutput, stateOutputSize);
    if (!no_print && !was_done && v_.done)
    { // src/game_logic.cpp:216, 224-238; the two speed messages can never fire there: vx, vy are zeroed before they are tested
        if (v_.landed_successfully) std::cout << "GameLogicCpp: Successful landing!" << std::endl;
        else
        {
            const double lpad_x1 = v_.landing_pad_center_x - lpad_w_ / 2.0;
            if (!(v_.x >= lpad_x1 && v_.x + lander_w_ <= lpad_x1 + lpad_w_)) std::cout << "GameLogicCpp: Crashed - outside landing area!" << std::endl;
        }
    }
    return v_.done != 0;
}

template <typename T> std::pair<std::vector<T>, bool> GameLogicCpp<T>::update(int action)
{
    std::vector<T> s(5);
    const bool done = update(action, s.data(), s.size());
    return {s, done};
}

template <typename T> T GameLogicCpp<T>::calculate_step_penalty(int action) const
{
    double p = 0.0;
    if (ll_sim_step_penalty(ctx_, action, &p) != LL_OK) raise("calculate_step_penalty");
    return static_cast<T>(p);
}

template <typename T> T GameLogicCpp<T>::calculate_terminal_penalty(int steps_taken) const
{
    double p = 0.0;
    if (ll_sim_terminal_penalty(ctx_, steps_taken, -1, &p, nullptr) != LL_OK) raise("calculate_terminal_penalty");
    return static_cast<T>(p);
}

template <typename T> T GameLogicCpp<T>::calculate_terminal_penalty(int steps_taken, size_t direction, std::array<T, 2>& landing_bonus_lr) const
{
    if (direction >= 2) throw std::invalid_argument("Direction must be 0 or 1"); // assert, src/game_logic.cpp:356
    double p = 0.0, bonus[2] = {static_cast<double>(landing_bonus_lr[0]), static_cast<double>(landing_bonus_lr[1])};
    if (ll_sim_terminal_penalty(ctx_, steps_taken, static_cast<int32_t>(direction), &p, bonus) != LL_OK) raise("calculate_terminal_penalty");
    landing_bonus_lr[0] = static_cast<T>(bonus[0]), landing_bonus_lr[1] = static_cast<T>(bonus[1]);
    return static_cast<T>(p);
}

template <typename T> void GameLogicCpp<T>::set(const ll_sim_view& members)
{
    if (ll_sim_set(ctx_, &members, &v_) != LL_OK) raise("set");
}

template class GameLogicCpp<float>;
template class GameLogicCpp<double>;
