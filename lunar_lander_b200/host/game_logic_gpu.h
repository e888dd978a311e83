// game_logic_gpu.h — GameLogicCpp<T> (src/game_logic.h:15-93) with the lander living on the GPU: every method is one
// ll_sim_* call of include/ll_gpu.h, which runs the episode kernel's own device functions on a single lander
// (csrc/ll_sim.cu), in T with the exact arithmetic.  Same constructor arguments, method names, return values, messages
// and failure behaviour as the reference class; the public data members (x, y, vx, vy, fuel, the flags,
// landing_pad_center_x, landing_pad_y, last_action) are read through view() and written through set(), because a plain
// member cannot mirror device state — the pybind module binds them as read/write properties of the same names.
#ifndef LL_HOST_GAME_LOGIC_GPU_H_
#define LL_HOST_GAME_LOGIC_GPU_H_

#include <array>
#include <cstddef>
#include <string>
#include <utility>
#include <vector>

#include "ll_gpu.h"

template <typename T> class GameLogicCpp
{
  public:
    bool no_print;

    // If config_file is empty the process-wide Config is used as loaded; otherwise it is loaded first
    // (std::runtime_error "FATAL: Could not load configuration from ..." on failure, src/game_logic.cpp:40-46).
    GameLogicCpp(bool no_print_flag = false, const std::string& config_file = "");
    ~GameLogicCpp();
    GameLogicCpp(const GameLogicCpp&)            = delete;
    GameLogicCpp& operator=(const GameLogicCpp&) = delete;

    void reset();
    void reset(T spad_x1_new, T lpad_x1_new);
    std::pair<std::vector<T>, bool> update(int action);
    bool update(int action, T* pStateOutput, size_t stateOutputSize);
    bool is_done() const { return v_.done != 0; }
    std::vector<T> get_state() const;
    void get_state(T* pOutputs, size_t outputsSize) const;
    T calculate_step_penalty(int action) const;
    T calculate_terminal_penalty(int steps_taken) const;
    T calculate_terminal_penalty(int steps_taken, size_t direction, std::array<T, 2>& landing_bonus_lr) const;

    const ll_sim_view& view() const { return v_; } // state after the last call
    void set(const ll_sim_view& members);          // writes the def_readwrite members (src/lunar_lander_cpp.cpp:27-38)

  private:
    [[noreturn]] void raise(const char* what) const;
    ll_ctx* ctx_ = nullptr;
    ll_sim_view v_{};
    double lander_w_ = 0.0, lpad_w_ = 0.0;
};

#endif
