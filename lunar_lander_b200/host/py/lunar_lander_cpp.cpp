// lunar_lander_cpp — the reference's pybind11 module of the same name (src/lunar_lander_cpp.cpp:16-107), built over the
// GPU-resident GameLogicCpp<T> of ../game_logic_gpu.h (C ABI: ll_sim_*).  Same classes (GameLogicCppFloat,
// GameLogicCppDouble), constructor defaults, member names (bound as read/write properties: the state lives on the
// device), method names, argument names, return types and error messages, plus generate_random_pad_positions, so
// `import lunar_lander_cpp` in main.py:24-28 works unchanged once this directory is on sys.path.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <optional>
#include <stdexcept>

#include "game_logic_gpu.h"

namespace py = pybind11;

namespace
{
    template <typename T> void bind_game_logic_class(py::module& m, const char* class_name)
    {
        using Class = GameLogicCpp<T>;
        py::class_<Class> c(m, class_name);
        c.def(py::init<bool, const std::string&>(), py::arg("no_print_flag") = false, py::arg("config_file") = "");
#define LL_RW_REAL(name)                                                                                            \
    c.def_property(#name, [](const Class& s) { return static_cast<T>(s.view().name); },                             \
                   [](Class& s, T v) { ll_sim_view w = s.view(); w.name = static_cast<double>(v); s.set(w); })
#define LL_RW_FLAG(name)                                                                                            \
    c.def_property(#name, [](const Class& s) { return s.view().name != 0; },                                        \
                   [](Class& s, bool v) { ll_sim_view w = s.view(); w.name = v ? 1 : 0; s.set(w); })
        LL_RW_REAL(x);
        LL_RW_REAL(y);
        LL_RW_REAL(vx);
        LL_RW_REAL(vy);
        LL_RW_REAL(fuel);
        LL_RW_FLAG(landed);
        LL_RW_FLAG(crashed);
        LL_RW_FLAG(landed_successfully);
        LL_RW_REAL(landing_pad_center_x);
        // landing_pad_y is a cached configuration value; the C ABI exposes it read-only
        c.def_property_readonly("landing_pad_y", [](const Class& s) { return static_cast<T>(s.view().landing_pad_y); });
        c.def_readwrite("no_print", &Class::no_print);
        c.def_property("last_action", [](const Class& s) { return static_cast<int>(s.view().last_action); },
                       [](Class& s, int v) { ll_sim_view w = s.view(); w.last_action = v; s.set(w); });
        c.def_property_readonly("time_step", [](const Class& s) { return static_cast<int>(s.view().time_step); }); // extra: private in the reference
#undef LL_RW_REAL
#undef LL_RW_FLAG
        c.def("reset", static_cast<void (Class::*)()>(&Class::reset), "Resets the game state to the initial configuration.");
        c.def("reset", static_cast<void (Class::*)(T, T)>(&Class::reset), py::arg("spad_x1"), py::arg("lpad_x1_new"),
              "Resets the game state with specific pad positions.");
        c.def("is_done", &Class::is_done, "Checks if the game episode has finished (landed or crashed).");
        c.def(
            "get_state",
            [](const Class& self, py::array_t<T> state_output) {
                py::buffer_info buf_info = state_output.request();
                if (buf_info.ndim != 1) { throw std::runtime_error("get_state: Output array must be 1-dimensional."); }
                self.get_state(static_cast<T*>(buf_info.ptr), static_cast<size_t>(buf_info.shape[0]));
            },
            py::arg("state_output"), "Fills the provided NumPy array with the current state vector.");
        c.def(
            "update",
            [](Class& self, int action, py::array_t<T> state_output) -> bool {
                py::buffer_info buf_info = state_output.request();
                if (buf_info.ndim != 1) { throw std::runtime_error("update: Output state array must be 1-dimensional."); }
                return self.update(action, static_cast<T*>(buf_info.ptr), static_cast<size_t>(buf_info.shape[0]));
            },
            py::arg("action"), py::arg("state_output"),
            "Performs one time step, fills the provided NumPy array with the new state, and returns the done flag (bool).");
        c.def("calculate_step_penalty", &Class::calculate_step_penalty, py::arg("action"),
              "Calculates the penalty applied at each step based on state and action.");
        c.def("calculate_terminal_penalty", static_cast<T (Class::*)(int) const>(&Class::calculate_terminal_penalty), py::arg("steps_taken"),
              "Calculates the terminal penalty/reward based on the final state and steps taken (original version).");
        c.def(
            "calculate_terminal_penalty_lr_bonus",
            [](const Class& self, int steps_taken, size_t direction) -> py::tuple {
                std::array<T, 2> landing_bonus_lr = {T(0.0), T(0.0)};
                T penalty                         = self.calculate_terminal_penalty(steps_taken, direction, landing_bonus_lr);
                return py::make_tuple(penalty, landing_bonus_lr);
            },
            py::arg("steps_taken"), py::arg("direction"),
            "Calculates terminal penalty, populating L/R bonus array. Returns (penalty, [bonus_L, bonus_R]).");
    }

    std::pair<int, int> generate_random_pad_positions(double screen_width, int spad_width, int lpad_width, std::optional<int> seed,
                                                      bool force_left_to_right, bool force_right_to_left)
    { // src/game_utils.cpp:12-74 behind ll_generate_random_pad_positions
        int32_t s = 0, l = 0;
        if (ll_generate_random_pad_positions(screen_width, spad_width, lpad_width, seed ? 1 : 0, seed ? *seed : 0, force_left_to_right ? 1 : 0,
                                             force_right_to_left ? 1 : 0, &s, &l) != LL_OK)
        {
            const char* why = ll_last_error(nullptr);
            throw std::runtime_error(std::string("generate_random_pad_positions: ") + (why ? why : "failed"));
        }
        return {s, l};
    }
} // namespace

PYBIND11_MODULE(lunar_lander_cpp, m)
{
    m.doc() = "pybind11 plugin for C++ Lunar Lander (B200: the lander is stepped on the GPU behind include/ll_gpu.h)";
    // the reference asserts on these (src/game_logic.cpp:277, :356); a Python caller gets an AssertionError instead of an abort
    py::register_exception_translator([](std::exception_ptr p) {
        try
        {
            if (p) std::rethrow_exception(p);
        }
        catch (const std::invalid_argument& e)
        {
            PyErr_SetString(PyExc_AssertionError, e.what());
        }
    });
    bind_game_logic_class<float>(m, "GameLogicCppFloat");
    bind_game_logic_class<double>(m, "GameLogicCppDouble");
    m.def("generate_random_pad_positions", &generate_random_pad_positions, py::arg("screen_width"), py::arg("spad_width"), py::arg("lpad_width"),
          py::arg("seed") = py::none(), py::arg("force_left_to_right") = false, py::arg("force_right_to_left") = false,
          "Generates random, non-overlapping positions for start and landing pads.");
}
