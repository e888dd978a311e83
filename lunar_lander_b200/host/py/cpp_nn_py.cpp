// cpp_nn_py — ma-libs' Python module as the reference uses it (mod_nn_play.py:19-56): ANN_MLP_GA_double / ANN_MLP_GA_float
// with SetName, Deserialize, Serialize, GetEpochs, GetNetworkSize, GetPopSize, GetTopPerformersSize and
// feedforwardIndex(inputs, member).  The network lives on the GPU behind include/ll_gpu.h (nn::ANN_MLP_GA<T> of
// ../ann_mlp_ga_gpu.h); feedforwardIndex is one ll_feedforward_index call: the exact-arithmetic forward pass of the
// episode kernel.  ma-libs' own source is absent: names and argument meaning follow the call sites.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <fstream>
#include <stdexcept>

#include "ann_mlp_ga_gpu.h"
#include "config_loader.h"

namespace py = pybind11;

namespace
{
    // The forward pass does not depend on the environment constants, but a context carries them: take them from the process-wide
    // Config when it is loaded, else from ./config.txt (the file main.py hands to GameLogicCpp, main.py:127), else the defaults of
    // the reference's config.txt.
    void ensure_config()
    {
        if (Config::has("GameCfg.pad_y1")) return;
        if (std::ifstream("config.txt").good() && Config::loadConfiguration("config.txt") && Config::has("GameCfg.pad_y1")) return;
        static const char* defaults[][2] = {
            {"GameCfg.pad_y1", "490"}, {"GameCfg.terrain_y", "50"}, {"GameCfg.max_vx", "0.5"}, {"GameCfg.max_vy", "2.0"}, {"PlanetCfg.g", "0.05"},
            {"PlanetCfg.mu_x", "0.01"}, {"PlanetCfg.mu_y", "0.01"}, {"LanderCfg.width", "65.5"}, {"LanderCfg.height", "54.2"},
            {"LanderCfg.max_fuel", "1000"}, {"GameCfg.lpad_x1", "560"}, {"GameCfg.lpad_width", "200"}, {"DisplayCfg.width", "960"},
            {"DisplayCfg.height", "540"}, {"GameCfg.spad_width", "80"}, {"GameCfg.x0", "52.0,433.8"}, {"GameCfg.v0", "0.0,0.0"},
            {"GameCfg.a0", "0.0,0.0"}, {"NNTrainingCfg.sp_dist_factor", "0.001"}, {"NNTrainingCfg.sp_action_factor", "0.01"},
            {"NNTrainingCfg.tp_steps_factor", "0.1"}, {"NNTrainingCfg.tp_dist_factor", "0.5"}, {"NNTrainingCfg.tp_landed_bonus", "0"},
            {"NNTrainingCfg.tp_landed_lr_bonus", "10000"}, {"NNTrainingCfg.tp_fuel_bonus_factor", "2"}, {"NNTrainingCfg.tp_crashed_penalty", "0"},
            {"NNTrainingCfg.tp_crash_v_mag_factor", "10"}, {"NNTrainingCfg.tp_no_fuel_penalty", "20"}};
        for (auto& kv : defaults)
            if (!Config::has(kv[0])) Config::set(kv[0], kv[1]);
    }

    template <typename T> void bind_net(py::module& m, const char* class_name)
    {
        using Net = nn::ANN_MLP_GA<T>;
        py::class_<Net>(m, class_name)
            .def(py::init([]() {
                ensure_config();
                auto* n = new Net();
                n->SetMode(LL_MODE_EXACT);
                return n;
            }))
            .def("SetName", [](Net& n, const std::string& name) { n.SetName(name.c_str()); }, py::arg("name"))
            .def("Deserialize", &Net::Deserialize, py::arg("path"))
            .def("Serialize", &Net::Serialize, py::arg("path"))
            .def("GetEpochs", &Net::GetEpochs)
            .def("GetNetworkSize",
                 [](const Net& n) {
                     if (!n.Context()) throw std::runtime_error("ANN_MLP_GA: the network has no shape yet (Deserialize first)");
                     return n.GetNetworkSize();
                 })
            .def("GetPopSize", &Net::GetPopSize)
            .def("GetTopPerformersSize", &Net::GetTopPerformersSize)
            .def(
                "feedforwardIndex",
                [](Net& n, py::array_t<T, py::array::c_style | py::array::forcecast> inputs, size_t member) -> size_t {
                    py::buffer_info b = inputs.request();
                    if (b.ndim != 1) throw std::runtime_error("feedforwardIndex: inputs must be a 1-dimensional array");
                    return n.feedforward(static_cast<const T*>(b.ptr), static_cast<size_t>(b.shape[0]), member);
                },
                py::arg("inputs"), py::arg("member") = 0);
    }
} // namespace

PYBIND11_MODULE(cpp_nn_py, m)
{
    m.doc() = "ANN_MLP_GA as mod_nn_play.py uses it, evaluated on the GPU behind include/ll_gpu.h";
    bind_net<double>(m, "ANN_MLP_GA_double");
    bind_net<float>(m, "ANN_MLP_GA_float");
}
