// ORACLE / TEST INFRASTRUCTURE — not part of the product.
// The product's device row layout, restated: per layer Wt[k][out_pad] (k-major, out_pad = out rounded up to 8) followed by b[out_pad]
// (lunar_lander_b200/csrc/ll_host.cpp ll_build_topo).  The float GA draws its Gaussians per device-order quad (ll_normal4_f32), so the
// checker needs the map device position -> canonical parameter index (per layer W[out][in] row-major then b[out]; -1 = padding).
#ifndef ORACLE_LL_DEVLAYOUT_H_
#define ORACLE_LL_DEVLAYOUT_H_

#include <cstddef>
#include <cstdint>
#include <vector>

inline std::vector<int> orc_dev2canon(const uint64_t* sizes, int n_sizes)
{
    std::vector<int> map;
    int canon = 0;
    for (int l = 0; l + 1 < n_sizes; ++l)
    {
        const int in = (int)sizes[l], out = (int)sizes[l + 1], op = (out + 7) & ~7;
        const size_t base = map.size();
        map.resize(base + (size_t)(in + 1) * op, -1);
        for (int k = 0; k < in; ++k)
            for (int o = 0; o < out; ++o) map[base + (size_t)k * op + o] = canon + o * in + k;
        for (int o = 0; o < out; ++o) map[base + (size_t)in * op + o] = canon + in * out + o;
        canon += in * out + out;
    }
    return map;
}

#endif
