// table_grid.cuh -- interval grid shared by the filter tables of tensornet.cu and et.cu.
//
// The distance filters are univariate functions of d built from ExpNormal RBFs (models/utils.py:339-344): their
// features have a constant width in u = exp(-alpha d), i.e. they are narrow at small d and broad near the cutoff,
// while the cutoff / activation factors vary on the scale of the cutoff itself.  A grid uniform in
//     S(d) = 1 - exp(-alpha d) + gamma d            (alpha = 5 / cutoff_upper, gamma = 0.05 / cutoff_upper)
// puts the intervals where the functions vary: 16 K intervals (K = num_rbf) reach the interpolation error of a
// 4096-interval grid uniform in d (|f| 1e-9, |df/dd| 6e-7 for the example hparams; tests/test_host.py), so a table
// is 3 MB instead of 25 MB -- and, because S compresses the far end where most neighbours of a condensed system
// are, the last ~24 intervals serve 50-75 % of all lookups: the large-system edge kernels keep those rows in shared
// memory (hot rows) and fetch only the rest through L2.  alpha = 0 selects the plain grid uniform in d
// (Gaussian smearing, cutoff_lower > 0).
#pragma once
#include <cuda_runtime.h>

struct TabGrid {
    float a2;      // -alpha * log2(e)
    float alpha;   // 0: uniform in d
    float gamma;
    float inv_h;   // n_int / S(cutoff_upper)
    int n_int;
};

static inline TabGrid make_tab_grid(int n_int, double hi, double alpha, double gamma) {
    TabGrid g;
    g.n_int = n_int;
    g.alpha = (float)alpha;
    g.gamma = (float)gamma;
    g.a2 = (float)(-alpha * 1.4426950408889634);
    const double T = alpha > 0 ? 1.0 - exp(-alpha * hi) + gamma * hi : hi;
    g.inv_h = n_int > 0 ? (float)((double)n_int / T) : 0.f;
    return g;
}

#ifdef __CUDACC__
// interval k, local coordinate s in [0,1) (clamped interval, s may exceed 1 only by rounding) and ds/dd
__device__ __forceinline__ void tab_locate(const TabGrid& g, float d, int& k, float& s, float& dsdd) {
    float sv;
    if (g.alpha > 0.f) {
        const float u = exp2f(g.a2 * d);
        sv = (1.0f - u + g.gamma * d) * g.inv_h;
        dsdd = (g.alpha * u + g.gamma) * g.inv_h;
    } else {
        sv = d * g.inv_h;
        dsdd = g.inv_h;
    }
    k = (int)sv;
    k = k > g.n_int - 1 ? g.n_int - 1 : k;
    s = sv - (float)k;
}
#endif
