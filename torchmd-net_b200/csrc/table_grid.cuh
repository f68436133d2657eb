// table_grid.cuh -- interval grid shared by the filter tables of tensornet.cu and et.cu.
//
// The distance filters are univariate functions of d built from ExpNormal RBFs (models/utils.py:339-344): their
// features have a constant width in u = exp(-alpha d), i.e. they are narrow at small d and broad near the cutoff,
// while the cutoff / activation factors vary on the scale of the cutoff itself.  A grid uniform in
//     t(d) = (u(d) - u(r_c)) + gamma (r_c - d)        (alpha = 5 / r_c, gamma = 0.05 / r_c; t(r_c) = 0)
// puts the intervals where the functions vary: max(1024, 32 K) intervals (K = num_rbf) give |f| errors of 1e-9 and
// |df/dd| errors of 1e-7 (tests/test_host.py) -- better than 4096 intervals uniform in d at a quarter of the size
// (6 MB instead of 25 MB per table at H = 128).  t is measured FROM THE CUTOFF so that fp32 evaluates it accurately
// where it is flat: near r_c both terms are small numbers and the rounding of t maps back to a distance error below
// one ulp of d.  (Measured from d = 0 the far end sits at t ~ 1 with an ulp of 6e-8 and a slope of 0.02 -- 3e-6 A of
// noise in d, which cost the 375-atom periodic fixture 5e-5 in the energy, gpurun r02a.)  Row 0 of a table is
// therefore the interval that ends at the cutoff.
// alpha = 0 selects the grid uniform in d (Gaussian smearing, cutoff_lower > 0), also counted from the cutoff.
#pragma once
#include <cuda_runtime.h>

struct TabGrid {
    float a2;      // -alpha * log2(e)
    float alpha;   // 0: uniform in d
    float gamma;
    float inv_h;   // n_int / t(0)
    float u_min;   // exp(-alpha r_c)
    float hi;      // r_c
    int n_int;
};

static inline TabGrid make_tab_grid(int n_int, double hi, double alpha, double gamma) {
    TabGrid g;
    g.n_int = n_int;
    g.alpha = (float)alpha;
    g.gamma = (float)gamma;
    g.a2 = (float)(-alpha * 1.4426950408889634);
    g.u_min = (float)exp(-alpha * hi);
    g.hi = (float)hi;
    const double T = alpha > 0 ? 1.0 - exp(-alpha * hi) + gamma * hi : hi;
    g.inv_h = n_int > 0 ? (float)((double)n_int / T) : 0.f;
    return g;
}

#ifdef __CUDACC__
// interval k (0 = the interval that ends at the cutoff), local coordinate s in [0,1] and ds/dd (negative)
__device__ __forceinline__ void tab_locate(const TabGrid& g, float d, int& k, float& s, float& dsdd) {
    float q;
    if (g.alpha > 0.f) {
        // a2 * d lies in [-7.3, 0]: the bare MUFU.EX2 (what exp2f issues inside the normal range, without its four
        // instructions of denormal-range handling -- these loops are issue-bound); same bits
        float u;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(u) : "f"(g.a2 * d));
        q = ((u - g.u_min) + g.gamma * (g.hi - d)) * g.inv_h;
        dsdd = -(g.alpha * u + g.gamma) * g.inv_h;
    } else {
        q = (g.hi - d) * g.inv_h;
        dsdd = -g.inv_h;
    }
    k = (int)q;
    k = k > g.n_int - 1 ? g.n_int - 1 : (k < 0 ? 0 : k);
    s = q - (float)k;
}
#endif
