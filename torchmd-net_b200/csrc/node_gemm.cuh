// node_gemm.cuh -- dispatch of the plain node GEMMs: tcgen05 (tc_gemm.cuh) when the problem fits its constraints,
// the fp32 FFMA kernel (gemm.cuh) otherwise.  TMDN_TC_GEMM=0 forces the FFMA kernel (A/B measurements).
#pragma once
#include "gemm.cuh"
#include "tc_gemm.cuh"

static inline int node_gemm1(int epi, const GemmP& p, cudaStream_t stream, bool allow_tc = true) {
    static const int use_tc = [] { const char* e = getenv("TMDN_TC_GEMM"); return e ? atoi(e) : 1; }();
    // below ~1k rows the launch is a handful of tiles: gemm1 picks the latency-oriented small kernel (gemm.cuh)
    if (use_tc && allow_tc && p.M >= 1024 && tcgemm::supported(epi, p)) return tcgemm::launch(epi, p, stream);
    return gemm1(epi, p, stream);
}
