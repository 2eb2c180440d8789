// cg_pair_inst.cu -- one translation unit per (element type, quantile count) instance of the
// pair-parallel streaming kernel.  Built with
//   -DCG_INST_T=double|float -DCG_INST_NQ=1..4 -DCG_INST_NAME=cg_launch_pair_<d|f><nq>
#include <cuda_runtime.h>
#include "cg_pairstream.cuh"

cudaError_t CG_INST_NAME(const cg::PairArgs& a, int grid, cudaStream_t s) {
    return cg::launch_pair_tn<CG_INST_T, CG_INST_NQ>(a, grid, s);
}

#if defined(CG_TIMING) && defined(CG_TIMING_EXPORT)
extern "C" int cg_debug_timing(unsigned long long* out16, int reset) {
    if (cudaMemcpyFromSymbol(out16, cg::cg_timing, sizeof(unsigned long long) * 16) != cudaSuccess) return 1;
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(cg::cg_timing, z, sizeof(z)); }
    return 0;
}
#endif
