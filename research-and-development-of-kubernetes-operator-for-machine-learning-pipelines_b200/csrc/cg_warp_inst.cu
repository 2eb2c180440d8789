// cg_warp_inst.cu -- one translation unit per (element type, quantile count) instance of the
// streaming kernel, so the instances compile in parallel.  Built with
//   -DCG_INST_T=double|float -DCG_INST_NQ=1..4 -DCG_INST_NAME=cg_launch_warp_<d|f><nq>
#include <cuda_runtime.h>
#include "cg_warp.cuh"

cudaError_t CG_INST_NAME(const cg::WarpArgs& a, int grid, cudaStream_t s) {
    return cg::launch_warp_tn<CG_INST_T, CG_INST_NQ>(a, grid, s);
}

#if defined(CG_CHECKED) && defined(CG_CHECK_EXPORT)
// developer build only (see CG_CHECK in cg_warp.cuh): failed bounds checks of THIS instance since the last reset
extern "C" int cg_debug_check_failures(unsigned int* out8, int reset) {
    if (cudaMemcpyFromSymbol(out8, cg::cg_check_failures, sizeof(unsigned int) * 8) != cudaSuccess) return 1;
    if (reset) { unsigned int z[8] = {0}; cudaMemcpyToSymbol(cg::cg_check_failures, z, sizeof(z)); }
    return 0;
}
#endif
