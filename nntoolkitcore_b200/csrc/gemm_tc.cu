// gemm_tc.cu -- K2/K3, tensor-core path (NNTK_MATH_TF32): placeholder until the tcgen05 kernel lands.
// Returns 1 = "shape not eligible", which makes nntk_cuda_gemm run the fp32 FFMA kernel.
#include "common.cuh"
extern "C" int nntk_cuda_gemm_tf32(const nntk_gemm_args *g) { (void)g; return 1; }
