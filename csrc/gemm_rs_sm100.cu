// GEMM -> reduce-scatter in one kernel (mtb_gemm_bf16_rs, include/mtb200.h): the tcgen05 GEMM engine of
// gemm_sm100.cu compiled a second time with the epilogue that adds every tile into the owner rank's gradient slice
// over NVLink peer memory.  A separate translation unit so that the plain GEMM kernels' code generation is
// untouched by it (adding the branch to the shared epilogue cost the flagship step 7 us).
#define MTB_GEMM_RS_TU 1
#include "gemm_sm100.cu"
