// placeholder until the tcgen05 GEMM lands: report "does not qualify" so the dispatcher uses SIMT
#include "common.cuh"
namespace acai {
int linear_tcgen05(const void*, int64_t, const void*, int64_t, const float*, const void*, int, int64_t, void*, int,
                   int64_t, int, int, int, int, cudaStream_t) { return ACAI_ERR_UNSUPPORTED; }
}
