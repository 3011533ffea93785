// tcgen05 / TMEM GEMM for the prefill linears (placeholder until the tensor-core path lands:
// returning B200_EUNSUPPORTED makes b200_linear fall through to the SIMT tiled GEMM).
#include "common.cuh"
namespace b200 {
int gemm_tcgen05_f16(const void*, const void*, const void*, void*, int, int, int, cudaStream_t) {
    return B200_EUNSUPPORTED;
}
}  // namespace b200
