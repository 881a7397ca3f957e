// tcgen05 GEMM path -- placeholder until the UMMA kernel lands.
#include "common.cuh"
#include "gemm.h"

int gemm_tc_launch(const float*, int, const float*, int, float*, int, long, int, int, GemmEpilogue, int, cudaStream_t) {
    return GEMM_TC_UNSUPPORTED;
}
