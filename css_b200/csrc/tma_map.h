// TMA tensor map over bf16 planes [planes][rows][ld] (gemm_tc.cu): box = BK columns x box_rows rows x 1 plane,
// SWIZZLE_128B for BK = 64, SWIZZLE_64B for BK = 32.
#pragma once
#include <cuda.h>
int css_make_plane_map(CUtensorMap* map, const void* base, long rows, int K, int ld, long plane_stride, int planes, int box_rows,
                       int BK);
