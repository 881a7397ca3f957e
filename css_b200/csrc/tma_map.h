// TMA tensor map over bf16 planes [planes][rows][ld] (gemm_tc.cu): box = BK columns x box_rows rows x 1 plane,
// SWIZZLE_128B for BK = 64, SWIZZLE_64B for BK = 32.
#pragma once
#include <cuda.h>
int css_make_plane_map(CUtensorMap* map, const void* base, long rows, int K, int ld, long plane_stride, int planes, int box_rows,
                       int BK);
// 3-D fp32 tensor map, no swizzle: dims / box in elements (innermost first), strides of dims 1 and 2 in bytes (multiples of 16).
// Out-of-bounds box elements read as zero.
int css_make_f32_map3(CUtensorMap* map, const void* base, const unsigned long long dims[3], const unsigned long long strides[2],
                      const unsigned box[3]);
