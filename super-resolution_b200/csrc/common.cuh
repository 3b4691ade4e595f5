// Host-side helpers shared by all translation units of libsrb200.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/srb200.h"

namespace srb {

int set_error(const char* msg);                         // returns 1
int set_cuda_error(const char* what, cudaError_t e);    // returns 1
int sm_count();

// 2-D bf16 tensor map: dims {inner, outer}, row stride in bytes, box {box_inner, box_outer}, 128B swizzle,
// out-of-bounds elements read as zero.  Returns 0 on success.
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t row_stride_bytes,
                      uint32_t box_inner, uint32_t box_outer);

inline int ceil_div(long long a, long long b) { return int((a + b - 1) / b); }

}  // namespace srb
