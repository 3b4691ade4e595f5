#include "common.cuh"

#include <map>
#include <mutex>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

namespace srb {

static thread_local char g_err[512] = "";

int set_error(const char* msg) {
  snprintf(g_err, sizeof(g_err), "CUDA srb200: %s", msg);
  return 1;
}
int set_cuda_error(const char* what, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "CUDA error in %s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  return 1;
}
const char* last_error() { return g_err; }

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("SRB_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
  return v != 0;
}

int sm_count() {
  static int cache[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  int n = cache[dev];
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cache[dev] = n;
  }
  return n;
}

void ensure_max_smem(const void* kernel, int bytes) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, int> done;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(mu);
  int& have = done[std::make_pair(kernel, dev)];
  if (have >= bytes) return;
  cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  have = bytes;
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t row_stride_bytes,
                      uint32_t box_inner, uint32_t box_outer) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return set_error("cuTensorMapEncodeTiled entry point not available");
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (row_stride_bytes & 15))
    return set_error("tensor map: base and row stride must be 16-byte aligned");
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[128];
    snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed with CUresult %d", int(r));
    return set_error(buf);
  }
  return 0;
}

int make_tmap_4d_bf16(CUtensorMap* out, const void* base, const uint64_t dims[4], const uint32_t box[4]) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return set_error("cuTensorMapEncodeTiled entry point not available");
  const uint64_t inner_bytes = dims[0] * 2, box_bytes = uint64_t(box[0]) * 2;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (inner_bytes & 15))
    return set_error("tensor map: base and innermost extent must be 16-byte aligned");
  if (box[0] > dims[0] || (box_bytes != 32 && box_bytes != 64 && box_bytes != 128))
    return set_error("tensor map (4d): innermost box must be 16, 32 or 64 bf16");
  cuuint64_t d[4] = {dims[0], dims[1], dims[2], dims[3]};
  cuuint64_t strides[3] = {inner_bytes, inner_bytes * dims[1], inner_bytes * dims[1] * dims[2]};
  cuuint32_t b[4] = {box[0], box[1], box[2], box[3]};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUtensorMapSwizzle sw = box_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                              : box_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), d, strides, b, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[128];
    snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled (4d) failed with CUresult %d", int(r));
    return set_error(buf);
  }
  return 0;
}

int make_tmap_3d_f32(CUtensorMap* out, const void* base, uint64_t W, uint64_t H, uint64_t B, uint32_t box_w,
                     uint32_t box_h) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return set_error("cuTensorMapEncodeTiled entry point not available");
  if ((reinterpret_cast<uintptr_t>(base) & 15) || ((W * 4) & 15) || ((uint64_t(box_w) * 4) & 15) || box_w > 256 || box_h > 256)
    return set_error("tensor map (3d f32): base, row pitch and box row must be multiples of 16 bytes, box <= 256");
  cuuint64_t d[3] = {W, H, B};
  cuuint64_t strides[2] = {W * 4, W * H * 4};
  cuuint32_t b[3] = {box_w, box_h, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(base), d, strides, b, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[128];
    snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled (3d f32) failed with CUresult %d", int(r));
    return set_error(buf);
  }
  return 0;
}

const char* last_error();

}  // namespace srb

extern "C" const char* srb_last_error(void) { return srb::last_error(); }
extern "C" int srb_abi_version(void) { return SRB_ABI_VERSION; }
