// Host-side helpers shared by all translation units of libsrb200.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <utility>

#include "../../include/srb200.h"

namespace srb {

int set_error(const char* msg);                         // returns 1
int set_cuda_error(const char* what, cudaError_t e);    // returns 1
int sm_count();                                         // SMs of the CURRENT device (cached per device)
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (current device, kernel): the attribute is per device, so a
// process-wide "done" flag would leave the second GPU of a multi-device process at the 48 KB default.
void ensure_max_smem(const void* kernel, int bytes);

// 2-D bf16 tensor map: dims {inner, outer}, row stride in bytes, box {box_inner, box_outer}, 128B swizzle,
// out-of-bounds elements read as zero.  Returns 0 on success.
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t row_stride_bytes,
                      uint32_t box_inner, uint32_t box_outer);
// 4-D tiled map over a dense bf16 [d3][d2][d1][d0] tensor (d0 innermost; the box takes 16/32/64 of its elements);
// swizzle = the innermost box extent in bytes (32B/64B/128B), out-of-bounds elements read as zero.
int make_tmap_4d_bf16(CUtensorMap* out, const void* base, const uint64_t dims[4], const uint32_t box[4]);
// 3-D tiled map over a dense fp32 [B][H][W] image stack, box = box_h rows x box_w columns of one image, no swizzle,
// out-of-bounds elements (halo) read as zero.
int make_tmap_3d_f32(CUtensorMap* out, const void* base, uint64_t W, uint64_t H, uint64_t B, uint32_t box_w,
                     uint32_t box_h);

inline int ceil_div(long long a, long long b) { return int((a + b - 1) / b); }

// Programmatic dependent launch: the kernel may be scheduled while its predecessor in the stream is still running
// (its CTAs move in as the predecessor's retire, the launch latency is hidden); it must execute pdl_wait() before it
// touches anything the predecessor wrote.  SRB_PDL=0 in the environment turns the attribute off.
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

// csrc/disc_mma.cu: warp-MMA 3x3 convolution / weight gradient with explicit memory pitches (see there)
int dmma_conv_pitched(const void* in_bf16, const float* in_f32, int cin, const void* w_packed, const float* bias,
                      int cout, void* out_bf16, float* out_f32, const void* ref, int B, int Hs, int Ws, int Hd,
                      int Wd, int Hsp, int Wsp, int Hdp, int Wdp, int stride, int transposed, int act, float slope,
                      void* stream);
int dmma_wgrad_pitched(const void* gz_bf16, const float* gz_f32, int cout, const void* a_bf16, int cin, float* dw,
                       float* db, int B, int Hi, int Wi, int Hip, int Wip, int Ho, int Wo, int stride, void* stream);

// csrc/conv_out_tc.cu: conv3.2 (64 -> 1 channel) forward + head and its fused backward, taps folded into the GEMM.
// c32_fwd_launch returns -1 when the geometry does not fit (the caller keeps the general path).
int c32_fwd_launch(const void* x, int B, int H, int W, const float* w, const float* bias, float inv_multiplier,
                   float power, float thres, int training, float* out, float* srs, float* z, cudaStream_t st);
// discriminator image end, input gradient (dl1_dgrad) on the tap-partial scheme; -1: geometry not covered
int d16_dgrad_launch(const void* gz, const float* w, float* gx, int C, int B, int H, int W, cudaStream_t st);
int c32_bwd_launch(const float* gz, const float* w, const void* act, void* gy, float* dw, float* db, int B, int H, int W,
                   float slope, cudaStream_t st);

}  // namespace srb
