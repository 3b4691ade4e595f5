// Weight re-layout kernels: fp32 OIHW (the state_dict layout, models.py:19) -> bf16 K-major tiles
// [tap][k-block][out-channel][64] consumed by the implicit-GEMM kernels.  Bandwidth-bound, one element/thread.
#include "common.cuh"

namespace srb {

__global__ void pack_fwd_kernel(const float* __restrict__ w, int cout, int cin, int shuffle, float scale,
                                __nv_bfloat16* __restrict__ wp) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long total = 9LL * cout * cin;
  if (idx >= total) return;
  const int k = int(idx % 64);
  long long r = idx / 64;
  const int n = int(r % cout); r /= cout;
  const int KB = cin / 64;
  const int kb = int(r % KB);
  const int t = int(r / KB);
  const int C = cout / 4;
  const int co = shuffle ? (4 * (n % C) + n / C) : n;
  wp[idx] = __float2bfloat16(scale * w[((size_t)co * cin + kb * 64 + k) * 9 + t]);
}

__global__ void pack_dgrad_kernel(const float* __restrict__ w, int cout, int cin, int ci_off, int n_out,
                                  int unshuffle, float scale, __nv_bfloat16* __restrict__ wp, int kb_dst,
                                  int kb_total) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long total = 9LL * cout * n_out;
  if (idx >= total) return;
  const int k = int(idx % 64);
  long long r = idx / 64;
  const int n = int(r % n_out); r /= n_out;
  const int KBs = cout / 64;
  const int kb = int(r % KBs);
  const int t = int(r / KBs);
  const int kk = kb * 64 + k;                 // channel of the incoming gradient (packed order)
  const int C = cout / 4;
  const int co = unshuffle ? (4 * (kk % C) + kk / C) : kk;
  const float v = scale * w[((size_t)co * cin + ci_off + n) * 9 + (8 - t)];
  wp[(((size_t)t * kb_total + kb_dst + kb) * n_out + n) * 64 + k] = __float2bfloat16(v);
}

}  // namespace srb

extern "C" int srb_pack_conv3x3_fwd(const float* w, int cout, int cin, int shuffle, float scale, void* wp,
                                    void* stream) {
  using namespace srb;
  if (cin % 64 || cout % 64 || (shuffle && cout % 256)) return set_error("pack_fwd: channels must be multiples of 64");
  const long long total = 9LL * cout * cin;
  pack_fwd_kernel<<<ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(w, cout, cin, shuffle, scale,
                                                                          (__nv_bfloat16*)wp);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : set_cuda_error("pack_fwd launch", e);
}

extern "C" int srb_pack_conv3x3_dgrad(const float* w, int cout, int cin, int ci_off, int n_out, int unshuffle,
                                      float scale, void* wp, int kb_dst, int kb_total, void* stream) {
  using namespace srb;
  if (cout % 64 || n_out % 64 || ci_off + n_out > cin || kb_dst + cout / 64 > kb_total)
    return set_error("pack_dgrad: bad channel arguments");
  const long long total = 9LL * cout * n_out;
  pack_dgrad_kernel<<<ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(
      w, cout, cin, ci_off, n_out, unshuffle, scale, (__nv_bfloat16*)wp, kb_dst, kb_total);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : set_cuda_error("pack_dgrad launch", e);
}
