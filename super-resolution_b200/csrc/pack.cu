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

// Table-driven form: one launch re-packs every convolution of a model (608 items for the default generator).
// blocks are assigned to items through an exclusive prefix of per-item block counts (binary search).
struct PackItem {
  const float* w; __nv_bfloat16* wp;
  int cout, cin, ci_off, n_out;
  int flag;      // fwd: shuffle; dgrad: unshuffle
  int kind;      // 0 forward pack, 1 dgrad pack
  int kb_dst, kb_total;
  float scale; int pad;
};

// One thread per (output channel, input channel) pair: it reads the pair's nine taps (36 contiguous bytes of the OIHW
// tensor) and writes one bf16 into each of the nine tap planes, where a warp covers 32 consecutive k (64 contiguous
// bytes).  The element-per-thread mapping it replaces read the fp32 weights with a 36-byte stride (9x sector traffic).
__global__ void pack_batch_kernel(const PackItem* __restrict__ items, const int* __restrict__ prefix, int n_items) {
  int lo = 0, hi = n_items - 1;
  while (lo < hi) {                               // last item whose first block <= blockIdx.x
    const int mid = (lo + hi + 1) >> 1;
    if (prefix[mid] <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const PackItem it = items[lo];
  const long long idx = (blockIdx.x - prefix[lo]) * (long long)blockDim.x + threadIdx.x;
  if (it.kind == 0) {
    const long long pairs = 1LL * it.cout * it.cin;
    if (idx >= pairs) return;
    const int k = int(idx % 64);
    long long r = idx / 64;
    const int n = int(r % it.cout);
    const int kb = int(r / it.cout);
    const int KB = it.cin / 64;
    const int C = it.cout / 4;
    const int co = it.flag ? (4 * (n % C) + n / C) : n;
    const float* src = it.w + ((size_t)co * it.cin + kb * 64 + k) * 9;
#pragma unroll
    for (int t = 0; t < 9; ++t)
      it.wp[(((size_t)t * KB + kb) * it.cout + n) * 64 + k] = __float2bfloat16(it.scale * __ldg(src + t));
  } else {
    const long long pairs = 1LL * it.cout * it.n_out;
    if (idx >= pairs) return;
    const int k = int(idx % 64);
    long long r = idx / 64;
    const int n = int(r % it.n_out);
    const int kb = int(r / it.n_out);
    const int kk = kb * 64 + k;
    const int C = it.cout / 4;
    const int co = it.flag ? (4 * (kk % C) + kk / C) : kk;
    const float* src = it.w + ((size_t)co * it.cin + it.ci_off + n) * 9;
#pragma unroll
    for (int t = 0; t < 9; ++t)
      it.wp[(((size_t)t * it.kb_total + it.kb_dst + kb) * it.n_out + n) * 64 + k] =
          __float2bfloat16(it.scale * __ldg(src + (8 - t)));
  }
}

}  // namespace srb

extern "C" int srb_pack_batch(const void* items_dev, const int* block_prefix_dev, int n_items, int total_blocks,
                              void* stream) {
  using namespace srb;
  if (!items_dev || !block_prefix_dev || n_items < 1 || total_blocks < 1) return set_error("pack_batch: bad arguments");
  pack_batch_kernel<<<total_blocks, 256, 0, (cudaStream_t)stream>>>((const PackItem*)items_dev, block_prefix_dev, n_items);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : set_cuda_error("pack_batch launch", e);
}

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
