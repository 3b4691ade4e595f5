// Evaluation post-processing: per-image statistics of a batch of (super-resolved) jet images in
// ONE pass on the device, instead of `.cpu().numpy()` plus a Python loop per image per statistic.
//
// Replaces, per batch, the collectors of evaluation/eval.py:323-335 (modes max, min, mean, nnz, sum, meannnz) and
// get_nth_hardest (eval.py:98-105: the n-th largest pixel of every image, duplicates counted, n = 1..nmax).
// One CTA per image; HBM-bound: H*W*4 bytes read per image, (6 + nmax)*4 written.
//   out[b] = { max, min, mean, nnz(> thres), sum, mean of the pixels > thres (0 if none), 1st, 2nd, ... nmax-th hardest }
// max / min / nnz / n-th hardest are exact; sum and the means are accumulated in fp64 and rounded once.
#include "common.cuh"

namespace srb {

constexpr int kStatsTop = 8;      // largest supported nmax
constexpr int kStatsFixed = 6;

__global__ void __launch_bounds__(256)
image_stats_kernel(const float* __restrict__ img, int HW, float thres, int nmax, float* __restrict__ out) {
  __shared__ float s_f[8][2];
  __shared__ double s_d[8][2];
  __shared__ int s_i[8];
  __shared__ float s_best[8];
  __shared__ int s_owner[8];
  const float* x = img + static_cast<size_t>(blockIdx.x) * HW;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float mx = -INFINITY, mn = INFINITY;
  double sum = 0.0, sum_nz = 0.0;
  int nnz = 0;
  float top[kStatsTop];                        // this thread's largest values, descending
#pragma unroll
  for (int k = 0; k < kStatsTop; ++k) top[k] = -INFINITY;
  for (int i = tid; i < HW; i += 256) {
    const float v = __ldg(x + i);
    mx = fmaxf(mx, v); mn = fminf(mn, v);
    sum += v;
    if (v > thres) { ++nnz; sum_nz += v; }
    if (v > top[kStatsTop - 1]) {              // insertion into the sorted list
      float c = v;
#pragma unroll
      for (int k = 0; k < kStatsTop; ++k) {
        const float t = top[k];
        if (c > t) { top[k] = c; c = t; }
      }
    }
  }
  // block reductions of the scalar statistics
  for (int o = 16; o; o >>= 1) {
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
    sum_nz += __shfl_xor_sync(0xffffffffu, sum_nz, o);
    nnz += __shfl_xor_sync(0xffffffffu, nnz, o);
  }
  if (lane == 0) { s_f[warp][0] = mx; s_f[warp][1] = mn; s_d[warp][0] = sum; s_d[warp][1] = sum_nz; s_i[warp] = nnz; }
  __syncthreads();
  float* o_ = out + static_cast<size_t>(blockIdx.x) * (kStatsFixed + nmax);
  if (tid == 0) {
    for (int w = 1; w < 8; ++w) {
      mx = fmaxf(mx, s_f[w][0]); mn = fminf(mn, s_f[w][1]); sum += s_d[w][0]; sum_nz += s_d[w][1]; nnz += s_i[w];
    }
    o_[0] = mx; o_[1] = mn; o_[2] = static_cast<float>(sum / HW); o_[3] = static_cast<float>(nnz);
    o_[4] = static_cast<float>(sum); o_[5] = nnz ? static_cast<float>(sum_nz / nnz) : 0.f;
  }
  // n-th hardest: nmax rounds of "largest head of all per-thread lists", the owner then advances its list
  int head = 0;
  for (int k = 0; k < nmax; ++k) {
    float cand = -INFINITY;
#pragma unroll
    for (int j = 0; j < kStatsTop; ++j) if (j == head) cand = top[j];
    float best = cand; int owner = tid;
    for (int o = 16; o; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oo = __shfl_xor_sync(0xffffffffu, owner, o);
      if (ob > best || (ob == best && oo < owner)) { best = ob; owner = oo; }
    }
    __syncthreads();                           // previous round's s_best / s_owner fully consumed
    if (lane == 0) { s_best[warp] = best; s_owner[warp] = owner; }
    __syncthreads();
    best = s_best[0]; owner = s_owner[0];
    for (int w = 1; w < 8; ++w)
      if (s_best[w] > best || (s_best[w] == best && s_owner[w] < owner)) { best = s_best[w]; owner = s_owner[w]; }
    if (tid == owner) ++head;
    if (tid == 0) o_[kStatsFixed + k] = best;
  }
}

}  // namespace srb

extern "C" int srb_image_stats(const float* img, int B, int HW, float thres, int nmax, float* out, void* stream) {
  using namespace srb;
  if (!img || !out) return set_error("image_stats: null pointer");
  if (B <= 0 || HW <= 0) return set_error("image_stats: bad sizes");
  if (nmax < 0 || nmax > kStatsTop) return set_error("image_stats: nmax must be 0..8");
  image_stats_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(img, HW, thres, nmax, out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("image_stats launch", e);
  return 0;
}
