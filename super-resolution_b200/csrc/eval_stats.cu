// Evaluation post-processing: statistics of a batch of (super-resolved) jet images on the
// device, instead of `.cpu().numpy()` plus a Python loop per image per statistic.
//
//   srb_image_stats   per-image collectors of evaluation/eval.py:323-335 (modes max, min, mean, nnz, sum, meannnz) and
//                     get_nth_hardest (eval.py:98-105: the n-th largest pixel of every image, duplicates counted,
//                     n = 1..nmax <= 32: the default eval_modes reach E_30, options/default.json), plus the has-NaN
//                     flag the reference skips batches on (eval.py:319-321).  One CTA per image.
//   srb_l1_metrics    calculate_metrics, eval.py:477-485: per image mean |SumPool(gen) - lr| and mean |gen - hr|.
//   srb_eval_rasters  SumRaster.add (utils.py:275-296, mode 'hitogram') and MeanImage.add (utils.py:299-328, mode
//                     'meanimg') of one batch, accumulated into the caller's running arrays.
// All HBM-bound: one read of the batch.  max / min / nnz / n-th hardest / counts are exact; sums are accumulated in fp64
// and rounded once.
#include "common.cuh"

namespace srb {

constexpr int kStatsTopMax = 32;  // largest supported nmax
constexpr int kStatsFixed = 7;    // max, min, mean, nnz, sum, meannnz, has_nan

// The image is staged in shared memory once; the n-th hardest rounds then repeatedly take the block-wide maximum and
// strike it out, so any nmax works with any number of pixels per thread (round 1 kept 8 candidates per thread).
__global__ void __launch_bounds__(256)
image_stats_kernel(const float* __restrict__ img, int HW, float thres, int nmax, float* __restrict__ out) {
  extern __shared__ float s_img[];
  __shared__ float s_f[8][2];
  __shared__ double s_d[8][2];
  __shared__ int s_i[8][2];
  __shared__ float s_best[8];
  __shared__ int s_idx[8];
  const float* x = img + static_cast<size_t>(blockIdx.x) * HW;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float mx = -INFINITY, mn = INFINITY;
  double sum = 0.0, sum_nz = 0.0;
  int nnz = 0, nan = 0;
  for (int i = tid; i < HW; i += 256) {
    const float v = __ldg(x + i);
    s_img[i] = v;
    nan |= (v != v);
    mx = fmaxf(mx, v); mn = fminf(mn, v);
    sum += v;
    if (v > thres) { ++nnz; sum_nz += v; }
  }
  for (int o = 16; o; o >>= 1) {
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
    sum_nz += __shfl_xor_sync(0xffffffffu, sum_nz, o);
    nnz += __shfl_xor_sync(0xffffffffu, nnz, o);
    nan |= __shfl_xor_sync(0xffffffffu, nan, o);
  }
  if (lane == 0) {
    s_f[warp][0] = mx; s_f[warp][1] = mn; s_d[warp][0] = sum; s_d[warp][1] = sum_nz; s_i[warp][0] = nnz; s_i[warp][1] = nan;
  }
  __syncthreads();
  float* o_ = out + static_cast<size_t>(blockIdx.x) * (kStatsFixed + nmax);
  if (tid == 0) {
    for (int w = 1; w < 8; ++w) {
      mx = fmaxf(mx, s_f[w][0]); mn = fminf(mn, s_f[w][1]); sum += s_d[w][0]; sum_nz += s_d[w][1];
      nnz += s_i[w][0]; nan |= s_i[w][1];
    }
    o_[0] = mx; o_[1] = mn; o_[2] = static_cast<float>(sum / HW); o_[3] = static_cast<float>(nnz);
    o_[4] = static_cast<float>(sum); o_[5] = nnz ? static_cast<float>(sum_nz / nnz) : 0.f;
    o_[6] = nan ? 1.f : 0.f;
  }
  // n-th hardest: nmax rounds of block-wide (max, lowest index); the winner is struck out of the staged image
  for (int k = 0; k < nmax; ++k) {
    float best = -INFINITY; int idx = 0x7fffffff;
    for (int i = tid; i < HW; i += 256) {
      const float v = s_img[i];
      if (v > best) { best = v; idx = i; }             // ascending i: ties keep the lowest index
    }
    for (int o = 16; o; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
      if (ob > best || (ob == best && oi < idx)) { best = ob; idx = oi; }
    }
    if (lane == 0) { s_best[warp] = best; s_idx[warp] = idx; }
    __syncthreads();
    best = s_best[0]; idx = s_idx[0];
    for (int w = 1; w < 8; ++w)
      if (s_best[w] > best || (s_best[w] == best && s_idx[w] < idx)) { best = s_best[w]; idx = s_idx[w]; }
    __syncthreads();                                   // everyone has read s_best / s_idx before the next round
    if (tid == 0) {
      o_[kStatsFixed + k] = best;
      if (idx < HW) s_img[idx] = -INFINITY;
    }
    __syncthreads();
  }
}

// One CTA per image: mean |SumPool_k(gen) - lr| over the LR pixels and mean |gen - hr| over the HR pixels.
__global__ void __launch_bounds__(256)
l1_metrics_kernel(const float* __restrict__ gen, const float* __restrict__ hr, const float* __restrict__ lr, int H, int W,
                  int k, float* __restrict__ out) {
  __shared__ double s_d[8][2];
  const int HW = H * W, h = H / k, w = W / k;
  const float* g = gen + static_cast<size_t>(blockIdx.x) * HW;
  const float* t = hr + static_cast<size_t>(blockIdx.x) * HW;
  const float* l = lr + static_cast<size_t>(blockIdx.x) * h * w;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double a_hr = 0.0, a_lr = 0.0;
  for (int i = tid; i < HW; i += 256) a_hr += fabsf(__ldg(g + i) - __ldg(t + i));
  for (int i = tid; i < h * w; i += 256) {
    const int y = i / w, x = i - y * w;
    float s = 0.f;
    for (int dy = 0; dy < k; ++dy)
      for (int dx = 0; dx < k; ++dx) s += __ldg(g + (y * k + dy) * W + x * k + dx);
    a_lr += fabsf(s - __ldg(l + i));
  }
  for (int o = 16; o; o >>= 1) {
    a_hr += __shfl_xor_sync(0xffffffffu, a_hr, o);
    a_lr += __shfl_xor_sync(0xffffffffu, a_lr, o);
  }
  if (lane == 0) { s_d[warp][0] = a_hr; s_d[warp][1] = a_lr; }
  __syncthreads();
  if (tid == 0) {
    for (int w_ = 1; w_ < 8; ++w_) { a_hr += s_d[w_][0]; a_lr += s_d[w_][1]; }
    out[2 * blockIdx.x] = static_cast<float>(a_lr / (h * w));
    out[2 * blockIdx.x + 1] = static_cast<float>(a_hr / HW);
  }
}

// One thread per pixel position, looping over the batch (coalesced across the warp):
//   cnt_r = #images with pixel > thr_raster  -> raster[(y % f) * f + (x % f)] += cnt_r               (SumRaster.add)
//   cnt_m = #images with pixel > thr_mean, clip = sum_b max(pixel, thr_mean)
//   mean_acc[p] += energy ? clip * cnt_m : cnt_m                                                    (MeanImage.add)
__global__ void __launch_bounds__(128)
eval_rasters_kernel(const float* __restrict__ img, int B, int H, int W, int f, float thr_raster, float thr_mean,
                    int energy, double* __restrict__ raster, double* __restrict__ mean_acc) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int HW = H * W;
  if (p >= HW) return;
  int cnt_r = 0, cnt_m = 0;
  double clip = 0.0;
  for (int b = 0; b < B; ++b) {
    const float v = __ldg(img + static_cast<size_t>(b) * HW + p);
    cnt_r += v > thr_raster;
    cnt_m += v > thr_mean;
    clip += fmaxf(v, thr_mean);
  }
  if (mean_acc) mean_acc[p] += energy ? clip * cnt_m : static_cast<double>(cnt_m);
  if (raster && cnt_r) {
    const int y = p / W, x = p - y * W;
    atomicAdd(raster + (y % f) * f + (x % f), static_cast<double>(cnt_r));
  }
}

}  // namespace srb

extern "C" int srb_image_stats(const float* img, int B, int HW, float thres, int nmax, float* out, void* stream) {
  using namespace srb;
  if (!img || !out) return set_error("image_stats: null pointer");
  if (B <= 0 || HW <= 0) return set_error("image_stats: bad sizes");
  if (nmax < 0 || nmax > kStatsTopMax) return set_error("image_stats: nmax must be 0..32");
  const size_t smem = static_cast<size_t>(HW) * 4;
  if (smem > 200 * 1024) return set_error("image_stats: image does not fit in shared memory (H*W <= 51200)");
  if (smem > 48 * 1024) ensure_max_smem(reinterpret_cast<const void*>(image_stats_kernel), 200 * 1024);
  image_stats_kernel<<<B, 256, smem, (cudaStream_t)stream>>>(img, HW, thres, nmax, out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("image_stats launch", e);
  return 0;
}

extern "C" int srb_l1_metrics(const float* gen, const float* hr, const float* lr, int B, int H, int W, int factor,
                              float* out, void* stream) {
  using namespace srb;
  if (!gen || !hr || !lr || !out) return set_error("l1_metrics: null pointer");
  if (B <= 0 || H <= 0 || W <= 0 || factor <= 0 || H % factor || W % factor) return set_error("l1_metrics: bad sizes");
  l1_metrics_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(gen, hr, lr, H, W, factor, out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("l1_metrics launch", e);
  return 0;
}

extern "C" int srb_eval_rasters(const float* img, int B, int H, int W, int factor, float thr_raster, float thr_mean,
                                int energy, double* raster, double* mean_acc, void* stream) {
  using namespace srb;
  if (!img || (!raster && !mean_acc)) return set_error("eval_rasters: null pointer");
  if (B <= 0 || H <= 0 || W <= 0 || factor <= 0) return set_error("eval_rasters: bad sizes");
  eval_rasters_kernel<<<ceil_div(1LL * H * W, 128), 128, 0, (cudaStream_t)stream>>>(img, B, H, W, factor, thr_raster,
                                                                                    thr_mean, energy, raster, mean_acc);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("eval_rasters launch", e);
  return 0;
}
