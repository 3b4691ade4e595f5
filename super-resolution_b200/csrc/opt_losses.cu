// Optional generator loss terms of esrgan.train (all behind lambda = 0 in options/default.json):
//   hist  esrgan.py:538-545  DiffableHistogram (models.py:308-342) of the positive pixels of generated / ground truth,
//                            compared with KLD_hist (utils.py:90-113)
//   nnz   esrgan.py:530-533  mse( sum softgreater(gen, 0, 50000), count(gt > 0) ) per image (utils.py:259-261)
//   mask  esrgan.py:534-537  L1( nnz_mask(gen), nnz_mask(gt) ), nnz_mask = sigmoid(5e4 x) (utils.py:271-272)
//   hit   esrgan.py:546-550  mse of the "hitograms" (utils.py:264-268): mean soft hit indicator per position inside the
//                            factor x factor cells
// Each is a fused reduction that also produces d loss / d generated (added into the generator step's gradient image),
// so none of them builds an autograd graph or materialises the [bins x pixels] tensor the reference's histogram does.
// HBM-bound: one or two reads of the batch.  (The sorted-Wasserstein / Sinkhorn term of esrgan.py:539-542 lives in
// csrc/sinkhorn.cu.)
#include "common.cuh"

namespace srb {

constexpr int kHistMaxBins = 64;

__device__ __forceinline__ float sigm(float z) { return 1.f / (1.f + expf(-z)); }

__device__ __forceinline__ float block_sum_f(float v, float* red) {      // red: >= 32 floats of shared memory
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float s = 0.f;
  if (warp == 0) {
    s = lane < (int)((blockDim.x + 31) >> 5) ? red[lane] : 0.f;
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  }
  return s;                                                               // valid in thread 0
}

// ---------------------------------------------------------------- DiffableHistogram over the positive entries
// hist[j] += sum_{x_i > 0} sigmoid(sigma (x_i - c_j + d_j / 2)) - sigmoid(sigma (x_i - c_j - d_j / 2))
__global__ void __launch_bounds__(256)
soft_hist_kernel(const float* __restrict__ x, long long n, const float* __restrict__ centers,
                 const float* __restrict__ deltas, int nb, float sigma, float* __restrict__ hist) {
  __shared__ float s_c[kHistMaxBins], s_d[kHistMaxBins], s_h[kHistMaxBins];
  if (threadIdx.x < nb) { s_c[threadIdx.x] = centers[threadIdx.x]; s_d[threadIdx.x] = 0.5f * deltas[threadIdx.x]; s_h[threadIdx.x] = 0.f; }
  __syncthreads();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = __ldg(x + i);
    if (!(v > 0.f)) continue;                       // jet images are > 99 % zeros: almost every thread skips
    for (int j = 0; j < nb; ++j) {
      const float t = sigm(sigma * (v - s_c[j] + s_d[j])) - sigm(sigma * (v - s_c[j] - s_d[j]));
      if (t != 0.f) atomicAdd(&s_h[j], t);
    }
  }
  __syncthreads();
  if (threadIdx.x < nb && s_h[threadIdx.x] != 0.f) atomicAdd(hist + threadIdx.x, s_h[threadIdx.x]);
}

// KLD_hist.forward(q = generated histogram, p = ground-truth histogram), utils.py:103-113, and d loss / d q.
//   P_j = p_j b_j / N_p;  Q_j = (q_j + 1e-6) b_j / N_q  (N_q = sum q before the epsilon);  loss = sum_j P_j (log P_j - log Q_j) / mean(b)
__global__ void kld_hist_kernel(const float* __restrict__ q, const float* __restrict__ p, const float* __restrict__ bins,
                                int nb, float* __restrict__ loss, float* __restrict__ dq) {
  __shared__ float red[32];
  __shared__ float s_tot[4];
  const int j = threadIdx.x;
  const float qj = j < nb ? q[j] : 0.f, pj = j < nb ? p[j] : 0.f, bj = j < nb ? bins[j] : 0.f;
  float t;
  t = block_sum_f(pj, red); if (threadIdx.x == 0) s_tot[0] = t;
  t = block_sum_f(qj, red); if (threadIdx.x == 0) s_tot[1] = t;
  t = block_sum_f(bj, red); if (threadIdx.x == 0) s_tot[2] = t;
  __syncthreads();
  const float Np = s_tot[0], Nq = s_tot[1], bmean = s_tot[2] / nb;
  const float P = pj * bj / Np;
  const float logQ = logf((qj + 1e-6f) * bj / Nq);
  const float term = (j < nb && P > 0.f) ? P * (logf(P) - logQ) : 0.f;       // nn.KLDivLoss: 0 where the target is 0
  t = block_sum_f(term, red); if (threadIdx.x == 0) *loss = t / bmean;
  t = block_sum_f(j < nb ? P : 0.f, red); if (threadIdx.x == 0) s_tot[3] = t;
  __syncthreads();
  if (j < nb) dq[j] = (-P / (qj + 1e-6f) + s_tot[3] / Nq) / bmean;
}

// grad_i (+)= w * sum_j dq_j * sigma * (s'(sigma (x_i - c_j + d_j/2)) - s'(sigma (x_i - c_j - d_j/2))) for x_i > 0
__global__ void __launch_bounds__(256)
soft_hist_bwd_kernel(const float* __restrict__ x, long long n, const float* __restrict__ centers,
                     const float* __restrict__ deltas, int nb, float sigma, const float* __restrict__ dq, float w,
                     int accumulate, float* __restrict__ grad) {
  __shared__ float s_c[kHistMaxBins], s_d[kHistMaxBins], s_q[kHistMaxBins];
  if (threadIdx.x < nb) { s_c[threadIdx.x] = centers[threadIdx.x]; s_d[threadIdx.x] = 0.5f * deltas[threadIdx.x]; s_q[threadIdx.x] = dq[threadIdx.x]; }
  __syncthreads();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = __ldg(x + i);
    float g = 0.f;
    if (v > 0.f) {
      for (int j = 0; j < nb; ++j) {
        const float a = sigm(sigma * (v - s_c[j] + s_d[j])), b = sigm(sigma * (v - s_c[j] - s_d[j]));
        g += s_q[j] * sigma * (a * (1.f - a) - b * (1.f - b));
      }
      g *= w;
    }
    if (accumulate) { if (g != 0.f) grad[i] += g; }
    else grad[i] = g;
  }
}

// ---------------------------------------------------------------- nnz loss: one CTA per image, then a finishing pass
// s_b = sum_i sigmoid(sigma x_i), t_b = #(gt_i > 0);  loss = mean_b (s_b - t_b)^2;  d/dx_i = (2 / B)(s_b - t_b) sigma s'(sigma x_i)
__global__ void __launch_bounds__(256)
nnz_sum_kernel(const float* __restrict__ gen, const float* __restrict__ gt, int HW, float sigma, float* __restrict__ diff) {
  __shared__ float red[32];
  const float* g = gen + static_cast<size_t>(blockIdx.x) * HW;
  const float* t = gt + static_cast<size_t>(blockIdx.x) * HW;
  float s = 0.f, c = 0.f;
  for (int i = threadIdx.x; i < HW; i += 256) { s += sigm(sigma * __ldg(g + i)); c += __ldg(t + i) > 0.f ? 1.f : 0.f; }
  const float S = block_sum_f(s, red);
  const float Cn = block_sum_f(c, red);
  if (threadIdx.x == 0) diff[blockIdx.x] = S - Cn;
}
__global__ void __launch_bounds__(256)
nnz_finish_kernel(const float* __restrict__ gen, const float* __restrict__ diff, int B, int HW, float sigma, float w,
                  int accumulate, float* __restrict__ loss, float* __restrict__ grad) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    float a = 0.f;
    for (int b = 0; b < B; ++b) a += diff[b] * diff[b];
    *loss = a / B;
  }
  if (!grad || i >= 1LL * B * HW) return;
  const int b = int(i / HW);
  const float a = sigm(sigma * __ldg(gen + i));
  const float g = w * (2.f / B) * diff[b] * sigma * a * (1.f - a);
  grad[i] = accumulate ? grad[i] + g : g;
}

// ---------------------------------------------------------------- mask loss: mean |sigmoid(s g) - sigmoid(s t)|
__global__ void __launch_bounds__(256)
mask_loss_kernel(const float* __restrict__ gen, const float* __restrict__ gt, long long n, float sigma, float w,
                 int accumulate, float* __restrict__ acc, float* __restrict__ grad) {
  __shared__ float red[32];
  const long long stride = (long long)gridDim.x * blockDim.x;
  float a = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) {
    const float mg = sigm(sigma * __ldg(gen + i)), mt = sigm(sigma * __ldg(gt + i));
    const float d = mg - mt;
    a += fabsf(d);
    if (grad) {
      const float g = w * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) * sigma * mg * (1.f - mg) / float(n);
      grad[i] = accumulate ? grad[i] + g : g;
    }
  }
  a = block_sum_f(a, red);
  if (threadIdx.x == 0) atomicAdd(acc, a / float(n));
}

// ---------------------------------------------------------------- hitogram loss
// hit[(y % f), (x % f)] = mean over images and cells of sigmoid(sig (pixel - thr));  loss = mean_{f x f} (hit_gen - hit_gt)^2
__global__ void __launch_bounds__(256)
hit_sum_kernel(const float* __restrict__ gen, const float* __restrict__ gt, int B, int H, int W, int f, float thr,
               float sig, float* __restrict__ work) {                    // work[0 .. f*f): gen sums, [f*f .. 2 f*f): gt sums
  extern __shared__ float sh[];
  for (int i = threadIdx.x; i < 2 * f * f; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  const long long n = 1LL * B * H * W, stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) {
    const int x = int(i % W), y = int((i / W) % H);
    const int c = (y % f) * f + (x % f);
    atomicAdd(&sh[c], sigm(sig * (__ldg(gen + i) - thr)));
    atomicAdd(&sh[f * f + c], sigm(sig * (__ldg(gt + i) - thr)));
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * f * f; i += blockDim.x) atomicAdd(work + i, sh[i]);
}
__global__ void __launch_bounds__(256)
hit_finish_kernel(const float* __restrict__ gen, const float* __restrict__ work, int B, int H, int W, int f, float thr,
                  float sig, float w, int accumulate, float* __restrict__ loss, float* __restrict__ grad) {
  const float cells = float(B) * (H / f) * (W / f);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    float a = 0.f;
    for (int c = 0; c < f * f; ++c) { const float d = (work[c] - work[f * f + c]) / cells; a += d * d; }
    *loss = a / (f * f);
  }
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (!grad || i >= 1LL * B * H * W) return;
  const int x = int(i % W), y = int((i / W) % H);
  const int c = (y % f) * f + (x % f);
  const float d = (work[c] - work[f * f + c]) / cells;
  const float a = sigm(sig * (__ldg(gen + i) - thr));
  const float g = w * (2.f / (f * f)) * d * sig * a * (1.f - a) / cells;
  grad[i] = accumulate ? grad[i] + g : g;
}

static int grid_for(long long n, int block) {
  long long b = (n + block - 1) / block;
  const long long cap = 16LL * sm_count();
  return int(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace srb

using namespace srb;
#define SRB_OL_CHECK(name)                                             \
  do {                                                                 \
    cudaError_t e_ = cudaGetLastError();                               \
    if (e_ != cudaSuccess) return set_cuda_error(name " launch", e_);  \
  } while (0)

extern "C" int srb_soft_hist(const float* x, long long n, const float* centers, const float* deltas, int nbins,
                             float sigma, float* hist, void* stream) {
  if (!x || !centers || !deltas || !hist) return set_error("soft_hist: null pointer");
  if (nbins < 1 || nbins > kHistMaxBins || n <= 0) return set_error("soft_hist: 1..64 bins, n > 0");
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(hist, 0, sizeof(float) * nbins, st);
  soft_hist_kernel<<<grid_for(n, 256), 256, 0, st>>>(x, n, centers, deltas, nbins, sigma, hist);
  SRB_OL_CHECK("soft_hist");
  return 0;
}

extern "C" int srb_kld_hist(const float* q_hist, const float* p_hist, const float* binsizes, int nbins, float* loss,
                            float* dq, void* stream) {
  if (!q_hist || !p_hist || !binsizes || !loss || !dq) return set_error("kld_hist: null pointer");
  if (nbins < 1 || nbins > kHistMaxBins) return set_error("kld_hist: 1..64 bins");
  kld_hist_kernel<<<1, 64, 0, (cudaStream_t)stream>>>(q_hist, p_hist, binsizes, nbins, loss, dq);
  SRB_OL_CHECK("kld_hist");
  return 0;
}

extern "C" int srb_soft_hist_bwd(const float* x, long long n, const float* centers, const float* deltas, int nbins,
                                 float sigma, const float* dq, float w, int accumulate, float* grad, void* stream) {
  if (!x || !centers || !deltas || !dq || !grad) return set_error("soft_hist_bwd: null pointer");
  if (nbins < 1 || nbins > kHistMaxBins || n <= 0) return set_error("soft_hist_bwd: 1..64 bins, n > 0");
  soft_hist_bwd_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(x, n, centers, deltas, nbins, sigma, dq, w,
                                                                          accumulate, grad);
  SRB_OL_CHECK("soft_hist_bwd");
  return 0;
}

extern "C" int srb_nnz_loss(const float* gen, const float* gt, int B, int HW, float sigma, float w, int accumulate,
                            float* per_image, float* loss, float* grad, void* stream) {
  if (!gen || !gt || !per_image || !loss) return set_error("nnz_loss: null pointer");
  if (B <= 0 || HW <= 0) return set_error("nnz_loss: bad sizes");
  cudaStream_t st = (cudaStream_t)stream;
  nnz_sum_kernel<<<B, 256, 0, st>>>(gen, gt, HW, sigma, per_image);
  SRB_OL_CHECK("nnz_sum");
  nnz_finish_kernel<<<ceil_div(1LL * B * HW, 256), 256, 0, st>>>(gen, per_image, B, HW, sigma, w, accumulate, loss, grad);
  SRB_OL_CHECK("nnz_finish");
  return 0;
}

extern "C" int srb_mask_loss(const float* gen, const float* gt, long long n, float sigma, float w, int accumulate,
                             float* loss, float* grad, void* stream) {
  if (!gen || !gt || !loss) return set_error("mask_loss: null pointer");
  if (n <= 0) return set_error("mask_loss: bad size");
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(loss, 0, sizeof(float), st);
  mask_loss_kernel<<<grid_for(n, 256), 256, 0, st>>>(gen, gt, n, sigma, w, accumulate, loss, grad);
  SRB_OL_CHECK("mask_loss");
  return 0;
}

extern "C" int srb_hit_loss(const float* gen, const float* gt, int B, int H, int W, int factor, float thr, float sig,
                            float w, int accumulate, float* work, float* loss, float* grad, void* stream) {
  if (!gen || !gt || !work || !loss) return set_error("hit_loss: null pointer");
  if (B <= 0 || factor <= 0 || H % factor || W % factor || factor > 32) return set_error("hit_loss: bad sizes");
  if (!(sig > 0.f)) return set_error("hit_loss: the soft hitogram needs sig > 0 (utils.py:265)");
  cudaStream_t st = (cudaStream_t)stream;
  const int ff = factor * factor;
  cudaMemsetAsync(work, 0, sizeof(float) * 2 * ff, st);
  const long long n = 1LL * B * H * W;
  hit_sum_kernel<<<grid_for(n, 256), 256, sizeof(float) * 2 * ff, st>>>(gen, gt, B, H, W, factor, thr, sig, work);
  SRB_OL_CHECK("hit_sum");
  hit_finish_kernel<<<ceil_div(n, 256), 256, 0, st>>>(gen, work, B, H, W, factor, thr, sig, w, accumulate, loss, grad);
  SRB_OL_CHECK("hit_finish");
  return 0;
}
