// K6/K7/K8: loss reductions of the GAN step and the fused Adam update.  All bandwidth/latency-bound:
// coalesced, vectorised where the data is large (Adam: 28 B/param), warp-shuffle reductions for the scalars.
//   srb_adam_step        torch.optim.Adam as used at esrgan.py:299,305,555,624 (coupled L2, no amsgrad), one launch
//                        over a flat parameter buffer, with a device-side skip predicate replacing the host sync of
//                        esrgan.py:623 (`if loss_D.item() > d_threshold`).
//   srb_pixel_losses     esrgan.py:487,490 (+ SumPool2d of 471): L1 on the batch-mean image and L1 on the pooled LR
//                        image, and their gradient w.r.t. the generated image, in three small kernels.
//   srb_ragan            esrgan.py:507-508 / 578-588: relativistic-average BCE-with-logits, loss and both gradients.
//   srb_gp_fwd / _bwd    esrgan.py:604-605: ((||g_b||_2 - 1)^2).mean() * lambda/2 and its gradient.
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace srb {

__device__ __forceinline__ float warp_sum(float v) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// block-wide sum, result valid in thread 0 (blockDim.x <= 1024)
__device__ __forceinline__ float block_sum(float v) {
  __shared__ float red[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float s = 0.f;
  if (warp == 0) {
    s = lane < (int)((blockDim.x + 31) >> 5) ? red[lane] : 0.f;
    s = warp_sum(s);
  }
  return s;
}

// ------------------------------------------------------------------ Adam
// state[0] = step count (float), advanced only when the update is applied; state[1] = skip flag of this step;
// state[2] = learning rate, used when the host passes lr < 0 (so a captured CUDA graph follows lr schedules:
// esrgan.py:308-310,666-667 lower lr_g / lr_d with ReduceLROnPlateau).
__global__ void adam_tick_kernel(float* state, const float* gate_value, float gate_threshold) {
  const bool skip = gate_value && !(*gate_value > gate_threshold);
  state[1] = skip ? 1.f : 0.f;
  if (!skip) state[0] += 1.f;
}

__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, long long n, float lr, float b1, float b2, float eps, float wd,
                            float grad_scale, const float* __restrict__ state) {
  if (state[1] != 0.f) return;                       // gated off (discriminator already "too good")
  const float step = state[0];
  if (lr < 0.f) lr = state[2];
  const float bc1 = 1.f - powf(b1, step), bc2 = 1.f - powf(b2, step);
  const float step_size = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* pa = &pp.x; const float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float gr = ga[k] * grad_scale;
      if (wd != 0.f) gr += wd * pa[k];
      ma[k] = b1 * ma[k] + (1.f - b1) * gr;
      va[k] = b2 * va[k] + (1.f - b2) * gr * gr;
      pa[k] -= step_size * ma[k] / (sqrtf(va[k]) * inv_sqrt_bc2 + eps);
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  // tail
  for (long long i = (n4 << 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) {
    float gr = g[i] * grad_scale;
    if (wd != 0.f) gr += wd * p[i];
    const float mi = b1 * m[i] + (1.f - b1) * gr;
    const float vi = b2 * v[i] + (1.f - b2) * gr * gr;
    m[i] = mi; v[i] = vi;
    p[i] -= step_size * mi / (sqrtf(vi) * inv_sqrt_bc2 + eps);
  }
}

// ------------------------------------------------------------------ pixel losses
// A: per HR pixel p, m[p] = mean_b gen - mean_b gt; accumulates sum |m| into acc[0].
// H W pixels alone are 25 blocks at 80x80: the batch is split over the kPixMeanSplit blocks of a thread-block cluster
// (each sums its images in order, eight images' loads in flight per thread) and rank 0 adds the partial sums through
// distributed shared memory in rank order - deterministic, one launch.
constexpr int kPixMeanSplit = 8;
__global__ void __cluster_dims__(1, kPixMeanSplit, 1) __launch_bounds__(256)
pix_mean_kernel(const float* __restrict__ gen, const float* __restrict__ gt, int B, int HW,
                float* __restrict__ msign, float* __restrict__ acc) {
  __shared__ float part[256];
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank();
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int per = (B + kPixMeanSplit - 1) / kPixMeanSplit;
  const int b0 = rank * per, b1 = min(B, b0 + per);
  float s = 0.f;
  if (p < HW) {
#pragma unroll 8
    for (int b = b0; b < b1; ++b) s += gen[(size_t)b * HW + p] - gt[(size_t)b * HW + p];
  }
  part[threadIdx.x] = s;
  cluster.sync();
  float a = 0.f;
  if (rank == 0 && p < HW) {
    s = 0.f;
#pragma unroll
    for (int r = 0; r < kPixMeanSplit; ++r) s += cluster.map_shared_rank(part, r)[threadIdx.x];
    s /= B;
    msign[p] = s > 0.f ? 1.f : (s < 0.f ? -1.f : 0.f);
    a = fabsf(s);
  }
  cluster.sync();                                    // the other ranks' shared memory stays alive until rank 0 has read it
  if (rank != 0) return;
  a = block_sum(a);
  if (threadIdx.x == 0) atomicAdd(acc, a);
}
// B: per LR pixel, d = sum of k x k block of gen - lr; sign to dsign; sum |d| into acc[1]
// lr_power != 1 (opt.scaling_power, esrgan.py:472-473,490 for k = 1): d = pool^p - lr^p and dsign carries the whole
// derivative factor sign(d) * p * pool^(p-1) (torch.pow semantics, including its infinities at pool == 0 for p < 1).
__global__ void pix_lr_kernel(const float* __restrict__ gen, const float* __restrict__ lr, int B, int H, int W, int k,
                              float lr_power, float* __restrict__ dsign, float* __restrict__ acc) {
  const int h = H / k, w = W / k;
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  float a = 0.f;
  if (i < 1LL * B * h * w) {
    const int xo = int(i % w), yo = int((i / w) % h), b = int(i / (1LL * w * h));
    const float* g = gen + ((size_t)b * H + yo * k) * W + xo * k;
    float s = 0.f;
    for (int r = 0; r < k; ++r)
      for (int c = 0; c < k; ++c) s += g[r * W + c];
    float dfac = 1.f;
    if (lr_power != 1.f) {
      dfac = lr_power * powf(s, lr_power - 1.f);
      s = powf(s, lr_power) - powf(lr[i], lr_power);
    } else {
      s -= lr[i];
    }
    dsign[i] = (s > 0.f ? 1.f : (s < 0.f ? -1.f : 0.f)) * dfac;
    a = fabsf(s);
  }
  a = block_sum(a);
  if (threadIdx.x == 0) atomicAdd(acc + 1, a);
}
// C: grad[b,p] (+)= c_hr * msign[p] + c_lr * dsign[b, p/k]
__global__ void pix_grad_kernel(const float* __restrict__ msign, const float* __restrict__ dsign, int B, int H, int W,
                                int k, float c_hr, float c_lr, int accumulate, float* __restrict__ grad) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= 1LL * B * H * W) return;
  const int x = int(i % W), y = int((i / W) % H), b = int(i / (1LL * W * H));
  const int h = H / k, w = W / k;
  float g = c_hr * msign[y * W + x];
  if (y / k < h && x / k < w) g += c_lr * dsign[((size_t)b * h + y / k) * w + x / k];
  grad[i] = accumulate ? grad[i] + g : g;
}

// ------------------------------------------------------------------ relativistic average GAN loss
// L = 0.5 * [ mean softplus(-(eps + u - mean_b v)) + mean softplus(eps + v - mean_b u) ];  one block.
__device__ __forceinline__ float softplus(float x) { return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x))); }
__device__ __forceinline__ float sigmoidf(float x) { return 1.f / (1.f + expf(-x)); }

__global__ void ragan_kernel(const float* __restrict__ u, const float* __restrict__ v, int B, int n, float eps,
                             float* __restrict__ loss, float* __restrict__ gu, float* __restrict__ gv) {
  extern __shared__ float sm[];           // mean_u[n], mean_v[n], s1[n] = sum_b sig(-x1), s2[n] = sum_b sig(x2)
  float* mu = sm; float* mv = sm + n; float* s1 = sm + 2 * n; float* s2 = sm + 3 * n;
  for (int j = threadIdx.x; j < n; j += blockDim.x) {
    float a = 0.f, c = 0.f;
    for (int b = 0; b < B; ++b) { a += u[(size_t)b * n + j]; c += v[(size_t)b * n + j]; }
    mu[j] = a / B; mv[j] = c / B;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < n; j += blockDim.x) {
    float a = 0.f, c = 0.f;
    for (int b = 0; b < B; ++b) {
      a += sigmoidf(-(eps + u[(size_t)b * n + j] - mv[j]));
      c += sigmoidf(eps + v[(size_t)b * n + j] - mu[j]);
    }
    s1[j] = a; s2[j] = c;
  }
  __syncthreads();
  const float scale = 0.5f / (float(B) * n);
  float acc = 0.f;
  for (int i = threadIdx.x; i < B * n; i += blockDim.x) {
    const int j = i % n;
    const float x1 = eps + u[i] - mv[j], x2 = eps + v[i] - mu[j];
    acc += softplus(-x1) + softplus(x2);
    if (gu) gu[i] = scale * (-sigmoidf(-x1) - s2[j] / B);
    if (gv) gv[i] = scale * (sigmoidf(x2) + s1[j] / B);
  }
  acc = block_sum(acc);
  if (threadIdx.x == 0) *loss = acc * scale;
}

// Same result with the batch reductions spread over the whole block: G = blockDim / n thread groups each sum a slice
// of the batch per patch position (n <= blockDim; the prediction map of the patch discriminator has n = 25).
__global__ void __launch_bounds__(1024)
ragan_wide_kernel(const float* __restrict__ u, const float* __restrict__ v, int B, int n, float eps,
                  float* __restrict__ loss, float* __restrict__ gu, float* __restrict__ gv) {
  extern __shared__ float sm[];           // mu[n], mv[n], s1[n], s2[n], part[2][G][n]
  const int G = blockDim.x / n;
  float* mu = sm; float* mv = sm + n; float* s1 = sm + 2 * n; float* s2 = sm + 3 * n;
  float* pa = sm + 4 * n; float* pc = pa + G * n;
  const int g = threadIdx.x / n, j = threadIdx.x - g * n;
  const bool act = g < G;
  float a = 0.f, c = 0.f;
  if (act) for (int b = g; b < B; b += G) { a += u[(size_t)b * n + j]; c += v[(size_t)b * n + j]; }
  if (act) { pa[g * n + j] = a; pc[g * n + j] = c; }
  __syncthreads();
  if (threadIdx.x < n) {
    float x = 0.f, y = 0.f;
    for (int k = 0; k < G; ++k) { x += pa[k * n + threadIdx.x]; y += pc[k * n + threadIdx.x]; }
    mu[threadIdx.x] = x / B; mv[threadIdx.x] = y / B;
  }
  __syncthreads();
  a = 0.f; c = 0.f;
  if (act) for (int b = g; b < B; b += G) {
    a += sigmoidf(-(eps + u[(size_t)b * n + j] - mv[j]));
    c += sigmoidf(eps + v[(size_t)b * n + j] - mu[j]);
  }
  if (act) { pa[g * n + j] = a; pc[g * n + j] = c; }
  __syncthreads();
  if (threadIdx.x < n) {
    float x = 0.f, y = 0.f;
    for (int k = 0; k < G; ++k) { x += pa[k * n + threadIdx.x]; y += pc[k * n + threadIdx.x]; }
    s1[threadIdx.x] = x; s2[threadIdx.x] = y;
  }
  __syncthreads();
  const float scale = 0.5f / (float(B) * n);
  float acc = 0.f;
  for (int i = threadIdx.x; i < B * n; i += blockDim.x) {
    const int jj = i % n;
    const float x1 = eps + u[i] - mv[jj], x2 = eps + v[i] - mu[jj];
    acc += softplus(-x1) + softplus(x2);
    if (gu) gu[i] = scale * (-sigmoidf(-x1) - s2[jj] / B);
    if (gv) gv[i] = scale * (sigmoidf(x2) + s1[jj] / B);
  }
  acc = block_sum(acc);
  if (threadIdx.x == 0) *loss = acc * scale;
}

// ------------------------------------------------------------------ gradient penalty
__global__ void gp_fwd_kernel(const float* __restrict__ g, int B, int n, float coef, float* __restrict__ norms,
                              float* __restrict__ loss) {
  const int b = blockIdx.x;
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { const float v = g[(size_t)b * n + i]; s += v * v; }
  s = block_sum(s);
  if (threadIdx.x == 0) {
    const float nr = sqrtf(s);
    norms[b] = nr;
    atomicAdd(loss, coef * (nr - 1.f) * (nr - 1.f) / B);
  }
}
__global__ void gp_bwd_kernel(const float* __restrict__ g, const float* __restrict__ norms, int B, int n, float coef,
                              const float* __restrict__ gloss, float* __restrict__ gg) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= 1LL * B * n) return;
  const float nr = norms[i / n];
  const float f = nr > 0.f ? 2.f * coef / B * (nr - 1.f) / nr : 0.f;
  gg[i] = (gloss ? *gloss : 1.f) * f * g[i];
}

}  // namespace srb

using namespace srb;
#define SRB_CHK(name)                                                  \
  do {                                                                 \
    cudaError_t e_ = cudaGetLastError();                               \
    if (e_ != cudaSuccess) return set_cuda_error(name " launch", e_);  \
  } while (0)

/* state: device float[2] = {step count, skip flag}.  gate_value: device scalar or NULL; the update (and the step
 * count) is skipped unless *gate_value > gate_threshold. */
extern "C" int srb_adam_step(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                             float eps, float weight_decay, float grad_scale, float* state, const float* gate_value,
                             float gate_threshold, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
       reinterpret_cast<uintptr_t>(v)) & 15)
    return set_error("adam: buffers must be 16-byte aligned");
  adam_tick_kernel<<<1, 1, 0, st>>>(state, gate_value, gate_threshold);
  SRB_CHK("adam_tick");
  long long blocks = (n / 4 + 255) / 256;
  const long long cap = 8LL * sm_count();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  adam_kernel<<<(int)blocks, 256, 0, st>>>(p, g, m, v, n, lr, b1, b2, eps, weight_decay, grad_scale, state);
  SRB_CHK("adam");
  return 0;
}

/* acc: device float[2], zeroed here; afterwards loss_hr = acc[0] / HW, loss_lr = acc[1] / (B*h*w).
 * grad (fp32 [B,1,H,W]) (+)= c_hr * d loss_hr + c_lr * d loss_lr, the 1/N factors already folded in. */
extern "C" int srb_pixel_losses(const float* gen, const float* gt, const float* lr, int B, int H, int W, int k,
                                float w_hr, float w_lr, int accumulate, float* msign, float* dsign, float* acc,
                                float* grad, void* stream) {
  return srb_pixel_losses_pow(gen, gt, lr, B, H, W, k, w_hr, w_lr, 1.f, accumulate, msign, dsign, acc, grad, stream);
}

extern "C" int srb_pixel_losses_pow(const float* gen, const float* gt, const float* lr, int B, int H, int W, int k,
                                    float w_hr, float w_lr, float lr_power, int accumulate, float* msign, float* dsign,
                                    float* acc, float* grad, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const int HW = H * W, h = H / k, w = W / k;
  cudaMemsetAsync(acc, 0, 2 * sizeof(float), st);
  pix_mean_kernel<<<dim3(ceil_div(HW, 256), kPixMeanSplit), 256, 0, st>>>(gen, gt, B, HW, msign, acc);
  SRB_CHK("pix_mean");
  pix_lr_kernel<<<ceil_div(1LL * B * h * w, 256), 256, 0, st>>>(gen, lr, B, H, W, k, lr_power, dsign, acc);
  SRB_CHK("pix_lr");
  if (grad) {
    pix_grad_kernel<<<ceil_div(1LL * B * HW, 256), 256, 0, st>>>(msign, dsign, B, H, W, k, w_hr / (float(B) * HW),
                                                                w_lr / (float(B) * h * w), accumulate, grad);
    SRB_CHK("pix_grad");
  }
  return 0;
}

extern "C" int srb_ragan(const float* u, const float* v, int B, int n, float eps, float* loss, float* gu, float* gv,
                         void* stream) {
  if (n <= 512) {
    const int G = 1024 / n;
    ragan_wide_kernel<<<1, 1024, (4 * n + 2 * G * n) * sizeof(float), (cudaStream_t)stream>>>(u, v, B, n, eps, loss, gu, gv);
  } else {
    ragan_kernel<<<1, 256, 4 * n * sizeof(float), (cudaStream_t)stream>>>(u, v, B, n, eps, loss, gu, gv);
  }
  SRB_CHK("ragan");
  return 0;
}

extern "C" int srb_gp_fwd(const float* g, int B, int n, float coef, float* norms, float* loss, void* stream) {
  cudaMemsetAsync(loss, 0, sizeof(float), (cudaStream_t)stream);
  gp_fwd_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(g, B, n, coef, norms, loss);
  SRB_CHK("gp_fwd");
  return 0;
}
extern "C" int srb_gp_bwd(const float* g, const float* norms, int B, int n, float coef, const float* gloss, float* gg,
                          void* stream) {
  gp_bwd_kernel<<<ceil_div(1LL * B * n, 256), 256, 0, (cudaStream_t)stream>>>(g, norms, B, n, coef, gloss, gg);
  SRB_CHK("gp_bwd");
  return 0;
}
