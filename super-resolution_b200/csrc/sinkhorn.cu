// Sorted-Wasserstein / Sinkhorn generator loss term of esrgan.train (lambda_wasser = 0 in
// options/default.json):
//   esrgan.py:539-542   gen_sort / real_sort = per-image ascending sort of the generated / true pixels;
//                       loss = WasserDist(cut_smaller(gen_sort)[..., None], cut_smaller(real_sort)[..., None])
//   utils.py:241-255    cut_smaller: keep the last P columns, P = the largest per-image count of entries > 0 (all of them
//                       when that count is 0)
//   utils.py:147-238    SinkhornDistance(eps, 100, 'sum'): C_ij = |x_i - y_j|^2, uniform marginals, log-domain updates
//                         u_i <- eps (log(mu + 1e-8) - LSE_j((-C_ij + u_i + v_j) / eps)) + u_i
//                         v_j <- eps (log(nu + 1e-8) - LSE_i((-C_ij + u_i + v_j) / eps)) + v_j
//                       until mean_b sum_i |u - u_prev| < 0.1 (a host sync per iteration in the reference), then
//                       sum_b sum_ij pi_ij C_ij, pi = exp((-C + u + v) / eps).
// Here: no cost matrix, no autograd graph, no host sync.  One CTA owns one image; C_ij is recomputed from the sorted
// values held in shared memory.  Every Sinkhorn iteration is one launch whose CTAs leave at once when the previous
// launch's error sum says "converged" (the stopping test is a batch mean, so it needs a grid-wide value: the launch
// boundary provides it, and the fixed launch count keeps the term capturable in a CUDA graph).  The duals of every
// iteration are kept ([iter][B][P]); the backward pass walks them in reverse with the analytic adjoint of the update
//   u' = eps log mu - eps LSE_j((v_j - C_ij) / eps),   v' = eps log nu - eps LSE_i((u'_i - C_ij) / eps)
// (the "+ u" / "+ v" of the reference cancels the dual's own contribution inside the LSE, so d u' / d u = 0):
//   T_ij = softmax_i((u'_i - C_ij) / eps), S_ij = softmax_j((v_j - C_ij) / eps)
//   gu_i <- gu'_i - sum_j gv'_j T_ij;  gC_ij = gv'_j T_ij + gu_i S_ij;  gv_j <- - sum_i gu_i S_ij;  gu <- 0
// and d loss / d x_i = sum_j gC_ij 2 (x_i - y_j), scattered back through the sort permutation.
#include <math.h>

#include "common.cuh"

namespace srb {

constexpr int kSinkThreads = 1024;
constexpr int kSinkMaxIter = 1024;

// workspace header (ints), followed by the float arrays laid out by sink_layout()
struct SinkLayout {
  size_t cnt, err, cost, xs, ys, xi, uh, vh, gu, gv, gx, total;     // byte offsets
};
static SinkLayout sink_layout(int B, int HW, int cap1, int cap2, int max_iter) {
  SinkLayout L{};
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t r = o; o += (bytes + 255) & ~size_t(255); return r; };
  L.cnt = take(8 * sizeof(int));                                     // [0] P1 count, [1] P2 count, [2] nits, [3] overflow
  L.err = take(sizeof(float) * max_iter);
  L.cost = take(sizeof(float) * B);
  L.xs = take(sizeof(float) * B * HW);
  L.ys = take(sizeof(float) * B * HW);
  L.xi = take(sizeof(int) * B * HW);
  L.uh = take(sizeof(float) * (size_t)(max_iter + 1) * B * cap1);
  L.vh = take(sizeof(float) * (size_t)(max_iter + 1) * B * cap2);
  L.gu = take(sizeof(float) * B * cap1);
  L.gv = take(sizeof(float) * B * cap2);
  L.gx = take(sizeof(float) * B * cap1);
  L.total = o;
  return L;
}

struct SinkP {
  int B, HW, cap1, cap2, max_iter;
  float eps, thresh;
  int* cnt; float* err; float* cost;
  float* xs; float* ys; int* xi;
  float* uh; float* vh; float* gu; float* gv; float* gx;
};

__device__ __forceinline__ float sk_warp_max(float v) {
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float sk_warp_sum(float v) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float sk_block_sum(float v, float* red) {   // result valid in every thread
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = sk_warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float s = lane < (int)(blockDim.x >> 5) ? red[lane] : 0.f;
  return sk_warp_sum(s);
}

// number of points of a cloud: the largest per-image count of positive entries, or everything when that is 0
__device__ __forceinline__ int sk_points(int count, int HW) { return count > 0 ? count : HW; }

// ---------------------------------------------------------------- per-image ascending sort (bitonic, shared memory)
// grid (B, 2): y == 0 sorts the generated image (values + source indices), y == 1 the ground truth (values only).
__global__ void __launch_bounds__(kSinkThreads)
sink_sort_kernel(const float* __restrict__ gen, const float* __restrict__ truth, SinkP p, int n2) {
  extern __shared__ __align__(16) unsigned char sk_smem[];
  float* key = reinterpret_cast<float*>(sk_smem);
  unsigned short* idx = reinterpret_cast<unsigned short*>(key + n2);
  const int b = blockIdx.x, which = blockIdx.y;
  const float* src = (which == 0 ? gen : truth) + static_cast<size_t>(b) * p.HW;
  int pos = 0;
  for (int i = threadIdx.x; i < n2; i += blockDim.x) {
    const float v = i < p.HW ? src[i] : INFINITY;
    key[i] = v;
    idx[i] = static_cast<unsigned short>(i);
    pos += (i < p.HW && v > 0.f) ? 1 : 0;
  }
  __shared__ float red[32];
  const int count = static_cast<int>(sk_block_sum(static_cast<float>(pos), red) + 0.5f);
  if (threadIdx.x == 0) atomicMax(&p.cnt[which], count);
  __syncthreads();
  for (int k = 2; k <= n2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n2; i += blockDim.x) {
        const int l = i ^ j;
        if (l > i) {
          const float a = key[i], c = key[l];
          const bool up = (i & k) == 0;
          if (up ? (a > c) : (a < c)) {
            key[i] = c; key[l] = a;
            const unsigned short t = idx[i]; idx[i] = idx[l]; idx[l] = t;
          }
        }
      }
      __syncthreads();
    }
  }
  float* dst = (which == 0 ? p.xs : p.ys) + static_cast<size_t>(b) * p.HW;
  for (int i = threadIdx.x; i < p.HW; i += blockDim.x) {
    dst[i] = key[i];
    if (which == 0) p.xi[static_cast<size_t>(b) * p.HW + i] = idx[i];
  }
}

// ---------------------------------------------------------------- shared-memory staging of one image's clouds
struct SinkTile {
  float *x, *a, *b1, *y, *v, *c, *d;      // P1-sized: x, a, b1;  P2-sized: y, v, c, d
  int P1, P2;
};
__device__ __forceinline__ bool sink_setup(const SinkP& p, unsigned char* smem, SinkTile& t, int n1, int n2) {
  t.P1 = sk_points(p.cnt[0], p.HW);
  t.P2 = sk_points(p.cnt[1], p.HW);
  if (t.P1 > p.cap1 || t.P2 > p.cap2) {
    if (blockIdx.x == 0 && threadIdx.x == 0) p.cnt[3] = 1;
    return false;
  }
  float* f = reinterpret_cast<float*>(smem);
  t.x = f; t.a = f + p.cap1; t.b1 = f + 2 * p.cap1;
  float* g = f + n1 * p.cap1;
  t.y = g; t.v = g + p.cap2; t.c = g + 2 * p.cap2; t.d = g + 3 * p.cap2;
  (void)n2;
  const int b = blockIdx.x;
  const float* xs = p.xs + static_cast<size_t>(b) * p.HW + (p.HW - t.P1);
  const float* ys = p.ys + static_cast<size_t>(b) * p.HW + (p.HW - t.P2);
  for (int i = threadIdx.x; i < t.P1; i += blockDim.x) t.x[i] = xs[i];
  for (int j = threadIdx.x; j < t.P2; j += blockDim.x) t.y[j] = ys[j];
  return true;
}

// ---------------------------------------------------------------- one Sinkhorn iteration (utils.py:196-204)
__global__ void __launch_bounds__(kSinkThreads)
sink_iter_kernel(SinkP p, int it) {
  extern __shared__ __align__(16) unsigned char sk_smem[];
  __shared__ float red[32];
  if (it > 0 && !(p.err[it - 1] / p.B >= p.thresh)) return;          // converged (or never ran): utils.py:203-204
  SinkTile t;
  if (!sink_setup(p, sk_smem, t, 2, 2)) return;
  const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  if (b == 0 && threadIdx.x == 0) p.cnt[2] = it + 1;
  const float* u_old = p.uh + (static_cast<size_t>(it) * p.B + b) * p.cap1;
  const float* v_old = p.vh + (static_cast<size_t>(it) * p.B + b) * p.cap2;
  float* u_new = p.uh + (static_cast<size_t>(it + 1) * p.B + b) * p.cap1;
  float* v_new = p.vh + (static_cast<size_t>(it + 1) * p.B + b) * p.cap2;
  float* u = t.a;
  for (int i = threadIdx.x; i < t.P1; i += blockDim.x) u[i] = u_old[i];
  for (int j = threadIdx.x; j < t.P2; j += blockDim.x) t.v[j] = v_old[j];
  __syncthreads();
  const float eps = p.eps;
  const float lmu = logf(1.0f / t.P1 + 1e-8f), lnu = logf(1.0f / t.P2 + 1e-8f);
  // u update: one thread per row, two passes over the columns
  float err = 0.f;
  for (int i = threadIdx.x; i < t.P1; i += blockDim.x) {
    const float xi = t.x[i], ui = u[i];
    float m = -INFINITY;
    for (int j = 0; j < t.P2; ++j) {
      const float dd = xi - t.y[j];
      m = fmaxf(m, (-(dd * dd) + ui + t.v[j]) / eps);
    }
    float s = 0.f;
    for (int j = 0; j < t.P2; ++j) {
      const float dd = xi - t.y[j];
      s += expf((-(dd * dd) + ui + t.v[j]) / eps - m);
    }
    const float un = eps * (lmu - (m + logf(s))) + ui;
    err += fabsf(un - ui);
    u_new[i] = un;
  }
  err = sk_block_sum(err, red);                                       // (includes the barrier that orders u reads/writes)
  if (threadIdx.x == 0) atomicAdd(&p.err[it], err);
  for (int i = threadIdx.x; i < t.P1; i += blockDim.x) u[i] = u_new[i];
  __syncthreads();
  // v update: one warp per column, lanes stride over the rows
  for (int j = warp; j < t.P2; j += nwarp) {
    const float yj = t.y[j], vj = t.v[j];
    float m = -INFINITY;
    for (int i = lane; i < t.P1; i += 32) {
      const float dd = t.x[i] - yj;
      m = fmaxf(m, (-(dd * dd) + u[i] + vj) / eps);
    }
    m = sk_warp_max(m);
    float s = 0.f;
    for (int i = lane; i < t.P1; i += 32) {
      const float dd = t.x[i] - yj;
      s += expf((-(dd * dd) + u[i] + vj) / eps - m);
    }
    s = sk_warp_sum(s);
    if (lane == 0) v_new[j] = eps * (lnu - (m + logf(s))) + vj;
  }
}

// ---------------------------------------------------------------- transport cost and the adjoints it starts
// cost_b = sum_ij pi_ij C_ij;  gu_i = sum_j pi C / eps;  gv_j = sum_i pi C / eps;  gx_i = sum_j pi (1 - C / eps) 2 (x_i - y_j)
__global__ void __launch_bounds__(kSinkThreads)
sink_cost_kernel(SinkP p, int want_grad) {
  extern __shared__ __align__(16) unsigned char sk_smem[];
  __shared__ float red[32];
  SinkTile t;
  if (!sink_setup(p, sk_smem, t, 2, 2)) {
    if (threadIdx.x == 0) p.cost[blockIdx.x] = NAN;
    return;
  }
  const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int nits = p.cnt[2];
  const float* U = p.uh + (static_cast<size_t>(nits) * p.B + b) * p.cap1;
  const float* V = p.vh + (static_cast<size_t>(nits) * p.B + b) * p.cap2;
  float* u = t.a;
  for (int i = threadIdx.x; i < t.P1; i += blockDim.x) u[i] = U[i];
  for (int j = threadIdx.x; j < t.P2; j += blockDim.x) t.v[j] = V[j];
  __syncthreads();
  const float eps = p.eps;
  float cost = 0.f;
  for (int i = threadIdx.x; i < t.P1; i += blockDim.x) {
    const float xi = t.x[i], ui = u[i];
    float gu = 0.f, gx = 0.f;
    for (int j = 0; j < t.P2; ++j) {
      const float dd = xi - t.y[j], c = dd * dd;
      const float pi = expf((-c + ui + t.v[j]) / eps);
      cost += pi * c;
      gu += pi * c;
      gx += pi * (1.f - c / eps) * 2.f * dd;
    }
    if (want_grad) {
      p.gu[static_cast<size_t>(b) * p.cap1 + i] = gu / eps;
      p.gx[static_cast<size_t>(b) * p.cap1 + i] = gx;
    }
  }
  cost = sk_block_sum(cost, red);
  if (threadIdx.x == 0) p.cost[b] = cost;
  if (!want_grad) return;
  for (int j = warp; j < t.P2; j += nwarp) {
    const float yj = t.y[j], vj = t.v[j];
    float g = 0.f;
    for (int i = lane; i < t.P1; i += 32) {
      const float dd = t.x[i] - yj, c = dd * dd;
      g += expf((-c + u[i] + vj) / eps) * c;
    }
    g = sk_warp_sum(g);
    if (lane == 0) p.gv[static_cast<size_t>(b) * p.cap2 + j] = g / eps;
  }
}

// ---------------------------------------------------------------- adjoint of one Sinkhorn iteration
__global__ void __launch_bounds__(kSinkThreads)
sink_bwd_kernel(SinkP p, int it) {
  extern __shared__ __align__(16) unsigned char sk_smem[];
  const int nits = p.cnt[2];
  if (it >= nits) return;
  SinkTile t;
  if (!sink_setup(p, sk_smem, t, 3, 4)) return;
  const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const float* un = p.uh + (static_cast<size_t>(it + 1) * p.B + b) * p.cap1;      // u' of this iteration
  const float* vo = p.vh + (static_cast<size_t>(it) * p.B + b) * p.cap2;          // v it started from
  float* gu_g = p.gu + static_cast<size_t>(b) * p.cap1;
  float* gv_g = p.gv + static_cast<size_t>(b) * p.cap2;
  float* gx_g = p.gx + static_cast<size_t>(b) * p.cap1;
  float* a = t.a;         // u', later the row LSE
  float* gut = t.b1;      // gu_i after the v' adjoint
  float* gv = t.c;        // gv'_j, later the new gv_j
  float* lc = t.d;        // column LSE
  const bool last = it == nits - 1;                       // only the final duals feed the cost directly
  for (int i = threadIdx.x; i < t.P1; i += blockDim.x) { a[i] = un[i]; gut[i] = last ? gu_g[i] : 0.f; }
  for (int j = threadIdx.x; j < t.P2; j += blockDim.x) { t.v[j] = vo[j]; gv[j] = gv_g[j]; }
  __syncthreads();
  const float eps = p.eps;
  // A: column LSE of (u'_i - C_ij) / eps
  for (int j = warp; j < t.P2; j += nwarp) {
    const float yj = t.y[j];
    float m = -INFINITY;
    for (int i = lane; i < t.P1; i += 32) {
      const float dd = t.x[i] - yj;
      m = fmaxf(m, (a[i] - dd * dd) / eps);
    }
    m = sk_warp_max(m);
    float s = 0.f;
    for (int i = lane; i < t.P1; i += 32) {
      const float dd = t.x[i] - yj;
      s += expf((a[i] - dd * dd) / eps - m);
    }
    s = sk_warp_sum(s);
    if (lane == 0) lc[j] = m + logf(s);
  }
  __syncthreads();
  // B + C: per row, the v' adjoint (T), then the u' adjoint (S)
  for (int i = threadIdx.x; i < t.P1; i += blockDim.x) {
    const float xi = t.x[i], ui = a[i];
    float acc = 0.f, gx = 0.f, m = -INFINITY;
    for (int j = 0; j < t.P2; ++j) {
      const float dd = xi - t.y[j], c = dd * dd;
      const float w = gv[j] * expf((ui - c) / eps - lc[j]);
      acc += w;
      gx += w * 2.f * dd;
      m = fmaxf(m, (t.v[j] - c) / eps);
    }
    const float g = gut[i] - acc;
    float s = 0.f, sx = 0.f;
    for (int j = 0; j < t.P2; ++j) {
      const float dd = xi - t.y[j], c = dd * dd;
      const float e = expf((t.v[j] - c) / eps - m);
      s += e;
      sx += e * 2.f * dd;
    }
    gx += g * sx / s;
    gx_g[i] += gx;
    gut[i] = g;
    a[i] = m + logf(s);                                   // row LSE (this thread's own element: no hazard)
  }
  __syncthreads();
  // D: gv_j = - sum_i gu_i S_ij
  for (int j = warp; j < t.P2; j += nwarp) {
    const float yj = t.y[j], vj = t.v[j];
    float g = 0.f;
    for (int i = lane; i < t.P1; i += 32) {
      const float dd = t.x[i] - yj;
      g += gut[i] * expf((vj - dd * dd) / eps - a[i]);
    }
    g = sk_warp_sum(g);
    if (lane == 0) gv_g[j] = -g;
  }
}

// ---------------------------------------------------------------- loss value and the scatter through the sort
__global__ void __launch_bounds__(256)
sink_finish_kernel(SinkP p, float w, float* __restrict__ loss, float* __restrict__ grad, int* __restrict__ info) {
  const int b = blockIdx.x;
  const int P1 = sk_points(p.cnt[0], p.HW), P2 = sk_points(p.cnt[1], p.HW);
  if (b == 0 && threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < p.B; ++i) s += p.cost[i];           // 'sum' reduction over the batch, in order
    *loss = s;
    if (info) { info[0] = P1; info[1] = P2; info[2] = p.cnt[2]; info[3] = p.cnt[3]; }
  }
  if (!grad || P1 > p.cap1 || P2 > p.cap2) return;
  const int* xi = p.xi + static_cast<size_t>(b) * p.HW + (p.HW - P1);
  const float* gx = p.gx + static_cast<size_t>(b) * p.cap1;
  for (int i = threadIdx.x; i < P1; i += blockDim.x) grad[static_cast<size_t>(b) * p.HW + xi[i]] += w * gx[i];
}

}  // namespace srb

extern "C" size_t srb_sinkhorn_workspace_bytes(int B, int HW, int cap1, int cap2, int max_iter) {
  if (B < 1 || HW < 1 || cap1 < 1 || cap2 < 1 || max_iter < 1) return 0;
  return srb::sink_layout(B, HW, cap1, cap2, max_iter).total;
}

extern "C" int srb_sinkhorn_loss(const float* gen, const float* truth, int B, int HW, float eps, int max_iter,
                                 float thresh, int cap1, int cap2, float w, void* workspace, size_t workspace_bytes,
                                 float* loss, float* grad, int* info, void* stream) {
  using namespace srb;
  if (!gen || !truth || !workspace || !loss) return set_error("sinkhorn_loss: null pointer");
  if (B < 1 || HW < 2 || HW > 32768) return set_error("sinkhorn_loss: 2 <= H*W <= 32768 (the per-image sort runs in shared memory)");
  if (max_iter < 1 || max_iter > kSinkMaxIter) return set_error("sinkhorn_loss: 1 <= max_iter <= 1024");
  if (!(eps > 0.f) || !(thresh > 0.f)) return set_error("sinkhorn_loss: eps and the stopping threshold must be positive");
  if (cap1 < 1 || cap2 < 1 || cap1 > HW || cap2 > HW) return set_error("sinkhorn_loss: 1 <= capacity <= H*W");
  const size_t smem_bwd = sizeof(float) * (3 * size_t(cap1) + 4 * size_t(cap2));
  if (smem_bwd > 220 * 1024)
    return set_error("sinkhorn_loss: 12 * cap1 + 16 * cap2 bytes of shared memory per image exceed 220 KB");
  const SinkLayout L = sink_layout(B, HW, cap1, cap2, max_iter);
  if (workspace_bytes < L.total) return set_error("sinkhorn_loss: workspace too small");
  if (reinterpret_cast<uintptr_t>(workspace) & 255) return set_error("sinkhorn_loss: workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  SinkP p{};
  p.B = B; p.HW = HW; p.cap1 = cap1; p.cap2 = cap2; p.max_iter = max_iter; p.eps = eps; p.thresh = thresh;
  p.cnt = reinterpret_cast<int*>(ws + L.cnt); p.err = reinterpret_cast<float*>(ws + L.err);
  p.cost = reinterpret_cast<float*>(ws + L.cost);
  p.xs = reinterpret_cast<float*>(ws + L.xs); p.ys = reinterpret_cast<float*>(ws + L.ys);
  p.xi = reinterpret_cast<int*>(ws + L.xi);
  p.uh = reinterpret_cast<float*>(ws + L.uh); p.vh = reinterpret_cast<float*>(ws + L.vh);
  p.gu = reinterpret_cast<float*>(ws + L.gu); p.gv = reinterpret_cast<float*>(ws + L.gv);
  p.gx = reinterpret_cast<float*>(ws + L.gx);
  // header + error sums + costs are contiguous at the front; the duals start from zero (utils.py:190-191)
  cudaMemsetAsync(ws, 0, L.xs, st);
  cudaMemsetAsync(p.uh, 0, sizeof(float) * size_t(B) * cap1, st);
  cudaMemsetAsync(p.vh, 0, sizeof(float) * size_t(B) * cap2, st);
  int n2 = 2; while (n2 < HW) n2 <<= 1;
  const size_t smem_sort = size_t(n2) * 6;
  const size_t smem_it = sizeof(float) * (2 * size_t(cap1) + 4 * size_t(cap2));
  ensure_max_smem(reinterpret_cast<const void*>(sink_sort_kernel), (int)smem_sort);
  ensure_max_smem(reinterpret_cast<const void*>(sink_iter_kernel), (int)smem_it);
  ensure_max_smem(reinterpret_cast<const void*>(sink_cost_kernel), (int)smem_it);
  ensure_max_smem(reinterpret_cast<const void*>(sink_bwd_kernel), (int)smem_bwd);
  sink_sort_kernel<<<dim3(B, 2), kSinkThreads, smem_sort, st>>>(gen, truth, p, n2);
  for (int it = 0; it < max_iter; ++it) sink_iter_kernel<<<B, kSinkThreads, smem_it, st>>>(p, it);
  sink_cost_kernel<<<B, kSinkThreads, smem_it, st>>>(p, grad ? 1 : 0);
  if (grad)
    for (int it = max_iter - 1; it >= 0; --it) sink_bwd_kernel<<<B, kSinkThreads, smem_bwd, st>>>(p, it);
  sink_finish_kernel<<<B, 256, 0, st>>>(p, w, loss, grad, info);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("sinkhorn_loss launch", e);
  return 0;
}
