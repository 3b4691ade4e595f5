// K5: small-channel 3x3 convolutions of the discriminator stack (models.py:140-174), fp32 NCHW, stride 1 or 2,
// padding 1 - forward, input gradient and weight gradient as three independent linear kernels so that the
// Python autograd Functions can compose them for the double backward the gradient penalty needs
// (esrgan.py:598-606).  Plus the elementwise LeakyReLU pair and SumPool2d (models.py:297-305).
// Channel counts are 1/16/32/64: latency/bandwidth-bound CUDA-core kernels, not tensor-core work.
#include "common.cuh"

namespace srb {

// y[b,o,yo,xo] = bias[o] + sum_{i,ky,kx} w[o,i,ky,kx] * x[b,i,yo*s+ky-1,xo*s+kx-1]
// block = (32 x-positions) x (8 output channels); weights for the 8 channels staged in smem.
template <int OCB>
__global__ void dconv_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                 float* __restrict__ y, int B, int Cin, int H, int W, int Cout, int s, int Ho, int Wo) {
  extern __shared__ float ws[];                 // [OCB][Cin*9]
  const int ocb = blockIdx.y * OCB;
  const int b = blockIdx.z;
  const int K = Cin * 9;
  for (int i = threadIdx.x; i < OCB * K; i += blockDim.x) {
    const int o = ocb + i / K;
    ws[i] = o < Cout ? w[static_cast<size_t>(o) * K + i % K] : 0.f;
  }
  __syncthreads();
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= Ho * Wo) return;
  const int yo = pix / Wo, xo = pix % Wo;
  float acc[OCB];
#pragma unroll
  for (int o = 0; o < OCB; ++o) acc[o] = (bias && ocb + o < Cout) ? bias[ocb + o] : 0.f;
  const float* xb = x + static_cast<size_t>(b) * Cin * H * W;
  for (int i = 0; i < Cin; ++i) {
    float v[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int yy = yo * s + t / 3 - 1, xx = xo * s + t % 3 - 1;
      v[t] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(xb + (static_cast<size_t>(i) * H + yy) * W + xx) : 0.f;
    }
#pragma unroll
    for (int o = 0; o < OCB; ++o) {
      const float* wr = ws + o * K + i * 9;
#pragma unroll
      for (int t = 0; t < 9; ++t) acc[o] += v[t] * wr[t];
    }
  }
#pragma unroll
  for (int o = 0; o < OCB; ++o)
    if (ocb + o < Cout) y[((static_cast<size_t>(b) * Cout + ocb + o) * Ho + yo) * Wo + xo] = acc[o];
}

// gx[b,i,y,x] = sum_{o,ky,kx} w[o,i,ky,kx] * gy[b,o,(y+1-ky)/s,(x+1-kx)/s]   (when divisible and in range)
template <int ICB>
__global__ void dconv_dgrad_kernel(const float* __restrict__ gy, const float* __restrict__ w, float* __restrict__ gx,
                                   int B, int Cin, int H, int W, int Cout, int s, int Ho, int Wo) {
  extern __shared__ float ws[];                 // [Cout][ICB][9]
  const int icb = blockIdx.y * ICB;
  const int b = blockIdx.z;
  for (int idx = threadIdx.x; idx < Cout * ICB * 9; idx += blockDim.x) {
    const int t = idx % 9, ii = (idx / 9) % ICB, o = idx / (9 * ICB);
    ws[idx] = (icb + ii < Cin) ? w[(static_cast<size_t>(o) * Cin + icb + ii) * 9 + t] : 0.f;
  }
  __syncthreads();
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= H * W) return;
  const int yy = pix / W, xx = pix % W;
  float acc[ICB];
#pragma unroll
  for (int i = 0; i < ICB; ++i) acc[i] = 0.f;
  const float* gb = gy + static_cast<size_t>(b) * Cout * Ho * Wo;
  int oy[3], ox[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int ny = yy + 1 - k, nx = xx + 1 - k;
    oy[k] = (ny >= 0 && ny % s == 0 && ny / s < Ho) ? ny / s : -1;
    ox[k] = (nx >= 0 && nx % s == 0 && nx / s < Wo) ? nx / s : -1;
  }
  for (int o = 0; o < Cout; ++o) {
    float v[9];
#pragma unroll
    for (int t = 0; t < 9; ++t)
      v[t] = (oy[t / 3] >= 0 && ox[t % 3] >= 0) ? __ldg(gb + (static_cast<size_t>(o) * Ho + oy[t / 3]) * Wo + ox[t % 3]) : 0.f;
#pragma unroll
    for (int i = 0; i < ICB; ++i) {
      const float* wr = ws + (o * ICB + i) * 9;
#pragma unroll
      for (int t = 0; t < 9; ++t) acc[i] += v[t] * wr[t];
    }
  }
#pragma unroll
  for (int i = 0; i < ICB; ++i)
    if (icb + i < Cin) gx[((static_cast<size_t>(b) * Cin + icb + i) * H + yy) * W + xx] = acc[i];
}

// dw[o,i,t] += sum_{b,yo,xo} gy[b,o,yo,xo] * x[b,i,yo*s+ky-1,xo*s+kx-1];  db[o] += sum gy
// grid = (Cout*Cin, chunks); block reduces its chunk of (b, pixel) pairs and issues 9 (+1) atomics.
__global__ void dconv_wgrad_kernel(const float* __restrict__ gy, const float* __restrict__ x, float* __restrict__ dw,
                                   float* __restrict__ db, int B, int Cin, int H, int W, int Cout, int s, int Ho,
                                   int Wo, long long per_chunk) {
  const int o = blockIdx.x / Cin, i = blockIdx.x % Cin;
  const long long total = 1LL * B * Ho * Wo;
  const long long e0 = blockIdx.y * per_chunk;
  long long e1 = e0 + per_chunk; if (e1 > total) e1 = total;
  float acc[9], bs = 0.f;
#pragma unroll
  for (int t = 0; t < 9; ++t) acc[t] = 0.f;
  for (long long e = e0 + threadIdx.x; e < e1; e += blockDim.x) {
    const int xo = int(e % Wo); const int yo = int((e / Wo) % Ho); const int b = int(e / (1LL * Wo * Ho));
    const float g = __ldg(gy + ((static_cast<size_t>(b) * Cout + o) * Ho + yo) * Wo + xo);
    bs += g;
    if (g == 0.f) continue;
    const float* xb = x + (static_cast<size_t>(b) * Cin + i) * H * W;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int yy = yo * s + t / 3 - 1, xx = xo * s + t % 3 - 1;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) acc[t] += g * __ldg(xb + static_cast<size_t>(yy) * W + xx);
    }
  }
  __shared__ float red[10][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int t = 0; t < 10; ++t) {
    float v = t < 9 ? acc[t] : bs;
    for (int k = 16; k; k >>= 1) v += __shfl_xor_sync(0xffffffffu, v, k);
    if (lane == 0) red[t][warp] = v;
  }
  __syncthreads();
  if (threadIdx.x < 10) {
    float v = 0.f;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) v += red[threadIdx.x][k];
    if (threadIdx.x < 9) atomicAdd(dw + (static_cast<size_t>(o) * Cin + i) * 9 + threadIdx.x, v);
    else if (db && i == 0) atomicAdd(db + o, v);
  }
}

__global__ void lrelu_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long n, float slope) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i < n) { const float v = x[i]; y[i] = v > 0.f ? v : v * slope; }
}
// gx = g * (ref > 0 ? 1 : slope); ref may be the LeakyReLU input or output (same sign for slope > 0)
__global__ void lrelu_bwd_kernel(const float* __restrict__ g, const float* __restrict__ ref, float* __restrict__ gx,
                                 long long n, float slope) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i < n) gx[i] = ref[i] > 0.f ? g[i] : g[i] * slope;
}

// SumPool2d: y[b,c,yo,xo] = sum of the k x k block (models.py:304-305: k*k*AvgPool2d(k)); bwd broadcasts.
__global__ void sumpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int BC, int H, int W, int k) {
  const int Ho = H / k, Wo = W / k;
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= 1LL * BC * Ho * Wo) return;
  const int xo = int(i % Wo); const int yo = int((i / Wo) % Ho); const int bc = int(i / (1LL * Wo * Ho));
  const float* p = x + (static_cast<size_t>(bc) * H + yo * k) * W + xo * k;
  float s = 0.f;
  for (int a = 0; a < k; ++a)
    for (int c = 0; c < k; ++c) s += p[a * W + c];
  y[i] = s;
}
__global__ void sumpool_bwd_kernel(const float* __restrict__ gy, float* __restrict__ gx, int BC, int H, int W, int k) {
  const int Ho = H / k, Wo = W / k;
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= 1LL * BC * H * W) return;
  const int xx = int(i % W); const int yy = int((i / W) % H); const int bc = int(i / (1LL * W * H));
  const int yo = yy / k, xo = xx / k;
  gx[i] = (yo < Ho && xo < Wo) ? gy[(static_cast<size_t>(bc) * Ho + yo) * Wo + xo] : 0.f;
}

}  // namespace srb

using namespace srb;
#define SRB_CHECK_LAUNCH(name)                                         \
  do {                                                                 \
    cudaError_t e_ = cudaGetLastError();                               \
    if (e_ != cudaSuccess) return set_cuda_error(name " launch", e_);  \
  } while (0)

namespace srb {
int dconv_fwd_fast(const float* x, const float* w, const float* bias, float* y, int B, int Cin, int H, int W, int Cout,
                   int stride, int act, float slope, cudaStream_t st);
int dconv_dgrad_fast(const float* gy, const float* w, float* gx, int B, int Cin, int H, int W, int Cout, int stride,
                     cudaStream_t st);
int dconv_wgrad_fast(const float* gy, const float* x, float* dw, float* db, int B, int Cin, int H, int W, int Cout,
                     int stride, cudaStream_t st);
static int g_dconv_impl = 0;     // 0: register-tiled kernels (disc_fast.cu); 1: the simple kernels above (cross-check)
}  // namespace srb
extern "C" int srb_set_dconv_impl(int impl) { srb::g_dconv_impl = impl; return 0; }

extern "C" int srb_dconv_fwd(const float* x, const float* w, const float* bias, float* y, int B, int Cin, int H, int W,
                             int Cout, int stride, int act, float slope, void* stream) {
  if (stride != 1 && stride != 2) return set_error("dconv: stride must be 1 or 2");
  if (g_dconv_impl == 0) return dconv_fwd_fast(x, w, bias, y, B, Cin, H, W, Cout, stride, act, slope, (cudaStream_t)stream);
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  constexpr int OCB = 8;
  dim3 grid(ceil_div(Ho * Wo, 128), ceil_div(Cout, OCB), B);
  const size_t smem = sizeof(float) * OCB * Cin * 9;
  dconv_fwd_kernel<OCB><<<grid, 128, smem, (cudaStream_t)stream>>>(x, w, bias, y, B, Cin, H, W, Cout, stride, Ho, Wo);
  SRB_CHECK_LAUNCH("dconv_fwd");
  if (act) {
    const long long n = 1LL * B * Cout * Ho * Wo;
    lrelu_fwd_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(y, y, n, slope);
    SRB_CHECK_LAUNCH("lrelu_fwd");
  }
  return 0;
}

extern "C" int srb_dconv_dgrad(const float* gy, const float* w, float* gx, int B, int Cin, int H, int W, int Cout,
                               int stride, void* stream) {
  if (stride != 1 && stride != 2) return set_error("dconv: stride must be 1 or 2");
  if (g_dconv_impl == 0) return dconv_dgrad_fast(gy, w, gx, B, Cin, H, W, Cout, stride, (cudaStream_t)stream);
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  constexpr int ICB = 8;
  dim3 grid(ceil_div(H * W, 128), ceil_div(Cin, ICB), B);
  const size_t smem = sizeof(float) * Cout * ICB * 9;
  dconv_dgrad_kernel<ICB><<<grid, 128, smem, (cudaStream_t)stream>>>(gy, w, gx, B, Cin, H, W, Cout, stride, Ho, Wo);
  SRB_CHECK_LAUNCH("dconv_dgrad");
  return 0;
}

extern "C" int srb_dconv_wgrad(const float* gy, const float* x, float* dw, float* db, int B, int Cin, int H, int W,
                               int Cout, int stride, void* stream) {
  if (stride != 1 && stride != 2) return set_error("dconv: stride must be 1 or 2");
  if (g_dconv_impl == 0) return dconv_wgrad_fast(gy, x, dw, db, B, Cin, H, W, Cout, stride, (cudaStream_t)stream);
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  cudaMemsetAsync(dw, 0, sizeof(float) * Cout * Cin * 9, (cudaStream_t)stream);
  if (db) cudaMemsetAsync(db, 0, sizeof(float) * Cout, (cudaStream_t)stream);
  const long long total = 1LL * B * Ho * Wo;
  int chunks = ceil_div(4LL * sm_count(), 1LL * Cout * Cin);
  if (chunks < 1) chunks = 1;
  const long long max_chunks = (total + 1023) / 1024;
  if (chunks > max_chunks) chunks = (int)max_chunks;
  const long long per = (total + chunks - 1) / chunks;
  dim3 grid(Cout * Cin, chunks);
  dconv_wgrad_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(gy, x, dw, db, B, Cin, H, W, Cout, stride, Ho, Wo, per);
  SRB_CHECK_LAUNCH("dconv_wgrad");
  return 0;
}

extern "C" int srb_lrelu_fwd(const float* x, float* y, long long n, float slope, void* stream) {
  lrelu_fwd_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(x, y, n, slope);
  SRB_CHECK_LAUNCH("lrelu_fwd");
  return 0;
}
extern "C" int srb_lrelu_bwd(const float* g, const float* ref, float* gx, long long n, float slope, void* stream) {
  lrelu_bwd_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(g, ref, gx, n, slope);
  SRB_CHECK_LAUNCH("lrelu_bwd");
  return 0;
}
extern "C" int srb_sumpool_fwd(const float* x, float* y, int BC, int H, int W, int k, void* stream) {
  sumpool_fwd_kernel<<<ceil_div(1LL * BC * (H / k) * (W / k), 256), 256, 0, (cudaStream_t)stream>>>(x, y, BC, H, W, k);
  SRB_CHECK_LAUNCH("sumpool_fwd");
  return 0;
}
extern "C" int srb_sumpool_bwd(const float* gy, float* gx, int BC, int H, int W, int k, void* stream) {
  sumpool_bwd_kernel<<<ceil_div(1LL * BC * H * W, 256), 256, 0, (cudaStream_t)stream>>>(gy, gx, BC, H, W, k);
  SRB_CHECK_LAUNCH("sumpool_bwd");
  return 0;
}
