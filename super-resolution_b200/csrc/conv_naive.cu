// CUDA-core restatement of K1/K3 on the same PNHWC / packed-weight layouts.  Not a fallback for speed: it is the
// on-device cross-check for the tcgen05 kernels at sizes the CPU oracle cannot reach.
#include "common.cuh"
#include "sm100.cuh"

namespace srb {

struct NaiveConvParams {
  int R, P, pitch, H, W, KB, cout, x_cstride, x_coff;
  const __nv_bfloat16* x; const __nv_bfloat16* w; const float* bias;
  __nv_bfloat16* y; int y_cstride, y_coff, store_mode, y_pitch, y_P;
  float alpha; int act; float slope;
  const __nv_bfloat16* mask; int mask_cstride, mask_coff;
  const __nv_bfloat16* res1; int res1_cstride, res1_coff; float res1_scale;
  const __nv_bfloat16* res2; int res2_cstride, res2_coff; float res2_scale;
  int res_f32; float* y32; int y32_cstride, y32_coff;
  const __nv_bfloat16* res3; int res3_cstride, res3_coff; float res3_scale;
  const float* chan_scale;
};

__device__ __forceinline__ float naive_res(const void* base, long long i, bool f32) {
  return f32 ? static_cast<const float*>(base)[i] : __bfloat162float(static_cast<const __nv_bfloat16*>(base)[i]);
}

__global__ void conv3x3_naive_kernel(const NaiveConvParams p) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int n = int(idx % p.cout);
  const long long q = idx / p.cout;
  if (q >= p.R) return;
  const int r = int(q % p.P);
  const int py = r / p.pitch, px = r - py * p.pitch;
  if (px >= p.W || py >= p.H) return;
  float acc = 0.f;
  for (int t = 0; t < 9; ++t) {
    const long long qs = q + (t / 3 - 1) * p.pitch + (t % 3 - 1);
    if (qs < 0 || qs >= p.R) continue;                       // what TMA zero-fill does in the tcgen05 kernel
    const __nv_bfloat16* xr = p.x + qs * p.x_cstride + p.x_coff;
    for (int kb = 0; kb < p.KB; ++kb) {
      const __nv_bfloat16* wr = p.w + ((size_t)(t * p.KB + kb) * p.cout + n) * 64;
      for (int k = 0; k < 64; ++k) acc += __bfloat162float(xr[kb * 64 + k]) * __bfloat162float(wr[k]);
    }
  }
  float v = acc;
  if (p.bias) v += p.bias[n];
  if (p.chan_scale) v *= p.chan_scale[(q / p.P) * p.cout + n];
  v *= p.alpha;
  if (p.res1) v += p.res1_scale * naive_res(p.res1, q * p.res1_cstride + p.res1_coff + n, p.res_f32 & 1);
  if (p.res2) v += p.res2_scale * naive_res(p.res2, q * p.res2_cstride + p.res2_coff + n, p.res_f32 & 2);
  if (p.res3) v += p.res3_scale * __bfloat162float(p.res3[q * p.res3_cstride + p.res3_coff + n]);
  if (p.act == SRB_ACT_LRELU) v = v > 0.f ? v : v * p.slope;
  else if (p.act == SRB_ACT_LRELU_MASK)
    v *= __bfloat162float(p.mask[q * p.mask_cstride + p.mask_coff + n]) > 0.f ? 1.f : p.slope;
  if (p.store_mode == SRB_STORE_SHUFFLE) {
    const int Cy = p.cout / 4;
    const int sub = n / Cy, c = n % Cy;                      // packed channel order (2i+j)*Cy + c
    const long long img = q / p.P;
    const long long yrow = img * p.y_P + (long long)(2 * py + (sub >> 1)) * p.y_pitch + (2 * px + (sub & 1));
    p.y[yrow * p.y_cstride + p.y_coff + c] = __float2bfloat16(v);
  } else {
    p.y[q * p.y_cstride + p.y_coff + n] = __float2bfloat16(v);
    if (p.y32) p.y32[q * p.y32_cstride + p.y32_coff + n] = v;
  }
}

int conv3x3_naive_launch(const srb_conv3x3_args* a, cudaStream_t stream) {
  if (a->cin % 64 || a->cout % 64) return set_error("conv3x3(naive): channels must be multiples of 64");
  if (a->store_mode == SRB_STORE_SHUFFLE && a->cout % 256)
    return set_error("conv3x3(naive): shuffle store needs cout % 256 == 0");
  NaiveConvParams p{};
  p.pitch = a->W + 1; p.P = (a->H + 1) * (a->W + 1); p.R = a->batch * p.P; p.H = a->H; p.W = a->W;
  p.KB = a->cin / 64; p.cout = a->cout; p.x_cstride = a->x_cstride; p.x_coff = a->x_coff;
  p.x = (const __nv_bfloat16*)a->x; p.w = (const __nv_bfloat16*)a->w; p.bias = a->bias;
  p.y = (__nv_bfloat16*)a->y; p.y_cstride = a->y_cstride; p.y_coff = a->y_coff; p.store_mode = a->store_mode;
  if (a->store_mode == SRB_STORE_SHUFFLE) { p.y_pitch = 2 * a->W + 1; p.y_P = (2 * a->H + 1) * (2 * a->W + 1); }
  else { p.y_pitch = p.pitch; p.y_P = p.P; }
  p.alpha = a->alpha; p.act = a->act; p.slope = a->slope;
  p.mask = (const __nv_bfloat16*)a->mask; p.mask_cstride = a->mask_cstride; p.mask_coff = a->mask_coff;
  p.res1 = (const __nv_bfloat16*)a->res1; p.res1_cstride = a->res1_cstride; p.res1_coff = a->res1_coff; p.res1_scale = a->res1_scale;
  p.res2 = (const __nv_bfloat16*)a->res2; p.res2_cstride = a->res2_cstride; p.res2_coff = a->res2_coff; p.res2_scale = a->res2_scale;
  p.res_f32 = a->res_f32; p.y32 = a->y32; p.y32_cstride = a->y32_cstride; p.y32_coff = a->y32_coff;
  p.res3 = (const __nv_bfloat16*)a->res3; p.res3_cstride = a->res3_cstride; p.res3_coff = a->res3_coff; p.res3_scale = a->res3_scale;
  p.chan_scale = a->chan_scale;
  if ((p.y32 || p.chan_scale) && a->store_mode != SRB_STORE_PLAIN)
    return set_error("conv3x3(naive): y32 / chan_scale need a plain store");
  const long long total = 1LL * p.R * p.cout;
  conv3x3_naive_kernel<<<ceil_div(total, 256), 256, 0, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("conv3x3_naive launch", e);
  return 0;
}

// ------------------------------------------------------------------------------ weight gradient (naive)
struct NaiveWgradParams {
  int R, P, pitch, H, W, cin, cout, x_cstride, x_coff, gy_cstride, gy_coff;
  const __nv_bfloat16* x; const __nv_bfloat16* gy;
  float* dw; float* dbias; float alpha; int accumulate; int unshuffle;
};

// one block per (ci, co): threads stride over pixels, block-reduce the 9 taps (+ bias sum)
__global__ void wgrad3x3_naive_kernel(const NaiveWgradParams p) {
  const int ci = blockIdx.x;
  const int co = blockIdx.y;      // gy channel (packed order)
  float acc[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) acc[t] = 0.f;
  float bsum = 0.f;
  for (long long q = threadIdx.x; q < p.R; q += blockDim.x) {
    const float g = __bfloat162float(p.gy[q * p.gy_cstride + p.gy_coff + co]);
    if (g == 0.f) continue;
    bsum += g;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const long long qs = q + (t / 3 - 1) * p.pitch + (t % 3 - 1);
      if (qs < 0 || qs >= p.R) continue;
      acc[t] += g * __bfloat162float(p.x[qs * p.x_cstride + p.x_coff + ci]);
    }
  }
  __shared__ float red[10][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int t = 0; t < 10; ++t) {
    float v = t < 9 ? acc[t] : bsum;
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[t][warp] = v;
  }
  __syncthreads();
  if (threadIdx.x < 10) {
    float v = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) v += red[threadIdx.x][w];
    v *= p.alpha;
    const int cq = p.cout / 4;
    const int co_dst = p.unshuffle ? (4 * (co % cq) + co / cq) : co;
    if (threadIdx.x < 9) {
      float* d = p.dw + ((size_t)co_dst * p.cin + ci) * 9 + threadIdx.x;
      *d = p.accumulate ? *d + v : v;
    } else if (p.dbias && ci == 0) {
      float* d = p.dbias + co_dst;
      *d = p.accumulate ? *d + v : v;
    }
  }
}

int wgrad3x3_naive_launch(const srb_wgrad3x3_args* a, cudaStream_t stream) {
  NaiveWgradParams p{};
  p.pitch = a->W + 1; p.P = (a->H + 1) * (a->W + 1); p.R = a->batch * p.P; p.H = a->H; p.W = a->W;
  p.cin = a->cin; p.cout = a->cout; p.x_cstride = a->x_cstride; p.x_coff = a->x_coff;
  p.gy_cstride = a->gy_cstride; p.gy_coff = a->gy_coff;
  p.x = (const __nv_bfloat16*)a->x; p.gy = (const __nv_bfloat16*)a->gy;
  p.dw = a->dw; p.dbias = a->dbias; p.alpha = a->alpha; p.accumulate = a->accumulate; p.unshuffle = a->unshuffle;
  dim3 grid(a->cin, a->cout);
  wgrad3x3_naive_kernel<<<grid, 256, 0, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("wgrad3x3_naive launch", e);
  return 0;
}

}  // namespace srb
