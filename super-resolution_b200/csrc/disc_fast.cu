// K5 (tuned): register-tiled fp32 kernels for the discriminator's small-channel 3x3 convolutions.
//   dconv_tile_kernel  forward conv (stride 1|2) and, with transposed/flipped weight staging, the stride-1 input
//                      gradient: each thread owns 4 consecutive output pixels x 8 channels (288 FMA per 18 loads);
//                      bias and LeakyReLU fused for the forward.
//   dconv_dgrad_s2_kernel  stride-2 input gradient by parity: a thread owns two 2x2 input quads x 8 channels; every
//                      one of the 9 taps is used exactly once per quad (no wasted taps, no divisibility tests).
//   dconv_wgrad_tile_kernel  weight gradient: a block stages a band of gy and x in shared memory and each thread
//                      accumulates TO x 9 filter taps for one input channel over the band's pixels, looping over its
//                      share of (image, band) tiles before a single round of atomics.
#include "common.cuh"

namespace srb {

constexpr int kOCB = 8;     // channels per thread
constexpr int kPX = 4;      // pixels per thread along x

// MODE 0: forward   y[b,o,yo,xo] = act(bias[o] + sum_{i,t} w[o,i,t] x[b,i,yo*S+ky-1,xo*S+kx-1])
// MODE 1: stride-1 input gradient  gx[b,i,y,x] = sum_{o,t} w[o,i,8-t'] ...  (same loop, weights staged transposed)
template <int S, int MODE>
__global__ void __launch_bounds__(128)
dconv_tile_kernel(const float* __restrict__ in, const float* __restrict__ w, const float* __restrict__ bias,
                  float* __restrict__ out, int B, int Cred, int H, int W, int Cout_total, int Cin_total, int Ho, int Wo,
                  int act, float slope) {
  // Cred = reduction channels (Cin for MODE 0, Cout for MODE 1); output channels = Cout_total (0) / Cin_total (1)
  extern __shared__ float ws[];                      // [Cred][9][kOCB]
  const int n_out = MODE == 0 ? Cout_total : Cin_total;
  const int ocb = blockIdx.y * kOCB;
  const int b = blockIdx.z;
  for (int idx = threadIdx.x; idx < Cred * 9 * kOCB; idx += blockDim.x) {
    const int oc = idx % kOCB, t = (idx / kOCB) % 9, r = idx / (9 * kOCB);
    float v = 0.f;
    if (ocb + oc < n_out) {
      if (MODE == 0) v = w[(static_cast<size_t>(ocb + oc) * Cin_total + r) * 9 + t];
      else v = w[(static_cast<size_t>(r) * Cin_total + ocb + oc) * 9 + (8 - t)];
    }
    ws[idx] = v;
  }
  __syncthreads();
  const int WoQ = (Wo + kPX - 1) / kPX;
  const int item = blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= Ho * WoQ) return;
  const int yo = item / WoQ, xo0 = (item % WoQ) * kPX;
  constexpr int NC = (kPX - 1) * S + 3;              // input columns needed
  float acc[kPX][kOCB];
#pragma unroll
  for (int p = 0; p < kPX; ++p)
#pragma unroll
    for (int o = 0; o < kOCB; ++o) acc[p][o] = (MODE == 0 && bias && ocb + o < n_out) ? __ldg(bias + ocb + o) : 0.f;
  const float* ib = in + static_cast<size_t>(b) * Cred * H * W;
  const int y0 = yo * S - 1, x0 = xo0 * S - 1;
  for (int c = 0; c < Cred; ++c) {
    float v[3][NC];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int yy = y0 + r;
      const bool rok = yy >= 0 && yy < H;
      const float* rowp = ib + (static_cast<size_t>(c) * H + (rok ? yy : 0)) * W;
#pragma unroll
      for (int q = 0; q < NC; ++q) {
        const int xx = x0 + q;
        v[r][q] = (rok && xx >= 0 && xx < W) ? __ldg(rowp + xx) : 0.f;
      }
    }
    const float4* wc = reinterpret_cast<const float4*>(ws + c * 9 * kOCB);
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float4 wa = wc[2 * t], wb = wc[2 * t + 1];
      const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
      for (int p = 0; p < kPX; ++p) {
        const float val = v[t / 3][p * S + t % 3];
#pragma unroll
        for (int o = 0; o < kOCB; ++o) acc[p][o] += val * wv[o];
      }
    }
  }
#pragma unroll
  for (int o = 0; o < kOCB; ++o) {
    if (ocb + o >= n_out) continue;
    float* op = out + ((static_cast<size_t>(b) * n_out + ocb + o) * Ho + yo) * Wo + xo0;
#pragma unroll
    for (int p = 0; p < kPX; ++p) {
      if (xo0 + p < Wo) {
        float r = acc[p][o];
        if (MODE == 0 && act) r = r > 0.f ? r : r * slope;
        op[p] = r;
      }
    }
  }
}

// stride-2 input gradient.  gx[b,i,y,x] = sum_o sum_{ky,kx: (y+1-ky),(x+1-kx) even} w[o,i,ky,kx] gy[b,o,(y+1-ky)/2,(x+1-kx)/2]
// A thread owns input rows 2a, 2a+1 and columns 2c..2c+3 (two quads) for 8 input channels.
__global__ void __launch_bounds__(128)
dconv_dgrad_s2_kernel(const float* __restrict__ gy, const float* __restrict__ w, float* __restrict__ gx, int B, int Cin,
                      int H, int W, int Cout, int Ho, int Wo) {
  extern __shared__ float ws[];                      // [Cout][9][kOCB]
  const int icb = blockIdx.y * kOCB;
  const int b = blockIdx.z;
  for (int idx = threadIdx.x; idx < Cout * 9 * kOCB; idx += blockDim.x) {
    const int ic = idx % kOCB, t = (idx / kOCB) % 9, o = idx / (9 * kOCB);
    ws[idx] = (icb + ic < Cin) ? w[(static_cast<size_t>(o) * Cin + icb + ic) * 9 + t] : 0.f;
  }
  __syncthreads();
  const int QW = (W + 3) / 4, QH = (H + 1) / 2;
  const int item = blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= QH * QW) return;
  const int a = item / QW, c0 = (item % QW) * 2;     // gy row a, gy columns c0, c0+1, c0+2
  float acc[2][4][kOCB];
#pragma unroll
  for (int r = 0; r < 2; ++r)
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int i = 0; i < kOCB; ++i) acc[r][q][i] = 0.f;
  const float* gb = gy + static_cast<size_t>(b) * Cout * Ho * Wo;
  for (int o = 0; o < Cout; ++o) {
    float g[2][3];
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        const int yy = a + r, xx = c0 + q;
        g[r][q] = (yy < Ho && xx < Wo) ? __ldg(gb + (static_cast<size_t>(o) * Ho + yy) * Wo + xx) : 0.f;
      }
    const float4* wc = reinterpret_cast<const float4*>(ws + o * 9 * kOCB);
    float wv[9][kOCB];
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float4 wa = wc[2 * t], wb = wc[2 * t + 1];
      wv[t][0] = wa.x; wv[t][1] = wa.y; wv[t][2] = wa.z; wv[t][3] = wa.w;
      wv[t][4] = wb.x; wv[t][5] = wb.y; wv[t][6] = wb.z; wv[t][7] = wb.w;
    }
#pragma unroll
    for (int qd = 0; qd < 2; ++qd) {                 // quad index: gy columns c0+qd (same) and c0+qd+1 (next)
#pragma unroll
      for (int i = 0; i < kOCB; ++i) {
        const float g00 = g[0][qd], g01 = g[0][qd + 1], g10 = g[1][qd], g11 = g[1][qd + 1];
        // even row 2a: ky = 1
        acc[0][2 * qd][i] += g00 * wv[4][i];                                   // even col: kx = 1
        acc[0][2 * qd + 1][i] += g01 * wv[3][i] + g00 * wv[5][i];              // odd col: kx = 0 (next), 2 (same)
        // odd row 2a+1: ky = 0 (next gy row), ky = 2 (same gy row)
        acc[1][2 * qd][i] += g10 * wv[1][i] + g00 * wv[7][i];
        acc[1][2 * qd + 1][i] += g11 * wv[0][i] + g10 * wv[2][i] + g01 * wv[6][i] + g00 * wv[8][i];
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kOCB; ++i) {
    if (icb + i >= Cin) continue;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int yy = 2 * a + r;
      if (yy >= H) continue;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int xx = 2 * c0 + q;
        if (xx < W) gx[((static_cast<size_t>(b) * Cin + icb + i) * H + yy) * W + xx] = acc[r][q][i];
      }
    }
  }
}

// weight gradient, tiled.  Block threads = pairs * lanes, pair = (o-tile of TO outputs, input channel i).
// Tiles = (image b, band of TY output rows).  smem: gy band [Cout][TY*Wo], x band [Cin][(TY-1)*S+3][W+2] (zero halo).
template <int TO, int S>
__global__ void dconv_wgrad_tile_kernel(const float* __restrict__ gy, const float* __restrict__ x,
                                        float* __restrict__ dw, float* __restrict__ db, int B, int Cin, int H, int W,
                                        int Cout, int Ho, int Wo, int TY, int lanes, int tiles_per_block) {
  extern __shared__ float sm[];
  const int n_ot = Cout / TO;
  const int pairs = n_ot * Cin;
  const int XR = (TY - 1) * S + 3, XW = W + 2;
  const int XS = (XR * XW) | 1;                      // odd channel stride: spreads channels over the smem banks
  const int gy_sz = Cout * TY * Wo;
  float* sgy = sm;                                   // [Cout][TY*Wo]
  float* sx = sm + gy_sz;                            // [Cin][XR][XW]
  const int pair = threadIdx.x % pairs, lane = threadIdx.x / pairs;
  const int ot = pair / Cin, ci = pair % Cin;        // lanes of a warp vary ci fastest
  const int bands = (Ho + TY - 1) / TY;
  const long long n_tiles = 1LL * B * bands;
  float acc[TO][9];
#pragma unroll
  for (int o = 0; o < TO; ++o)
#pragma unroll
    for (int t = 0; t < 9; ++t) acc[o][t] = 0.f;
  float bsum[TO];
#pragma unroll
  for (int o = 0; o < TO; ++o) bsum[o] = 0.f;

  for (int k = 0; k < tiles_per_block; ++k) {
    const long long tile = 1LL * blockIdx.x * tiles_per_block + k;
    if (tile >= n_tiles) break;
    const int b = int(tile / bands), band = int(tile % bands);
    const int yo0 = band * TY;
    const int rows = min(TY, Ho - yo0);
    __syncthreads();
    for (int idx = threadIdx.x; idx < gy_sz; idx += blockDim.x) {
      const int o = idx / (TY * Wo), r = (idx / Wo) % TY, c = idx % Wo;
      sgy[idx] = r < rows ? __ldg(gy + ((static_cast<size_t>(b) * Cout + o) * Ho + yo0 + r) * Wo + c) : 0.f;
    }
    const int yi0 = yo0 * S - 1;
    for (int idx = threadIdx.x; idx < Cin * XR * XW; idx += blockDim.x) {
      const int i = idx / (XR * XW), r = (idx / XW) % XR, c = idx % XW;
      const int yy = yi0 + r, xx = c - 1;
      sx[i * XS + r * XW + c] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(x + ((static_cast<size_t>(b) * Cin + i) * H + yy) * W + xx)
                                                         : 0.f;
    }
    __syncthreads();
    if (lane < lanes) {
      const float* xg = sx + ci * XS;
      const int npix = rows * Wo;
      for (int pidx = lane; pidx < npix; pidx += lanes) {
        const int r = pidx / Wo, c = pidx % Wo;
        float g[TO];
        bool any = false;
#pragma unroll
        for (int o = 0; o < TO; ++o) { g[o] = sgy[(ot * TO + o) * TY * Wo + pidx]; any |= g[o] != 0.f; }
        if (!any) continue;
        const float* xp = xg + (r * S) * XW + c * S;
        float xv[9];
#pragma unroll
        for (int t = 0; t < 9; ++t) xv[t] = xp[(t / 3) * XW + t % 3];
#pragma unroll
        for (int o = 0; o < TO; ++o) {
          if (ci == 0) bsum[o] += g[o];
#pragma unroll
          for (int t = 0; t < 9; ++t) acc[o][t] += g[o] * xv[t];
        }
      }
    }
  }
  if (lane < lanes) {
#pragma unroll
    for (int o = 0; o < TO; ++o) {
      float* d = dw + (static_cast<size_t>(ot * TO + o) * Cin + ci) * 9;
#pragma unroll
      for (int t = 0; t < 9; ++t)
        if (acc[o][t] != 0.f) atomicAdd(d + t, acc[o][t]);
      if (db && ci == 0 && bsum[o] != 0.f) atomicAdd(db + ot * TO + o, bsum[o]);
    }
  }
}

// ------------------------------------------------------------------------------ launchers
int dconv_fwd_fast(const float* x, const float* w, const float* bias, float* y, int B, int Cin, int H, int W, int Cout,
                   int stride, int act, float slope, cudaStream_t st) {
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  const int WoQ = (Wo + kPX - 1) / kPX;
  dim3 grid(ceil_div(Ho * WoQ, 128), ceil_div(Cout, kOCB), B);
  const size_t smem = sizeof(float) * Cin * 9 * kOCB;
  if (stride == 1)
    dconv_tile_kernel<1, 0><<<grid, 128, smem, st>>>(x, w, bias, y, B, Cin, H, W, Cout, Cin, Ho, Wo, act, slope);
  else
    dconv_tile_kernel<2, 0><<<grid, 128, smem, st>>>(x, w, bias, y, B, Cin, H, W, Cout, Cin, Ho, Wo, act, slope);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : set_cuda_error("dconv_fwd launch", e);
}

int dconv_dgrad_fast(const float* gy, const float* w, float* gx, int B, int Cin, int H, int W, int Cout, int stride,
                     cudaStream_t st) {
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  const size_t smem = sizeof(float) * Cout * 9 * kOCB;
  if (stride == 1) {
    const int WQ = (W + kPX - 1) / kPX;
    dim3 grid(ceil_div(H * WQ, 128), ceil_div(Cin, kOCB), B);
    dconv_tile_kernel<1, 1><<<grid, 128, smem, st>>>(gy, w, nullptr, gx, B, Cout, Ho, Wo, Cout, Cin, H, W, 0, 0.f);
  } else {
    const int QW = (W + 3) / 4, QH = (H + 1) / 2;
    dim3 grid(ceil_div(QH * QW, 128), ceil_div(Cin, kOCB), B);
    dconv_dgrad_s2_kernel<<<grid, 128, smem, st>>>(gy, w, gx, B, Cin, H, W, Cout, Ho, Wo);
  }
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : set_cuda_error("dconv_dgrad launch", e);
}

template <int TO, int S>
static int wgrad_launch(const float* gy, const float* x, float* dw, float* db, int B, int Cin, int H, int W, int Cout,
                        int Ho, int Wo, cudaStream_t st) {
  const int pairs = (Cout / TO) * Cin;
  if (pairs > 1024) return set_error("dconv_wgrad: too many channel pairs for one block");
  int lanes = 256 / pairs; if (lanes < 1) lanes = 1;
  const int threads = ((pairs * lanes + 31) / 32) * 32;
  // band height: keep the staged tile under ~96 KB
  int TY = Ho;
  auto bytes = [&](int ty) { return sizeof(float) * (size_t(Cout) * ty * Wo + (size_t(Cin) * ((((ty - 1) * S + 3) * (W + 2)) | 1))); };
  while (TY > 1 && bytes(TY) > 96 * 1024) TY = (TY + 1) / 2;
  const size_t smem = bytes(TY);
  if (smem > 200 * 1024) return set_error("dconv_wgrad: tile does not fit shared memory");
  const long long n_tiles = 1LL * B * ((Ho + TY - 1) / TY);
  long long blocks = 2LL * sm_count(); if (blocks > n_tiles) blocks = n_tiles;
  const int tpb = int((n_tiles + blocks - 1) / blocks);
  blocks = (n_tiles + tpb - 1) / tpb;
  auto kern = dconv_wgrad_tile_kernel<TO, S>;
  ensure_max_smem(reinterpret_cast<const void*>(kern), 200 * 1024);
  kern<<<(int)blocks, threads, smem, st>>>(gy, x, dw, db, B, Cin, H, W, Cout, Ho, Wo, TY, lanes, tpb);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : set_cuda_error("dconv_wgrad launch", e);
}

int dconv_wgrad_fast(const float* gy, const float* x, float* dw, float* db, int B, int Cin, int H, int W, int Cout,
                     int stride, cudaStream_t st) {
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  cudaMemsetAsync(dw, 0, sizeof(float) * Cout * Cin * 9, st);
  if (db) cudaMemsetAsync(db, 0, sizeof(float) * Cout, st);
  if (Cout % 8 == 0) {
    return stride == 1 ? wgrad_launch<8, 1>(gy, x, dw, db, B, Cin, H, W, Cout, Ho, Wo, st)
                       : wgrad_launch<8, 2>(gy, x, dw, db, B, Cin, H, W, Cout, Ho, Wo, st);
  }
  return stride == 1 ? wgrad_launch<1, 1>(gy, x, dw, db, B, Cin, H, W, Cout, Ho, Wo, st)
                     : wgrad_launch<1, 2>(gy, x, dw, db, B, Cin, H, W, Cout, Ho, Wo, st);
}

}  // namespace srb
