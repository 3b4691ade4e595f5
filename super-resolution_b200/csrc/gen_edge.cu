// K4: the bandwidth-bound ends of the generator.
//   conv_in : models.py:122-123  x -> multiplier * x**power -> conv1 (channels -> F), fp32 NCHW image in,
//             PNHWC bf16 out (written into slot 0 of the first dense block's growth buffer).
//   conv_out: models.py:99,131-135 + out() 114-118: conv3.2 (F -> channels) / multiplier, then the output head
//             (hardshrink / relu / **(1/power)); writes the image `out`, the side output `srs` and the
//             pre-head value z (kept for backward).
//   Their backward kernels, and the PixelShuffle adjoint ("unshuffle") used by the upsampling backward.
// All are CUDA-core kernels: K = 9*channels or N = channels (1 by default) cannot feed a tensor core.
#include <stdlib.h>

#include "common.cuh"
#include "sm100.cuh"

namespace srb {

// ---------------------------------------------------------------- conv_in forward
// one thread = one pixel x 8 output channels (16-byte store)
__global__ void conv_in_fwd_kernel(const float* __restrict__ img, int B, int Cimg, int H, int W, float multiplier,
                                   float power, const float* __restrict__ w, const float* __restrict__ bias, int F,
                                   __nv_bfloat16* __restrict__ y, int y_cstride, int y_coff,
                                   __nv_bfloat16* __restrict__ y2, int y2_cstride, int y2_coff) {
  const int groups = F / 8;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long total = 1LL * B * H * W * groups;
  if (idx >= total) return;
  const int g = int(idx % groups);
  long long r = idx / groups;
  const int px = int(r % W); r /= W;
  const int py = int(r % H);
  const int b = int(r / H);
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = bias[g * 8 + i];
  for (int c = 0; c < Cimg; ++c) {
    const float* ip = img + (static_cast<size_t>(b) * Cimg + c) * H * W;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int yy = py + t / 3 - 1, xx = px + t % 3 - 1;
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      float v = ip[yy * W + xx];
      if (power != 1.f) v = powf(v, power);
      v *= multiplier;
      if (v == 0.f) continue;                                   // jet images are >99% zeros
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] += v * __ldg(w + (static_cast<size_t>(g * 8 + i) * Cimg + c) * 9 + t);
    }
  }
  uint4 o;
  o.x = pack_bf16(acc[0], acc[1]); o.y = pack_bf16(acc[2], acc[3]);
  o.z = pack_bf16(acc[4], acc[5]); o.w = pack_bf16(acc[6], acc[7]);
  const size_t row = (static_cast<size_t>(b) * (H + 1) + py) * (W + 1) + px;
  *reinterpret_cast<uint4*>(y + row * y_cstride + y_coff + g * 8) = o;
  if (y2) *reinterpret_cast<uint4*>(y2 + row * y2_cstride + y2_coff + g * 8) = o;
}

// ---------------------------------------------------------------- conv_in backward (weights only)
// dW[f][c][t] += sum_p g[p][f] * xin[p + s_t][c];  db[f] += sum_p g[p][f].
// One warp walks a contiguous range of pixels with a running (b, y, x) counter (no per-pixel divisions); a lane owns
// one channel pair of the 64-wide slice blockIdx.y, so a pixel's gradient row is one coalesced 128-byte load, four
// pixels' loads in flight per warp.  The image taps are warp-uniform: zero taps (jet images are >99 % zeros) skip their
// FMAs.  Warps of a block meet in shared memory, blocks in global atomics.
template <int CC>                                      // image channels (1..4): sizes the accumulator registers
__global__ void __launch_bounds__(256)
conv_in_bwd_kernel(const float* __restrict__ img, int B, int Cimg, int H, int W, float multiplier,
                   float power, const __nv_bfloat16* __restrict__ g, int g_cstride, int g_coff,
                   const __nv_bfloat16* __restrict__ g2, int g2_cstride, int g2_coff, int F,
                   float* __restrict__ dw, float* __restrict__ db, int pix_per_warp) {
  __shared__ float red[37 * 64];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int fl = lane * 2, f = blockIdx.y * 64 + fl;
  const bool fok = f < F;
  for (int i = threadIdx.x; i < 37 * 64; i += 256) red[i] = 0.f;
  __syncthreads();
  const long long npix = 1LL * B * H * W;
  const long long p0 = (1LL * blockIdx.x * 8 + warp) * pix_per_warp;
  long long p1 = p0 + pix_per_warp; if (p1 > npix) p1 = npix;
  float acc[CC][9][2];
#pragma unroll
  for (int c = 0; c < CC; ++c)
#pragma unroll
    for (int t = 0; t < 9; ++t) acc[c][t][0] = acc[c][t][1] = 0.f;
  float bs0 = 0.f, bs1 = 0.f;
  if (p0 < p1) {
    int px = int(p0 % W), py = int((p0 / W) % H), b = int(p0 / (1LL * W * H));
    const int n = int(p1 - p0);
    for (int i0 = 0; i0 < n; i0 += 4) {
      uint32_t gw[4], gw2[4];
      int bb[4], yy[4], xx[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        bb[j] = b; yy[j] = py; xx[j] = px;
        const size_t row = (static_cast<size_t>(b) * (H + 1) + py) * (W + 1) + px;
        const bool ok = fok && i0 + j < n;
        gw[j] = ok ? __ldg(reinterpret_cast<const uint32_t*>(g + row * g_cstride + g_coff + f)) : 0u;
        gw2[j] = (ok && g2) ? __ldg(reinterpret_cast<const uint32_t*>(g2 + row * g2_cstride + g2_coff + f)) : 0u;
        if (++px == W) { px = 0; if (++py == H) { py = 0; ++b; } }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (i0 + j >= n) break;
        const float g0 = bf16lo(gw[j]) + bf16lo(gw2[j]), g1 = bf16hi(gw[j]) + bf16hi(gw2[j]);
        bs0 += g0; bs1 += g1;
#pragma unroll
        for (int c = 0; c < CC; ++c) {
          const float* ip = img + (static_cast<size_t>(bb[j]) * Cimg + c) * H * W;
#pragma unroll
          for (int t = 0; t < 9; ++t) {
            const int y2 = yy[j] + t / 3 - 1, x2 = xx[j] + t % 3 - 1;
            if (y2 < 0 || y2 >= H || x2 < 0 || x2 >= W) continue;
            float v = __ldg(ip + y2 * W + x2);
            if (v == 0.f && power > 0.f) continue;                 // warp-uniform: every lane reads the same tap
            if (power != 1.f) v = powf(v, power);
            v *= multiplier;
            acc[c][t][0] += g0 * v; acc[c][t][1] += g1 * v;
          }
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < CC; ++c) {
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      if (acc[c][t][0] != 0.f) atomicAdd(&red[(c * 9 + t) * 64 + fl], acc[c][t][0]);
      if (acc[c][t][1] != 0.f) atomicAdd(&red[(c * 9 + t) * 64 + fl + 1], acc[c][t][1]);
    }
  }
  atomicAdd(&red[36 * 64 + fl], bs0);
  atomicAdd(&red[36 * 64 + fl + 1], bs1);
  __syncthreads();
  for (int i = threadIdx.x; i < 37 * 64; i += 256) {
    const int k = i >> 6, fo = blockIdx.y * 64 + (i & 63);
    const float s = red[i];
    if (fo >= F || s == 0.f) continue;
    if (k == 36) atomicAdd(db + fo, s);
    else if (k < CC * 9) atomicAdd(dw + (static_cast<size_t>(fo) * Cimg + k / 9) * 9 + k % 9, s);
  }
}

// ---------------------------------------------------------------- conv_out forward + head
__device__ __forceinline__ float hardshrink(float v, float l) { return (v > l || v < -l) ? v : 0.f; }

// one warp = one output pixel (all image channels, normally 1); lane = channel pair of the 64-wide input
__global__ void __launch_bounds__(256)
conv_out_fwd_kernel(const __nv_bfloat16* __restrict__ x, int x_cstride, int x_coff, int B, int H, int W,
                    int F, const float* __restrict__ w, const float* __restrict__ bias, int Cimg,
                    float inv_multiplier, float power, float thres, int training,
                    float* __restrict__ out, float* __restrict__ srs, float* __restrict__ z) {
  // Eight lanes per output pixel, each lane owns 8 consecutive channels of every 64-channel block: one tap of one
  // pixel is a single 128-byte row segment read with 16-byte loads (the 3x3 neighbourhood is re-read from L1/L2).
  // Weights are staged once per block as fp32 [oc][t][c].
  extern __shared__ float wsm[];
  for (int i = threadIdx.x; i < Cimg * 9 * F; i += blockDim.x) {
    const int c = i % F, t = (i / F) % 9, oc = i / (9 * F);
    wsm[i] = w[(static_cast<size_t>(oc) * F + c) * 9 + t];
  }
  __syncthreads();
  const int sub = threadIdx.x & 7;
  const long long pix = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 3;
  const long long npix = 1LL * B * H * W;
  const bool live = pix < npix;
  const long long pc = live ? pix : npix - 1;
  const int px = int(pc % W); const int py = int((pc / W) % H); const int b = int(pc / (1LL * W * H));
  const int pitch = W + 1;
  const long long row = (static_cast<long long>(b) * (H + 1) + py) * pitch + px;
  const long long R = 1LL * B * (H + 1) * pitch;
  for (int oc = 0; oc < Cimg; ++oc) {
    float acc = 0.f;
    for (int c0 = sub * 8; c0 < F; c0 += 64) {
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const long long rs = row + (t / 3 - 1) * pitch + (t % 3 - 1);
        if (rs < 0 || rs >= R) continue;                 // zero rows / columns of PNHWC give the padding
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(x + rs * x_cstride + x_coff + c0));
        const float4 w0 = *reinterpret_cast<const float4*>(wsm + (oc * 9 + t) * F + c0);
        const float4 w1 = *reinterpret_cast<const float4*>(wsm + (oc * 9 + t) * F + c0 + 4);
        acc += bf16lo(v.x) * w0.x + bf16hi(v.x) * w0.y + bf16lo(v.y) * w0.z + bf16hi(v.y) * w0.w +
               bf16lo(v.z) * w1.x + bf16hi(v.z) * w1.y + bf16lo(v.w) * w1.z + bf16hi(v.w) * w1.w;
      }
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    if (sub == 0 && live) {
      const float zz = (acc + bias[oc]) * inv_multiplier;
      const size_t oi = ((static_cast<size_t>(b) * Cimg + oc) * H + py) * W + px;
      if (z) z[oi] = zz;
      // models.py:132  srs = out(z, power): hardshrink(z or relu(z), thres**power)
      const float lam_p = powf(thres, power);
      float s = training ? zz : fmaxf(zz, 0.f);
      srs[oi] = hardshrink(s, lam_p);
      float o2 = zz;
      if (power != 1.f) o2 = powf(fmaxf(zz, 0.f), 1.f / power);      // models.py:133-134
      if (!training) o2 = fmaxf(o2, 0.f);
      out[oi] = hardshrink(o2, thres);                               // models.py:135
    }
  }
}

// gz[b][c][y][x] = g_out * d out/dz + g_srs * d srs/dz   (training-mode head, models.py:114-116,131-135)
__global__ void head_bwd_kernel(const float* __restrict__ z, const float* __restrict__ g_out,
                                const float* __restrict__ g_srs, long long n, float inv_multiplier, float power,
                                float thres, int training, float* __restrict__ gz) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float zz = z[i];
  float g = 0.f;
  if (g_srs) {
    const float lam_p = powf(thres, power);
    const float s = training ? zz : fmaxf(zz, 0.f);
    const bool pass = (s > lam_p || s < -lam_p) && (training || zz > 0.f);
    if (pass) g += g_srs[i];
  }
  if (g_out) {
    float o2 = zz, d = 1.f;
    if (power != 1.f) {
      if (zz > 0.f) { o2 = powf(zz, 1.f / power); d = o2 / (power * zz); } else { o2 = 0.f; d = 0.f; }
    }
    if (!training && o2 <= 0.f) d = 0.f;
    if (o2 > thres || o2 < -thres) g += g_out[i] * d;
  }
  gz[i] = g * inv_multiplier;     // z = conv / multiplier
}

// conv_out input gradient: gy[p][c] = (sum_t w[oc][c][t] * gz[p - s_t]) * lrelu'(act[p][c]); PNHWC bf16 out.
// one thread = one pixel x 8 channels
__global__ void conv_out_dgrad_kernel(const float* __restrict__ gz, int B, int Cimg, int H, int W,
                                      const float* __restrict__ w, int F, const __nv_bfloat16* __restrict__ act,
                                      int act_cstride, int act_coff, float slope, __nv_bfloat16* __restrict__ gy,
                                      int gy_cstride, int gy_coff) {
  const int groups = F / 8;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long total = 1LL * B * H * W * groups;
  if (idx >= total) return;
  const int g = int(idx % groups);
  long long r = idx / groups;
  const int px = int(r % W); r /= W;
  const int py = int(r % H);
  const int b = int(r / H);
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int oc = 0; oc < Cimg; ++oc) {
    const float* gp = gz + (static_cast<size_t>(b) * Cimg + oc) * H * W;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int yy = py - (t / 3 - 1), xx = px - (t % 3 - 1);
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      const float v = gp[yy * W + xx];
      if (v == 0.f) continue;
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] += v * __ldg(w + (static_cast<size_t>(oc) * F + g * 8 + i) * 9 + t);
    }
  }
  const size_t row = (static_cast<size_t>(b) * (H + 1) + py) * (W + 1) + px;
  const uint4 a = *reinterpret_cast<const uint4*>(act + row * act_cstride + act_coff + g * 8);
  const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
  uint32_t ow[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float m0 = bf16lo(aw[j]) > 0.f ? 1.f : slope, m1 = bf16hi(aw[j]) > 0.f ? 1.f : slope;
    ow[j] = pack_bf16(acc[2 * j] * m0, acc[2 * j + 1] * m1);
  }
  *reinterpret_cast<uint4*>(gy + row * gy_cstride + gy_coff + g * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
}

// conv_out weight gradient: dw[oc][c][t] += sum_p gz[p] * act[p + s_t][c]; db[oc] += sum gz.
// block = 256 threads: lane pair-of-channels (32) x 8 pixel lanes; each block reduces a pixel chunk.
__global__ void conv_out_wgrad_kernel(const float* __restrict__ gz, int B, int Cimg, int H, int W,
                                      const __nv_bfloat16* __restrict__ act, int act_cstride, int act_coff, int F,
                                      float* __restrict__ dw, float* __restrict__ db, int pix_per_block) {
  const int cp = threadIdx.x & 31, pl = threadIdx.x >> 5;
  const int c0 = blockIdx.y * 64 + cp * 2;
  const long long npix = 1LL * B * H * W;
  const long long p0 = 1LL * blockIdx.x * pix_per_block;
  long long p1 = p0 + pix_per_block; if (p1 > npix) p1 = npix;
  const int pitch = W + 1;
  const long long R = 1LL * B * (H + 1) * pitch;
  for (int oc = 0; oc < Cimg; ++oc) {
    float a0[9], a1[9], bs = 0.f;
#pragma unroll
    for (int t = 0; t < 9; ++t) { a0[t] = 0.f; a1[t] = 0.f; }
    for (long long pidx = p0 + pl; pidx < p1; pidx += 8) {
      const int px = int(pidx % W); const int py = int((pidx / W) % H); const int b = int(pidx / (1LL * W * H));
      const float gv = gz[((static_cast<size_t>(b) * Cimg + oc) * H + py) * W + px];
      if (gv == 0.f) continue;
      bs += gv;
      const long long row = (static_cast<long long>(b) * (H + 1) + py) * pitch + px;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const long long rs = row + (t / 3 - 1) * pitch + (t % 3 - 1);
        if (rs < 0 || rs >= R) continue;
        const uint32_t v = *reinterpret_cast<const uint32_t*>(act + rs * act_cstride + act_coff + c0);
        a0[t] += gv * bf16lo(v); a1[t] += gv * bf16hi(v);
      }
    }
    __shared__ float red[8][64][9];
    __shared__ float redb[8];
#pragma unroll
    for (int t = 0; t < 9; ++t) { red[pl][cp * 2][t] = a0[t]; red[pl][cp * 2 + 1][t] = a1[t]; }
    if (cp == 0) redb[pl] = bs;
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * 9; i += blockDim.x) {
      const int c = i / 9, t = i % 9;
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) s += red[k][c][t];
      if (s != 0.f && blockIdx.y * 64 + c < F) atomicAdd(dw + (static_cast<size_t>(oc) * F + blockIdx.y * 64 + c) * 9 + t, s);
    }
    if (threadIdx.x == 0 && blockIdx.y == 0) {
      float s = 0.f;
      for (int k = 0; k < 8; ++k) s += redb[k];
      atomicAdd(db + oc, s);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------- PixelShuffle adjoint
// src: PNHWC (2H,2W) with C channels (gradient at the shuffled resolution, already masked);
// dst: PNHWC (H,W) with 4C channels in packed order n = (2i+j)*C + c.  One thread = 8 channels.
__global__ void unshuffle_kernel(const __nv_bfloat16* __restrict__ src, int s_cstride, int s_coff, int B, int H, int W,
                                 int Cc, __nv_bfloat16* __restrict__ dst, int d_cstride, int d_coff) {
  const int groups = Cc / 8;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long total = 1LL * B * H * W * 4 * groups;
  if (idx >= total) return;
  const int g = int(idx % groups);
  long long r = idx / groups;
  const int sub = int(r % 4); r /= 4;
  const int px = int(r % W); r /= W;
  const int py = int(r % H);
  const int b = int(r / H);
  const size_t srow = (static_cast<size_t>(b) * (2 * H + 1) + 2 * py + (sub >> 1)) * (2 * W + 1) + 2 * px + (sub & 1);
  const size_t drow = (static_cast<size_t>(b) * (H + 1) + py) * (W + 1) + px;
  *reinterpret_cast<uint4*>(dst + drow * d_cstride + d_coff + sub * Cc + g * 8) =
      *reinterpret_cast<const uint4*>(src + srow * s_cstride + s_coff + g * 8);
}

// ---------------------------------------------------------------- Dropout2d backward helper
// x: PNHWC slot of C channels (the gradient w.r.t. a dense block's output).  copy <- x (unmasked, for the identity
// paths), x <- x * m[image][channel] in place (what the fifth convolution's gradients see, models.py:39-40).
// One thread = 8 channels of one valid pixel.
__global__ void chan_mask_split_kernel(__nv_bfloat16* __restrict__ x, int x_cstride, int x_coff,
                                       __nv_bfloat16* __restrict__ copy, int c_cstride, int c_coff,
                                       const float* __restrict__ m, int B, int H, int W, int Cc) {
  const int groups = Cc / 8;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long total = 1LL * B * H * W * groups;
  if (idx >= total) return;
  const int g = int(idx % groups);
  long long r = idx / groups;
  const int px = int(r % W); r /= W;
  const int py = int(r % H);
  const int b = int(r / H);
  const size_t row = (static_cast<size_t>(b) * (H + 1) + py) * (W + 1) + px;
  uint4* xp = reinterpret_cast<uint4*>(x + row * x_cstride + x_coff + g * 8);
  const uint4 v = *xp;
  *reinterpret_cast<uint4*>(copy + row * c_cstride + c_coff + g * 8) = v;
  const float4 m0 = __ldg(reinterpret_cast<const float4*>(m + static_cast<size_t>(b) * Cc + g * 8));
  const float4 m1 = __ldg(reinterpret_cast<const float4*>(m + static_cast<size_t>(b) * Cc + g * 8) + 1);
  uint4 o;
  o.x = pack_bf16(bf16lo(v.x) * m0.x, bf16hi(v.x) * m0.y);
  o.y = pack_bf16(bf16lo(v.y) * m0.z, bf16hi(v.y) * m0.w);
  o.z = pack_bf16(bf16lo(v.z) * m1.x, bf16hi(v.z) * m1.y);
  o.w = pack_bf16(bf16lo(v.w) * m1.z, bf16hi(v.w) * m1.w);
  *xp = o;
}

}  // namespace srb

// ================================================================== C ABI
using namespace srb;
#define SRB_LAUNCH_CHECK(name)                                         \
  do {                                                                 \
    cudaError_t e_ = cudaGetLastError();                               \
    if (e_ != cudaSuccess) return set_cuda_error(name " launch", e_);  \
  } while (0)

extern "C" int srb_conv_in_fwd(const float* img, int B, int Cimg, int H, int W, float multiplier, float power,
                               const float* w, const float* bias, int F, void* y, int y_cstride, int y_coff, void* y2,
                               int y2_cstride, int y2_coff, void* stream) {
  if (F % 8 || (y_cstride % 8) || (y_coff % 8)) return set_error("conv_in: F and channel offsets must be multiples of 8");
  const long long total = 1LL * B * H * W * (F / 8);
  conv_in_fwd_kernel<<<ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(
      img, B, Cimg, H, W, multiplier, power, w, bias, F, (__nv_bfloat16*)y, y_cstride, y_coff, (__nv_bfloat16*)y2,
      y2_cstride, y2_coff);
  SRB_LAUNCH_CHECK("conv_in_fwd");
  return 0;
}

extern "C" int srb_conv_in_bwd(const float* img, int B, int Cimg, int H, int W, float multiplier, float power,
                               const void* g, int g_cstride, int g_coff, const void* g2, int g2_cstride, int g2_coff,
                               int F, float* dw, float* db, void* stream) {
  if (Cimg > 4) return set_error("conv_in_bwd: at most 4 image channels");
  cudaMemsetAsync(dw, 0, sizeof(float) * F * Cimg * 9, (cudaStream_t)stream);
  cudaMemsetAsync(db, 0, sizeof(float) * F, (cudaStream_t)stream);
  const long long npix = 1LL * B * H * W;
  if (F % 2) return set_error("conv_in_bwd: F must be even");
  const long long warps = 3LL * sm_count() * 8;          // 80 registers per thread (one image channel): 3 blocks per SM
  int ppw = ceil_div(npix, warps); if (ppw < 16) ppw = 16;
  dim3 grid(ceil_div(npix, 8LL * ppw), (F + 63) / 64);
#define SRB_CIB(CC) conv_in_bwd_kernel<CC><<<grid, 256, 0, (cudaStream_t)stream>>>(img, B, Cimg, H, W, multiplier, power, \
      (const __nv_bfloat16*)g, g_cstride, g_coff, (const __nv_bfloat16*)g2, g2_cstride, g2_coff, F, dw, db, ppw)
  if (Cimg == 1) SRB_CIB(1); else if (Cimg == 2) SRB_CIB(2); else if (Cimg == 3) SRB_CIB(3); else SRB_CIB(4);
#undef SRB_CIB
  SRB_LAUNCH_CHECK("conv_in_bwd");
  return 0;
}

// ---- conv3.2 on the warp-MMA kernels of csrc/disc_mma.cu (single image channel, 64 dense feature channels) ----
// bf16 weight packs in the layouts those kernels read:  forward [9][8][64] (row 0 = the one output channel),
// input gradient [9][64][16] (the 1-channel fp32 gradient is column 0; taps flipped).
__device__ __nv_bfloat16 g_c32_wfwd[9 * 8 * 64];
__device__ __nv_bfloat16 g_c32_wdg[9 * 64 * 16];

__global__ void conv_out_pack_kernel(const float* __restrict__ w) {     // w: fp32 [1][64][3][3]
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 9 * 8 * 64 + 9 * 64 * 16; i += gridDim.x * blockDim.x) {
    if (i < 9 * 8 * 64) {
      const int k = i % 64, r = (i / 64) % 8, t = i / (64 * 8);
      g_c32_wfwd[i] = __float2bfloat16(r == 0 ? w[k * 9 + t] : 0.f);
    } else {
      const int j = i - 9 * 8 * 64;
      const int k = j % 16, ci = (j / 16) % 64, t = j / (16 * 64);
      g_c32_wdg[j] = __float2bfloat16(k == 0 ? w[ci * 9 + (8 - t)] : 0.f);
    }
  }
}

// z = raw / multiplier and the output head (models.py:131-135), in place on `out` which holds the raw convolution
__global__ void head_fwd_kernel(float* __restrict__ out, float* __restrict__ srs, float* __restrict__ z, long long n,
                                float inv_multiplier, float power, float thres, int training) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float zz = out[i] * inv_multiplier;
  if (z) z[i] = zz;
  const float lam_p = powf(thres, power);
  const float s = training ? zz : fmaxf(zz, 0.f);
  srs[i] = hardshrink(s, lam_p);
  float o2 = zz;
  if (power != 1.f) o2 = powf(fmaxf(zz, 0.f), 1.f / power);
  if (!training) o2 = fmaxf(o2, 0.f);
  out[i] = hardshrink(o2, thres);
}

// SRB_CONV_OUT_GENERAL=1: conv3.2 on the general warp-MMA kernels of disc_mma.cu (one product per tap) - the round-1 path,
// kept for A/B timing and as the fall-back for geometries conv_out_tc.cu does not take
static bool conv_out_general_path() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("SRB_CONV_OUT_GENERAL"); v = (e && atoi(e) != 0) ? 1 : 0; }
  return v == 1;
}
static bool conv_out_tc_ok(int Cimg, int F, int cstride, int coff, int W) {
  return Cimg == 1 && F == 64 && cstride == 64 && coff == 0 && W >= 8;
}

extern "C" int srb_conv_out_fwd(const void* x, int x_cstride, int x_coff, int B, int H, int W, int F, const float* w,
                                const float* bias, int Cimg, float multiplier, float power, float thres, int training,
                                float* out, float* srs, float* z, void* stream) {
  if (F % 64) return set_error("conv_out: F must be a multiple of 64");
  if ((x_cstride % 8) || (x_coff % 8)) return set_error("conv_out: channel stride/offset must keep 16-byte alignment");
  if (conv_out_tc_ok(Cimg, F, x_cstride, x_coff, W)) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!conv_out_general_path()) {          // taps-in-the-GEMM kernel with the output head fused (csrc/conv_out_tc.cu)
      const int rc = c32_fwd_launch(x, B, H, W, w, bias, 1.f / multiplier, power, thres, training, out, srs, z, st);
      if (rc >= 0) return rc;
    }
    conv_out_pack_kernel<<<8, 256, 0, st>>>(w);
    SRB_LAUNCH_CHECK("conv_out_pack");
    void* wp = nullptr;
    cudaGetSymbolAddress(&wp, g_c32_wfwd);
    if (dmma_conv_pitched(x, nullptr, 64, wp, bias, 1, nullptr, out, nullptr, B, H, W, H, W, H + 1, W + 1, H, W, 1, 0, 0,
                          0.f, stream))
      return 1;
    const long long n = 1LL * B * H * W;
    head_fwd_kernel<<<ceil_div(n, 256), 256, 0, st>>>(out, srs, z, n, 1.f / multiplier, power, thres, training);
    SRB_LAUNCH_CHECK("head_fwd");
    return 0;
  }
  const long long threads = 1LL * B * H * W * 8;
  const size_t wsm = sizeof(float) * Cimg * 9 * F;
  if (wsm > 48 * 1024) return set_error("conv_out: weights do not fit in shared memory");
  conv_out_fwd_kernel<<<ceil_div(threads, 256), 256, wsm, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, x_cstride, x_coff, B, H, W, F, w, bias, Cimg, 1.f / multiplier, power, thres, training,
      out, srs, z);
  SRB_LAUNCH_CHECK("conv_out_fwd");
  return 0;
}

extern "C" int srb_conv_out_bwd(const float* z, const float* g_out, const float* g_srs, int B, int Cimg, int H, int W,
                                float multiplier, float power, float thres, int training, const float* w, int F,
                                const void* act, int act_cstride, int act_coff, float slope, float* gz_scratch,
                                void* gy, int gy_cstride, int gy_coff, float* dw, float* db, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const long long n = 1LL * B * Cimg * H * W;
  head_bwd_kernel<<<ceil_div(n, 256), 256, 0, st>>>(z, g_out, g_srs, n, 1.f / multiplier, power, thres, training,
                                                    gz_scratch);
  SRB_LAUNCH_CHECK("head_bwd");
  if (conv_out_tc_ok(Cimg, F, act_cstride, act_coff, W) && gy_cstride == 64 && gy_coff == 0) {
    if (!conv_out_general_path()) {          // input gradient + weight gradient in one pass over the activations
      cudaMemsetAsync(dw, 0, sizeof(float) * F * 9, st);
      cudaMemsetAsync(db, 0, sizeof(float), st);
      return c32_bwd_launch(gz_scratch, w, act, gy, dw, db, B, H, W, slope, st);
    }
    conv_out_pack_kernel<<<8, 256, 0, st>>>(w);
    SRB_LAUNCH_CHECK("conv_out_pack");
    void* wp = nullptr;
    cudaGetSymbolAddress(&wp, g_c32_wdg);
    // input gradient = 3x3 convolution of the 1-channel gradient with the flipped taps, times LeakyReLU'(act)
    if (dmma_conv_pitched(nullptr, gz_scratch, 1, wp, nullptr, 64, gy, nullptr, act, B, H, W, H, W, H, W, H + 1, W + 1, 1,
                          0, 2, slope, stream))
      return 1;
    cudaMemsetAsync(dw, 0, sizeof(float) * F * 9, st);
    cudaMemsetAsync(db, 0, sizeof(float), st);
    return dmma_wgrad_pitched(nullptr, gz_scratch, 1, act, 64, dw, db, B, H, W, H + 1, W + 1, H, W, 1, stream);
  }
  const long long total = 1LL * B * H * W * (F / 8);
  conv_out_dgrad_kernel<<<ceil_div(total, 256), 256, 0, st>>>(gz_scratch, B, Cimg, H, W, w, F,
                                                              (const __nv_bfloat16*)act, act_cstride, act_coff, slope,
                                                              (__nv_bfloat16*)gy, gy_cstride, gy_coff);
  SRB_LAUNCH_CHECK("conv_out_dgrad");
  cudaMemsetAsync(dw, 0, sizeof(float) * Cimg * F * 9, st);
  cudaMemsetAsync(db, 0, sizeof(float) * Cimg, st);
  const long long npix = 1LL * B * H * W;
  int ppb = ceil_div(npix, 4 * sm_count()); if (ppb < 64) ppb = 64;
  dim3 grid(ceil_div(npix, ppb), F / 64);
  conv_out_wgrad_kernel<<<grid, 256, 0, st>>>(gz_scratch, B, Cimg, H, W, (const __nv_bfloat16*)act, act_cstride,
                                              act_coff, F, dw, db, ppb);
  SRB_LAUNCH_CHECK("conv_out_wgrad");
  return 0;
}

extern "C" int srb_unshuffle2(const void* src, int s_cstride, int s_coff, int B, int H, int W, int C, void* dst,
                              int d_cstride, int d_coff, void* stream) {
  if (C % 8) return set_error("unshuffle: C must be a multiple of 8");
  const long long total = 1LL * B * H * W * 4 * (C / 8);
  unshuffle_kernel<<<ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)src, s_cstride, s_coff, B, H, W, C, (__nv_bfloat16*)dst, d_cstride, d_coff);
  SRB_LAUNCH_CHECK("unshuffle");
  return 0;
}

extern "C" int srb_chan_mask_split(void* x, int x_cstride, int x_coff, void* copy, int c_cstride, int c_coff,
                                   const float* mask, int B, int H, int W, int C, void* stream) {
  if (!x || !copy || !mask) return set_error("chan_mask_split: null pointer");
  if (C % 8 || x_cstride % 8 || c_cstride % 8 || x_coff % 8 || c_coff % 8)
    return set_error("chan_mask_split: channels / strides / offsets must be multiples of 8");
  const long long total = 1LL * B * H * W * (C / 8);
  chan_mask_split_kernel<<<ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(
      (__nv_bfloat16*)x, x_cstride, x_coff, (__nv_bfloat16*)copy, c_cstride, c_coff, mask, B, H, W, C);
  SRB_LAUNCH_CHECK("chan_mask_split");
  return 0;
}
