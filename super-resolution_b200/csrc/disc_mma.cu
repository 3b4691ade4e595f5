// K5 (tensor-core path): the discriminator's small-channel 3x3 convolutions (models.py:140-146,168) and both
// halves of their backward on warp-level bf16 MMA (mma.sync m16n8k16, fp32 accumulate).
//
// Why not tcgen05 here: the layers are 16/32/64 channels wide on 80x80 ... 5x5 maps (23 MMAC per image for the whole
// stack); they are bound by HBM traffic and launch count, not by the tensor pipe, and the UMMA tile (128 x N x 16 from
// 128B-swizzled smem) would be mostly padding.  The warp-level MMA lets one fragment be ANY 16 output pixels: rows of
// an ldmatrix are per-lane shared-memory addresses, so 3x3 taps, stride 2, the four phases of the stride-2 input
// gradient and image borders are all just address arithmetic on a zero-padded shared-memory tile.
//
// Layout: activations bf16 NHWC [B][H][W][C] (C in {16,32,64}; no padding in global memory), images fp32 [B][H][W].
//   dmma_conv_kernel<KS,NT,MODE>   MODE 0: y = epi(conv_s(x, w) + b)       (forward, stride 1|2; also the stride-1
//                                          input gradient when given the flipped/transposed weight pack)
//                                  MODE 1: stride-2 input gradient (transposed conv) by output phase
//                                  epi: bias, LeakyReLU, or multiply by LeakyReLU'(ref) (the mask of the layer below)
//   dmma_wgrad_kernel<KS,MT>       dW[co][ci][t] += sum_pixels gz[p][co] * a[p*s + t][ci];  db[co] += sum gz
//   dl1_* kernels                  the 1-channel image end of the stack (Cin = 1) on CUDA cores in fp32
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "sm100.cuh"

namespace srb {

__device__ __forceinline__ uint32_t dsmem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2(uint32_t addr, uint32_t& r0, uint32_t& r1) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
               "{%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t dpack_bf16(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float dbf_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float dbf_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

struct DmmaConvP {
  const __nv_bfloat16* in;     // bf16 NHWC source [B][Hs][Ws][cin]          (or in_f32 for a 1-channel fp32 source)
  const float* in_f32;
  const __nv_bfloat16* w;      // bf16 [9][NT*8][KS*16], zero padded
  const float* bias;           // fp32 [cout] or null
  __nv_bfloat16* out;          // bf16 NHWC [B][Hd][Wd][cout] or null
  float* out_f32;              // fp32 [B][Hd][Wd] (cout == 1) or null
  const __nv_bfloat16* ref;    // act == 2: mask reference, geometry of out
  int B, Hs, Ws, Hd, Wd;       // source / destination maps
  int Hsp, Wsp, Hdp, Wdp;      // their pitches in memory (rows, pixels per row); PNHWC buffers carry one zero row/column
  int cin, cout;               // real channel counts (<= KS*16, <= NT*8)
  int stride;                  // MODE 0
  int TH, IMGS;                // tile = IMGS images x TH rows (MODE 0: destination rows; MODE 1: phase-grid rows)
  int nst, tile_bytes;         // ring of nst source-tile buffers, tile_bytes apart (multiple of 1024)
  uint32_t mg_img, mg_w;       // ceil(2^32 / (TH*Wg)), ceil(2^32 / Wg): exact division of fragment-row indices (< 2^16)
  int tiles_y, n_tiles;
  int act;                     // 0 none, 1 LeakyReLU(slope), 2 multiply by (ref > 0 ? 1 : slope)
  float slope;
};

constexpr int kDmmaThreads = 256;
constexpr int kDmmaMaxStages = 4;

// Source tiles arrive by TMA: one 4-D box [IMGS][TR][TC][cin] per tile (halo included, zero-filled outside the image
// and past the last image), dense in shared memory with the hardware swizzle of the pixel size (32/64/128 B), which
// keeps the ldmatrix reads of 8 neighbouring pixels conflict-free.  p.nst tile buffers form a ring: the load of tile
// i + nst - 1 is issued when tile i - 1 has been consumed.
template <int KS, int NT, int MODE>
__global__ void __launch_bounds__(kDmmaThreads)
dmma_conv_kernel(const DmmaConvP p, const __grid_constant__ CUtensorMap tmap) {
  constexpr int PP = KS * 32 + 16;                    // weight row pitch (16 B skew: conflict-free ldmatrix)
  constexpr int P = KS * 32;                          // source pixel pitch (swizzled)
  constexpr int SH = KS == 1 ? 2 : KS == 2 ? 1 : 0;   // swizzle: 16-byte chunk index ^= (pixel >> SH) & SWM
  constexpr int SWM = 2 * KS - 1;
  constexpr int NROWS = NT * 8;
  extern __shared__ __align__(1024) uint8_t dsm[];
  uint8_t* in_s0 = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dsm) + 1023) & ~uintptr_t(1023));
  uint8_t* w_s = in_s0 + p.nst * p.tile_bytes;
  uint8_t* stage_s = w_s + 9 * NROWS * PP;                       // per-warp epilogue staging (NT >= 2)
  uint64_t* bars = reinterpret_cast<uint64_t*>(stage_s + (NT >= 2 ? (kDmmaThreads / 32) * 32 * (NT * 16 + 16) : 0));
  const int S = MODE == 0 ? p.stride : 1;
  const int Wg = MODE == 0 ? p.Wd : p.Wd / 2;         // width of the fragment grid (destination or phase grid)
  const int Hg = MODE == 0 ? p.Hd : p.Hd / 2;
  const int TR = MODE == 0 ? p.TH * S + 2 : p.TH + 1;
  const int TC = MODE == 0 ? Wg * S + 2 : Wg + 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool tma = p.in_f32 == nullptr;

  uint64_t* empty = bars + kDmmaMaxStages;             // "tile consumed by every warp" (TMA path only)
  if (threadIdx.x == 0 && tma) {
    tma_prefetch_desc(&tmap);
    for (int i = 0; i < p.nst; ++i) { mbar_init(&bars[i], 1); mbar_init(&empty[i], kDmmaThreads / 32); }
    fence_mbar_init();
  }
  // weights -> shared memory (once per CTA)
  {
    constexpr int CH = KS * 2;                        // 16-byte chunks per row
    const uint4* src = reinterpret_cast<const uint4*>(p.w);
    for (int i = threadIdx.x; i < 9 * NROWS * CH; i += kDmmaThreads) {
      const int row = i / CH, c = i - row * CH;
      *reinterpret_cast<uint4*>(w_s + row * PP + c * 16) = __ldg(src + i);
    }
  }
  const int M_img = p.TH * Wg;
  const int M = p.IMGS * M_img;                       // fragment rows per tile (per phase for MODE 1)
  const int nfrag = (M + 15) >> 4;
  const int npair = (nfrag + 1) >> 1;
  constexpr int NPH = MODE == 0 ? 1 : 4;

  auto load_tile = [&](int tile, int slot) {
    const int ty_ = tile % p.tiles_y;
    const int img0_ = (tile / p.tiles_y) * p.IMGS;
    const int gy0 = MODE == 0 ? ty_ * p.TH * S - 1 : ty_ * p.TH;
    const int gx0 = MODE == 0 ? -1 : 0;
    uint8_t* dst = in_s0 + slot * p.tile_bytes;
    if (tma) {
      if (threadIdx.x == 0) {
        mbar_expect_tx(&bars[slot], static_cast<uint32_t>(p.IMGS * TR * TC * P));
        tma_load_4d(dst, &tmap, 0, gx0, gy0, img0_, &bars[slot]);
      }
    } else {                                          // single fp32 channel (tiny layers): convert on the way in
      constexpr int CH = KS * 2;
      const int total = p.IMGS * TR * TC * CH;
      for (int i = threadIdx.x; i < total; i += kDmmaThreads) {
        const int n = i / CH, c = i - n * CH;
        int r = n;
        const int tc = r % TC; r /= TC;
        const int tr = r % TR;
        const int im = r / TR;
        const int gy = gy0 + tr, gx = gx0 + tc, b = img0_ + im;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (c == 0 && b < p.B && gy >= 0 && gy < p.Hs && gx >= 0 && gx < p.Ws)
          v.x = dpack_bf16(__ldg(p.in_f32 + (static_cast<size_t>(b) * p.Hs + gy) * p.Ws + gx), 0.f);
        *reinterpret_cast<uint4*>(dst + n * P + ((c ^ ((n >> SH) & SWM)) << 4)) = v;
      }
    }
  };
  float bia[NT][2];                                   // this lane's accumulator columns: n*8 + 2*(lane&3), +1
#pragma unroll
  for (int n = 0; n < NT; ++n) {
    const int col = n * 8 + (lane & 3) * 2;
    bia[n][0] = (NT >= 2 && p.bias) ? __ldg(p.bias + col) : 0.f;
    bia[n][1] = (NT >= 2 && p.bias) ? __ldg(p.bias + col + 1) : 0.f;
  }
  __syncthreads();                                    // barriers initialised (weights: covered by the loop's barrier)
  for (int k = 0; k < p.nst - 1; ++k) {
    const int t = blockIdx.x + k * gridDim.x;
    if (t < p.n_tiles) load_tile(t, k);
  }
  int k = 0;
  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++k) {
    const int ty = tile % p.tiles_y;
    const int img0 = (tile / p.tiles_y) * p.IMGS;
    const int y0 = ty * p.TH;                         // first fragment-grid row of the tile
    const int slot = k % p.nst;
    if (tma) {
      // The warps of a CTA share nothing but the source tiles: no block-wide barrier per tile.  Every warp arrives on the
      // tile's "empty" barrier when it is done with it; only the thread that issues the TMA waits for tile k-1 to be
      // consumed before its buffer takes tile k + nst - 1, the other warps run ahead by up to nst - 1 tiles.
      if (threadIdx.x == 0) {
        const int t = tile + (p.nst - 1) * gridDim.x;
        if (t < p.n_tiles) {
          if (k > 0) mbar_wait(&empty[(k - 1) % p.nst], ((k - 1) / p.nst) & 1, 0x31);
          load_tile(t, (k + p.nst - 1) % p.nst);
        }
      }
      __syncwarp();
      mbar_wait(&bars[slot], (k / p.nst) & 1, 0x30);
    } else {
      __syncthreads();                                // tile k-1 fully consumed: its buffer takes tile k + nst - 1
      const int t = tile + (p.nst - 1) * gridDim.x;
      if (t < p.n_tiles) load_tile(t, (k + p.nst - 1) % p.nst);
    }
    const uint32_t tile_u32 = dsmem_u32(in_s0 + slot * p.tile_bytes);

    for (int it = warp; it < NPH * npair; it += kDmmaThreads / 32) {
      const int ph = MODE == 0 ? 0 : it / npair;
      const int pr = MODE == 0 ? it : it - ph * npair;
      const int py = ph >> 1, px = ph & 1;
      float acc[2][NT][4];                            // NT >= 2: starts from the bias (the epilogue adds nothing)
#pragma unroll
      for (int f = 0; f < 2; ++f)
#pragma unroll
        for (int n = 0; n < NT; ++n)
#pragma unroll
          for (int i = 0; i < 4; ++i) acc[f][n][i] = NT >= 2 ? bia[n][i & 1] : 0.f;
      // per-lane tile pixel (at tap 0,0) of the fragment row this lane addresses, for the two fragments
      int nbase[2];
#pragma unroll
      for (int f = 0; f < 2; ++f) {
        int m = (pr * 2 + f) * 16 + (lane & 15);
        if (m >= M) m = 0;
        const int im = __umulhi(m, p.mg_img);
        const int rem = m - im * M_img;
        const int yl = __umulhi(rem, p.mg_w), x = rem - yl * Wg;
        nbase[f] = (im * TR + yl * S) * TC + x * S;
      }
      const int hi = lane >> 4;
      const bool two = (pr * 2 + 1) < nfrag;
      // destination pixel of staging row `lane` of the epilogue (-1: nothing to store).  It does not depend on the
      // accumulators: with the mask epilogue its reference row goes to L1 now, behind the MMAs, instead of being a
      // ~1 us dependent load after them (these layers are latency-bound: 16 warps per SM).
      int mypix = -1;
      if constexpr (NT >= 2) {
        const int m = pr * 32 + lane;
        if (m < M && (lane < 16 || two)) {
          const int im = __umulhi(m, p.mg_img);
          const int rem = m - im * M_img;
          const int yl = __umulhi(rem, p.mg_w), x = rem - yl * Wg;
          const int b = img0 + im, yg = y0 + yl;
          if (b < p.B && yg < Hg) {
            const int yd = MODE == 0 ? yg : 2 * yg + py;
            const int xd = MODE == 0 ? x : 2 * x + px;
            mypix = (b * p.Hdp + yd) * p.Wdp + xd;
          }
        }
        if (p.act == 2 && mypix >= 0)
          asm volatile("prefetch.global.L1 [%0];" ::"l"(reinterpret_cast<const uint4*>(p.ref) +
                                                        static_cast<size_t>(static_cast<uint32_t>(mypix)) * NT));
      }
      const uint32_t wlane = dsmem_u32(w_s) + ((lane & 7) + ((lane >> 4) << 3)) * PP + ((lane >> 3) & 1) * 16;
      const uint32_t wlane2 = dsmem_u32(w_s) + (lane & 7) * PP + ((lane >> 3) & 1) * 16;   // x2 form (NT == 1)
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
        if (MODE == 1 && ((ky ^ py) & 1) == 0) continue;        // needs ky == py + 1 (mod 2)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          if (MODE == 1 && ((kx ^ px) & 1) == 0) continue;
          int noff;
          if (MODE == 0) noff = ky * TC + kx;
          else noff = ((py == 1 && ky == 0) ? TC : 0) + ((px == 1 && kx == 0) ? 1 : 0);
          const int wt = (ky * 3 + kx) * NROWS * PP;
          const int n0 = nbase[0] + noff, n1 = nbase[1] + noff;
          const uint32_t row0 = tile_u32 + n0 * P, row1 = tile_u32 + n1 * P;
          const int sw0 = (n0 >> SH) & SWM, sw1 = (n1 >> SH) & SWM;
#pragma unroll
          for (int kk = 0; kk < KS; ++kk) {
            uint32_t a0[4], a1[4];
            ldsm_x4(row0 + (((kk * 2 + hi) ^ sw0) << 4), a0);
            if (two) ldsm_x4(row1 + (((kk * 2 + hi) ^ sw1) << 4), a1);
            if (NT == 1) {
              uint32_t b0, b1;
              ldsm_x2(wlane2 + wt + kk * 32, b0, b1);
              mma_bf16_16816(acc[0][0], a0, b0, b1);
              if (two) mma_bf16_16816(acc[1][0], a1, b0, b1);
            } else {
#pragma unroll
              for (int np = 0; np < NT / 2; ++np) {
                uint32_t b[4];
                ldsm_x4(wlane + wt + np * 16 * PP + kk * 32, b);
                mma_bf16_16816(acc[0][2 * np], a0, b[0], b[1]);
                mma_bf16_16816(acc[0][2 * np + 1], a0, b[2], b[3]);
                if (two) {
                  mma_bf16_16816(acc[1][2 * np], a1, b[0], b[1]);
                  mma_bf16_16816(acc[1][2 * np + 1], a1, b[2], b[3]);
                }
              }
            }
          }
        }
      }
      const int g = lane >> 2, t2 = (lane & 3) * 2;
      if constexpr (NT >= 2) {
        // ---- epilogue through a per-warp staging tile: fragments are scattered 4-byte pieces, global memory wants
        // whole pixel rows.  Values go to shared memory as bf16 [32 rows][COUT] (+16 B skew), then each lane moves
        // 16-byte chunks: consecutive lanes cover consecutive channels of consecutive pixels (coalesced stores and,
        // for the mask epilogue, coalesced reads of the reference activation).
        constexpr int SP = NT * 16 + 16;
        constexpr int RPI = 32 / NT;                  // staging rows covered by the warp per 16-byte pass
        const uint32_t stg = dsmem_u32(stage_s) + warp * (32 * SP);
        if (p.act == 1) {                             // LeakyReLU, slope in (0, 1): max(v, slope * v)
#pragma unroll
          for (int f = 0; f < 2; ++f)
#pragma unroll
            for (int n = 0; n < NT; ++n)
#pragma unroll
              for (int i = 0; i < 4; ++i) acc[f][n][i] = fmaxf(acc[f][n][i], acc[f][n][i] * p.slope);
        }
        {
          const uint32_t mine = stg + g * SP + t2 * 2;
#pragma unroll
          for (int f = 0; f < 2; ++f)
#pragma unroll
            for (int hrow = 0; hrow < 2; ++hrow)
#pragma unroll
              for (int n = 0; n < NT; ++n)
                asm volatile("st.shared.b32 [%0], %1;" ::"r"(mine + (f * 16 + hrow * 8) * SP + n * 16),
                             "r"(dpack_bf16(acc[f][n][hrow * 2], acc[f][n][hrow * 2 + 1])) : "memory");
        }
        __syncwarp();
        const int part = lane % NT, r0 = lane / NT;
        const bool masked = p.act == 2;
#pragma unroll
        for (int i = 0; i < NT; ++i) {
          const int r = r0 + i * RPI;
          const int pix = __shfl_sync(0xffffffffu, mypix, r);
          if (pix < 0) continue;
          uint4 v;
          asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                       : "r"(stg + r * SP + part * 16));
          const size_t off = static_cast<size_t>(static_cast<uint32_t>(pix) * static_cast<uint32_t>(NT)) + part;   // in 16-byte units
          if (masked) {
            const uint4 m4 = __ldg(reinterpret_cast<const uint4*>(p.ref) + off);
            const uint32_t mw[4] = {m4.x, m4.y, m4.z, m4.w};
            uint32_t vw[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int q = 0; q < 4; ++q)
              vw[q] = dpack_bf16(dbf_lo(vw[q]) * (dbf_lo(mw[q]) > 0.f ? 1.f : p.slope),
                                 dbf_hi(vw[q]) * (dbf_hi(mw[q]) > 0.f ? 1.f : p.slope));
            v = make_uint4(vw[0], vw[1], vw[2], vw[3]);
          }
          *(reinterpret_cast<uint4*>(p.out) + off) = v;
        }
        __syncwarp();
      } else {
      // ---- direct epilogue (single output channel): rows g and g+8 of each fragment, columns nt*8 + 2t, +1
#pragma unroll
      for (int f = 0; f < 2; ++f) {
        if (f == 1 && !two) break;
#pragma unroll
        for (int hrow = 0; hrow < 2; ++hrow) {
          const int m = (pr * 2 + f) * 16 + g + hrow * 8;
          if (m >= M) continue;
          const int im = m / M_img;
          const int rem = m - im * M_img;
          const int yl = rem / Wg, x = rem - yl * Wg;
          const int b = img0 + im, yg = y0 + yl;
          if (b >= p.B || yg >= Hg) continue;
          const int yd = MODE == 0 ? yg : 2 * yg + py;
          const int xd = MODE == 0 ? x : 2 * x + px;
          const size_t pix = (static_cast<size_t>(b) * p.Hdp + yd) * p.Wdp + xd;
#pragma unroll
          for (int n = 0; n < NT; ++n) {
            const int col = n * 8 + t2;
            if (col >= p.cout) continue;
            float v0 = acc[f][n][hrow * 2], v1 = acc[f][n][hrow * 2 + 1];
            if (p.bias) {
              v0 += __ldg(p.bias + col);
              if (col + 1 < p.cout) v1 += __ldg(p.bias + col + 1);
            }
            if (p.act == 1) {
              v0 = v0 > 0.f ? v0 : v0 * p.slope;
              v1 = v1 > 0.f ? v1 : v1 * p.slope;
            } else if (p.act == 2) {
              if (p.cout == 1) {
                v0 *= __bfloat162float(p.ref[pix]) > 0.f ? 1.f : p.slope;
              } else {
                const uint32_t r = __ldg(reinterpret_cast<const uint32_t*>(p.ref + pix * p.cout + col));
                v0 *= dbf_lo(r) > 0.f ? 1.f : p.slope;
                v1 *= dbf_hi(r) > 0.f ? 1.f : p.slope;
              }
            }
            if (p.cout == 1) {
              if (p.out_f32) p.out_f32[pix] = v0;
              if (p.out) p.out[pix] = __float2bfloat16(v0);
            } else {
              *reinterpret_cast<uint32_t*>(p.out + pix * p.cout + col) = dpack_bf16(v0, v1);
            }
          }
        }
      }
      }
    }
    if (tma) {                                        // this warp is done with the tile (its ldmatrix reads are complete)
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
    }
  }
}

// ---------------------------------------------------------------------------------------------- weight gradient
struct DmmaWgradP {
  const __nv_bfloat16* gz;     // bf16 NHWC [B][Ho][Wo][cout]   (or gz_f32 for a 1-channel fp32 gradient)
  const float* gz_f32;
  const __nv_bfloat16* a;      // bf16 NHWC [B][Hi][Wi][cin]
  float* dw;                   // fp32 [cout][cin][3][3], accumulated with atomics
  float* db;                   // fp32 [cout] or null
  int B, Hi, Wi, Ho, Wo, cin, cout, stride;
  int co_off;                  // first output channel handled by this launch (MT*16 channels from there)
  int TH, IMGS, tiles_y, n_tiles;
  int Hip, Wip;                // pitches of a in memory (rows, pixels per row)
  int nst, a_bytes, g_bytes;   // ring of nst stages; activation / gradient tile sizes (multiples of 1024)
};

constexpr int kDwgThreads = 288;   // nine warps: one per filter tap
constexpr int kDwgMaxStages = 3;

// Both operand tiles arrive by TMA into a ring of p.nst stages (see dmma_conv_kernel): the activation tile with its
// halo as a [IMGS][TR][TC][cin] box, the gradient tile as a [IMGS][TH][Wo][MT*16] box whose rows are the fragment-row
// index m; both dense with the hardware swizzle of their pixel size.
template <int KS, int MT>
__global__ void __launch_bounds__(kDwgThreads)
dmma_wgrad_kernel(const DmmaWgradP p, const __grid_constant__ CUtensorMap tmap_a,
                  const __grid_constant__ CUtensorMap tmap_g) {
  constexpr int PA = KS * 32, PG = MT * 32;
  constexpr int SHA = KS == 1 ? 2 : KS == 2 ? 1 : 0, SWA = 2 * KS - 1;
  constexpr int SHG = MT == 1 ? 2 : MT == 2 ? 1 : 0, SWG = 2 * MT - 1;
  constexpr int NT = KS * 2;
  extern __shared__ __align__(1024) uint8_t dsm[];
  uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dsm) + 1023) & ~uintptr_t(1023));
  const int stage_bytes = p.a_bytes + p.g_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + p.nst * stage_bytes);
  int* ntab = reinterpret_cast<int*>(bars + kDwgMaxStages);      // [M16] activation-tile pixel of fragment row m, tap (0,0)
  const int S = p.stride;
  const int TR = p.TH * S + 2, TC = p.Wo * S + 2;
  const int M_img = p.TH * p.Wo;
  const int M = p.IMGS * M_img;
  const int M16 = (M + 15) & ~15;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ky = warp / 3, kx = warp - ky * 3;
  const bool g_tma = p.gz_f32 == nullptr;
  // the tile geometry is the same for every tile: do the index arithmetic once (rows >= M carry a zero gradient,
  // any valid pixel will do)
  for (int m = threadIdx.x; m < M16; m += kDwgThreads) {
    const int mm = m < M ? m : 0;
    const int im = mm / M_img;
    const int rem = mm - im * M_img;
    const int yl = rem / p.Wo, x = rem - yl * p.Wo;
    ntab[m] = (im * TR + yl * S) * TC + x * S;
  }

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_a);
    if (g_tma) tma_prefetch_desc(&tmap_g);
    for (int i = 0; i < p.nst; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  // fragment rows M .. M16-1 of every gradient buffer stay zero (TMA never writes them)
  for (int i = threadIdx.x; i < p.nst * (M16 - M) * (PG / 16); i += kDwgThreads) {
    const int st = i / ((M16 - M) * (PG / 16)), r = i - st * ((M16 - M) * (PG / 16));
    *reinterpret_cast<uint4*>(ring + st * stage_bytes + p.a_bytes + M * PG + r * 16) = make_uint4(0, 0, 0, 0);
  }

  float acc[MT][NT][4];
#pragma unroll
  for (int m = 0; m < MT; ++m)
#pragma unroll
    for (int n = 0; n < NT; ++n)
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[m][n][i] = 0.f;
  float db0 = 0.f, db1 = 0.f;

  auto load_tile = [&](int tile, int slot) {
    const int ty_ = tile % p.tiles_y;
    const int img0_ = (tile / p.tiles_y) * p.IMGS;
    const int y0_ = ty_ * p.TH;
    uint8_t* a_dst = ring + slot * stage_bytes;
    uint8_t* g_dst = a_dst + p.a_bytes;
    if (threadIdx.x == 0) {
      mbar_expect_tx(&bars[slot], static_cast<uint32_t>(p.IMGS * TR * TC * PA + (g_tma ? M * PG : 0)));
      tma_load_4d(a_dst, &tmap_a, 0, -1, y0_ * S - 1, img0_, &bars[slot]);
      if (g_tma) tma_load_4d(g_dst, &tmap_g, p.co_off, 0, y0_, img0_, &bars[slot]);
    }
    if (!g_tma) {                                      // single fp32 gradient channel (last layer): convert on the way in
      constexpr int CH = MT * 2;
      for (int i = threadIdx.x; i < M * CH; i += kDwgThreads) {
        const int m = i / CH, c = i - m * CH;
        const int im = m / M_img;
        const int rem = m - im * M_img;
        const int yl = rem / p.Wo, x = rem - yl * p.Wo;
        const int b = img0_ + im, y = y0_ + yl;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (c == 0 && b < p.B && y < p.Ho) v.x = dpack_bf16(__ldg(p.gz_f32 + (static_cast<size_t>(b) * p.Ho + y) * p.Wo + x), 0.f);
        *reinterpret_cast<uint4*>(g_dst + m * PG + ((c ^ ((m >> SHG) & SWG)) << 4)) = v;
      }
    }
  };
  __syncthreads();
  for (int k = 0; k < p.nst - 1; ++k) {
    const int t = blockIdx.x + k * gridDim.x;
    if (t < p.n_tiles) load_tile(t, k);
  }
  int k = 0;
  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++k) {
    const int slot = k % p.nst;
    __syncthreads();                                  // tile k-1 fully consumed: its stage takes tile k + nst - 1
    {
      const int t = tile + (p.nst - 1) * gridDim.x;
      if (t < p.n_tiles) load_tile(t, (k + p.nst - 1) % p.nst);
    }
    mbar_wait(&bars[slot], (k / p.nst) & 1, 0x31);
    const uint32_t a_u32 = dsmem_u32(ring + slot * stage_bytes);
    const uint32_t g_u32 = a_u32 + p.a_bytes;
    if (p.db) {                                       // bias gradient: every warp sums its share of the rows
      constexpr int NCP = MT * 8;                     // 4-byte column pairs per row
      const int cp = lane % NCP, ro = lane / NCP;
      for (int m = warp * (32 / NCP) + ro; m < M; m += 9 * (32 / NCP)) {
        uint32_t v;
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v)
                     : "r"(g_u32 + m * PG + (((cp >> 2) ^ ((m >> SHG) & SWG)) << 4) + (cp & 3) * 4));
        db0 += dbf_lo(v); db1 += dbf_hi(v);
      }
    }
    const int kA = (lane & 7) + ((lane >> 4) << 3);
    const int kB = (lane & 7) + (((lane >> 3) & 1) << 3);
    for (int ch = 0; ch < M16 / 16; ++ch) {
      const int n = ntab[ch * 16 + kB] + ky * TC + kx;
      const uint32_t brow = a_u32 + n * PA;
      const int swb = (n >> SHA) & SWA;
      const int r = ch * 16 + kA;
      const uint32_t arow = g_u32 + r * PG;
      const int swa = (r >> SHG) & SWG;
      uint32_t af[MT][4];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) ldsm_x4_t(arow + (((mt * 2 + ((lane >> 3) & 1)) ^ swa) << 4), af[mt]);
#pragma unroll
      for (int np = 0; np < KS; ++np) {
        uint32_t b[4];
        ldsm_x4_t(brow + (((np * 2 + (lane >> 4)) ^ swb) << 4), b);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          mma_bf16_16816(acc[mt][2 * np], af[mt], b[0], b[1]);
          mma_bf16_16816(acc[mt][2 * np + 1], af[mt], b[2], b[3]);
        }
      }
    }
  }
  // ---- flush: dw[co][ci][tap]
  const int g = lane >> 2, t2 = (lane & 3) * 2;
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int n = 0; n < NT; ++n)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int co = p.co_off + mt * 16 + g + (i >> 1) * 8;
        const int ci = n * 8 + t2 + (i & 1);
        if (co < p.cout && ci < p.cin) atomicAdd(p.dw + (static_cast<size_t>(co) * p.cin + ci) * 9 + warp, acc[mt][n][i]);
      }
  if (p.db) {
    constexpr int NCP = MT * 8;
#pragma unroll
    for (int o = NCP; o < 32; o <<= 1) {                // fold the row groups of the warp together
      db0 += __shfl_xor_sync(0xffffffffu, db0, o);
      db1 += __shfl_xor_sync(0xffffffffu, db1, o);
    }
    const int co = p.co_off + lane * 2;
    if (lane < NCP && co < p.cout) {
      atomicAdd(p.db + co, db0);
      if (co + 1 < p.cout) atomicAdd(p.db + co + 1, db1);
    }
  }
}

// ---------------------------------------------------------------------------------------------- image end (Cin = 1)
// y[b][p][c] = epi(bias[c] + sum_t w[c][t] * img[b][p + t]);  one thread per pixel, all C channels (C <= 64, C % 8 == 0)
template <int C>
__global__ void __launch_bounds__(256)
dl1_fwd_kernel(const float* __restrict__ img, const float* __restrict__ w, const float* __restrict__ bias,
               __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ ref, int B, int H, int W, int act,
               float slope) {
  __shared__ float ws[9][C];
  __shared__ float bs[C];
  for (int i = threadIdx.x; i < 9 * C; i += blockDim.x) ws[i % 9][i / 9] = w[i];     // w is [C][1][3][3]
  for (int i = threadIdx.x; i < C; i += blockDim.x) bs[i] = bias ? bias[i] : 0.f;
  __syncthreads();
  const long long total = 1LL * B * H * W;
  for (long long pix = blockIdx.x * (long long)blockDim.x + threadIdx.x; pix < total; pix += (long long)gridDim.x * blockDim.x) {
    const int x = int(pix % W);
    const int y = int((pix / W) % H);
    const float* ip = img + pix;
    float v[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int yy = y + t / 3 - 1, xx = x + t % 3 - 1;
      v[t] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(ip + (t / 3 - 1) * W + (t % 3 - 1)) : 0.f;
    }
    uint4* op = reinterpret_cast<uint4*>(out + pix * C);
    const uint4* rp = reinterpret_cast<const uint4*>(ref + pix * C);
#pragma unroll
    for (int c8 = 0; c8 < C / 8; ++c8) {
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float s = bs[c8 * 8 + j];
#pragma unroll
        for (int t = 0; t < 9; ++t) s += v[t] * ws[t][c8 * 8 + j];
        o[j] = s;
      }
      if (act == 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = o[j] > 0.f ? o[j] : o[j] * slope;
      } else if (act == 2) {
        const uint4 r = __ldg(rp + c8);
        const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          o[2 * j] *= dbf_lo(rw[j]) > 0.f ? 1.f : slope;
          o[2 * j + 1] *= dbf_hi(rw[j]) > 0.f ? 1.f : slope;
        }
      }
      uint4 ov;
      ov.x = dpack_bf16(o[0], o[1]); ov.y = dpack_bf16(o[2], o[3]); ov.z = dpack_bf16(o[4], o[5]); ov.w = dpack_bf16(o[6], o[7]);
      op[c8] = ov;
    }
  }
}

// gx[b][p] = sum_t sum_c gz[b][p - t][c] * w[c][t]   (fp32 image gradient out)
template <int C>
__global__ void __launch_bounds__(256)
dl1_dgrad_kernel(const __nv_bfloat16* __restrict__ gz, const float* __restrict__ w, float* __restrict__ gx, int B,
                 int H, int W) {
  __shared__ float ws[9][C];
  for (int i = threadIdx.x; i < 9 * C; i += blockDim.x) ws[i % 9][i / 9] = w[i];
  __syncthreads();
  const long long total = 1LL * B * H * W;
  for (long long pix = blockIdx.x * (long long)blockDim.x + threadIdx.x; pix < total; pix += (long long)gridDim.x * blockDim.x) {
    const int x = int(pix % W);
    const int y = int((pix / W) % H);
    float s = 0.f;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      // output pixel q = p - (t - center) contributed w[t] * img[p] to y[q]  ->  gx[p] += gz[q] * w[t]
      const int yy = y - (t / 3 - 1), xx = x - (t % 3 - 1);
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      const uint4* gp = reinterpret_cast<const uint4*>(gz + (pix - (t / 3 - 1) * W - (t % 3 - 1)) * C);
#pragma unroll
      for (int c8 = 0; c8 < C / 8; ++c8) {
        const uint4 r = __ldg(gp + c8);
        const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) s += dbf_lo(rw[j]) * ws[t][c8 * 8 + 2 * j] + dbf_hi(rw[j]) * ws[t][c8 * 8 + 2 * j + 1];
      }
    }
    gx[pix] = s;
  }
}

// dw[c][t] += sum_p gz[p][c] * img[p + t];  db[c] += sum_p gz[p][c].   Block = one image band; thread = (c, t) pair
// (+ C extra threads for the bias), both operands staged in shared memory.
template <int C>
__global__ void __launch_bounds__(C * 10)
dl1_wgrad_kernel(const __nv_bfloat16* __restrict__ gz, const float* __restrict__ img, float* __restrict__ dw,
                 float* __restrict__ db, int B, int H, int W, int TH) {
  extern __shared__ __align__(16) uint8_t dsm[];
  const int TR = TH + 2, TC = W + 2;
  float* im_s = reinterpret_cast<float*>(dsm);                          // [TR][TC]
  __nv_bfloat16* g_s = reinterpret_cast<__nv_bfloat16*>(im_s + ((TR * TC + 3) & ~3));          // [TH*W][C]
  const int bands = (H + TH - 1) / TH;
  const int c = threadIdx.x % C, t = threadIdx.x / C;                   // t == 9: bias thread
  const int dy = t / 3, dx = t % 3;
  float acc = 0.f;
  for (int item = blockIdx.x; item < B * bands; item += gridDim.x) {
    const int b = item / bands, y0 = (item % bands) * TH;
    const int rows = min(TH, H - y0);
    __syncthreads();
    for (int i = threadIdx.x; i < TR * TC; i += blockDim.x) {
      const int r = i / TC, cc = i % TC;
      const int gy = y0 - 1 + r, gxx = cc - 1;
      im_s[i] = (gy >= 0 && gy < H && gxx >= 0 && gxx < W) ? __ldg(img + (static_cast<size_t>(b) * H + gy) * W + gxx) : 0.f;
    }
    const uint4* src = reinterpret_cast<const uint4*>(gz + (static_cast<size_t>(b) * H + y0) * W * C);
    for (int i = threadIdx.x; i < rows * W * C / 8; i += blockDim.x) reinterpret_cast<uint4*>(g_s)[i] = __ldg(src + i);
    __syncthreads();
    if (t < 9) {
      for (int r = 0; r < rows; ++r) {
        const float* ip = im_s + (r + dy) * TC + dx;
        const __nv_bfloat16* gp = g_s + r * W * C + c;
        for (int x = 0; x < W; ++x) acc += __bfloat162float(gp[x * C]) * ip[x];
      }
    } else {
      for (int i = 0; i < rows * W; ++i) acc += __bfloat162float(g_s[i * C + c]);
    }
  }
  if (t < 9) atomicAdd(dw + c * 9 + t, acc);
  else if (db) atomicAdd(db + c, acc);
}

// Tensor-core form of the same sums for C == 16 and W % 16 == 0:  D[c][n] = sum_p gz[p][c] * Bm[p][n]  with
// Bm[p][0..8] = img[p + tap n], Bm[p][9] = 1 (the bias gradient), as m16n8k16 MMAs over 16-pixel chunks of an image row
// (two n8 tiles: taps 0-7 | tap 8, ones).  The fp32 image is split into bf16 hi + lo parts (two MMAs), which keeps
// the products exact to 2^-17; gz is bf16 already.  Tiles (one image band: gz rows by 4-D TMA, the image rows with
// their halo by 3-D TMA from column x = -4, zero-filled at the borders) go through a ring of p.nst stages; each warp takes every 8th chunk.
struct Dl1WgradP {
  float* dw; float* db;
  int B, H, W, TH, bands, n_tiles, nst, g_bytes, stage_bytes, TCp;
};

__global__ void __launch_bounds__(256)
dl1_wgrad_tc_kernel(const Dl1WgradP p, const __grid_constant__ CUtensorMap tmap_g,
                    const __grid_constant__ CUtensorMap tmap_img) {
  extern __shared__ __align__(1024) uint8_t dsm[];
  uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dsm) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + p.nst * p.stage_bytes);
  float* red = reinterpret_cast<float*>(bars + 8);                    // [16][10]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int TR = p.TH + 2;
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_g);
    tma_prefetch_desc(&tmap_img);
    for (int i = 0; i < p.nst; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  if (threadIdx.x < 160) red[threadIdx.x] = 0.f;
  auto load_tile = [&](int tile, int slot) {
    if (threadIdx.x != 0) return;
    const int b = tile / p.bands, y0 = (tile - b * p.bands) * p.TH;
    uint8_t* dst = ring + slot * p.stage_bytes;
    mbar_expect_tx(&bars[slot], static_cast<uint32_t>(p.TH * p.W * 32 + TR * p.TCp * 4));
    tma_load_4d(dst, &tmap_g, 0, 0, y0, b, &bars[slot]);
    tma_load_3d(dst + p.g_bytes, &tmap_img, -4, y0 - 1, b, &bars[slot]);   // shared-memory column j holds x = j - 4
  };
  __syncthreads();
  for (int k = 0; k < p.nst - 1; ++k) {
    const int tt = blockIdx.x + k * gridDim.x;
    if (tt < p.n_tiles) load_tile(tt, k);
  }
  float acc1[4] = {0.f, 0.f, 0.f, 0.f}, acc2[4] = {0.f, 0.f, 0.f, 0.f};
  // image-tile word offset of this lane's B rows: tap g for the first n-tile, tap 8 (lanes with g == 0) for the second
  const int off1 = (g / 3) * p.TCp + (g % 3) + 3 + 2 * t;
  const int off2 = 2 * p.TCp + 2 + 3 + 2 * t;
  const int kA = (lane & 7) + ((lane >> 4) << 3);
  const int chunks_x = p.W >> 4;
  int k = 0;
  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++k) {
    const int slot = k % p.nst;
    __syncthreads();
    {
      const int tt = tile + (p.nst - 1) * gridDim.x;
      if (tt < p.n_tiles) load_tile(tt, (k + p.nst - 1) % p.nst);
    }
    mbar_wait(&bars[slot], (k / p.nst) & 1, 0x32);
    const uint32_t g_u32 = dsmem_u32(ring + slot * p.stage_bytes);
    const float* im = reinterpret_cast<const float*>(ring + slot * p.stage_bytes + p.g_bytes);
    for (int ch = warp; ch < p.TH * chunks_x; ch += 8) {
      const int r = ch / chunks_x, x0 = (ch - r * chunks_x) << 4;
      uint32_t af[4];
      {
        const int n = ch * 16 + kA;
        ldsm_x4_t(g_u32 + n * 32 + ((((lane >> 3) & 1) ^ ((n >> 2) & 1)) << 4), af);
      }
      const float* row = im + r * p.TCp + x0;
      float v[4] = {row[off1], row[off1 + 1], row[off1 + 8], row[off1 + 9]};
      uint32_t hi0 = dpack_bf16(v[0], v[1]), hi1 = dpack_bf16(v[2], v[3]);
      uint32_t lo0 = dpack_bf16(v[0] - dbf_lo(hi0), v[1] - dbf_hi(hi0));
      uint32_t lo1 = dpack_bf16(v[2] - dbf_lo(hi1), v[3] - dbf_hi(hi1));
      mma_bf16_16816(acc1, af, hi0, hi1);
      mma_bf16_16816(acc1, af, lo0, lo1);
      if (g == 0) {
        v[0] = row[off2]; v[1] = row[off2 + 1]; v[2] = row[off2 + 8]; v[3] = row[off2 + 9];
        hi0 = dpack_bf16(v[0], v[1]); hi1 = dpack_bf16(v[2], v[3]);
        lo0 = dpack_bf16(v[0] - dbf_lo(hi0), v[1] - dbf_hi(hi0));
        lo1 = dpack_bf16(v[2] - dbf_lo(hi1), v[3] - dbf_hi(hi1));
      } else {
        hi0 = hi1 = g == 1 ? 0x3f803f80u : 0u;          // bf16 1.0 pairs: column 9 sums gz (bias gradient)
        lo0 = lo1 = 0u;
      }
      mma_bf16_16816(acc2, af, hi0, hi1);
      mma_bf16_16816(acc2, af, lo0, lo1);
    }
  }
  // accumulator (row g | g+8, columns 2t, 2t+1) -> red[c][n];  second tile: column 0 = tap 8, column 1 = bias
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = g + (i >> 1) * 8, n = 2 * t + (i & 1);
    atomicAdd(&red[c * 10 + n], acc1[i]);
    if (t == 0) atomicAdd(&red[c * 10 + 8 + (i & 1)], acc2[i]);
  }
  __syncthreads();
  if (threadIdx.x < 160) {
    const int c = threadIdx.x / 10, n = threadIdx.x - c * 10;
    if (n < 9) atomicAdd(p.dw + c * 9 + n, red[threadIdx.x]);
    else if (p.db) atomicAdd(p.db + c, red[threadIdx.x]);
  }
}

// ---------------------------------------------------------------------------------------------- weight packs
// One launch packs every layer of a discriminator: fp32 OIHW (flat parameter buffer) -> bf16 [9][rows][kpad].
//   kind 0 (forward, and MODE 1 input gradient of stride-2 layers uses kind 2):  wp[t][co][ci] = w[co][ci][t]
//   kind 1 (stride-1 input gradient as a forward conv over gz):                  wp[t][ci][co] = w[co][ci][8 - t]
//   kind 2 (stride-2 input gradient, transposed form):                           wp[t][ci][co] = w[co][ci][t]
struct DPackItem { long long src_off; long long dst_off; int cout, cin, rows, kpad, kind; };
struct DPackTable { int n; DPackItem it[24]; };

__global__ void dpack_kernel(const float* __restrict__ flat, __nv_bfloat16* __restrict__ dst, const DPackTable tab) {
  const DPackItem e = tab.it[blockIdx.y];
  const int total = 9 * e.rows * e.kpad;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int k = i % e.kpad;
    const int r = (i / e.kpad) % e.rows;
    const int t = i / (e.kpad * e.rows);
    float v = 0.f;
    if (e.kind == 0) {
      if (r < e.cout && k < e.cin) v = flat[e.src_off + (static_cast<long long>(r) * e.cin + k) * 9 + t];
    } else {
      const int tt = e.kind == 1 ? 8 - t : t;
      if (r < e.cin && k < e.cout) v = flat[e.src_off + (static_cast<long long>(k) * e.cin + r) * 9 + tt];
    }
    dst[e.dst_off + i] = __float2bfloat16(v);
  }
}

// ---------------------------------------------------------------------------------------------- host side
static int pad16(int c) { return (c + 15) & ~15; }
static int pad8(int c) { return (c + 7) & ~7; }

static void dmma_tile_plan(int Hg, int Wg, int B, size_t bytes_per_src_pixel_row /*unused*/, int* TH, int* IMGS) {
  const int target = 256;                              // at most 8 warps x one 32-pixel fragment pair
  if (Hg * Wg >= 128) {
    int th = target / Wg;
    if (th < 1) th = 1;
    if (th > Hg) th = Hg;
    const int nt = (Hg + th - 1) / th;
    th = (Hg + nt - 1) / nt;
    *TH = th; *IMGS = 1;
  } else {
    int n = target / (Hg * Wg);
    if (n > B) n = B;
    if (n < 1) n = 1;
    *TH = Hg; *IMGS = n;
  }
  // small layers: enough tiles to put two on every SM matter more than full-size tiles
  auto tiles = [&] { return ((Hg + *TH - 1) / *TH) * ((B + *IMGS - 1) / *IMGS); };
  while (tiles() < 2 * sm_count() && (*IMGS) * (*TH) * Wg > 64) {
    if (*IMGS > 1) *IMGS = (*IMGS + 1) / 2; else if (*TH > 1) *TH = (*TH + 1) / 2; else break;
  }
}

template <int KS, int NT, int MODE>
static int dmma_conv_launch_t(DmmaConvP& p, cudaStream_t st) {
  constexpr int PP = KS * 32 + 16, P = KS * 32;
  const int S = MODE == 0 ? p.stride : 1;
  const int Wg = MODE == 0 ? p.Wd : p.Wd / 2, Hg = MODE == 0 ? p.Hd : p.Hd / 2;
  // shared memory = weights + per-warp epilogue staging + barriers + a ring of source-tile buffers (1 KB aligned).
  // First try to stay under 108 KB (two CTAs per SM) with at least 96 fragment rows per tile and two buffers; layers
  // whose weights alone are too large for that (64 channels: 83 KB) run one CTA per SM with up to 200 KB.  Whatever
  // room is left under the cap goes into more buffers (deeper prefetch), up to kDmmaMaxStages.
  const size_t fixed = size_t(9) * NT * 8 * PP + (NT >= 2 ? size_t(kDmmaThreads / 32) * 32 * (NT * 16 + 16) : 0) +
                       2 * kDmmaMaxStages * 8 + 1024;
  int TH = 0, IMGS = 0, TR = 0, TC = 0, nst = 2;
  size_t smem = 0, tile_bytes = 0;
  bool ok = false;
  for (int attempt = 0; attempt < 2 && !ok; ++attempt) {
    const size_t cap = attempt == 0 ? 108 * 1024 : 200 * 1024;
    dmma_tile_plan(Hg, Wg, p.B, 0, &TH, &IMGS);
    const int whole = Hg * Wg * (p.B < IMGS ? p.B : IMGS);
    for (;;) {
      TR = MODE == 0 ? TH * S + 2 : TH + 1; TC = MODE == 0 ? Wg * S + 2 : Wg + 1;
      tile_bytes = (size_t(IMGS) * TR * TC * P + 1023) & ~size_t(1023);
      smem = fixed + 2 * tile_bytes;
      const int M = IMGS * TH * Wg;
      if (attempt == 0 && M < 96 && M < whole) break;
      if (smem <= cap && IMGS <= 256 && TR <= 256 && TC <= 256) {
        ok = true;
        // deeper ring only while it does not cost a resident CTA
        auto ctas = [](size_t sm) { int c = int((224 * 1024) / (sm + 1024)); return c > 4 ? 4 : c; };
        nst = 2;
        while (nst < kDmmaMaxStages && fixed + (nst + 1) * tile_bytes <= cap &&
               ctas(fixed + (nst + 1) * tile_bytes) == ctas(smem)) ++nst;
        smem = fixed + nst * tile_bytes;
        break;
      }
      if (IMGS > 1) --IMGS; else if (TH > 1) TH = (TH + 1) / 2; else break;
    }
  }
  if (!ok) return set_error("dmma_conv: tile does not fit in shared memory");
  p.TH = TH; p.IMGS = IMGS; p.nst = nst; p.tile_bytes = int(tile_bytes);
  if (Wg < 2) return set_error("dmma_conv: map too narrow");
  p.mg_img = uint32_t(((1ull << 32) + uint64_t(TH * Wg) - 1) / uint64_t(TH * Wg));
  p.mg_w = uint32_t(((1ull << 32) + uint64_t(Wg) - 1) / uint64_t(Wg));
  p.tiles_y = (Hg + TH - 1) / TH;
  p.n_tiles = p.tiles_y * ((p.B + IMGS - 1) / IMGS);
  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  if (!p.in_f32) {
    const uint64_t dims[4] = {uint64_t(p.cin), uint64_t(p.Wsp), uint64_t(p.Hsp), uint64_t(p.B)};
    const uint32_t box[4] = {uint32_t(p.cin), uint32_t(TC), uint32_t(TR), uint32_t(IMGS)};
    if (make_tmap_4d_bf16(&tmap, p.in, dims, box)) return 1;
  }
  ensure_max_smem(reinterpret_cast<const void*>(dmma_conv_kernel<KS, NT, MODE>), 201 * 1024);
  int per_sm = int((224 * 1024) / (smem + 1024)); if (per_sm < 1) per_sm = 1; if (per_sm > 4) per_sm = 4;
  int grid = sm_count() * per_sm; if (grid > p.n_tiles) grid = p.n_tiles;
  dmma_conv_kernel<KS, NT, MODE><<<grid, kDmmaThreads, smem, st>>>(p, tmap);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("dmma_conv launch", e);
  return 0;
}

template <int KS, int MODE>
static int dmma_conv_launch_n(DmmaConvP& p, cudaStream_t st) {
  const int nt = pad8(p.cout) / 8;
  switch (nt) {
    case 1: return dmma_conv_launch_t<KS, 1, MODE>(p, st);
    case 2: return dmma_conv_launch_t<KS, 2, MODE>(p, st);
    case 4: return dmma_conv_launch_t<KS, 4, MODE>(p, st);
    case 8: return dmma_conv_launch_t<KS, 8, MODE>(p, st);
  }
  return set_error("dmma_conv: output channels must be 1, 16, 32 or 64");
}

template <int KS, int MT>
static int dmma_wgrad_launch_t(DmmaWgradP& p, cudaStream_t st) {
  constexpr int PA = KS * 32, PG = MT * 32;
  const size_t fixed = kDwgMaxStages * 8 + 1024 + 272 * 4;      // barriers, alignment slack, pixel table (M16 <= 272)
  int TH = 0, IMGS = 0, TR = 0, TC = 0, nst = 2;
  size_t a_bytes = 0, g_bytes = 0, smem = 0;
  bool ok = false;
  // two stages under 72 KB (three CTAs per SM) if the tile keeps >= 96 rows, else whatever fits in 200 KB
  for (int attempt = 0; attempt < 2 && !ok; ++attempt) {
    const size_t cap = attempt == 0 ? 72 * 1024 : 200 * 1024;
    dmma_tile_plan(p.Ho, p.Wo, p.B, 0, &TH, &IMGS);
    const int whole = p.Ho * p.Wo * (p.B < IMGS ? p.B : IMGS);
    for (;;) {
      TR = TH * p.stride + 2; TC = p.Wo * p.stride + 2;
      const int M = IMGS * TH * p.Wo, M16 = (M + 15) & ~15;
      a_bytes = (size_t(IMGS) * TR * TC * PA + 1023) & ~size_t(1023);
      g_bytes = (size_t(M16) * PG + 1023) & ~size_t(1023);
      smem = fixed + 2 * (a_bytes + g_bytes);
      if (attempt == 0 && M < 96 && M < whole) break;
      if (smem <= cap && IMGS <= 256 && TR <= 256 && TC <= 256 && TH <= 256) {
        ok = true;
        auto ctas = [](size_t sm) { int c = int((224 * 1024) / (sm + 1024)); return c > 3 ? 3 : c; };
        nst = 2;
        while (nst < kDwgMaxStages && fixed + (nst + 1) * (a_bytes + g_bytes) <= cap &&
               ctas(fixed + (nst + 1) * (a_bytes + g_bytes)) == ctas(smem)) ++nst;
        smem = fixed + nst * (a_bytes + g_bytes);
        break;
      }
      if (IMGS > 1) --IMGS; else if (TH > 1) TH = (TH + 1) / 2; else break;
    }
  }
  if (!ok) return set_error("dmma_wgrad: tile does not fit in shared memory");
  if (((IMGS * TH * p.Wo + 15) & ~15) > 272) return set_error("dmma_wgrad: map wider than 256 pixels");
  p.TH = TH; p.IMGS = IMGS; p.nst = nst; p.a_bytes = int(a_bytes); p.g_bytes = int(g_bytes);
  p.tiles_y = (p.Ho + TH - 1) / TH;
  p.n_tiles = p.tiles_y * ((p.B + IMGS - 1) / IMGS);
  CUtensorMap tmap_a, tmap_g;
  memset(&tmap_g, 0, sizeof(tmap_g));
  {
    const uint64_t dims[4] = {uint64_t(p.cin), uint64_t(p.Wip), uint64_t(p.Hip), uint64_t(p.B)};
    const uint32_t box[4] = {uint32_t(p.cin), uint32_t(TC), uint32_t(TR), uint32_t(IMGS)};
    if (make_tmap_4d_bf16(&tmap_a, p.a, dims, box)) return 1;
  }
  if (!p.gz_f32) {
    const uint64_t dims[4] = {uint64_t(p.cout), uint64_t(p.Wo), uint64_t(p.Ho), uint64_t(p.B)};
    const uint32_t box[4] = {uint32_t(MT * 16), uint32_t(p.Wo), uint32_t(TH), uint32_t(IMGS)};
    if (make_tmap_4d_bf16(&tmap_g, p.gz, dims, box)) return 1;
  }
  ensure_max_smem(reinterpret_cast<const void*>(dmma_wgrad_kernel<KS, MT>), 201 * 1024);
  int per_sm = int((224 * 1024) / (smem + 1024)); if (per_sm < 1) per_sm = 1; if (per_sm > 3) per_sm = 3;
  // Every CTA ends with cout*cin*9 atomic adds, so a small layer wants few CTAs:  time(g) ~ tiles/g * t_tile +
  // g * atomics / rate  is least at  g = sqrt(tiles * t_tile * rate / atomics)   (t_tile ~ 1.5 us, rate ~ 250 / ns).
  const double atomics = double(MT * 16) * p.cin * 9;
  int grid = int(sqrt(double(p.n_tiles) * 1500.0 * 250.0 / atomics));
  if (grid < 16) grid = 16;
  if (grid > sm_count() * per_sm) grid = sm_count() * per_sm;
  if (grid > p.n_tiles) grid = p.n_tiles;
  dmma_wgrad_kernel<KS, MT><<<grid, kDwgThreads, smem, st>>>(p, tmap_a, tmap_g);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("dmma_wgrad launch", e);
  return 0;
}

}  // namespace srb

using namespace srb;

namespace srb {
// The C entry point with explicit memory pitches for the source (Hsp rows x Wsp pixels per image) and for the
// destination / mask reference (Hdp x Wdp): lets the generator's PNHWC buffers (one zero row and column per image)
// be read and written in place.  Used by csrc/gen_edge.cu for conv3.2.
int dmma_conv_pitched(const void* in_bf16, const float* in_f32, int cin, const void* w_packed, const float* bias,
                      int cout, void* out_bf16, float* out_f32, const void* ref, int B, int Hs, int Ws, int Hd,
                      int Wd, int Hsp, int Wsp, int Hdp, int Wdp, int stride, int transposed, int act, float slope,
                      void* stream) {
  if ((!in_bf16 && !in_f32) || !w_packed || (!out_bf16 && !out_f32)) return set_error("dmma_conv: null pointer");
  if (in_f32 && cin != 1) return set_error("dmma_conv: fp32 source must be single-channel");
  if (!in_f32 && cin != 16 && cin != 32 && cin != 64) return set_error("dmma_conv: input channels must be 16, 32 or 64");
  if (cout != 1 && cout != 16 && cout != 32 && cout != 64) return set_error("dmma_conv: output channels must be 1, 16, 32 or 64");
  if (cout != 1 && !out_bf16) return set_error("dmma_conv: bf16 output required");
  if (act == 2 && !ref) return set_error("dmma_conv: mask epilogue without reference");
  if (transposed && ((Hd & 1) || (Wd & 1) || Hs * 2 != Hd || Ws * 2 != Wd)) return set_error("dmma_conv: transposed form needs even maps");
  if (!transposed && (stride < 1 || stride > 2 || (Hs - 1) / stride + 1 != Hd || (Ws - 1) / stride + 1 != Wd))
    return set_error("dmma_conv: geometry mismatch");
  DmmaConvP p{};
  p.in = (const __nv_bfloat16*)in_bf16; p.in_f32 = in_f32; p.w = (const __nv_bfloat16*)w_packed; p.bias = bias;
  p.out = (__nv_bfloat16*)out_bf16; p.out_f32 = out_f32; p.ref = (const __nv_bfloat16*)ref;
  p.B = B; p.Hs = Hs; p.Ws = Ws; p.Hd = Hd; p.Wd = Wd; p.cin = cin; p.cout = cout; p.stride = stride;
  p.Hsp = Hsp; p.Wsp = Wsp; p.Hdp = Hdp; p.Wdp = Wdp;
  if (Hsp < Hs || Wsp < Ws || Hdp < Hd || Wdp < Wd) return set_error("dmma_conv: pitch smaller than the map");
  if (in_f32 && (Hsp != Hs || Wsp != Ws)) return set_error("dmma_conv: fp32 source must be dense");
  p.act = act; p.slope = slope;
  cudaStream_t st = (cudaStream_t)stream;
  const int ks = pad16(cin) / 16;
  if (transposed) {
    switch (ks) {
      case 1: return dmma_conv_launch_n<1, 1>(p, st);
      case 2: return dmma_conv_launch_n<2, 1>(p, st);
      case 4: return dmma_conv_launch_n<4, 1>(p, st);
    }
  } else {
    switch (ks) {
      case 1: return dmma_conv_launch_n<1, 0>(p, st);
      case 2: return dmma_conv_launch_n<2, 0>(p, st);
      case 4: return dmma_conv_launch_n<4, 0>(p, st);
    }
  }
  return set_error("dmma_conv: unsupported channel count");
}

int dmma_wgrad_pitched(const void* gz_bf16, const float* gz_f32, int cout, const void* a_bf16, int cin, float* dw,
                       float* db, int B, int Hi, int Wi, int Hip, int Wip, int Ho, int Wo, int stride, void* stream) {
  if ((!gz_bf16 && !gz_f32) || !a_bf16 || !dw) return set_error("dmma_wgrad: null pointer");
  if (gz_f32 && cout != 1) return set_error("dmma_wgrad: fp32 gradient must be single-channel");
  if (cin != 16 && cin != 32 && cin != 64) return set_error("dmma_wgrad: input channels must be 16, 32 or 64");
  if (cout != 1 && cout != 16 && cout != 32 && cout != 64) return set_error("dmma_wgrad: output channels must be 1, 16, 32 or 64");
  if (stride < 1 || stride > 2 || (Hi - 1) / stride + 1 != Ho || (Wi - 1) / stride + 1 != Wo) return set_error("dmma_wgrad: geometry mismatch");
  DmmaWgradP p{};
  p.gz = (const __nv_bfloat16*)gz_bf16; p.gz_f32 = gz_f32; p.a = (const __nv_bfloat16*)a_bf16; p.dw = dw; p.db = db;
  p.B = B; p.Hi = Hi; p.Wi = Wi; p.Ho = Ho; p.Wo = Wo; p.cin = cin; p.cout = cout; p.stride = stride;
  p.Hip = Hip; p.Wip = Wip;
  if (Hip < Hi || Wip < Wi) return set_error("dmma_wgrad: pitch smaller than the map");
  cudaStream_t st = (cudaStream_t)stream;
  const int ks = cin / 16, mt = pad16(cout) / 16;
#define SRB_WG(K, M) if (ks == K && mt == M) return dmma_wgrad_launch_t<K, M>(p, st);
  SRB_WG(1, 1) SRB_WG(1, 2) SRB_WG(1, 4) SRB_WG(2, 1) SRB_WG(2, 2) SRB_WG(2, 4) SRB_WG(4, 1) SRB_WG(4, 2)
#undef SRB_WG
  if (ks == 4 && mt == 4) {      // 64 x 64: two passes of 32 output channels (keeps the accumulators in registers)
    if (dmma_wgrad_launch_t<4, 2>(p, st)) return 1;
    p.co_off = 32;
    return dmma_wgrad_launch_t<4, 2>(p, st);
  }
  return set_error("dmma_wgrad: unsupported channel count");
}
}  // namespace srb

extern "C" int srb_dmma_conv(const void* in_bf16, const float* in_f32, int cin, const void* w_packed, const float* bias,
                             int cout, void* out_bf16, float* out_f32, const void* ref, int B, int Hs, int Ws, int Hd,
                             int Wd, int stride, int transposed, int act, float slope, void* stream) {
  return dmma_conv_pitched(in_bf16, in_f32, cin, w_packed, bias, cout, out_bf16, out_f32, ref, B, Hs, Ws, Hd, Wd, Hs, Ws,
                           Hd, Wd, stride, transposed, act, slope, stream);
}

extern "C" int srb_dmma_wgrad(const void* gz_bf16, const float* gz_f32, int cout, const void* a_bf16, int cin, float* dw,
                              float* db, int B, int Hi, int Wi, int Ho, int Wo, int stride, void* stream) {
  return dmma_wgrad_pitched(gz_bf16, gz_f32, cout, a_bf16, cin, dw, db, B, Hi, Wi, Hi, Wi, Ho, Wo, stride, stream);
}

extern "C" int srb_dl1_fwd(const float* img, const float* w, const float* bias, void* out_bf16, const void* ref, int C,
                           int B, int H, int W, int act, float slope, void* stream) {
  if (!img || !w || !out_bf16) return set_error("dl1_fwd: null pointer");
  if (act == 2 && !ref) return set_error("dl1_fwd: mask epilogue without reference");
  const long long total = 1LL * B * H * W;
  int grid = ceil_div(total, 256); if (grid > sm_count() * 8) grid = sm_count() * 8;
  cudaStream_t st = (cudaStream_t)stream;
#define SRB_L1(CC) if (C == CC) dl1_fwd_kernel<CC><<<grid, 256, 0, st>>>(img, w, bias, (__nv_bfloat16*)out_bf16, (const __nv_bfloat16*)ref, B, H, W, act, slope);
  SRB_L1(16) else SRB_L1(32) else SRB_L1(64) else return set_error("dl1_fwd: channels must be 16, 32 or 64");
#undef SRB_L1
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("dl1_fwd launch", e);
  return 0;
}

extern "C" int srb_dl1_dgrad(const void* gz_bf16, const float* w, float* gx, int C, int B, int H, int W, void* stream) {
  if (!gz_bf16 || !w || !gx) return set_error("dl1_dgrad: null pointer");
  {
    static int general = -1;          // SRB_DL1_GENERAL=1: the CUDA-core kernel below (A/B timing, fall-back)
    if (general < 0) { const char* e = getenv("SRB_DL1_GENERAL"); general = (e && atoi(e) != 0) ? 1 : 0; }
    if (!general) {
      const int rc = d16_dgrad_launch(gz_bf16, w, gx, C, B, H, W, (cudaStream_t)stream);
      if (rc >= 0) return rc;
    }
  }
  const long long total = 1LL * B * H * W;
  int grid = ceil_div(total, 256); if (grid > sm_count() * 8) grid = sm_count() * 8;
  cudaStream_t st = (cudaStream_t)stream;
#define SRB_L1(CC) if (C == CC) dl1_dgrad_kernel<CC><<<grid, 256, 0, st>>>((const __nv_bfloat16*)gz_bf16, w, gx, B, H, W);
  SRB_L1(16) else SRB_L1(32) else SRB_L1(64) else return set_error("dl1_dgrad: channels must be 16, 32 or 64");
#undef SRB_L1
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("dl1_dgrad launch", e);
  return 0;
}

extern "C" int srb_dl1_wgrad(const void* gz_bf16, const float* img, float* dw, float* db, int C, int B, int H, int W,
                             void* stream) {
  if (!gz_bf16 || !img || !dw) return set_error("dl1_wgrad: null pointer");
  if (C == 16 && W % 16 == 0 && W + 8 <= 256) {          // tensor-core kernel
    Dl1WgradP p{};
    p.dw = dw; p.db = db; p.B = B; p.H = H; p.W = W;
    p.TH = H < 8 ? H : 8;
    p.bands = (H + p.TH - 1) / p.TH;
    p.n_tiles = B * p.bands;
    p.TCp = W + 8;                                       // columns x = -4 .. W+3: TMA box starts must be 16-byte aligned
    p.g_bytes = (p.TH * W * 32 + 1023) & ~1023;
    p.stage_bytes = p.g_bytes + (((p.TH + 2) * p.TCp * 4 + 1023) & ~1023);
    p.nst = 3;
    const size_t smem = 1024 + size_t(p.nst) * p.stage_bytes + 64 + 160 * 4;
    CUtensorMap tg, ti;
    const uint64_t dims[4] = {16, uint64_t(W), uint64_t(H), uint64_t(B)};
    const uint32_t box[4] = {16, uint32_t(W), uint32_t(p.TH), 1};
    if (make_tmap_4d_bf16(&tg, gz_bf16, dims, box)) return 1;
    if (make_tmap_3d_f32(&ti, img, W, H, B, p.TCp, p.TH + 2)) return 1;
    ensure_max_smem(reinterpret_cast<const void*>(dl1_wgrad_tc_kernel), 100 * 1024);
    if (smem > 100 * 1024) return set_error("dl1_wgrad: band does not fit in shared memory");
    int per_sm = int((224 * 1024) / (smem + 1024)); if (per_sm < 1) per_sm = 1; if (per_sm > 4) per_sm = 4;
    int grid = sm_count() * per_sm; if (grid > p.n_tiles) grid = p.n_tiles;
    dl1_wgrad_tc_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(p, tg, ti);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_cuda_error("dl1_wgrad launch", e);
    return 0;
  }
  int TH = 8; if (TH > H) TH = H;
  while (TH > 1 && size_t(TH + 2) * (W + 2) * 4 + 16 + size_t(TH) * W * C * 2 > 96 * 1024) TH = (TH + 1) / 2;
  const size_t smem = size_t(TH + 2) * (W + 2) * 4 + 16 + size_t(TH) * W * C * 2;
  const int bands = (H + TH - 1) / TH;
  int grid = B * bands; if (grid > sm_count() * 4) grid = sm_count() * 4;
  cudaStream_t st = (cudaStream_t)stream;
#define SRB_L1(CC) if (C == CC) { \
    ensure_max_smem(reinterpret_cast<const void*>(dl1_wgrad_kernel<CC>), 96 * 1024); \
    dl1_wgrad_kernel<CC><<<grid, CC * 10, smem, st>>>((const __nv_bfloat16*)gz_bf16, img, dw, db, B, H, W, TH); }
  SRB_L1(16) else SRB_L1(32) else SRB_L1(64) else return set_error("dl1_wgrad: channels must be 16, 32 or 64");
#undef SRB_L1
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("dl1_wgrad launch", e);
  return 0;
}

extern "C" int srb_dpack(const float* flat_params, void* dst_bf16, int n_items, const long long* src_off,
                         const long long* dst_off, const int* cout, const int* cin, const int* kind, void* stream) {
  if (!flat_params || !dst_bf16 || n_items < 1 || n_items > 24) return set_error("dpack: bad arguments");
  DPackTable tab{};
  tab.n = n_items;
  for (int i = 0; i < n_items; ++i) {
    DPackItem& e = tab.it[i];
    e.src_off = src_off[i]; e.dst_off = dst_off[i]; e.cout = cout[i]; e.cin = cin[i]; e.kind = kind[i];
    if (kind[i] == 0) { e.rows = pad8(cout[i]); e.kpad = pad16(cin[i]); }
    else { e.rows = pad8(cin[i]); e.kpad = pad16(cout[i]); }
  }
  dim3 grid(8, n_items);
  dpack_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(flat_params, (__nv_bfloat16*)dst_bf16, tab);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("dpack launch", e);
  return 0;
}
