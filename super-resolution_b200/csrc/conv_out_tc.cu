// conv3.2 of the generator (models.py:99,131: nn.Conv2d(64, 1, 3, padding 1) at HR geometry) and its backward, with the
// nine taps folded into a GEMM dimension.
//
// A 64 -> 1 channel 3x3 convolution is 7 MFLOP per image against 0.8 MB of bf16 activations: HBM-bound by two orders
// of magnitude.  The general warp-MMA kernels (disc_mma.cu) treat every tap as its own K = 64 (forward) or K = 16
// (backward, one real channel) product and re-read the shared-memory tile nine times; here each activation row is read
// from shared memory ONCE per product:
//   forward   d[q][t] = sum_c act[q][c] w[c][t]         [16 pixels x 64] x [64 x 16 (9 taps)]   per-pixel tap partials,
//             out[p]  = bias + sum_t d[p + s_t][t]       gathered from a shared-memory plane per tap (flat-shift: the
//                                                        zero pad row / column of PNHWC supplies the image border)
//   backward  G[p][t]  = gz[p - s_t]                     nine fp32 neighbours of the 1-channel gradient, bf16
//             gy[p][c] = (sum_t G[p][t] w[c][t]) LeakyReLU'(act[p][c])     [16 x 16] x [16 x 64]
//             dw[c][t] = sum_p act[p][c] G[p][t]         [64 x pixels] x [pixels x 16]: same tile, same G
//             db       = sum_p gz[p]
//   so the input gradient and the weight gradient share ONE pass over the activations (read 128 B, write 128 B per pixel).
// Every warp runs its own TMA ring over 16-pixel groups (2 KB boxes, 128B swizzle, read with ldmatrix): no block-wide
// barrier in the backward kernel, two per tile in the forward one (around the tap gather).
// The output head (models.py:114-118,131-135) is applied in the forward kernel's gather phase.
#include "common.cuh"
#include "sm100.cuh"

namespace srb {

constexpr int kC32Threads = 256;
constexpr int kC32Warps = kC32Threads / 32;
constexpr int kC32FwdStages = 3;
constexpr int kC32BwdStages = 4;
constexpr int kC32RB = 16;                 // image rows per forward tile (+ one halo row above and below)
constexpr int kC32GPitch = 48;             // bytes per pixel row of the G tile: 16 bf16 + 16 B skew (conflict-free ldmatrix)

struct C32FwdP {
  const float* w; const float* bias;
  float* out; float* srs; float* z;
  int B, H, W, pitch, P;                   // P = (H + 1) * pitch flat rows per image
  int blocks_per_img, n_tiles, tr_max;     // tr_max = (kC32RB + 2) * pitch floats per tap plane
  float inv_multiplier, power, thres; int training;
};

__global__ void __launch_bounds__(kC32Threads, 2)
c32_fwd_kernel(const __grid_constant__ CUtensorMap tmap, const C32FwdP p) {
  extern __shared__ __align__(1024) uint8_t c32_smem[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(c32_smem) + 1023) & ~uintptr_t(1023));
  uint8_t* ring = base;                                                            // [warp][stage][2048]
  __nv_bfloat16* w_s = reinterpret_cast<__nv_bfloat16*>(ring + kC32Warps * kC32FwdStages * 2048);   // [16 taps][64]
  float* ds = reinterpret_cast<float*>(w_s + 16 * 64);                             // [9][tr_max]
  uint64_t* bars = reinterpret_cast<uint64_t*>(ds + 9 * p.tr_max);                 // [warp][stage]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int i = threadIdx.x; i < 16 * 64; i += kC32Threads) {
    const int t = i >> 6, c = i & 63;
    w_s[i] = __float2bfloat16(t < 9 ? p.w[c * 9 + t] : 0.f);
  }
  if (lane == 0) {
    if (warp == 0) tma_prefetch_desc(&tmap);
    for (int s = 0; s < kC32FwdStages; ++s) mbar_init(&bars[warp * kC32FwdStages + s], 1);
    fence_mbar_init();
  }
  __syncthreads();
  // B fragments of w [K = 64 channels][N = 16 taps], stored [tap][channel]: constant for the whole kernel
  uint32_t wb[4][4];
#pragma unroll
  for (int kk = 0; kk < 4; ++kk)
    ldsm_x4(smem_u32(w_s) + ((lane & 7) + ((lane >> 4) << 3)) * 128 + kk * 32 + ((lane >> 3) & 1) * 16, wb[kk]);

  uint8_t* my_ring = ring + warp * kC32FwdStages * 2048;
  uint64_t* my_bars = bars + warp * kC32FwdStages;
  uint32_t issued = 0, consumed = 0;
  const float bias = p.bias ? p.bias[0] : 0.f;
  const float lam_p = powf(p.thres, p.power);

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
    const int b = tile / p.blocks_per_img, yb = tile - b * p.blocks_per_img;
    const int y0 = yb * kC32RB;
    const int rows_here = min(kC32RB, p.H - y0);
    const int TR = (rows_here + 2) * p.pitch;
    const int flat0 = b * p.P + (y0 - 1) * p.pitch;            // may be negative (first tile): TMA zero-fills
    const int ngroups = (TR + 15) >> 4;
    const int n_my = (ngroups - warp + kC32Warps - 1) / kC32Warps;
    auto issue = [&](int k) {
      if (lane == 0) {
        const uint32_t s = issued % kC32FwdStages;
        mbar_expect_tx(&my_bars[s], 2048);
        tma_load_2d(my_ring + s * 2048, &tmap, 0, flat0 + (warp + kC32Warps * k) * 16, &my_bars[s]);
      }
      ++issued;
    };
    for (int k = 0; k < kC32FwdStages - 1 && k < n_my; ++k) issue(k);
    for (int k = 0; k < n_my; ++k) {
      __syncwarp();                                            // group k-1's ldmatrix reads are done in every lane
      if (k + kC32FwdStages - 1 < n_my) issue(k + kC32FwdStages - 1);
      const uint32_t s = consumed % kC32FwdStages;
      mbar_wait(&my_bars[s], (consumed / kC32FwdStages) & 1, 0x41);
      ++consumed;
      const uint32_t ta = smem_u32(my_ring + s * 2048);
      float acc[2][4];
#pragma unroll
      for (int n = 0; n < 2; ++n)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[n][e] = 0.f;
      const int row = lane & 15, hi = lane >> 4;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        uint32_t a[4];
        ldsm_x4(ta + row * 128 + (((kk * 2 + hi) ^ (row & 7)) << 4), a);
        mma_bf16_16816(acc[0], a, wb[kk][0], wb[kk][1]);
        mma_bf16_16816(acc[1], a, wb[kk][2], wb[kk][3]);
      }
      const int r0 = (warp + kC32Warps * k) * 16 + (lane >> 2), r1 = r0 + 8;
      const int t0 = (lane & 3) * 2;
      if (r0 < TR) { ds[t0 * p.tr_max + r0] = acc[0][0]; ds[(t0 + 1) * p.tr_max + r0] = acc[0][1]; }
      if (r1 < TR) { ds[t0 * p.tr_max + r1] = acc[0][2]; ds[(t0 + 1) * p.tr_max + r1] = acc[0][3]; }
      if (t0 == 0) {
        if (r0 < TR) ds[8 * p.tr_max + r0] = acc[1][0];
        if (r1 < TR) ds[8 * p.tr_max + r1] = acc[1][2];
      }
    }
    __syncthreads();
    // gather the nine tap partials of every output pixel, then the output head
    const int npx = rows_here * p.W;
    for (int i = threadIdx.x; i < npx; i += kC32Threads) {
      const int y = i / p.W, x = i - y * p.W;
      const int r = (y + 1) * p.pitch + x;
      float sacc = bias;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const int idx = r + (t / 3 - 1) * p.pitch + (t % 3 - 1);
        sacc += idx >= 0 ? ds[t * p.tr_max + idx] : 0.f;       // idx == -1: the pad pixel in front of the tile (zero)
      }
      const float zz = sacc * p.inv_multiplier;
      const size_t oi = (static_cast<size_t>(b) * p.H + y0 + y) * p.W + x;
      if (p.z) p.z[oi] = zz;
      const float sv = p.training ? zz : fmaxf(zz, 0.f);
      p.srs[oi] = hardshrink(sv, lam_p);
      float o2 = zz;
      if (p.power != 1.f) o2 = powf(fmaxf(zz, 0.f), 1.f / p.power);
      if (!p.training) o2 = fmaxf(o2, 0.f);
      p.out[oi] = hardshrink(o2, p.thres);
    }
    __syncthreads();
  }
}

struct C32BwdP {
  const float* gz; const float* w;
  __nv_bfloat16* gy;
  float* dw; float* db;
  int B, H, W, pitch, P;
  int R, n_groups;
  float slope;
};

__global__ void __launch_bounds__(kC32Threads, 2)
c32_bwd_kernel(const __grid_constant__ CUtensorMap tmap, const C32BwdP p) {
  extern __shared__ __align__(1024) uint8_t c32_smem[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(c32_smem) + 1023) & ~uintptr_t(1023));
  uint8_t* ring = base;                                                            // [warp][stage][2048]
  uint8_t* g_s = ring + kC32Warps * kC32BwdStages * 2048;                          // [warp][16][kC32GPitch]
  __nv_bfloat16* w2_s = reinterpret_cast<__nv_bfloat16*>(g_s + kC32Warps * 16 * kC32GPitch);   // [64 ch][16 taps]
  float* red = reinterpret_cast<float*>(w2_s + 64 * 16);                           // [64][16]
  uint64_t* bars = reinterpret_cast<uint64_t*>(red + 64 * 16);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int i = threadIdx.x; i < 64 * 16; i += kC32Threads) {
    const int c = i >> 4, t = i & 15;
    w2_s[i] = __float2bfloat16(t < 9 ? p.w[c * 9 + t] : 0.f);
    red[i] = 0.f;
  }
  if (lane == 0) {
    if (warp == 0) tma_prefetch_desc(&tmap);
    for (int s = 0; s < kC32BwdStages; ++s) mbar_init(&bars[warp * kC32BwdStages + s], 1);
    fence_mbar_init();
  }
  __syncthreads();
  // B fragments of w2 [K = 16 taps][N = 64 channels], stored [channel][tap]: wd[jj] covers channels 16 jj .. 16 jj + 15
  uint32_t wd[4][4];
#pragma unroll
  for (int jj = 0; jj < 4; ++jj)
    ldsm_x4(smem_u32(w2_s) + (16 * jj + (lane & 7) + ((lane >> 4) << 3)) * 32 + ((lane >> 3) & 1) * 16, wd[jj]);

  uint8_t* my_ring = ring + warp * kC32BwdStages * 2048;
  uint8_t* my_g = g_s + warp * 16 * kC32GPitch;
  uint64_t* my_bars = bars + warp * kC32BwdStages;
  const int gw = blockIdx.x * kC32Warps + warp, nw = gridDim.x * kC32Warps;
  const int n_my = gw < p.n_groups ? (p.n_groups - gw + nw - 1) / nw : 0;
  float accw[4][2][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) accw[i][j][e] = 0.f;
  float db_acc = 0.f;
  uint32_t issued = 0;
  auto issue = [&](int k) {
    if (lane == 0) {
      const uint32_t s = issued % kC32BwdStages;
      mbar_expect_tx(&my_bars[s], 2048);
      tma_load_2d(my_ring + s * 2048, &tmap, 0, (gw + nw * k) * 16, &my_bars[s]);
    }
    ++issued;
  };
  for (int k = 0; k < kC32BwdStages - 1 && k < n_my; ++k) issue(k);
  for (int k = 0; k < n_my; ++k) {
    if (k > 0) fence_proxy_async();                          // group k-1's shared-memory writes vs the TMA re-fill
    __syncwarp();
    if (k + kC32BwdStages - 1 < n_my) issue(k + kC32BwdStages - 1);
    const int q0 = (gw + nw * k) * 16;
    // ---- G tile: nine neighbours of the 1-channel gradient per pixel, bf16 [16 pixels][16]
    if (lane < 16) {
      const int q = q0 + lane;
      float g9[9];
#pragma unroll
      for (int t = 0; t < 9; ++t) g9[t] = 0.f;
      if (q < p.R) {
        const int b = q / p.P, rem = q - b * p.P;
        const int y = rem / p.pitch, x = rem - y * p.pitch;
        if (y < p.H && x < p.W) {
          const float* gp = p.gz + static_cast<size_t>(b) * p.H * p.W;
#pragma unroll
          for (int t = 0; t < 9; ++t) {
            const int yy = y - (t / 3 - 1), xx = x - (t % 3 - 1);
            if (yy >= 0 && yy < p.H && xx >= 0 && xx < p.W) g9[t] = __ldg(gp + yy * p.W + xx);
          }
          db_acc += g9[4];
        }
      }
      uint4 lo, hi4;
      lo.x = pack_bf16(g9[0], g9[1]); lo.y = pack_bf16(g9[2], g9[3]);
      lo.z = pack_bf16(g9[4], g9[5]); lo.w = pack_bf16(g9[6], g9[7]);
      hi4.x = pack_bf16(g9[8], 0.f); hi4.y = 0u; hi4.z = 0u; hi4.w = 0u;
      *reinterpret_cast<uint4*>(my_g + lane * kC32GPitch) = lo;
      *reinterpret_cast<uint4*>(my_g + lane * kC32GPitch + 16) = hi4;
    }
    __syncwarp();
    uint32_t ga[4], gb[4];
    {
      const uint32_t gaddr = smem_u32(my_g) + (lane & 15) * kC32GPitch + (lane >> 4) * 16;
      ldsm_x4(gaddr, ga);              // A fragment [16 pixels][16 taps]
      ldsm_x4_t(gaddr, gb);            // B fragments [K = 16 pixels][N = taps 0-7 | 8-15]
    }
    const uint32_t s = static_cast<uint32_t>(k) % kC32BwdStages;
    mbar_wait(&my_bars[s], (static_cast<uint32_t>(k) / kC32BwdStages) & 1, 0x42);
    const uint32_t ta = smem_u32(my_ring + s * 2048);
    // ---- weight gradient: A = act^T [16 channels][16 pixels] per m-tile (transposing ldmatrix), B = G
    {
      const int pix = (lane & 7) + ((lane >> 4) << 3), sub = (lane >> 3) & 1;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint32_t aw[4];
        ldsm_x4_t(ta + pix * 128 + (((2 * i + sub) ^ (pix & 7)) << 4), aw);
        mma_bf16_16816(accw[i][0], aw, gb[0], gb[1]);
        mma_bf16_16816(accw[i][1], aw, gb[2], gb[3]);
      }
    }
    __syncwarp();                                            // every lane's transposed reads are done: the tile may be overwritten
    // ---- input gradient, written over the activation tile in place (each thread replaces exactly the words it masks with)
    {
      const int row = lane & 15, hi = lane >> 4;
      const int r0 = lane >> 2, r1 = r0 + 8, cb = (lane & 3) * 4;
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        uint32_t am[4];
        ldsm_x4(ta + row * 128 + (((2 * jj + hi) ^ (row & 7)) << 4), am);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float d[4] = {0.f, 0.f, 0.f, 0.f};
          mma_bf16_16816(d, ga, wd[jj][2 * h], wd[jj][2 * h + 1]);
          const uint32_t m0 = am[2 * h], m1 = am[2 * h + 1];
          const uint32_t o0 = pack_bf16(d[0] * (bf16lo(m0) > 0.f ? 1.f : p.slope), d[1] * (bf16hi(m0) > 0.f ? 1.f : p.slope));
          const uint32_t o1 = pack_bf16(d[2] * (bf16lo(m1) > 0.f ? 1.f : p.slope), d[3] * (bf16hi(m1) > 0.f ? 1.f : p.slope));
          const int ch = 2 * jj + h;
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(ta + r0 * 128 + ((ch ^ (r0 & 7)) << 4) + cb), "r"(o0) : "memory");
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(ta + r1 * 128 + ((ch ^ (r1 & 7)) << 4) + cb), "r"(o1) : "memory");
        }
      }
    }
    __syncwarp();
    // ---- copy-out: 16 rows x 128 B, 16-byte pieces, consecutive lanes = consecutive addresses
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int i = lane + 32 * it, row = i >> 3, c = i & 7;
      uint4 v;
      asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                   : "r"(ta + row * 128 + ((c ^ (row & 7)) << 4)));
      const int q = q0 + row;
      if (q < p.R) *reinterpret_cast<uint4*>(p.gy + static_cast<size_t>(q) * 64 + c * 8) = v;
    }
  }
  // ---- reductions: warps -> shared memory -> global
  db_acc += __shfl_xor_sync(0xffffffffu, db_acc, 8);
  db_acc += __shfl_xor_sync(0xffffffffu, db_acc, 4);
  db_acc += __shfl_xor_sync(0xffffffffu, db_acc, 2);
  db_acc += __shfl_xor_sync(0xffffffffu, db_acc, 1);
  if (lane == 0 && db_acc != 0.f) atomicAdd(p.db, db_acc);
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int c = 16 * i + (lane >> 2) + ((e & 2) ? 8 : 0);
        const int t = 8 * j + (lane & 3) * 2 + (e & 1);
        if (t < 9 && accw[i][j][e] != 0.f) atomicAdd(&red[c * 16 + t], accw[i][j][e]);
      }
  __syncthreads();
  for (int i = threadIdx.x; i < 64 * 9; i += kC32Threads) {
    const int c = i / 9, t = i - c * 9;
    const float v = red[c * 16 + t];
    if (v != 0.f) atomicAdd(p.dw + i, v);
  }
}

// ---------------------------------------------------------------- image end of the discriminator: input gradient
// gx[p] = sum_t sum_c gz[p - s_t][c] w[c][t]   (models.py:142 for in_filters = 1; esrgan.py:494-508 through the generator's
// adversarial term and 602 for the gradient penalty): the same tap-partial scheme as c32_fwd_kernel over UNPADDED bf16
// NHWC [B][H][W][C] rows (C = 16 KS channels): d[q][t] = sum_c gz[q][c] w[c][t], then the nine planes are gathered with
// explicit border tests.  The fp32 weights enter as bf16 hi + lo (two k-steps per 16 channels against the same A
// fragment), so the products keep ~2^-17 relative accuracy like the CUDA-core kernel this replaces.
struct D16DgradP {
  const float* w;                          // fp32 [C][1][3][3]
  float* gx;                               // fp32 [B][H][W]
  int B, H, W, blocks_per_img, n_tiles, tr_max;
};

template <int KS>
__global__ void __launch_bounds__(kC32Threads, 2)
d16_dgrad_kernel(const __grid_constant__ CUtensorMap tmap, const D16DgradP p) {
  constexpr int C = KS * 16, P = KS * 32;               // channels, bytes per pixel row
  constexpr int SH = KS == 1 ? 2 : KS == 2 ? 1 : 0, SWM = 2 * KS - 1;      // TMA swizzle of a P-byte row (disc_mma.cu)
  constexpr int kStages = 4;
  constexpr int RPB = 16;                               // pixel rows per TMA box (larger boxes cost resident CTAs: measured slower)
  constexpr int kSlot = RPB * P;
  extern __shared__ __align__(1024) uint8_t c32_smem[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(c32_smem) + 1023) & ~uintptr_t(1023));
  uint8_t* ring = base;                                                            // [warp][stage][kSlot]
  __nv_bfloat16* w_s = reinterpret_cast<__nv_bfloat16*>(ring + kC32Warps * kStages * kSlot);   // [2 (hi, lo)][16 taps][C]
  float* ds = reinterpret_cast<float*>(w_s + 2 * 16 * C);                          // [9][tr_max]
  uint64_t* bars = reinterpret_cast<uint64_t*>(ds + 9 * p.tr_max);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int i = threadIdx.x; i < 16 * C; i += kC32Threads) {
    const int t = i / C, c = i - t * C;
    const float v = t < 9 ? p.w[c * 9 + t] : 0.f;
    const __nv_bfloat16 hi = __float2bfloat16(v);
    w_s[i] = hi;
    w_s[16 * C + i] = __float2bfloat16(v - __bfloat162float(hi));
  }
  if (lane == 0) {
    if (warp == 0) tma_prefetch_desc(&tmap);
    for (int s = 0; s < kStages; ++s) mbar_init(&bars[warp * kStages + s], 1);
    fence_mbar_init();
  }
  __syncthreads();
  uint32_t wb[2][KS][4];                                 // [hi | lo][k-step][b0 nt0, b1 nt0, b0 nt1, b1 nt1]
#pragma unroll
  for (int h = 0; h < 2; ++h)
#pragma unroll
    for (int kk = 0; kk < KS; ++kk)
      ldsm_x4(smem_u32(w_s + h * 16 * C) + ((lane & 7) + ((lane >> 4) << 3)) * P + kk * 32 + ((lane >> 3) & 1) * 16, wb[h][kk]);

  uint8_t* my_ring = ring + warp * kStages * kSlot;
  uint64_t* my_bars = bars + warp * kStages;
  uint32_t issued = 0, consumed = 0;
  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
    const int b = tile / p.blocks_per_img, yb = tile - b * p.blocks_per_img;
    const int y0 = yb * kC32RB;
    const int rows_here = min(kC32RB, p.H - y0);
    const int TR = (rows_here + 2) * p.W;
    const int flat0 = (b * p.H + y0 - 1) * p.W;               // halo rows may belong to a neighbour image: masked below
    const int ngroups = (TR + RPB - 1) / RPB;
    const int n_my = (ngroups - warp + kC32Warps - 1) / kC32Warps;
    auto issue = [&](int k) {
      if (lane == 0) {
        const uint32_t s = issued % kStages;
        mbar_expect_tx(&my_bars[s], kSlot);
        tma_load_4d(my_ring + s * kSlot, &tmap, 0, flat0 + (warp + kC32Warps * k) * RPB, 0, 0, &my_bars[s]);
      }
      ++issued;
    };
    for (int k = 0; k < kStages - 1 && k < n_my; ++k) issue(k);
    for (int k = 0; k < n_my; ++k) {
      __syncwarp();
      if (k + kStages - 1 < n_my) issue(k + kStages - 1);
      const uint32_t s = consumed % kStages;
      mbar_wait(&my_bars[s], (consumed / kStages) & 1, 0x43);
      ++consumed;
      const uint32_t ta = smem_u32(my_ring + s * kSlot);
      const int row = lane & 15, hi = lane >> 4;
#pragma unroll
      for (int f = 0; f < RPB / 16; ++f) {
        float acc[2][4];
#pragma unroll
        for (int n = 0; n < 2; ++n)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[n][e] = 0.f;
        const int br = f * 16 + row;                           // row inside the box
#pragma unroll
        for (int kk = 0; kk < KS; ++kk) {
          uint32_t a[4];
          ldsm_x4(ta + br * P + (((kk * 2 + hi) ^ ((br >> SH) & SWM)) << 4), a);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            mma_bf16_16816(acc[0], a, wb[h][kk][0], wb[h][kk][1]);
            mma_bf16_16816(acc[1], a, wb[h][kk][2], wb[h][kk][3]);
          }
        }
        const int r0 = (warp + kC32Warps * k) * RPB + f * 16 + (lane >> 2), r1 = r0 + 8;
        const int t0 = (lane & 3) * 2;
        if (r0 < TR) { ds[t0 * p.tr_max + r0] = acc[0][0]; ds[(t0 + 1) * p.tr_max + r0] = acc[0][1]; }
        if (r1 < TR) { ds[t0 * p.tr_max + r1] = acc[0][2]; ds[(t0 + 1) * p.tr_max + r1] = acc[0][3]; }
        if (t0 == 0) {
          if (r0 < TR) ds[8 * p.tr_max + r0] = acc[1][0];
          if (r1 < TR) ds[8 * p.tr_max + r1] = acc[1][2];
        }
      }
    }
    __syncthreads();
    const int npx = rows_here * p.W;
    for (int i = threadIdx.x; i < npx; i += kC32Threads) {
      const int y = i / p.W, x = i - y * p.W;
      float sacc = 0.f;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const int dy = t / 3 - 1, dx = t % 3 - 1;
        const int yy = y0 + y - dy, xx = x - dx;
        if (yy >= 0 && yy < p.H && xx >= 0 && xx < p.W) sacc += ds[t * p.tr_max + (y + 1 - dy) * p.W + xx];
      }
      p.gx[(static_cast<size_t>(b) * p.H + y0 + y) * p.W + x] = sacc;
    }
    __syncthreads();
  }
}

template <int KS>
static int d16_dgrad_launch_t(const void* gz, const float* w, float* gx, int B, int H, int W, cudaStream_t st) {
  constexpr int C = KS * 16, P = KS * 32;
  D16DgradP p{};
  p.w = w; p.gx = gx; p.B = B; p.H = H; p.W = W;
  p.blocks_per_img = (H + kC32RB - 1) / kC32RB;
  p.n_tiles = B * p.blocks_per_img;
  p.tr_max = (((kC32RB + 2) * W + 27) / 32) * 32 + 4;     // >= the band, == 4 (mod 32): tap planes 2i land 8 banks apart
  const long long R = 1LL * B * H * W;
  if (R > 0x7fffffffLL - 4096) return -1;
  const size_t smem = 1024 + size_t(kC32Warps) * 4 * 16 * P + 2 * 16 * C * 2 + size_t(9) * p.tr_max * 4 + kC32Warps * 4 * 8;
  if (smem > 150 * 1024) return -1;
  CUtensorMap tm;
  const uint64_t dims[4] = {uint64_t(C), uint64_t(R), 1, 1};
  const uint32_t box[4] = {uint32_t(C), 16, 1, 1};
  if (make_tmap_4d_bf16(&tm, gz, dims, box)) return 1;
  ensure_max_smem(reinterpret_cast<const void*>(d16_dgrad_kernel<KS>), (int)smem);
  int per_sm = int((224 * 1024) / (smem + 1024)); if (per_sm > 3) per_sm = 3; if (per_sm < 1) per_sm = 1;
  int grid = per_sm * sm_count(); if (grid > p.n_tiles) grid = p.n_tiles;
  d16_dgrad_kernel<KS><<<grid, kC32Threads, smem, st>>>(tm, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("dl1_dgrad (tensor-core) launch", e);
  return 0;
}

// returns -1 when the geometry is not covered (the caller keeps the CUDA-core kernel)
int d16_dgrad_launch(const void* gz, const float* w, float* gx, int C, int B, int H, int W, cudaStream_t st) {
  if (W < 4 || H < 1) return -1;
  if (C == 16) return d16_dgrad_launch_t<1>(gz, w, gx, B, H, W, st);
  if (C == 32) return d16_dgrad_launch_t<2>(gz, w, gx, B, H, W, st);
  if (C == 64) return d16_dgrad_launch_t<4>(gz, w, gx, B, H, W, st);
  return -1;
}

// x / act / gy: bf16 PNHWC [B][H+1][W+1][64] (channel stride 64, offset 0); fp32 images [B][1][H][W].
int c32_fwd_launch(const void* x, int B, int H, int W, const float* w, const float* bias, float inv_multiplier,
                   float power, float thres, int training, float* out, float* srs, float* z, cudaStream_t st) {
  C32FwdP p{};
  p.w = w; p.bias = bias; p.out = out; p.srs = srs; p.z = z;
  p.B = B; p.H = H; p.W = W; p.pitch = W + 1; p.P = (H + 1) * (W + 1);
  const long long R = 1LL * B * p.P;
  if (R > 0x7fffffffLL - 4096) return set_error("conv_out: batch too large for 32-bit row indices");
  p.blocks_per_img = (H + kC32RB - 1) / kC32RB;
  p.n_tiles = B * p.blocks_per_img;
  p.tr_max = (((kC32RB + 2) * p.pitch + 27) / 32) * 32 + 4;   // >= the band, == 4 (mod 32): conflict-free fragment stores
  p.inv_multiplier = inv_multiplier; p.power = power; p.thres = thres; p.training = training;
  const size_t smem = 1024 + size_t(kC32Warps) * kC32FwdStages * 2048 + 16 * 64 * 2 + size_t(9) * p.tr_max * 4 +
                      kC32Warps * kC32FwdStages * 8;
  if (smem > 110 * 1024) return -1;                        // very wide images: the caller keeps the general path
  CUtensorMap tm;
  if (make_tmap_2d_bf16(&tm, x, 64, (uint64_t)R, 128, 64, 16)) return 1;
  ensure_max_smem(reinterpret_cast<const void*>(c32_fwd_kernel), (int)smem);
  int grid = 2 * sm_count(); if (grid > p.n_tiles) grid = p.n_tiles;
  c32_fwd_kernel<<<grid, kC32Threads, smem, st>>>(tm, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("conv_out forward launch", e);
  return 0;
}

int c32_bwd_launch(const float* gz, const float* w, const void* act, void* gy, float* dw, float* db, int B, int H, int W,
                   float slope, cudaStream_t st) {
  C32BwdP p{};
  p.gz = gz; p.w = w; p.gy = static_cast<__nv_bfloat16*>(gy); p.dw = dw; p.db = db;
  p.B = B; p.H = H; p.W = W; p.pitch = W + 1; p.P = (H + 1) * (W + 1);
  const long long R = 1LL * B * p.P;
  if (R > 0x7fffffffLL - 4096) return set_error("conv_out: batch too large for 32-bit row indices");
  p.R = (int)R; p.n_groups = int((R + 15) / 16); p.slope = slope;
  const size_t smem = 1024 + size_t(kC32Warps) * kC32BwdStages * 2048 + kC32Warps * 16 * kC32GPitch + 64 * 16 * 2 +
                      64 * 16 * 4 + kC32Warps * kC32BwdStages * 8;
  CUtensorMap tm;
  if (make_tmap_2d_bf16(&tm, act, 64, (uint64_t)R, 128, 64, 16)) return 1;
  ensure_max_smem(reinterpret_cast<const void*>(c32_bwd_kernel), (int)smem);
  int grid = 2 * sm_count();
  const int need = (p.n_groups + kC32Warps - 1) / kC32Warps;
  if (grid > need) grid = need;
  c32_bwd_kernel<<<grid, kC32Threads, smem, st>>>(tm, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("conv_out backward launch", e);
  return 0;
}

}  // namespace srb
