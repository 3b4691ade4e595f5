// K1b: 3x3 stride-1 pad-1 convolution, "filter row per MMA" formulation (tcgen05 / TMEM / TMA, sm_100a).
// Same contract as conv3x3_igemm_kernel (conv_igemm.cu), which it replaces on the hot path (models.py:19,36-41,53,
// 67,87-89,97-98 and their input gradients).
//
// Why.  With Cout = 64 fixed by the model, a 128x64x16 MMA reads 4 KB of A + 2 KB of B from shared memory per 32
// tensor-pipe clocks = 192 B/clk against the 128 B/clk an SM delivers: the one-tap-per-MMA kernel is shared-memory
// bound at 2/3 of the tensor rate (tools/probe1.py rate measures 48 clk for that MMA).  The three taps of a filter
// row (dy fixed, dx = -1, 0, +1) read the SAME 128 pixels shifted by one row of the flat PNHWC matrix each, so this
// kernel issues them as ONE 128x192x16 MMA whose B operand is the 192 x 64 stack [W(dy,-1); W(dy,0); W(dy,+1)] and
// whose A operand is the window at shift dy*pitch - 1:
//     D_j[m] = sum_k W(dy, j-1)[., k] X[q0 + m + dy*pitch - 1, k]        (column block j = 0, 1, 2)
//     Y[q0 + m] = D_0[m] + D_1[m + 1] + D_2[m + 2]
// A is read once instead of three times (10 KB per 96 clk, below the shared-memory limit), a third of the MMA
// instructions are issued (one issuer thread is enough), and the row shifts of the final sum are two warp shuffles per
// accumulator column in the epilogue.  A 128-row MMA tile therefore yields 126 output rows: sub-tiles advance by 126.
//
// CTA tile: up to two sub-tiles (254 A rows + halo, ONE TMA window per 64-channel k-block) so that each 24 KB weight
// slot (a filter row, exactly the layout the N = 192 descriptor wants) feeds two MMAs sets.  Accumulators: 2 x 192 TMEM
// columns, each released as soon as the epilogue has it in registers, so the issuer of a sub-tile resumes while the
// epilogue warps are still in the global-memory part (the two issuers drift into a stagger of one TMEM read-out).
// Work split: sub-tile units in contiguous ranges per CTA (persistent grid), balanced to one sub-tile.
// Warp roles: warp 0 TMA producer, warps 1-2 MMA issuers (one per sub-tile), warps 3..10 epilogue (two per TMEM lane
// quarter, 32 of the 64 output channels each).
#include "common.cuh"
#include "sm100.cuh"

namespace srb {

constexpr int kRowThreads = 352;        // warp 0 TMA, warps 1-2 MMA issuers, warps 3-10 epilogue
constexpr int kRowMaxCout = 1024;       // bias staged in shared memory
constexpr int kRowEpiWarps = 8;
constexpr int kRowSub = 126;           // output rows per 128-row sub-tile
constexpr int kRowMaxBSlots = 8;       // weight ring: one slot = one filter row = 3 x (64 x 64) bf16 = 24 KB; as many
                                       // slots as fit beside the two halo windows (6 at 20x20, 4 at 80x80)
constexpr int kRowAStages = 2;

struct RowSched {
  int n_sub;          // sub-tiles along the flat row dimension
  long long units;    // n_sub * n_ntiles
  int mt_max;         // 1 or 2 sub-tiles per CTA tile
  int n_bslots;       // weight ring depth (>= 4)
};

__device__ __forceinline__ float shfl_down1(float v) { return __shfl_down_sync(0xffffffffu, v, 1); }

// tcgen05.mma with the descriptors given as their low words (start address field); the high word (SBO, version,
// swizzle) is the same for both K-major operands.
__device__ __forceinline__ void umma_bf16_lo(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}\n" ::"r"(d_tmem),
      "r"(a_lo | (1u << 16)), "r"(b_lo | (1u << 16)), "r"(hi), "r"(idesc), "r"(accumulate)
      : "memory");
}

// Row epilogue of the filter-row kernel: as igemm_store_row, with the bias read from shared memory, the mask already in
// registers and the pixel coordinates supplied by the caller.
__device__ __forceinline__ void row_store(const IgemmParams& p, float (&v)[32], const uint4 (&mk)[4], const float* bias32,
                                          int q, int img, int py, int px, int nt, int half) {
  size_t yrow;
  int ycoff = p.y_coff;
  if (p.store_mode == SRB_STORE_SHUFFLE) {
    const int cblocks = p.n_ntiles >> 2;  // 64-wide blocks per shuffled output channel group (Cy / 64)
    const int sub = nt / cblocks;         // (i,j) = (sub>>1, sub&1)
    yrow = static_cast<size_t>(img) * p.y_P + static_cast<size_t>(2 * py + (sub >> 1)) * p.y_pitch +
           (2 * px + (sub & 1));
    ycoff += (nt % cblocks) * 64;
  } else {
    yrow = static_cast<size_t>(q);
    ycoff += nt * 64;
  }
  const int c0 = nt * 64 + half * 32;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 b = reinterpret_cast<const float4*>(bias32)[i];
    v[4 * i] = (v[4 * i] + b.x) * p.alpha; v[4 * i + 1] = (v[4 * i + 1] + b.y) * p.alpha;
    v[4 * i + 2] = (v[4 * i + 2] + b.z) * p.alpha; v[4 * i + 3] = (v[4 * i + 3] + b.w) * p.alpha;
  }
  if (p.res1) {
    const uint4* rp = reinterpret_cast<const uint4*>(p.res1 + static_cast<size_t>(q) * p.res1_cstride +
                                                     p.res1_coff + c0);
    uint4 u[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) u[i] = __ldg(rp + i);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t w[4] = {u[i].x, u[i].y, u[i].z, u[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[8 * i + 2 * j] += p.res1_scale * bf16lo(w[j]);
        v[8 * i + 2 * j + 1] += p.res1_scale * bf16hi(w[j]);
      }
    }
  }
  if (p.res2) {
    const uint4* rp = reinterpret_cast<const uint4*>(p.res2 + static_cast<size_t>(q) * p.res2_cstride +
                                                     p.res2_coff + c0);
    uint4 u[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) u[i] = __ldg(rp + i);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t w[4] = {u[i].x, u[i].y, u[i].z, u[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[8 * i + 2 * j] += p.res2_scale * bf16lo(w[j]);
        v[8 * i + 2 * j + 1] += p.res2_scale * bf16hi(w[j]);
      }
    }
  }
  if (p.act == SRB_ACT_LRELU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * p.slope;
  } else if (p.act == SRB_ACT_LRELU_MASK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t w[4] = {mk[i].x, mk[i].y, mk[i].z, mk[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[8 * i + 2 * j] *= bf16lo(w[j]) > 0.f ? 1.f : p.slope;
        v[8 * i + 2 * j + 1] *= bf16hi(w[j]) > 0.f ? 1.f : p.slope;
      }
    }
  }
  uint4* yp = reinterpret_cast<uint4*>(p.y + yrow * p.y_cstride + ycoff + half * 32);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint4 o;
    o.x = pack_bf16(v[8 * i], v[8 * i + 1]);
    o.y = pack_bf16(v[8 * i + 2], v[8 * i + 3]);
    o.z = pack_bf16(v[8 * i + 4], v[8 * i + 5]);
    o.w = pack_bf16(v[8 * i + 6], v[8 * i + 7]);
    yp[i] = o;
  }
}

__global__ void __launch_bounds__(kRowThreads, 1)
conv3x3_row_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                   const IgemmParams p, const RowSched sc) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_smem = smem;
  uint8_t* b_smem = a_smem + kRowAStages * p.a_stage_bytes;
  float* xch = reinterpret_cast<float*>(b_smem + sc.n_bslots * kBSlotBytes);     // [2 parity][2 half][4 lg][64]
  float* bias_s = xch + 2 * 2 * 4 * 64;                                           // [kRowMaxCout]
  uint64_t* bars = reinterpret_cast<uint64_t*>(bias_s + kRowMaxCout);
  uint64_t* a_full = bars;                        // [kRowAStages]
  uint64_t* a_empty = a_full + kRowAStages;       // [kRowAStages]
  uint64_t* b_full = a_empty + kRowAStages;       // [kRowMaxBSlots]
  uint64_t* b_empty = b_full + kRowMaxBSlots;     // [kRowMaxBSlots]
  uint64_t* t_full = b_empty + kRowMaxBSlots;     // [2]  one per sub-tile accumulator
  uint64_t* t_empty = t_full + 2;                 // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const bool stamp = (p.dbg & 32) && blockIdx.x == 0;       // clock64 timeline of CTA 0 (tools/exp_row_stamps.py)
  if (stamp && threadIdx.x == 0) g_igemm_stamps[0] = clock64();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
    for (int i = 0; i < kRowAStages; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 2); }
    for (int i = 0; i < kRowMaxBSlots; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 2); }
    for (int i = 0; i < 2; ++i) { mbar_init(&t_full[i], 1); mbar_init(&t_empty[i], kRowEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (stamp && threadIdx.x == 0) g_igemm_stamps[1] = clock64();
  pdl_trigger();
  pdl_wait();          // everything above overlapped the previous kernel; its results are visible from here on
  if (stamp && threadIdx.x == 0) g_igemm_stamps[2] = clock64();

  // this CTA's contiguous range of (n-tile, sub-tile) units
  const long long u_begin = sc.units * blockIdx.x / gridDim.x;
  const long long u_end = sc.units * (blockIdx.x + 1) / gridDim.x;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {   // elect.sync: ptxas then keeps descriptors and barrier addresses in uniform registers
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0;
      int na = 0, nb = 0;                        // loads issued so far (timing experiments, p.dbg)
      for (long long u = u_begin; u < u_end;) {
        const int nt = static_cast<int>(u / sc.n_sub);
        const int s = static_cast<int>(u - static_cast<long long>(nt) * sc.n_sub);
        int cnt = sc.n_sub - s;
        if (cnt > sc.mt_max) cnt = sc.mt_max;
        if (cnt > u_end - u) cnt = static_cast<int>(u_end - u);
        const int q0 = s * kRowSub;
        for (int kb = 0; kb < p.KB; ++kb) {
          mbar_wait(&a_empty[sa], pa ^ 1, 1);
          if ((p.dbg & 2) && na >= kRowAStages) mbar_arrive(&a_full[sa]);
          else {
            mbar_expect_tx(&a_full[sa], p.a_stage_bytes);
            for (int bx = 0; bx < p.a_nbox; ++bx)
              tma_load_2d(a_smem + sa * p.a_stage_bytes + bx * p.a_box_rows * 128, &tmap_x, p.x_coff + kb * 64,
                          q0 - p.halo + bx * p.a_box_rows, &a_full[sa]);
          }
          ++na;
          if (++sa == kRowAStages) { sa = 0; pa ^= 1; }
          for (int dy = 0; dy < 3; ++dy) {
            mbar_wait(&b_empty[sb], pb ^ 1, 2);
            if ((p.dbg & 1) && nb >= sc.n_bslots) mbar_arrive(&b_full[sb]);
            else {
              mbar_expect_tx(&b_full[sb], kBSlotBytes);
              for (int dx = 0; dx < 3; ++dx)
                tma_load_2d(b_smem + sb * kBSlotBytes + dx * kBTileBytes, &tmap_w, 0,
                            ((dy * 3 + dx) * p.KB + kb) * p.cout + nt * 64, &b_full[sb]);
            }
            ++nb;
            if (++sb == (uint32_t)sc.n_bslots) { sb = 0; pb ^= 1; }
          }
        }
        u += cnt;
      }
    }
  } else if (warp <= 2) {
    // ------------------------------------------------------------------ MMA issuers: warp 1 -> sub-tile 0, warp 2 -> sub-tile 1
    // One thread cannot keep the tensor pipe busy (about 150 clk of waits, descriptor arithmetic and commits per MMA
    // against the 96 clk a 128x192x16 MMA takes), so each sub-tile has its own issuer with its own accumulator.  Both
    // walk the same ring positions and both arrive on the slot / stage "empty" barriers (count 2); an issuer whose
    // sub-tile does not exist in a tile (odd tail of the CTA's range) walks the rings without issuing.
    if (elect_one()) {   // elect.sync: ptxas then keeps descriptors and barrier addresses in uniform registers
      const int mt = warp - 1;
      constexpr uint32_t idesc = make_idesc_bf16(128, 192, 0, 0);
      constexpr uint32_t desc_hi32 = (1024u >> 4) | (1u << 14) | (2u << 29);   // SBO 1024, version 1, SWIZZLE_128B
      const uint32_t a0 = ((smem_u32(a_smem) & 0x3FFFFu) >> 4) + static_cast<uint32_t>(mt * kRowSub) * 8u;
      const uint32_t b0 = (smem_u32(b_smem) & 0x3FFFFu) >> 4;
      const uint32_t a_stage16 = p.a_stage_bytes >> 4;
      const uint32_t pitch16 = static_cast<uint32_t>(p.pitch) * 8u;      // one image row, in 16-byte units
      const uint32_t d_tmem = tmem_base + mt * 256;
      const bool no_mma = (p.dbg & 16) != 0;
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0, pt = 0;
      for (long long u = u_begin; u < u_end;) {
        const int nt = static_cast<int>(u / sc.n_sub);
        const int s = static_cast<int>(u - static_cast<long long>(nt) * sc.n_sub);
        int cnt = sc.n_sub - s;
        if (cnt > sc.mt_max) cnt = sc.mt_max;
        if (cnt > u_end - u) cnt = static_cast<int>(u_end - u);
        const bool mine = mt < cnt && !no_mma;
        if (mt < cnt) {
          mbar_wait(&t_empty[mt], pt ^ 1, 3);
          tc_fence_after();
        }
        uint32_t acc = 0;
        for (int kb = 0; kb < p.KB; ++kb) {
          mbar_wait(&a_full[sa], pa, 4);
          if (stamp && mt == 0 && kb == 0 && u == u_begin) g_igemm_stamps[3] = clock64();
          uint32_t row16 = a0 + sa * a_stage16;
#pragma unroll
          for (int dy = 0; dy < 3; ++dy) {
            mbar_wait(&b_full[sb], pb, 5);
            tc_fence_after();
            const uint32_t b_lo = b0 + sb * (kBSlotBytes >> 4);
            if (mine) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                umma_bf16_lo(d_tmem, row16 + k * 2, b_lo + k * 2, desc_hi32, idesc, acc);
                acc = 1;
              }
            }
            umma_commit(&b_empty[sb]);
            row16 += pitch16;
            if (++sb == (uint32_t)sc.n_bslots) { sb = 0; pb ^= 1; }
          }
          umma_commit(&a_empty[sa]);
          if (++sa == kRowAStages) { sa = 0; pa ^= 1; }
        }
        if (mt < cnt) {
          umma_commit(&t_full[mt]);
          pt ^= 1;
        }
        if (stamp && mt == 0) g_igemm_stamps[u == u_begin ? 4 : 8] = clock64();
        u += cnt;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps 3..10
    const int lg = warp & 3;                    // TMEM lane quarter this warp may access
    const int half = (warp - 3) >> 2;           // which 32 of the 64 output channels
    const int m = lg * 32 + lane;
    // bias of the whole layer in shared memory (a parameter, not produced by the previous kernel)
    for (int i = threadIdx.x - 96; i < p.cout; i += kRowEpiWarps * 32) bias_s[i] = p.bias ? __ldg(p.bias + i) : 0.f;
    asm volatile("bar.sync 3, 256;" ::: "memory");
    const float inv_pitch = 1.0f / static_cast<float>(p.pitch);
    uint32_t pt0 = 0, pt1 = 0, par = 0;
    for (long long u = u_begin; u < u_end;) {
      const int nt = static_cast<int>(u / sc.n_sub);
      const int s = static_cast<int>(u - static_cast<long long>(nt) * sc.n_sub);
      int cnt = sc.n_sub - s;
      if (cnt > sc.mt_max) cnt = sc.mt_max;
      if (cnt > u_end - u) cnt = static_cast<int>(u_end - u);
#pragma unroll 1
      for (int mt = 0; mt < cnt; ++mt) {
        // pixel of this thread's row: one warp-uniform division per sub-tile, the rest is a few adds
        const int qb = (s + mt) * kRowSub;
        int img = qb / p.P;
        int r = qb - img * p.P + m;
        while (r >= p.P) { r -= p.P; ++img; }
        int py = __float2int_rz(static_cast<float>(r) * inv_pitch);
        if (py * p.pitch > r) --py;
        else if ((py + 1) * p.pitch <= r) ++py;
        const int px = r - py * p.pitch;
        const int q = qb + m;
        const bool valid = (m < kRowSub) && (q < p.R) && (px < p.W) && (py < p.H) && !(p.dbg & 4);
        // the LeakyReLU' mask does not depend on the accumulator: fetch it before waiting for the MMAs
        uint4 mk[4];
        if (p.act == SRB_ACT_LRELU_MASK && valid) {
          const uint4* mp = reinterpret_cast<const uint4*>(p.mask + static_cast<size_t>(q) * p.mask_cstride +
                                                           p.mask_coff + nt * 64 + half * 32);
#pragma unroll
          for (int i = 0; i < 4; ++i) mk[i] = __ldg(mp + i);
        }
        mbar_wait(&t_full[mt], mt ? pt1 : pt0, 6);
        tc_fence_after();
        if (stamp && warp == 3 && lane == 0 && mt == 0 && u == u_begin) g_igemm_stamps[5] = clock64();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lg * 32) << 16) + mt * 256 + half * 32;
        float* my_x = xch + ((par * 2 + half) * 4 + lg) * 64;
        const float* nx_x = my_x + 64;            // the next lane quarter's slot (read only when lg < 3)
        if (p.dbg & 8) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&t_empty[mt]);
          continue;
        }
        float t[32];
        {
          uint32_t d2[32], d1[32];
          tmem_ld32(taddr + 128, d2);
          tmem_ld32(taddr + 64, d1);
          tmem_ld_wait();
          if (stamp && warp == 3 && lane == 0 && mt == 0 && u == u_begin) g_igemm_stamps[11] = clock64();
#pragma unroll
          for (int i = 0; i < 32; ++i) t[i] = __uint_as_float(d1[i]) + shfl_down1(__uint_as_float(d2[i]));
          if (lane == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              reinterpret_cast<float4*>(my_x)[i] = make_float4(__uint_as_float(d2[4 * i]), __uint_as_float(d2[4 * i + 1]),
                                                               __uint_as_float(d2[4 * i + 2]), __uint_as_float(d2[4 * i + 3]));
              reinterpret_cast<float4*>(my_x + 32)[i] = make_float4(t[4 * i], t[4 * i + 1], t[4 * i + 2], t[4 * i + 3]);
            }
          }
          if (stamp && warp == 3 && lane == 0 && mt == 0 && u == u_begin) g_igemm_stamps[12] = clock64();
          asm volatile("bar.sync %0, 128;" ::"r"(1 + half) : "memory");
          if (stamp && warp == 3 && lane == 0 && mt == 0 && u == u_begin) g_igemm_stamps[13] = clock64();
          if (lane == 31 && lg < 3) {
            // row 31 of this quarter: D_2[m + 1] is row 0 of the next quarter
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 x = reinterpret_cast<const float4*>(nx_x)[i];
              t[4 * i] = __uint_as_float(d1[4 * i]) + x.x;
              t[4 * i + 1] = __uint_as_float(d1[4 * i + 1]) + x.y;
              t[4 * i + 2] = __uint_as_float(d1[4 * i + 2]) + x.z;
              t[4 * i + 3] = __uint_as_float(d1[4 * i + 3]) + x.w;
            }
          }
        }
        float v[32];
        {
          uint32_t d0[32];
          tmem_ld32(taddr, d0);
          tmem_ld_wait();
          if (stamp && warp == 3 && lane == 0 && mt == 0 && u == u_begin) g_igemm_stamps[14] = clock64();
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(d0[i]) + shfl_down1(t[i]);
          if (lane == 31 && lg < 3) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 x = reinterpret_cast<const float4*>(nx_x + 32)[i];
              v[4 * i] = __uint_as_float(d0[4 * i]) + x.x;
              v[4 * i + 1] = __uint_as_float(d0[4 * i + 1]) + x.y;
              v[4 * i + 2] = __uint_as_float(d0[4 * i + 2]) + x.z;
              v[4 * i + 3] = __uint_as_float(d0[4 * i + 3]) + x.w;
            }
          }
        }
        // the accumulator is in registers: hand it back before the global-memory part of the epilogue
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[mt]);
        par ^= 1;
        if (stamp && warp == 3 && lane == 0 && mt == 0 && u == u_begin) g_igemm_stamps[6] = clock64();
        if (valid) row_store(p, v, mk, bias_s + nt * 64 + half * 32, q, img, py, px, nt, half);
      }
      pt0 ^= 1;
      if (cnt == 2) pt1 ^= 1;
      if (stamp && warp == 3 && lane == 0) g_igemm_stamps[u == u_begin ? 7 : 9] = clock64();
      u += cnt;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
  if (stamp && threadIdx.x == 0) g_igemm_stamps[10] = clock64();
}

// ---------------------------------------------------------------------------------- host side
static size_t row_smem_bytes(uint32_t a_stage_bytes, int n_bslots) {
  return 1024 + size_t(kRowAStages) * a_stage_bytes + size_t(n_bslots) * kBSlotBytes + (2 * 2 * 4 * 64 + kRowMaxCout) * sizeof(float) + 256;
}

static void row_window(int mt, int pitch, int* box_rows, int* nbox, uint32_t* stage_bytes) {
  const int win = kRowSub * (mt - 1) + 128 + 2 * pitch;       // A rows [q0 - pitch - 1, q0 + 126 (mt-1) + 127 + pitch - 1]
  const int win8 = (win + 7) & ~7;
  *nbox = (win8 + 255) / 256;
  *box_rows = (((win8 + *nbox - 1) / *nbox) + 7) & ~7;
  *stage_bytes = uint32_t(*nbox) * *box_rows * 128;
}

// returns 0 on success, 1 on error, -1 if the geometry does not fit (caller falls back to the one-tap kernel)
int conv3x3_row_launch(const srb_conv3x3_args* a, cudaStream_t stream) {
  if (a->cout > kRowMaxCout) return -1;
  const int pitch = a->W + 1, P = (a->H + 1) * (a->W + 1);
  const long long R = 1LL * a->batch * P;
  const int sms = sm_count();
  const int n_ntiles = a->cout / 64;
  const int n_sub = int((R + kRowSub - 1) / kRowSub);
  const long long units = 1LL * n_sub * n_ntiles;

  int mt_max = units >= 2LL * sms ? 2 : 1;
  { static int f = -1; if (f < 0) { const char* e = getenv("SRB_ROW_MT"); f = e ? atoi(e) : 0; } if (f == 1) mt_max = 1; }
  int box_rows, nbox; uint32_t a_stage_bytes;
  row_window(mt_max, pitch, &box_rows, &nbox, &a_stage_bytes);
  if (row_smem_bytes(a_stage_bytes, 4) > 227 * 1024 && mt_max == 2) {
    mt_max = 1;
    row_window(mt_max, pitch, &box_rows, &nbox, &a_stage_bytes);
  }
  if (row_smem_bytes(a_stage_bytes, 4) > 227 * 1024) return -1;
  int n_bslots = 4;
  while (n_bslots < kRowMaxBSlots && row_smem_bytes(a_stage_bytes, n_bslots + 1) <= 227 * 1024) ++n_bslots;
  const size_t smem = row_smem_bytes(a_stage_bytes, n_bslots);

  CUtensorMap tx, tw;
  if (make_tmap_2d_bf16(&tx, a->x, (uint64_t)a->x_cstride, (uint64_t)R, (uint64_t)a->x_cstride * 2, 64, box_rows))
    return 1;
  if (make_tmap_2d_bf16(&tw, a->w, 64, (uint64_t)9 * (a->cin / 64) * a->cout, 128, 64, 64)) return 1;

  IgemmParams p{};
  p.R = (int)R; p.P = P; p.pitch = pitch; p.H = a->H; p.W = a->W;
  p.KB = a->cin / 64; p.cout = a->cout; p.n_ntiles = n_ntiles;
  p.num_work = 0;
  p.x_coff = a->x_coff; p.halo = pitch + 1; p.a_box_rows = box_rows; p.a_nbox = nbox; p.a_stage_bytes = a_stage_bytes;
  p.bias = a->bias;
  p.y = (__nv_bfloat16*)a->y; p.y_cstride = a->y_cstride; p.y_coff = a->y_coff; p.store_mode = a->store_mode;
  if (a->store_mode == SRB_STORE_SHUFFLE) { p.y_pitch = 2 * a->W + 1; p.y_P = (2 * a->H + 1) * (2 * a->W + 1); }
  else { p.y_pitch = pitch; p.y_P = P; }
  p.alpha = a->alpha; p.act = a->act; p.slope = a->slope;
  p.mask = (const __nv_bfloat16*)a->mask; p.mask_cstride = a->mask_cstride; p.mask_coff = a->mask_coff;
  p.res1 = (const __nv_bfloat16*)a->res1; p.res1_cstride = a->res1_cstride; p.res1_coff = a->res1_coff; p.res1_scale = a->res1_scale;
  p.res2 = (const __nv_bfloat16*)a->res2; p.res2_cstride = a->res2_cstride; p.res2_coff = a->res2_coff; p.res2_scale = a->res2_scale;
  if (p.act == SRB_ACT_LRELU_MASK && !p.mask) return set_error("conv3x3: mask activation without mask tensor");
  {                       // timing experiments only (results are wrong on purpose): SRB_ROW_DBG bits, see the kernel
    static int dbg = -1;
    if (dbg < 0) { const char* e = getenv("SRB_ROW_DBG"); dbg = e ? atoi(e) : 0; }
    p.dbg = dbg;
  }

  RowSched sc{};
  sc.n_sub = n_sub; sc.units = units; sc.mt_max = mt_max; sc.n_bslots = n_bslots;
  const int grid = units < sms ? (int)units : sms;
  ensure_max_smem(reinterpret_cast<const void*>(conv3x3_row_kernel), 227 * 1024);
  cudaError_t e = launch_pdl(conv3x3_row_kernel, dim3(grid), dim3(kRowThreads), smem, stream, tx, tw, p, sc);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("conv3x3_row launch", e);
  return 0;
}

}  // namespace srb
