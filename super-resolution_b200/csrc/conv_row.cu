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
// columns.  The issuer runs (k-block, sub-tile, dy, k) so sub-tile 0 finishes one pass (12 MMAs) before sub-tile 1:
// the epilogue of each sub-tile overlaps the other's MMAs without a second accumulator buffer.
// Work split: sub-tile units in contiguous ranges per CTA (persistent grid), balanced to one sub-tile.
// Warp roles: warp 0 TMA producer, warp 1 MMA issuer, warps 2..9 epilogue (two per TMEM lane quarter, 32 of the 64
// output channels each).
#include "common.cuh"
#include "sm100.cuh"

namespace srb {

constexpr int kRowThreads = 320;
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

__global__ void __launch_bounds__(kRowThreads, 1)
conv3x3_row_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                   const IgemmParams p, const RowSched sc) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_smem = smem;
  uint8_t* b_smem = a_smem + kRowAStages * p.a_stage_bytes;
  float* xch = reinterpret_cast<float*>(b_smem + sc.n_bslots * kBSlotBytes);     // [2 parity][2 half][4 lg][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(xch + 2 * 2 * 4 * 64);
  uint64_t* a_full = bars;                        // [kRowAStages]
  uint64_t* a_empty = a_full + kRowAStages;       // [kRowAStages]
  uint64_t* b_full = a_empty + kRowAStages;       // [kRowMaxBSlots]
  uint64_t* b_empty = b_full + kRowMaxBSlots;     // [kRowMaxBSlots]
  uint64_t* t_full = b_empty + kRowMaxBSlots;     // [2]  one per sub-tile accumulator
  uint64_t* t_empty = t_full + 2;                 // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
    for (int i = 0; i < kRowAStages; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < kRowMaxBSlots; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&t_full[i], 1); mbar_init(&t_empty[i], kRowEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();          // everything above overlapped the previous kernel; its results are visible from here on

  // this CTA's contiguous range of (n-tile, sub-tile) units
  const long long u_begin = sc.units * blockIdx.x / gridDim.x;
  const long long u_end = sc.units * (blockIdx.x + 1) / gridDim.x;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0;
      for (long long u = u_begin; u < u_end;) {
        const int nt = static_cast<int>(u / sc.n_sub);
        const int s = static_cast<int>(u - static_cast<long long>(nt) * sc.n_sub);
        int cnt = sc.n_sub - s;
        if (cnt > sc.mt_max) cnt = sc.mt_max;
        if (cnt > u_end - u) cnt = static_cast<int>(u_end - u);
        const int q0 = s * kRowSub;
        for (int kb = 0; kb < p.KB; ++kb) {
          mbar_wait(&a_empty[sa], pa ^ 1, 1);
          mbar_expect_tx(&a_full[sa], p.a_stage_bytes);
          for (int bx = 0; bx < p.a_nbox; ++bx)
            tma_load_2d(a_smem + sa * p.a_stage_bytes + bx * p.a_box_rows * 128, &tmap_x, p.x_coff + kb * 64,
                        q0 - p.halo + bx * p.a_box_rows, &a_full[sa]);
          if (++sa == kRowAStages) { sa = 0; pa ^= 1; }
          for (int dy = 0; dy < 3; ++dy) {
            mbar_wait(&b_empty[sb], pb ^ 1, 2);
            mbar_expect_tx(&b_full[sb], kBSlotBytes);
            for (int dx = 0; dx < 3; ++dx)
              tma_load_2d(b_smem + sb * kBSlotBytes + dx * kBTileBytes, &tmap_w, 0,
                          ((dy * 3 + dx) * p.KB + kb) * p.cout + nt * 64, &b_full[sb]);
            if (++sb == (uint32_t)sc.n_bslots) { sb = 0; pb ^= 1; }
          }
        }
        u += cnt;
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 192, 0, 0);
      constexpr uint64_t desc_hi = (uint64_t(16 >> 4) << 16) | (uint64_t(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      const uint32_t a0 = (smem_u32(a_smem) & 0x3FFFFu) >> 4;
      const uint32_t b0 = (smem_u32(b_smem) & 0x3FFFFu) >> 4;
      const uint32_t a_stage16 = p.a_stage_bytes >> 4;
      const uint32_t pitch16 = static_cast<uint32_t>(p.pitch) * 8u;      // one image row, in 16-byte units
      uint32_t sa = 0, pa = 0, sb0 = 0, pb0 = 0;                         // sb0/pb0: ring position of this k-block's dy = 0
      uint32_t pt0 = 0, pt1 = 0;                                         // accumulator phases
      for (long long u = u_begin; u < u_end;) {
        const int nt = static_cast<int>(u / sc.n_sub);
        const int s = static_cast<int>(u - static_cast<long long>(nt) * sc.n_sub);
        int cnt = sc.n_sub - s;
        if (cnt > sc.mt_max) cnt = sc.mt_max;
        if (cnt > u_end - u) cnt = static_cast<int>(u_end - u);
        for (int kb = 0; kb < p.KB; ++kb) {
          mbar_wait(&a_full[sa], pa, 4);
          const uint32_t a_lo = a0 + sa * a_stage16;
#pragma unroll 1
          for (int mt = 0; mt < cnt; ++mt) {
            if (kb == 0) {
              mbar_wait(&t_empty[mt], (mt ? pt1 : pt0) ^ 1, 3);
            }
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + mt * 256;
            const bool last_pass = mt == cnt - 1;
            uint32_t row16 = a_lo + static_cast<uint32_t>(mt * kRowSub) * 8u;
            uint32_t sb = sb0, pb = pb0;
#pragma unroll 1
            for (int dy = 0; dy < 3; ++dy) {
              if (mt == 0) {
                mbar_wait(&b_full[sb], pb, 5);
                tc_fence_after();
              }
              const uint32_t b_lo = b0 + sb * (kBSlotBytes >> 4);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint64_t ad = desc_hi | static_cast<uint64_t>(row16 + k * 2);
                const uint64_t bd = desc_hi | static_cast<uint64_t>(b_lo + k * 2);
                umma_bf16(d_tmem, ad, bd, idesc, k != 0 ? 1u : static_cast<uint32_t>((kb | dy) != 0));
              }
              if (last_pass) umma_commit(&b_empty[sb]);
              row16 += pitch16;
              if (++sb == (uint32_t)sc.n_bslots) { sb = 0; pb ^= 1; }
            }
            if (kb == p.KB - 1) umma_commit(&t_full[mt]);
          }
          umma_commit(&a_empty[sa]);
          sb0 += 3;
          if (sb0 >= (uint32_t)sc.n_bslots) { sb0 -= sc.n_bslots; pb0 ^= 1; }
          if (++sa == kRowAStages) { sa = 0; pa ^= 1; }
        }
        pt0 ^= 1;
        if (cnt == 2) pt1 ^= 1;
        u += cnt;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps 2..9
    const int lg = warp & 3;                    // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;           // which 32 of the 64 output channels
    uint32_t pt0 = 0, pt1 = 0, par = 0;
    for (long long u = u_begin; u < u_end;) {
      const int nt = static_cast<int>(u / sc.n_sub);
      const int s = static_cast<int>(u - static_cast<long long>(nt) * sc.n_sub);
      int cnt = sc.n_sub - s;
      if (cnt > sc.mt_max) cnt = sc.mt_max;
      if (cnt > u_end - u) cnt = static_cast<int>(u_end - u);
#pragma unroll 1
      for (int mt = 0; mt < cnt; ++mt) {
        const int m = lg * 32 + lane;
        const int q = (s + mt) * kRowSub + m;
        const int r = q % p.P;
        const int py = r / p.pitch, px = r - py * p.pitch;
        const bool valid = (m < kRowSub) && (q < p.R) && (px < p.W) && (py < p.H);
        mbar_wait(&t_full[mt], mt ? pt1 : pt0, 6);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lg * 32) << 16) + mt * 256 + half * 32;
        float* my_x = xch + ((par * 2 + half) * 4 + lg) * 64;
        const float* nx_x = my_x + 64;            // the next lane quarter's slot (read only when lg < 3)
        float t[32];
        {
          uint32_t d2[32], d1[32];
          tmem_ld32(taddr + 128, d2);
          tmem_ld32(taddr + 64, d1);
          tmem_ld_wait();
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
          asm volatile("bar.sync %0, 128;" ::"r"(1 + half) : "memory");
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
        if (valid) igemm_store_row(p, v, q, py, px, nt, half);
      }
      pt0 ^= 1;
      if (cnt == 2) pt1 ^= 1;
      u += cnt;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ---------------------------------------------------------------------------------- host side
static size_t row_smem_bytes(uint32_t a_stage_bytes, int n_bslots) {
  return 1024 + size_t(kRowAStages) * a_stage_bytes + size_t(n_bslots) * kBSlotBytes + 2 * 2 * 4 * 64 * sizeof(float) + 256;
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
  const int pitch = a->W + 1, P = (a->H + 1) * (a->W + 1);
  const long long R = 1LL * a->batch * P;
  const int sms = sm_count();
  const int n_ntiles = a->cout / 64;
  const int n_sub = int((R + kRowSub - 1) / kRowSub);
  const long long units = 1LL * n_sub * n_ntiles;

  int mt_max = units >= 2LL * sms ? 2 : 1;
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

  RowSched sc{};
  sc.n_sub = n_sub; sc.units = units; sc.mt_max = mt_max; sc.n_bslots = n_bslots;
  const int grid = units < sms ? (int)units : sms;
  static bool attr = false;
  if (!attr) { cudaFuncSetAttribute(conv3x3_row_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024); attr = true; }
  cudaError_t e = launch_pdl(conv3x3_row_kernel, dim3(grid), dim3(kRowThreads), smem, stream, tx, tw, p, sc);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("conv3x3_row launch", e);
  return 0;
}

}  // namespace srb
