// K1d: chained 3x3 convolutions, "2 + 1 taps per filter row" formulation (tcgen05 / TMEM / TMA, sm_100a).
// Same contract and chaining protocol as conv3x3_chain_kernel (conv_chain.cu); opt-in (SRB_CONV_T21=1 or flags bit 20).
//
// Why.  A 128x64x16 MMA reads 6 KB of shared memory per 32 tensor clocks and is therefore shared-memory bound at 48 clk
// (DESIGN.md 4.1): nine of them per filter position cost 3 x 144 clk per k-step.  Here the taps dx = -1 and dx = 0 of a
// filter row are ONE 128x128x16 MMA (A read once for both: 8 KB per 64 clk, exactly the shared-memory rate) and dx = +1
// is a 128x64x16 MMA into the same accumulator: 112 clk per filter row and k-step instead of 144.
//     MMA 1: A = window rows at shift dy*pitch - 1,  B = [W(dy,-1); W(dy,0)]  ->  D0[m] (tap -1, in place), D1[m] (tap 0)
//     MMA 2: A = window rows at shift dy*pitch + 1,  B = W(dy,+1)             ->  D0[m] += ...
//     Y[q0 + m] = D0[m] + D1[m + 1]
// so a 128-row MMA tile yields 127 output rows (sub-tiles advance by 127) and the epilogue shifts D1 by one row: a
// shuffle inside a warp, one float4 exchange through shared memory across the four lane quarters.  Accumulators:
// 128 columns per sub-tile, two sub-tiles per CTA tile, double buffered = all 512 TMEM columns; one issuer per
// sub-tile (elect.sync issue is fast enough: 112 clk of tensor time per MMA pair and issuer).
// The tile -> CTA map rotates by one CTA per layer: with 254-row tiles the batch-512 LR geometry has 889 tiles = 6 x 148
// + 1, and a fixed map would make the same CTA carry the seventh tile in every layer of a chain.
#include "common.cuh"
#include "sm100.cuh"

namespace srb {

constexpr int kT21Sub = 127;                 // output rows per 128-row MMA sub-tile
constexpr int kT21Rows = 2 * kT21Sub;        // rows per CTA tile
constexpr int kT21Iss = 2;                   // issuer warps (one per sub-tile)
constexpr int kT21XchBytes = 2 * 2 * 4 * 32 * 4;   // [parity][half][lane quarter][32 floats]

__global__ void __launch_bounds__(kIgemmThreads, 1)
conv3x3_chain21_kernel(const __grid_constant__ ChainParams cp) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int NB = kBSlots;
  const uint32_t a_stage_bytes = cp.L[0].a_stage_bytes;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_smem = smem;
  uint8_t* b_smem = a_smem + kAStages * a_stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_smem + NB * kBSlotBytes);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + kAStages;
  uint64_t* b_full = a_empty + kAStages;
  uint64_t* b_empty = b_full + kBSlots;
  uint64_t* t_full = b_empty + kBSlots;
  uint64_t* t_empty = t_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);
  float* xch = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 512;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&cp.tx[0]);
    tma_prefetch_desc(&cp.tw[0]);
    for (int i = 0; i < kAStages; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], kT21Iss); }
    for (int i = 0; i < kBSlots; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], kT21Iss); }
    for (int i = 0; i < 2; ++i) { mbar_init(&t_full[i], kT21Iss); mbar_init(&t_empty[i], kEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();

  const int n_layers = cp.n_layers;
  const int n_mtiles = cp.n_mtiles;
  const int num_work = cp.L[0].num_work;
  const int n_ntiles = cp.L[0].n_ntiles;
  const int G = gridDim.x;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0;
      for (int l = 0; l < n_layers; ++l) {
        const IgemmParams& p = cp.L[l];
        const CUtensorMap* tmx = &cp.tx[l];
        const CUtensorMap* tmw = &cp.tw[l];
        if (l + 1 < n_layers) { tma_prefetch_desc(&cp.tx[l + 1]); tma_prefetch_desc(&cp.tw[l + 1]); }
        const int KB = p.KB, x_coff = p.x_coff, cout = p.cout;
        for (int work = (blockIdx.x + l) % G; work < num_work; work += G) {
          const int nt = work % n_ntiles;
          const int mt_idx = work / n_ntiles;
          const int q0 = mt_idx * kT21Rows;
          if (l > 0) {
            const int* f = cp.flags + static_cast<size_t>(l - 1) * n_mtiles;
            const int t0 = mt_idx > 0 ? mt_idx - 1 : 0;
            const int t1 = mt_idx + 1 < n_mtiles ? mt_idx + 1 : n_mtiles - 1;
            for (int t = t0; t <= t1; ++t) chain_wait_flag(f + t, cp.target, l);
            fence_proxy_async_all();
          }
          for (int kb = 0; kb < KB; ++kb) {
            mbar_wait(&a_empty[sa], pa ^ 1, 1);
            mbar_expect_tx(&a_full[sa], a_stage_bytes);
            for (int bx = 0; bx < p.a_nbox; ++bx)
              tma_load_2d(a_smem + sa * a_stage_bytes + bx * p.a_box_rows * 128, tmx, x_coff + kb * 64,
                          q0 - p.halo + bx * p.a_box_rows, &a_full[sa]);
            if (++sa == kAStages) { sa = 0; pa ^= 1; }
            for (int dy = 0; dy < 3; ++dy) {
              mbar_wait(&b_empty[sb], pb ^ 1, 2);
              mbar_expect_tx(&b_full[sb], kBSlotBytes);
              for (int dx = 0; dx < 3; ++dx)
                tma_load_2d(b_smem + sb * kBSlotBytes + dx * kBTileBytes, tmw, 0,
                            ((dy * 3 + dx) * KB + kb) * cout + nt * 64, &b_full[sb]);
              if (++sb == NB) { sb = 0; pb ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp <= kT21Iss) {
    // ------------------------------------------------------------------ MMA issuers: warp 1 + mt owns sub-tile mt
    if (elect_one()) {
      const int mt = warp - 1;
      constexpr uint32_t idesc128 = make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc64 = make_idesc_bf16(128, 64, 0, 0);
      constexpr uint64_t desc_hi = (uint64_t(16 >> 4) << 16) | (uint64_t(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      const uint32_t a0 = ((smem_u32(a_smem) & 0x3FFFFu) >> 4) + static_cast<uint32_t>(mt * kT21Sub) * 8u;
      const uint32_t b0 = (smem_u32(b_smem) & 0x3FFFFu) >> 4;
      const uint32_t a_stage16 = a_stage_bytes >> 4;
      const int halo = cp.L[0].halo, pitch = cp.L[0].pitch;
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0, as = 0, pt = 0;
      for (int l = 0; l < n_layers; ++l) {
        const int KB = cp.L[l].KB;
        for (int work = (blockIdx.x + l) % G; work < num_work; work += G) {
          mbar_wait(&t_empty[as], pt ^ 1, 3);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * 256 + mt * 128;
          for (int kb = 0; kb < KB; ++kb) {
            mbar_wait(&a_full[sa], pa, 4);
            const uint32_t a_lo = a0 + sa * a_stage16;
            uint32_t row16 = static_cast<uint32_t>(halo - pitch - 1) * 8u;     // window row of tap (dy = -1, dx = -1)
#pragma unroll 1
            for (int dy = 0; dy < 3; ++dy) {
              mbar_wait(&b_full[sb], pb, 5);
              tc_fence_after();
              const uint32_t b_lo = b0 + sb * (kBSlotBytes >> 4);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint64_t ad1 = desc_hi | static_cast<uint64_t>(a_lo + row16 + k * 2);
                const uint64_t bd1 = desc_hi | static_cast<uint64_t>(b_lo + k * 2);
                umma_bf16(d_tmem, ad1, bd1, idesc128, static_cast<uint32_t>((kb | dy | k) != 0));
                const uint64_t ad2 = desc_hi | static_cast<uint64_t>(a_lo + row16 + 16 + k * 2);
                const uint64_t bd2 = desc_hi | static_cast<uint64_t>(b_lo + 2 * (kBTileBytes >> 4) + k * 2);
                umma_bf16(d_tmem, ad2, bd2, idesc64, 1u);
              }
              umma_commit(&b_empty[sb]);
              row16 += static_cast<uint32_t>(pitch) * 8u;
              if (++sb == NB) { sb = 0; pb ^= 1; }
            }
            umma_commit(&a_empty[sa]);
            if (kb == KB - 1) umma_commit(&t_full[as]);
            if (++sa == kAStages) { sa = 0; pa ^= 1; }
          }
          if (++as == 2) { as = 0; pt ^= 1; }
        }
      }
    }
  } else if (warp >= 5) {
    // ------------------------------------------------------------------ epilogue warps 5..12
    const int lg = warp & 3;                    // TMEM lane quarter
    const int half = (warp - 5) >> 2;           // which 32 of the 64 output channels
    const int m = lg * 32 + lane;
    uint32_t as = 0, pt = 0, par = 0;
    for (int l = 0; l < n_layers; ++l) {
      const IgemmParams& p = cp.L[l];
      const bool signal = l + 1 < n_layers;
      int* fl = cp.flags + static_cast<size_t>(l) * n_mtiles;
      float vs[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) vs[i] = 0.f;
      for (int work = (blockIdx.x + l) % G; work < num_work; work += G) {
        const int nt = work % n_ntiles;
        const int mt_idx = work / n_ntiles;
        if (p.act == SRB_ACT_LRELU_MASK) {
#pragma unroll
          for (int mt = 0; mt < 2; ++mt) {
            const int q = mt_idx * kT21Rows + mt * kT21Sub + m;
            if (q < p.R) {
              if (p.mask_bits) prefetch_l2(p.mask_bits + static_cast<size_t>(q) * p.maskb_cstride + p.maskb_coff + nt * 2 + half);
              else prefetch_l2(p.mask + static_cast<size_t>(q) * p.mask_cstride + p.mask_coff + nt * 64 + half * 32);
            }
          }
        }
        mbar_wait(&t_full[as], pt, 6);
        tc_fence_after();
#pragma unroll 1
        for (int mt = 0; mt < 2; ++mt) {
          const int q = mt_idx * kT21Rows + mt * kT21Sub + m;
          const int r = q % p.P;
          const int py = r / p.pitch, px = r - py * p.pitch;
          const bool valid = (m < kT21Sub) && (q < p.R) && (px < p.W) && (py < p.H);
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lg * 32) << 16) + as * 256 + mt * 128 + half * 32;
          float* my_x = xch + ((par * 2 + half) * 4 + lg) * 32;
          float v[32];
          {
            uint32_t d1[32];
            tmem_ld32(taddr + 64, d1);
            tmem_ld_wait();
            if (lane == 0) {                     // row 0 of this quarter = D1[m + 1] of the previous quarter's lane 31
#pragma unroll
              for (int i = 0; i < 8; ++i)
                reinterpret_cast<float4*>(my_x)[i] = make_float4(__uint_as_float(d1[4 * i]), __uint_as_float(d1[4 * i + 1]),
                                                                 __uint_as_float(d1[4 * i + 2]), __uint_as_float(d1[4 * i + 3]));
            }
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __shfl_down_sync(0xffffffffu, __uint_as_float(d1[i]), 1);
            asm volatile("bar.sync %0, 128;" ::"r"(1 + half) : "memory");
            if (lane == 31 && lg < 3) {
              const float4* nx = reinterpret_cast<const float4*>(my_x + 32);     // the next quarter's slot
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const float4 x = nx[i];
                v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
              }
            }
          }
          {
            uint32_t d0[32];
            tmem_ld32(taddr, d0);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] += __uint_as_float(d0[i]);
          }
          par ^= 1;
          if (valid) igemm_store_row<true, true>(p, v, q, py, px, nt, half);
          if (p.colsum && valid) {
#pragma unroll
            for (int i = 0; i < 32; ++i) vs[i] += v[i];
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[as]);
        if (signal) {
          fence_proxy_async_all();
          __threadfence();
          __syncwarp();
          if (lane == 0) atomicAdd(fl + mt_idx, 1);
        }
        if (++as == 2) { as = 0; pt ^= 1; }
      }
      if (p.colsum) atomicAdd(p.colsum + half * 32 + lane, warp_colsum32(vs, lane));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ---------------------------------------------------------------------------------- host side
// SRB_CONV_T21=1 (read once per process) turns the formulation on for every launch it fits.
bool conv_t21_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("SRB_CONV_T21"); v = (e && e[0] == '1') ? 1 : 0; }
  return v != 0;
}

// Plan: the 256-row geometry of conv_igemm.cu (same halo window, same rings) with 254-row tiles and the exchange buffer.
// Returns false when the geometry does not fit (image too wide for a two-sub-tile window).
bool conv_t21_plan(const srb_conv3x3_args* a, IgemmPlan* pl) {
  if (!igemm_geom(a, 2, pl)) return false;
  pl->rows_tile = kT21Rows;
  pl->n_mtiles = int((pl->R + kT21Rows - 1) / kT21Rows);
  pl->smem += kT21XchBytes;
  return pl->smem <= 227 * 1024;
}

int conv3x3_chain21_launch_kernel(const ChainParams& cp, const IgemmPlan& pl, int grid, cudaStream_t stream) {
  ensure_max_smem(reinterpret_cast<const void*>(conv3x3_chain21_kernel), 227 * 1024);
  cudaError_t e = launch_pdl(conv3x3_chain21_kernel, dim3(grid), dim3(kIgemmThreads), pl.smem, stream, cp);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("conv3x3_chain21 launch", e);
  return 0;
}

}  // namespace srb
