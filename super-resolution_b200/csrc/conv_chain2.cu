// K1p: the chained convolution kernel (conv_chain.cu) on CTA PAIRS: tcgen05.mma.cta_group::2.
//
// Why.  With Cout = 64 fixed by the model every MMA of conv_chain.cu is 128x64x16 and reads 4 KB of A + 2 KB of B from
// shared memory per 32 tensor clocks: 48 clocks at the 128 B/clk an SM delivers.  A CTA pair
// (two SMs of one TPC, thread-block cluster of 2) computes 256x64x16 per instruction: each CTA supplies its own 128
// rows of A and only HALF of the weight tile (32 of the 64 output channels), the tensor cores exchange the halves.
// Per CTA that is 4 KB + 1 KB per MMA (40 clocks on paper, 43 measured by the issue-rate probe,
// profiles/r01_s21_mma_rate_probe.txt), and the weights cross L2 -> shared memory once per pair instead of once per
// CTA (half the TMA writes of the weight ring into the same shared-memory pipe).
//
// What changes against conv_chain.cu (everything else - tile -> CTA map, flag protocol, epilogue - is that file's):
//   * CTA rank 0 of a pair (the leader) issues every MMA, for both CTAs; its four issuer threads keep their private
//     accumulators, which now exist at the same TMEM address in both CTAs (tcgen05.alloc.cta_group::2).
//   * Tiles: pair p works on items j = p, p + pairs, ...; item j = (M tiles 2 (j / n_ntiles) + {0, 1}, N tile
//     j % n_ntiles), the CTA of rank r taking the M tile + r.  With one N tile (cout = 64) that is the tile -> CTA map of
//     conv_chain.cu (CTA b: tiles b, b + gridDim.x, ...).  If the last M tile has no partner the peer runs a PHANTOM
//     tile: TMA zero-fills its rows (all out of bounds), nothing is stored, no flag is touched.
//   * full barriers live in the leader: both CTAs' TMA loads (cp.async.bulk.tensor ... .cta_group::2) complete their
//     bytes on the leader's a_full / b_full, which the leader's producer arms with the bytes of BOTH CTAs.
//   * empty / accumulator-ready barriers exist in both CTAs at the same offset and are signalled by ONE multicast
//     tcgen05.commit per issuer (mask 0b11).
//   * accumulator-drained (t_empty) lives in the leader: the epilogue warps of both CTAs arrive there (2 x kEpiWarps).
//   * release of a finished tile (flags[l][t]) is done by a warp of its own: the epilogue warps arrive on a CTA-local
//     mbarrier after their stores (st_full, release.cta) and go on with the next tile; warp 13 waits for the eight
//     arrivals, runs fence.proxy.async + __threadfence() (cumulative: it covers the stores that happen-before it through
//     the mbarrier, the same pattern as bar.sync + one thread's fence in a grid barrier) and adds kEpiWarps to the flag.
//     In conv_chain.cu every epilogue warp stalls ~1.6 us per tile in that fence - on the K = 64 / 128 layers, where
//     the epilogue is the slow side (tools/exp_pair.py, exp_pair_epi.py), that is time of the launch.
//   * the bias of a layer is staged in shared memory once per layer and warp (igemm_store_row's sbias).
//   * set-up and tear-down are cluster barriers: no CTA may signal a peer barrier before it is initialised, and no CTA
//     may free its tensor memory / exit while the pair's MMAs still read its shared memory.
// With cout > 64 a pair takes two M tiles of the SAME 64-wide output slice (the two CTAs must want the same weights).
// Results are bit-identical to conv_chain.cu (same products, same accumulation order per accumulator).
#include <stdlib.h>

#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "sm100.cuh"

namespace srb {

#ifndef SRB_CONV_PAIR_DEFAULT
#define SRB_CONV_PAIR_DEFAULT 1
#endif

constexpr int kPairBSlots = 6;                        // weight ring: one slot = this CTA's half of a filter row
constexpr int kPairBTileBytes = 32 * 64 * 2;          // 32 output channels x 64 input channels, bf16
constexpr int kPairBSlotBytes = 3 * kPairBTileBytes;
constexpr int kPairThreads = kIgemmThreads + 32;      // + warp 13: flag signaller (see "release" below)
constexpr int kPairTailBytes = 256 + kEpiWarps * 32 * 4 + 256;   // barriers | per-warp bias copies | slack

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// address of the same shared-memory location in CTA `rank` of the cluster (shared::cluster window)
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
  uint32_t d;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(d) : "r"(addr), "r"(rank));
  return d;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// wait with cluster-scope acquire: the arrivals come from the peer CTA's threads
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity, uint32_t tag) {
  uint32_t spins = 0;
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) break;
    if (++spins > SRB_WAIT_LIMIT) {
      atomicExch(&g_srb_fault, 0x80000000u | (tag << 16) | (blockIdx.x & 0xffffu));
      __trap();
    }
  }
}
// TMA load into THIS CTA's shared memory whose bytes complete on a barrier of the pair's leader
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* m, int c0, int c1, uint32_t leader_bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(leader_bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs once all MMAs this thread issued have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3))
               : "memory");
}

template <int MT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairThreads, 1)
conv3x3_chain_pair_kernel(const __grid_constant__ ChainParams cp) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int NB = kPairBSlots;
  const uint32_t a_stage_bytes = cp.L[0].a_stage_bytes;     // geometry is common to all layers
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_smem = smem;
  uint8_t* b_smem = a_smem + kAStages * a_stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_smem + NB * kPairBSlotBytes);
  uint64_t* a_full = bars;                      // used in the leader only
  uint64_t* a_empty = a_full + kAStages;
  uint64_t* b_full = a_empty + kAStages;        // used in the leader only
  uint64_t* b_empty = b_full + NB;
  uint64_t* t_full = b_empty + NB;
  uint64_t* t_empty = t_full + 2;               // used in the leader only
  uint64_t* st_full = t_empty + 2;              // tile stored by the eight epilogue warps -> signaller
  uint64_t* st_empty = st_full + 2;             // signaller done with the slot
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(st_empty + 2);
  float* sbias = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [kEpiWarps][32]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  constexpr uint32_t kTmemCols = 512;
  constexpr int NACC = kIssuers / MT;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&cp.tx[0]);
    tma_prefetch_desc(&cp.tw[0]);
    for (int i = 0; i < kAStages; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], kIssuers); }
    for (int i = 0; i < NB; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], kIssuers); }
    for (int i = 0; i < 2; ++i) { mbar_init(&t_full[i], kIssuers); mbar_init(&t_empty[i], 2 * kEpiWarps); }
    for (int i = 0; i < 2; ++i) { mbar_init(&st_full[i], kEpiWarps); mbar_init(&st_empty[i], 1); }
    fence_mbar_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncwarp();
  cluster_sync_all();       // both CTAs' barriers initialised and tensor memory allocated
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();

  const int rows_per_tile = 128 * MT;
  const int n_layers = cp.n_layers;
  const int n_mtiles = cp.n_mtiles;
  const int n_ntiles = cp.L[0].n_ntiles;
  // pair items: item j = (pair of M tiles j / n_ntiles, N tile j % n_ntiles); pair p takes items p, p + pairs, ...
  // (n_ntiles == 1: CTA b works on tiles b, b + gridDim.x, ... exactly as in conv_chain.cu)
  const int n_items = ((n_mtiles + 1) >> 1) * n_ntiles;
  const int j_first = static_cast<int>(blockIdx.x >> 1), j_step = static_cast<int>(gridDim.x >> 1);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    if (elect_one()) {
      const uint32_t a_full_l = mapa_u32(smem_u32(a_full), 0), b_full_l = mapa_u32(smem_u32(b_full), 0);
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0;
      for (int l = 0; l < n_layers; ++l) {
        const IgemmParams& p = cp.L[l];
        const CUtensorMap* tmx = &cp.tx[l];
        const CUtensorMap* tmw = &cp.tw[l];
        if (l + 1 < n_layers) { tma_prefetch_desc(&cp.tx[l + 1]); tma_prefetch_desc(&cp.tw[l + 1]); }
        const int KB = p.KB, x_coff = p.x_coff, cout = p.cout;
        for (int j = j_first; j < n_items; j += j_step) {
          const int nt = j % n_ntiles;
          const int mt_idx = 2 * (j / n_ntiles) + static_cast<int>(rank);
          const bool phantom = mt_idx >= n_mtiles;
          const int q0 = mt_idx * rows_per_tile;
          if (l > 0 && !phantom && !(p.dbg & 16)) {
            const int* f = cp.flags + static_cast<size_t>(l - 1) * n_mtiles;
            const int t0 = mt_idx > 0 ? mt_idx - 1 : 0;
            const int t1 = mt_idx + 1 < n_mtiles ? mt_idx + 1 : n_mtiles - 1;
            for (int t = t0; t <= t1; ++t) chain_wait_flag(f + t, cp.target, l);
            fence_proxy_async_all();
          }
          for (int kb = 0; kb < KB; ++kb) {
            mbar_wait(&a_empty[sa], pa ^ 1, 1);
            if (leader) mbar_expect_tx(&a_full[sa], 2 * a_stage_bytes);
            for (int bx = 0; bx < p.a_nbox; ++bx)
              tma_load_2d_pair(a_smem + sa * a_stage_bytes + bx * p.a_box_rows * 128, tmx, x_coff + kb * 64,
                               q0 - p.halo + bx * p.a_box_rows, a_full_l + sa * 8);
            if (++sa == kAStages) { sa = 0; pa ^= 1; }
            for (int dy = 0; dy < 3; ++dy) {
              mbar_wait(&b_empty[sb], pb ^ 1, 2);
              if (leader) mbar_expect_tx(&b_full[sb], 2 * kPairBSlotBytes);
              for (int dx = 0; dx < 3; ++dx)
                tma_load_2d_pair(b_smem + sb * kPairBSlotBytes + dx * kPairBTileBytes, tmw, 0,
                                 ((dy * 3 + dx) * KB + kb) * cout + nt * 64 + static_cast<int>(rank) * 32, b_full_l + sb * 8);
              if (++sb == NB) { sb = 0; pb ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp <= kIssuers) {
    // ------------------------------------------------------------------ MMA issuers (leader CTA only)
    if (leader && elect_one()) {
      const int iss = warp - 1;
      const int mt = MT == 4 ? iss : (MT == 2 ? (iss >> 1) : 0);
      const int k0 = MT == 4 ? 0 : (MT == 2 ? 2 * (iss & 1) : iss);
      constexpr int NK = MT == 4 ? 4 : (MT == 2 ? 2 : 1);
      constexpr uint32_t idesc = make_idesc_bf16(256, 64, 0, 0);
      constexpr uint64_t desc_hi = (uint64_t(16 >> 4) << 16) | (uint64_t(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      const uint32_t a0 = ((smem_u32(a_smem) & 0x3FFFFu) >> 4) + mt * 1024 + k0 * 2;
      const uint32_t b0 = ((smem_u32(b_smem) & 0x3FFFFu) >> 4) + k0 * 2;
      const uint32_t a_stage16 = a_stage_bytes >> 4;
      const int halo = cp.L[0].halo, pitch = cp.L[0].pitch;
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0, as = 0, pt = 0;
      for (int l = 0; l < n_layers; ++l) {
        const int KB = cp.L[l].KB;
        for (int j = j_first; j < n_items; j += j_step) {
          mbar_wait_cluster(&t_empty[as], pt ^ 1, 3);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * 256 + iss * 64;
          for (int kb = 0; kb < KB; ++kb) {
            mbar_wait(&a_full[sa], pa, 4);
            const uint32_t a_lo = a0 + sa * a_stage16;
            uint32_t row16 = static_cast<uint32_t>(halo - pitch - 1) * 8u;
#pragma unroll 1
            for (int dy = 0; dy < 3; ++dy) {
              mbar_wait(&b_full[sb], pb, 5);
              tc_fence_after();
              const uint32_t b_lo = b0 + sb * (kPairBSlotBytes >> 4);
#pragma unroll
              for (int dx = 0; dx < 3; ++dx) {
#pragma unroll
                for (int k = 0; k < NK; ++k) {
                  const uint64_t ad = desc_hi | static_cast<uint64_t>(a_lo + row16 + dx * 8 + k * 2);
                  const uint64_t bd = desc_hi | static_cast<uint64_t>(b_lo + dx * (kPairBTileBytes >> 4) + k * 2);
                  umma_bf16_pair(d_tmem, ad, bd, idesc, (dx | k) != 0 ? 1u : static_cast<uint32_t>((kb | dy) != 0));
                }
              }
              umma_commit_pair(&b_empty[sb]);
              row16 += static_cast<uint32_t>(pitch) * 8u;
              if (++sb == NB) { sb = 0; pb ^= 1; }
            }
            umma_commit_pair(&a_empty[sa]);
            if (kb == KB - 1) umma_commit_pair(&t_full[as]);
            if (++sa == kAStages) { sa = 0; pa ^= 1; }
          }
          if (++as == 2) { as = 0; pt ^= 1; }
        }
      }
    }
  } else if (warp < 5 + kEpiWarps) {
    // ------------------------------------------------------------------ epilogue warps 5..12 (both CTAs)
    const int lg = warp & 3;
    const int half = (warp - 5) >> 2;
    uint32_t ct = 0, cs = 0;                     // tiles done / tiles handed to the signaller: ring slot = count & 1,
                                                 // phase parity = (count >> 1) & 1
    for (int l = 0; l < n_layers; ++l) {
      const IgemmParams& p = cp.L[l];
      const bool signal = l + 1 < n_layers && !(p.dbg & 16);
      float vs[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) vs[i] = 0.f;
      if (j_first < n_items)
        igemm_prefetch_mask<MT>(p, 2 * (j_first / n_ntiles) + static_cast<int>(rank), j_first % n_ntiles, lg, lane, half);
      for (int j = j_first; j < n_items; j += j_step) {
        const int nt = j % n_ntiles;
        const int mt_idx = 2 * (j / n_ntiles) + static_cast<int>(rank);
        const bool phantom = mt_idx >= n_mtiles;
        {
          const int nj = j + j_step;      // (a phantom tile's rows are >= R: the prefetch skips them)
          if (nj < n_items) igemm_prefetch_mask<MT>(p, 2 * (nj / n_ntiles) + static_cast<int>(rank), nj % n_ntiles, lg, lane, half);
        }
        if (p.bias && (j == j_first || n_ntiles > 1)) {   // once per layer (and N tile): this warp's 32 bias values
          __syncwarp();
          sbias[(warp - 5) * 32 + lane] = __ldg(p.bias + nt * 64 + half * 32 + lane);
          __syncwarp();
        }
        const uint32_t as = ct & 1u, pt = (ct >> 1) & 1u;
        mbar_wait(&t_full[as], pt, 6);
        tc_fence_after();
#pragma unroll 1
        for (int mt = 0; mt < MT; ++mt) {
          const int q = mt_idx * rows_per_tile + mt * 128 + lg * 32 + lane;
          const int r = q % p.P;
          const int py = r / p.pitch, px = r - py * p.pitch;
          const bool valid = !phantom && (q < p.R) && (px < p.W) && (py < p.H) && !(p.dbg & 4);
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lg * 32) << 16) + as * 256 + mt * (NACC * 64);
          float v[32];
          if (p.dbg & 32) {       // timing experiment: no accumulator read-out
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = 0.f;
          } else {
            uint32_t acc[32];
            tmem_ld32(taddr + half * 32, acc);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(acc[i]);
#pragma unroll
            for (int j = 1; j < NACC; ++j) {
              tmem_ld32(taddr + j * 64 + half * 32, acc);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] += __uint_as_float(acc[i]);
            }
          }
          if (valid) {
            const float4* sb4 = p.bias ? reinterpret_cast<const float4*>(sbias + (warp - 5) * 32) : nullptr;
            if (p.ext) igemm_store_row<true, true>(p, v, q, py, px, nt, half, sb4);
            else igemm_store_row<true, false>(p, v, q, py, px, nt, half, sb4);
          }
          if (p.colsum && valid) {
#pragma unroll
            for (int i = 0; i < 32; ++i) vs[i] += v[i];
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&t_empty[as]), 0));
        if (signal && !phantom) {                // hand the tile to the signaller warp (release.cta); no fence here
          fence_proxy_async_all();
          __syncwarp();
          if (lane == 0) {
            mbar_wait(&st_empty[cs & 1u], ((cs >> 1) & 1u) ^ 1u, 7);
            mbar_arrive(&st_full[cs & 1u]);
          }
          ++cs;
        }
        ++ct;
      }
      if (p.colsum) atomicAdd(p.colsum + half * 32 + lane, warp_colsum32(vs, lane));
    }
  } else {
    // ------------------------------------------------------------------ warp 13: flag signaller (both CTAs)
    if (elect_one()) {
      uint32_t ss = 0, ps = 0;
      for (int l = 0; l + 1 < n_layers; ++l) {
        if (cp.L[l].dbg & 16) continue;
        int* fl = cp.flags + static_cast<size_t>(l) * n_mtiles;
        for (int j = j_first; j < n_items; j += j_step) {
          const int mt_idx = 2 * (j / n_ntiles) + static_cast<int>(rank);
          if (mt_idx >= n_mtiles) continue;      // phantom tile
          mbar_wait(&st_full[ss], ps, 8);        // the eight epilogue warps have stored their parts of the tile
          fence_proxy_async_all();
          __threadfence();
          atomicAdd(fl + mt_idx, kEpiWarps);
          mbar_arrive(&st_empty[ss]);
          if (++ss == 2) { ss = 0; ps ^= 1; }
        }
      }
    }
  }

  tc_fence_before();
  __syncwarp();
  cluster_sync_all();       // the pair's MMAs read both CTAs' shared memory and write both CTAs' tensor memory
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

// ---------------------------------------------------------------------------------- host side
// SRB_CONV_PAIR=0 / 1 (read once per process) switches the pair kernel off / on for every eligible chained launch;
// srb_conv3x3_args.flags bit 22 forces it, bit 23 forbids it (tests run both on the same inputs).
bool conv_pair_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("SRB_CONV_PAIR"); v = e ? (e[0] == '1' ? 1 : 0) : SRB_CONV_PAIR_DEFAULT; }
  return v != 0;
}

static size_t chain_pair_smem_bytes(uint32_t a_stage_bytes) {
  return 1024 + size_t(kAStages) * a_stage_bytes + size_t(kPairBSlots) * kPairBSlotBytes + kPairTailBytes;
}

bool conv_pair_eligible(const srb_conv3x3_args* a, int n, const IgemmPlan& pl) {
  if (pl.n_mtiles < 2) return false;
  for (int l = 0; l < n; ++l)
    if ((a[l].flags >> 1) & 0x7fff & ~(4 | 16 | 32)) return false;   // timing-experiment switches: only "no epilogue I/O" (4),
                                                                     // "no flags" (16), "no accumulator read-out" (32) exist here
  return chain_pair_smem_bytes(pl.a_stage_bytes) <= 227 * 1024;
}

// 2 x cudaOccupancyMaxActiveClusters of the kernel at this shared-memory size, cached per (device, MT, smem)
static int pair_max_ctas(int MT, size_t smem) {
  static std::mutex mu;
  static std::map<std::tuple<int, int, size_t>, int> cache;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_tuple(dev, MT, smem);
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  const void* k = MT == 4 ? reinterpret_cast<const void*>(conv3x3_chain_pair_kernel<4>)
                : MT == 2 ? reinterpret_cast<const void*>(conv3x3_chain_pair_kernel<2>)
                          : reinterpret_cast<const void*>(conv3x3_chain_pair_kernel<1>);
  ensure_max_smem(k, 227 * 1024);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(sm_count() & ~1); cfg.blockDim = dim3(kPairThreads); cfg.dynamicSmemBytes = smem;
  int clusters = 0;
  cudaError_t e = cudaOccupancyMaxActiveClusters(&clusters, k, &cfg);
  int v = (sm_count() & ~1);
  if (e == cudaSuccess && clusters > 0 && 2 * clusters < v) v = 2 * clusters;
  if (e != cudaSuccess) cudaGetLastError();
  cache[key] = v;
  return v;
}

// cp: filled by conv3x3_chain_launch (tensor maps of the weights still with 64-row boxes; re-encoded here)
int conv3x3_chain_pair_launch(ChainParams& cp, const srb_conv3x3_args* a, int n, const IgemmPlan& pl, cudaStream_t stream) {
  for (int l = 0; l < n; ++l)
    if (make_tmap_2d_bf16(&cp.tw[l], a[l].w, 64, (uint64_t)9 * (a[l].cin / 64) * a[l].cout, 128, 64, 32)) return 1;
  const size_t smem = chain_pair_smem_bytes(pl.a_stage_bytes);
  // every CTA of the grid must be able to be resident at once (flag waits between tiles of different CTAs): at most as
  // many pairs as the device can hold clusters of this kernel (74 on a 148-SM B200: both SMs of every TPC)
  const int sms = pair_max_ctas(pl.MT, smem);
  if (sms < 2) return set_error("conv3x3_chain (pair): no resident cluster possible");
  const int want = ((cp.n_mtiles + 1) >> 1) * cp.L[0].n_ntiles * 2;
  const int grid = want < sms ? want : sms;
  cudaError_t e;
  if (pl.MT == 4) {
    ensure_max_smem(reinterpret_cast<const void*>(conv3x3_chain_pair_kernel<4>), 227 * 1024);
    e = launch_pdl(conv3x3_chain_pair_kernel<4>, dim3(grid), dim3(kPairThreads), smem, stream, cp);
  } else if (pl.MT == 2) {
    ensure_max_smem(reinterpret_cast<const void*>(conv3x3_chain_pair_kernel<2>), 227 * 1024);
    e = launch_pdl(conv3x3_chain_pair_kernel<2>, dim3(grid), dim3(kPairThreads), smem, stream, cp);
  } else {
    ensure_max_smem(reinterpret_cast<const void*>(conv3x3_chain_pair_kernel<1>), 227 * 1024);
    e = launch_pdl(conv3x3_chain_pair_kernel<1>, dim3(grid), dim3(kPairThreads), smem, stream, cp);
  }
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("conv3x3_chain (pair) launch", e);
  return 0;
}

}  // namespace srb
