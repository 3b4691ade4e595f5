// K1c: a CHAIN of 3x3 convolutions over the same pixel grid in ONE persistent launch.
//
// Why.  A DenseResidualBlock (models.py:34-41) is five convolutions whose inputs are each other's outputs; launched one
// by one (conv_igemm.cu) every launch pays a fill/drain bubble of ~8 us (first TMA round trip, last tile's epilogue,
// grid drain + dependent launch) against 13 us of tensor-pipe work per 64 input channels at batch 512 - 18 % of a dense
// block.  The dependency between consecutive layers is only LOCAL: tile t of layer l reads the rows of tiles t-1, t, t+1
// (halo = one image row + 1 pixel < one tile) of the layers before it.  So all layers run in one launch with the same
// static tile -> CTA map, and a layer's tile starts as soon as the three neighbouring tiles of the previous layer are
// complete - normally several tiles earlier, so the MMA stream never drains at a layer boundary.
//
// Protocol.  flags[l][t] counts epilogue warps that finished tile t of layer l (zero on entry, caller's workspace).
//   writer (each epilogue warp): global stores; fence.proxy.async (generic -> async proxy, the consumer is a TMA load);
//     __threadfence(); __syncwarp(); lane 0: atomicAdd(flags[l][t], 1).
//   reader (TMA producer thread) before the first load of tile t of layer l > 0: ld.acquire.gpu of flags[l-1][t-1..t+1]
//     until each equals kEpiWarps * n_ntiles; fence.proxy.async; TMA loads.
//   Older layers are covered transitively: every CTA runs its layers in order, so "layer l-1 complete on tile t" implies
//   all earlier layers are complete there.  The same in-order argument covers write-after-read on ring buffers (the
//   inference growth-buffer ring, the gradient buffers): whoever overwrites rows of tile t at layer l has waited for the
//   neighbours' layer l-1, which is after every earlier read of those rows.
//   Residual / mask rows are read by the thread that wrote them (same tile -> same CTA -> same warp/lane) with
//   ld.global.cg, never through the non-coherent path.
// Progress.  grid <= SM count and one CTA per SM by shared-memory footprint, so every CTA gets an SM of its own as soon
// as nothing else holds it.  While another stream's kernel occupies SMs (NCCL channel CTAs of the bucketed all-reduce,
// the discriminator streams) some CTAs start late; their neighbours spin in chain_wait_flag until they arrive.  That is
// a delay, not a deadlock: a waiting CTA only ever waits for tiles of the PREVIOUS layer, foreign kernels never wait for
// this one, and a late CTA needs no SM a spinning CTA holds.  The spin limit (2^26 polls of >= 40 ns, several seconds)
// exists only so that a protocol bug surfaces as a fault word + trap instead of a hung GPU.
//
// Everything else (operand staging, four MMA issuers with private TMEM accumulators, epilogue) is conv_igemm.cu's.
#include "common.cuh"
#include "sm100.cuh"

namespace srb {

constexpr int kChainMax = SRB_CONV_CHAIN_MAX;

struct alignas(64) ChainParams {
  CUtensorMap tx[kChainMax];
  CUtensorMap tw[kChainMax];
  IgemmParams L[kChainMax];
  int n_layers;
  int n_mtiles;
  int target;          // kEpiWarps * n_ntiles
  int* flags;          // [n_layers][n_mtiles]
};

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

__device__ __forceinline__ void chain_wait_flag(const int* f, int target, int layer) {
  uint32_t spins = 0;
  while (ld_acquire_gpu(f) < target) {
    __nanosleep(40);
    if (++spins > (1u << 26)) {
      atomicExch(&g_srb_fault, 0xC0000000u | (static_cast<uint32_t>(layer & 0xff) << 16) | (blockIdx.x & 0xffffu));
      __trap();
    }
  }
}

template <int MT>
__global__ void __launch_bounds__(kIgemmThreads, 1)
conv3x3_chain_kernel(const __grid_constant__ ChainParams cp) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int NB = MT == 4 ? 3 : kBSlots;     // weight ring slots (MT == 4: two 512-row windows leave room for three)
  const uint32_t a_stage_bytes = cp.L[0].a_stage_bytes;     // geometry is common to all layers
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

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr uint32_t kTmemCols = 512;
  constexpr int NACC = kIssuers / MT;           // accumulators that sum to one 128-row sub-tile

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&cp.tx[0]);
    tma_prefetch_desc(&cp.tw[0]);
    for (int i = 0; i < kAStages; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], kIssuers); }
    for (int i = 0; i < kBSlots; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], kIssuers); }
    for (int i = 0; i < 2; ++i) { mbar_init(&t_full[i], kIssuers); mbar_init(&t_empty[i], kEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();

  const int rows_per_tile = 128 * MT;
  const int n_layers = cp.n_layers;
  const int n_mtiles = cp.n_mtiles;
  const int num_work = cp.L[0].num_work;       // same for every layer (same geometry, same cout)
  const int n_ntiles = cp.L[0].n_ntiles;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {   // elect.sync: ptxas then keeps descriptors and barrier addresses in uniform registers
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0;
      for (int l = 0; l < n_layers; ++l) {
        const IgemmParams& p = cp.L[l];
        const CUtensorMap* tmx = &cp.tx[l];
        const CUtensorMap* tmw = &cp.tw[l];
        if (l + 1 < n_layers) { tma_prefetch_desc(&cp.tx[l + 1]); tma_prefetch_desc(&cp.tw[l + 1]); }
        const int KB = p.KB, x_coff = p.x_coff, cout = p.cout;
        for (int work = blockIdx.x; work < num_work; work += gridDim.x) {
          const int nt = work % n_ntiles;
          const int mt_idx = work / n_ntiles;
          const int q0 = mt_idx * rows_per_tile;
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
  } else if (warp <= kIssuers) {
    // ------------------------------------------------------------------ MMA issuers (as conv3x3_igemm_kernel)
    if (elect_one()) {   // elect.sync: ptxas then keeps descriptors and barrier addresses in uniform registers
      const int iss = warp - 1;
      const int mt = MT == 4 ? iss : (MT == 2 ? (iss >> 1) : 0);
      const int k0 = MT == 4 ? 0 : (MT == 2 ? 2 * (iss & 1) : iss);
      constexpr int NK = MT == 4 ? 4 : (MT == 2 ? 2 : 1);
      constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
      constexpr uint64_t desc_hi = (uint64_t(16 >> 4) << 16) | (uint64_t(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      const uint32_t a0 = ((smem_u32(a_smem) & 0x3FFFFu) >> 4) + mt * 1024 + k0 * 2;
      const uint32_t b0 = ((smem_u32(b_smem) & 0x3FFFFu) >> 4) + k0 * 2;
      const uint32_t a_stage16 = a_stage_bytes >> 4;
      const int halo = cp.L[0].halo, pitch = cp.L[0].pitch;
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0, as = 0, pt = 0;
      for (int l = 0; l < n_layers; ++l) {
        const int KB = cp.L[l].KB;
        for (int work = blockIdx.x; work < num_work; work += gridDim.x) {
          mbar_wait(&t_empty[as], pt ^ 1, 3);
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
              const uint32_t b_lo = b0 + sb * (kBSlotBytes >> 4);
#pragma unroll
              for (int dx = 0; dx < 3; ++dx) {
#pragma unroll
                for (int k = 0; k < NK; ++k) {
                  const uint64_t ad = desc_hi | static_cast<uint64_t>(a_lo + row16 + dx * 8 + k * 2);
                  const uint64_t bd = desc_hi | static_cast<uint64_t>(b_lo + dx * (kBTileBytes >> 4) + k * 2);
                  umma_bf16(d_tmem, ad, bd, idesc, (dx | k) != 0 ? 1u : static_cast<uint32_t>((kb | dy) != 0));
                }
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
  } else {
    // ------------------------------------------------------------------ epilogue warps 5..12
    const int lg = warp & 3;
    const int half = (warp - 5) >> 2;
    uint32_t as = 0, pt = 0;
    for (int l = 0; l < n_layers; ++l) {
      const IgemmParams& p = cp.L[l];
      const bool signal = l + 1 < n_layers;
      int* fl = cp.flags + static_cast<size_t>(l) * n_mtiles;
      float vs[32];                             // p.colsum: per-column sums over the rows this lane stored
#pragma unroll
      for (int i = 0; i < 32; ++i) vs[i] = 0.f;
      // LeakyReLU' mask rows are prefetched into L2 one tile ahead: on the small-K layers the epilogue is the slow side,
      // so a prefetch issued when the tile's own turn comes has no lead over the loads that need it
      if (static_cast<int>(blockIdx.x) < num_work)
        igemm_prefetch_mask<MT>(p, blockIdx.x / n_ntiles, blockIdx.x % n_ntiles, lg, lane, half);
      for (int work = blockIdx.x; work < num_work; work += gridDim.x) {
        const int nt = work % n_ntiles;
        const int mt_idx = work / n_ntiles;
        {
          const int nw = work + gridDim.x;
          if (nw < num_work) igemm_prefetch_mask<MT>(p, nw / n_ntiles, nw % n_ntiles, lg, lane, half);
        }
        mbar_wait(&t_full[as], pt, 6);
        tc_fence_after();
#pragma unroll 1
        for (int mt = 0; mt < MT; ++mt) {
          const int q = mt_idx * rows_per_tile + mt * 128 + lg * 32 + lane;
          const int r = q % p.P;
          const int py = r / p.pitch, px = r - py * p.pitch;
          const bool valid = (q < p.R) && (px < p.W) && (py < p.H);
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lg * 32) << 16) + as * 256 + mt * (NACC * 64);
          float v[32];
          {
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
            if (p.ext) igemm_store_row<true, true>(p, v, q, py, px, nt, half);
            else igemm_store_row<true, false>(p, v, q, py, px, nt, half);
          }
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
      if (p.colsum) atomicAdd(p.colsum + half * 32 + lane, warp_colsum32(vs, lane));   // one butterfly per layer
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
bool conv_t21_enabled();                                                                       // conv_chain21.cu
bool conv_t21_plan(const srb_conv3x3_args* a, IgemmPlan* pl);
int conv3x3_chain21_launch_kernel(const ChainParams& cp, const IgemmPlan& pl, int grid, cudaStream_t stream);

bool conv_pair_enabled();                                                                      // conv_chain2.cu
bool conv_pair_eligible(const srb_conv3x3_args* a, int n, const IgemmPlan& pl);
int conv3x3_chain_pair_launch(ChainParams& cp, const srb_conv3x3_args* a, int n, const IgemmPlan& pl, cudaStream_t stream);

size_t conv3x3_chain_flag_bytes(int batch, int H, int W, int n) {
  // sized for the finer tiling (128 rows) so the caller need not know which one the plan picks
  const long long R = 1LL * batch * (H + 1) * (W + 1);
  return static_cast<size_t>(n) * static_cast<size_t>((R + 127) / 128) * sizeof(int);
}

int conv3x3_chain_launch(const srb_conv3x3_args* a, int n, int* flags, size_t flag_bytes, cudaStream_t stream) {
  if (n < 1 || n > kChainMax) return set_error("conv3x3_chain: 1..SRB_CONV_CHAIN_MAX layers");
  if (!flags && n > 1) return set_error("conv3x3_chain: flag workspace missing");
  for (int l = 0; l < n; ++l) {
    if (!a[l].x || !a[l].w || !a[l].y) return set_error("conv3x3_chain: null pointer");
    if (int rc = igemm_check_args(&a[l])) return rc;
    if (a[l].batch != a[0].batch || a[l].H != a[0].H || a[l].W != a[0].W || a[l].cout != a[0].cout)
      return set_error("conv3x3_chain: all layers must share batch, H, W and cout");
    if (a[l].store_mode != SRB_STORE_PLAIN && n > 1) return set_error("conv3x3_chain: plain stores only");
    if (a[l].impl != SRB_IMPL_TCGEN05) return set_error("conv3x3_chain: tcgen05 kernels only");
  }
  // flags bit 20 forces, bit 21 forbids the 2+1-tap kernel (conv_chain21.cu); SRB_CONV_T21=1 makes it the default
  IgemmPlan pl;
  const bool want21 = (a[0].flags & 0x100000) || (conv_t21_enabled() && !(a[0].flags & 0x200000));
  const bool t21 = want21 && conv_t21_plan(&a[0], &pl);
  if (!t21)
    if (int rc = igemm_make_plan(&a[0], &pl)) return rc;
  if (n > 1 && flag_bytes < static_cast<size_t>(n) * pl.n_mtiles * sizeof(int))
    return set_error("conv3x3_chain: flag workspace too small (srb_conv3x3_chain_flag_bytes)");
  if (n > 1 && pl.rows_tile < a[0].W + 2)   // a tile's halo (one image row + 1 pixel) must stay inside its two neighbours
    return set_error("conv3x3_chain: image row wider than a tile");
  ChainParams cp{};
  for (int l = 0; l < n; ++l)
    if (int rc = igemm_fill(&a[l], pl, &cp.L[l], &cp.tx[l], &cp.tw[l])) return rc;
  cp.n_layers = n; cp.n_mtiles = pl.n_mtiles; cp.target = kEpiWarps * cp.L[0].n_ntiles; cp.flags = flags;
  const int sms = sm_count();
  const int grid = cp.L[0].num_work < sms ? cp.L[0].num_work : sms;
  if (t21) return conv3x3_chain21_launch_kernel(cp, pl, grid, stream);
  // flags bit 22 forces, bit 23 forbids the CTA-pair kernel (conv_chain2.cu); SRB_CONV_PAIR sets the default
  if (((a[0].flags & 0x400000) || (conv_pair_enabled() && !(a[0].flags & 0x800000))) && conv_pair_eligible(a, n, pl))
    return conv3x3_chain_pair_launch(cp, a, n, pl, stream);
  cudaError_t e;
  if (pl.MT == 4) {
    ensure_max_smem(reinterpret_cast<const void*>(conv3x3_chain_kernel<4>), 227 * 1024);
    e = launch_pdl(conv3x3_chain_kernel<4>, dim3(grid), dim3(kIgemmThreads), pl.smem, stream, cp);
  } else if (pl.MT == 2) {
    ensure_max_smem(reinterpret_cast<const void*>(conv3x3_chain_kernel<2>), 227 * 1024);
    e = launch_pdl(conv3x3_chain_kernel<2>, dim3(grid), dim3(kIgemmThreads), pl.smem, stream, cp);
  } else {
    ensure_max_smem(reinterpret_cast<const void*>(conv3x3_chain_kernel<1>), 227 * 1024);
    e = launch_pdl(conv3x3_chain_kernel<1>, dim3(grid), dim3(kIgemmThreads), pl.smem, stream, cp);
  }
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("conv3x3_chain launch", e);
  return 0;
}

}  // namespace srb
