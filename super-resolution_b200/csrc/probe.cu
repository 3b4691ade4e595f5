// Bring-up microbenchmark (not on the product path): sustained tcgen05.mma issue rate for the operand shapes the
// conv kernels can use, with operands resident in shared memory (no loads in the timed loop).  Answers SURVEY.md
// section 7 "hard part 1": how far below peak is an N=64 tile, and does a CTA pair (cta_group::2) recover it.
#include "common.cuh"
#include "sm100.cuh"

namespace srb {

struct ProbeParams {
  int N;            // UMMA N (per instruction; for cta_group::2 the pair computes 256 x N)
  int iters;        // MMA instructions issued (K = 16 each)
  int a_row_shift;  // start row offset of the A window (exercises unaligned descriptor starts)
  int a_mn_major;   // 0: K-major A/B (forward conv shapes); 1: MN-major (wgrad shapes)
  int b_lbo;        // LBO for B when MN-major (128 = the one-pixel-shift trick)
  long long* cycles;  // [gridDim.x]
};

template <int CG>
__global__ void __launch_bounds__(128, 1) probe_mma_kernel(const ProbeParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t cta_rank = 0;
  if (CG == 2) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
  // operands: A window 64 KB (zero-filled), B 32 KB
  for (int i = threadIdx.x; i < (96 * 1024) / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  fence_proxy_async();
  if (warp == 0) {
    if (CG == 1) tmem_alloc(&tmem_slot, 512);
    else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
  }
  tc_fence_before();
  if (CG == 2) { asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  long long t0 = 0, t1 = 0;
  if (warp == 1 && cta_rank == 0) {
    const uint32_t idesc = make_idesc_bf16(CG == 2 ? 256 : 128, p.N, p.a_mn_major, p.a_mn_major);
    const uint32_t a_base = smem_u32(smem) + p.a_row_shift * 128;
    const uint32_t b_base = smem_u32(smem) + 64 * 1024;
    if (elect_one()) {
      t0 = clock64();
      for (int i = 0; i < p.iters; ++i) {
        const int k = i & 3;
        const int acc = (i >> 2) & 1;
        uint64_t ad, bd;
        if (p.a_row_shift < 0) {
          // replica of the conv kernel's issue order: 9 taps x 2 sub-tiles x 4 k-steps, 12 weight slots, pitch 21
          const int mt = (i >> 2) & 1, t = (i >> 3) % 9, slot = (i >> 3) % 12;
          const int variant = -p.a_row_shift;           // 1: all as in the kernel, 2: fixed A window, 3: fixed B slot
          const int shift = variant == 2 ? 0 : 22 + (t / 3 - 1) * 21 + (t % 3 - 1);
          const uint32_t aa = smem_u32(smem) + (shift + mt * 128) * 128 + k * 32;
          const uint32_t bb = smem_u32(smem) + 40 * 1024 + (variant == 3 ? 0 : slot) * 4096 + k * 32;
          ad = make_sw128_desc(aa, 16, 1024, 0);
          bd = make_sw128_desc(bb, 16, 1024, 0);
          if (CG == 1) umma_bf16(tmem_base + mt * 64, ad, bd, idesc, 1);
          continue;
        }
        if (p.a_mn_major) {
          ad = make_sw128_desc(a_base + k * 2048, 8192, 1024, 0);
          bd = make_sw128_desc(b_base + k * 2048, p.b_lbo, 1024, 0);
        } else {
          ad = make_sw128_desc(a_base + k * 32, 16, 1024, 0);
          bd = make_sw128_desc(b_base + k * 32, 16, 1024, 0);
        }
        if (CG == 1) umma_bf16(tmem_base + acc * 256, ad, bd, idesc, 1);
        else
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                       "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_base + acc * 256),
                       "l"(ad), "l"(bd), "r"(idesc), "r"(1u) : "memory");
      }
      if (CG == 1) umma_commit(&bar);
      else
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "h"((uint16_t)1) : "memory");
    }
    __syncwarp();
    mbar_wait(&bar, 0, 21);
    t1 = clock64();
    if (lane == 0) p.cycles[blockIdx.x] = t1 - t0;
  }
  tc_fence_before();
  if (CG == 2) { asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }
  else __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    if (CG == 1) tmem_dealloc(tmem_base, 512);
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

}  // namespace srb

// cycles_out: device array of gridDim entries (only leader CTAs write for cta_group 2).
extern "C" int srb_probe_mma_rate(int cta_group, int N, int iters, int a_row_shift, int mn_major, int b_lbo,
                                  int n_ctas, long long* cycles_out, void* stream) {
  using namespace srb;
  ProbeParams p{N, iters, a_row_shift, mn_major, b_lbo, cycles_out};
  const size_t smem = 97 * 1024 + 1024;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(n_ctas); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem; cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute at[1];
  cudaError_t e;
  if (cta_group == 2) {
    cudaFuncSetAttribute(probe_mma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, probe_mma_kernel<2>, p);
  } else {
    cudaFuncSetAttribute(probe_mma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    e = cudaLaunchKernelEx(&cfg, probe_mma_kernel<1>, p);
  }
  if (e != cudaSuccess) return set_cuda_error("probe launch", e);
  return 0;
}
