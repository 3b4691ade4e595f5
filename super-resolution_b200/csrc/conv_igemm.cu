// K1: 3x3 stride-1 pad-1 convolution as an implicit GEMM on tcgen05 tensor cores (sm_100a).
//
// Replaces the cuDNN/oneDNN convolutions behind models.py:19,36-41,53 (dense blocks incl. concat and
// residual adds), models.py:67,125-126 (conv2 + skip), models.py:87-89 (upsampling conv + LeakyReLU +
// PixelShuffle), models.py:97-98 (conv3.0 + LeakyReLU) and their input gradients.
//
// Formulation.  Activations are PNHWC bf16 (see include/srb200.h): the flat matrix X[R = B*(H+1)*(W+1), C]
// has zero rows/columns between images, so tap (dy,dx) of the convolution is the constant row shift
// s = dy*(W+1) + dx of X.  One CTA owns MT*128 consecutive flat rows (pixels) and one 64-wide slice of the
// output channels:
//   D[128*MT x 64] = sum over k-blocks kb (64 input channels) and taps t of  X[q0+s_t .. , kb] * Wp[t][kb]^T
// A operand: ONE TMA load per k-block of the halo window rows [q0-halo, q0+128*MT+halo) x 64 channels into
//   128B-swizzled shared memory; the nine taps are nine UMMA descriptors whose start address is offset by
//   (halo + s_t) rows into that window (no re-load, no im2col).
// B operand: per (k-block, tap) a 64x64 K-major weight tile (8 KB) streamed through a ring of slots.
// Accumulators: MT x 64 fp32 columns of TMEM, double buffered so the epilogue of tile i overlaps the MMAs
//   of tile i+1.  Persistent CTAs, static round-robin tile schedule.
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane), warps 2..5 = epilogue
//   (TMEM -> registers -> bias/scale/residual/activation -> bf16 global stores).
#include <stdlib.h>

#include "common.cuh"
#include "sm100.cuh"

namespace srb {

constexpr int kIgemmThreads = 416;   // warp 0 TMA producer, warps 1-4 MMA issuers, warps 5-12 epilogue
constexpr int kEpiWarps = 8;         // two per TMEM lane group: each takes 32 of a block's 64 columns, which doubles the
                                     // mask / residual loads in flight (the K = 64 input-gradient convs are bound by them)
constexpr int kIssuers = 4;
constexpr int kBSlots = 4;           // weight ring: one slot = the three taps of one filter row (3 x 8 KB)
constexpr int kAStages = 2;          // halo-window ring
constexpr int kBTileBytes = 64 * 64 * 2;
constexpr int kBSlotBytes = 3 * kBTileBytes;

struct IgemmParams {
  int R;            // total flat rows = batch * P
  int P, pitch, H, W;
  int KB;           // cin / 64
  int cout;
  int n_ntiles;     // cout / 64
  int num_work;     // n_mtiles * n_ntiles
  int x_coff;
  int halo;         // pitch + 1
  int a_box_rows, a_nbox;   // TMA boxes per halo window
  uint32_t a_stage_bytes;
  const float* bias;
  __nv_bfloat16* y; int y_cstride, y_coff;
  int store_mode;   // SRB_STORE_*
  int y_pitch, y_P; // geometry of y (== pitch,P for plain; 2W+1,(2H+1)(2W+1) for shuffle)
  float alpha;
  int act; float slope;
  const __nv_bfloat16* mask; int mask_cstride, mask_coff;
  const __nv_bfloat16* res1; int res1_cstride, res1_coff; float res1_scale;
  const __nv_bfloat16* res2; int res2_cstride, res2_coff; float res2_scale;
  uint32_t* sign_out; int sign_cstride, sign_coff;          // optional: (output > 0) bits, one word per 32 channels
  const uint32_t* mask_bits; int maskb_cstride, maskb_coff; // LRELU_MASK from sign words instead of `mask`
  float* colsum;    // optional [64]: per-channel sums of the stored outputs (bias gradient), n_ntiles == 1 only
  int res_f32;      // bit 0 / 1: res1 / res2 are fp32 streams
  float* y32; int y32_cstride, y32_coff;                    // optional fp32 copy of the output (plain store)
  const __nv_bfloat16* res3; int res3_cstride, res3_coff; float res3_scale;
  const float* chan_scale;                                  // optional [batch][cout] factor on (acc + bias)
  int ext;                                                  // any of the four fields above in use (EXT epilogue)
  int base_offset_mode;
  int dbg;          // timing experiments only (flags >> 1): 1 no weight reloads, 2 no activation reloads, 4 no epilogue I/O
};

__device__ __forceinline__ void load_row64(const __nv_bfloat16* p, uint4 (&v)[8]) {
  const uint4* q = reinterpret_cast<const uint4*>(p);
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __ldg(q + i);
}

// Per-row epilogue shared by the tcgen05 convolution kernels: v[32] = fp32 accumulators of flat row q, output channels
// [nt*64 + half*32, +32) -> bias, alpha, residual streams, LeakyReLU / LeakyReLU' mask -> bf16 store (plain or
// PixelShuffle-addressed).  Called for valid pixels only.
// COHERENT = true reads the residual / mask streams with ld.global.cg instead of the non-coherent path: the chained
// kernel (conv_chain.cu) reads rows that an earlier layer of the SAME launch wrote.
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// Sign words: row-major [pixel][cstride words] (cstride > 0) or plane-major [word][-cstride pixels] (cstride < 0).
// Plane-major makes a warp's 32 rows x one word a single 128-byte access instead of 32 sectors with 4 useful bytes.
__device__ __forceinline__ size_t sign_word_index(int q, int cstride, int word) {
  return cstride >= 0 ? static_cast<size_t>(q) * cstride + word
                      : static_cast<size_t>(word) * static_cast<size_t>(-cstride) + q;
}

// The LeakyReLU' mask rows of a tile do not depend on its accumulators: pull them into L2 before waiting for the MMAs
// (they are forward activations, long evicted by the time the input-gradient pass reads them).
template <int MT>
__device__ __forceinline__ void igemm_prefetch_mask(const IgemmParams& p, int mt_idx, int nt, int lg, int lane, int half) {
  if (p.act != SRB_ACT_LRELU_MASK) return;
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
    const int q = mt_idx * (128 * MT) + mt * 128 + lg * 32 + lane;
    if (q >= p.R) continue;
    if (p.mask_bits) prefetch_l2(p.mask_bits + sign_word_index(q, p.maskb_cstride, p.maskb_coff + nt * 2 + half));
    else prefetch_l2(p.mask + static_cast<size_t>(q) * p.mask_cstride + p.mask_coff + nt * 64 + half * 32);
  }
}

// Column sums of a warp's 32 rows x 32 columns (v[] = one row per lane): transposing butterfly, 31 shuffles; lane j
// returns the sum of column j.
__device__ __forceinline__ float warp_colsum32(const float (&v)[32], int lane) {
  float a[16];
  const bool b16 = lane & 16;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float keep = b16 ? v[i + 16] : v[i], send = b16 ? v[i] : v[i + 16];
    a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
  float b[8];
  const bool b8 = lane & 8;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float keep = b8 ? a[i + 8] : a[i], send = b8 ? a[i] : a[i + 8];
    b[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  float c[4];
  const bool b4 = lane & 4;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float keep = b4 ? b[i + 4] : b[i], send = b4 ? b[i] : b[i + 4];
    c[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  float d[2];
  const bool b2 = lane & 2;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float keep = b2 ? c[i + 2] : c[i], send = b2 ? c[i] : c[i + 2];
    d[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  }
  const bool b1 = lane & 1;
  const float keep = b1 ? d[1] : d[0], send = b1 ? d[0] : d[1];
  return keep + __shfl_xor_sync(0xffffffffu, send, 1);
}

template <bool COHERENT>
__device__ __forceinline__ uint4 ld_row16(const uint4* p) { return COHERENT ? __ldcg(p) : __ldg(p); }
// 32 bytes per thread in one instruction (LDG.256, sm_100): the per-row streams of the epilogue are 64 or 128 bytes per
// thread, each thread on its own row - half the load instructions for the LSU to sequence
template <bool COHERENT>
__device__ __forceinline__ void ld_row32(const void* p, uint32_t (&r)[8]) {
  if (COHERENT)
    asm volatile("ld.global.cg.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]),
                 "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p));
  else
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]),
                 "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p));
}
template <bool COHERENT>
__device__ __forceinline__ float4 ld_row16f(const float4* p) { return COHERENT ? __ldcg(p) : __ldg(p); }

// v[32] += scale * (32 consecutive channels of a residual row): bf16 or fp32 source
template <bool COHERENT>
__device__ __forceinline__ void add_residual32(float (&v)[32], const void* base, size_t q, int cstride, int coff, int c0,
                                               float scale, bool f32) {
  if (f32) {
    const float* rp = static_cast<const float*>(base) + q * cstride + coff + c0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t u[8];
      ld_row32<COHERENT>(rp + 8 * i, u);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[8 * i + j] += scale * __uint_as_float(u[j]);
    }
  } else {
    const __nv_bfloat16* rp = static_cast<const __nv_bfloat16*>(base) + q * cstride + coff + c0;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      uint32_t w[8];
      ld_row32<COHERENT>(rp + 16 * i, w);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[16 * i + 2 * j] += scale * bf16lo(w[j]);
        v[16 * i + 2 * j + 1] += scale * bf16hi(w[j]);
      }
    }
  }
}

// EXT = false is the round-1 epilogue (bf16 residuals only); EXT = true adds the fp32 residual streams, the third
// residual, the per-image channel factor and the fp32 output copy.  Two instantiations so that the common layers (four
// of the five convolutions of a dense block) keep the lean code: the K = 64 / 128 layers are epilogue-bound.
// sbias: optional shared-memory copy of the 32 bias values this thread's columns need (conv_chain2.cu stages it once per
// layer: the flag protocol's fences invalidate L1, so a global bias load per row is an L2 round trip on the critical path
// of the epilogue-bound K = 64 / 128 layers)
template <bool COHERENT = false, bool EXT = false>
__device__ __forceinline__ void igemm_store_row(const IgemmParams& p, float (&v)[32], int q, int py, int px, int nt,
                                                int half, const float4* sbias = nullptr) {
  const int n0 = nt * 64;
  size_t yrow;
  int ycoff = p.y_coff;
  if (p.store_mode == SRB_STORE_SHUFFLE) {
    const int img = q / p.P;
    const int cblocks = p.n_ntiles >> 2;  // 64-wide blocks per shuffled output channel group (Cy / 64)
    const int sub = nt / cblocks;         // (i,j) = (sub>>1, sub&1)
    yrow = static_cast<size_t>(img) * p.y_P + static_cast<size_t>(2 * py + (sub >> 1)) * p.y_pitch +
           (2 * px + (sub & 1));
    ycoff += (nt % cblocks) * 64;
  } else {
    yrow = static_cast<size_t>(q);
    ycoff += n0;
  }
  const int c0 = n0 + half * 32;
  if (sbias) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 b = sbias[i];
      v[4 * i] += b.x; v[4 * i + 1] += b.y; v[4 * i + 2] += b.z; v[4 * i + 3] += b.w;
    }
  } else if (p.bias) {                              // (parameter views: 16-byte alignment only, 128-bit loads)
    const float4* b4 = reinterpret_cast<const float4*>(p.bias + c0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 b = __ldg(b4 + i);
      v[4 * i] += b.x; v[4 * i + 1] += b.y; v[4 * i + 2] += b.z; v[4 * i + 3] += b.w;
    }
  }
  if (EXT && p.chan_scale) {
    const float4* cs4 = reinterpret_cast<const float4*>(p.chan_scale + static_cast<size_t>(q / p.P) * p.cout + c0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 c = __ldg(cs4 + i);
      v[4 * i] *= c.x; v[4 * i + 1] *= c.y; v[4 * i + 2] *= c.z; v[4 * i + 3] *= c.w;
    }
  }
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] *= p.alpha;
  if (p.res1) add_residual32<COHERENT>(v, p.res1, static_cast<size_t>(q), p.res1_cstride, p.res1_coff, c0, p.res1_scale,
                                       EXT && (p.res_f32 & 1));
  if (p.res2) add_residual32<COHERENT>(v, p.res2, static_cast<size_t>(q), p.res2_cstride, p.res2_coff, c0, p.res2_scale,
                                       EXT && (p.res_f32 & 2));
  if (EXT && p.res3) add_residual32<COHERENT>(v, p.res3, static_cast<size_t>(q), p.res3_cstride, p.res3_coff, c0,
                                              p.res3_scale, false);
  if (p.act == SRB_ACT_LRELU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * p.slope;
  } else if (p.act == SRB_ACT_LRELU_MASK && p.mask_bits) {
    const uint32_t* wp = p.mask_bits + sign_word_index(q, p.maskb_cstride, p.maskb_coff + (c0 >> 5));
    const uint32_t bits = COHERENT ? __ldcg(wp) : __ldg(wp);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] *= ((bits >> i) & 1u) ? 1.f : p.slope;
  } else if (p.act == SRB_ACT_LRELU_MASK) {
    const uint4* mp = reinterpret_cast<const uint4*>(p.mask + static_cast<size_t>(q) * p.mask_cstride +
                                                     p.mask_coff + c0);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint4 u = ld_row16<COHERENT>(mp + i);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[8 * i + 2 * j] *= bf16lo(w[j]) > 0.f ? 1.f : p.slope;
        v[8 * i + 2 * j + 1] *= bf16hi(w[j]) > 0.f ? 1.f : p.slope;
      }
    }
  }
  if (p.sign_out) {
    uint32_t bits = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) bits |= (v[i] > 0.f ? 1u : 0u) << i;
    p.sign_out[sign_word_index(q, p.sign_cstride, p.sign_coff + (c0 >> 5))] = bits;
  }
  if (EXT && p.y32) {
    float* y32 = p.y32 + static_cast<size_t>(q) * p.y32_cstride + p.y32_coff + c0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
      asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(y32 + 8 * i), "r"(__float_as_uint(v[8 * i])),
                   "r"(__float_as_uint(v[8 * i + 1])), "r"(__float_as_uint(v[8 * i + 2])), "r"(__float_as_uint(v[8 * i + 3])),
                   "r"(__float_as_uint(v[8 * i + 4])), "r"(__float_as_uint(v[8 * i + 5])), "r"(__float_as_uint(v[8 * i + 6])),
                   "r"(__float_as_uint(v[8 * i + 7])) : "memory");
  }
  // 64 bytes per row and thread: two 256-bit stores (STG.256, sm_100).  A warp's store instruction touches 32 rows either
  // way; 32 bytes per row make every touched sector a full one and halve the instructions the LSU has to sequence.
  __nv_bfloat16* yp = p.y + yrow * p.y_cstride + ycoff + half * 32;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = pack_bf16(v[16 * i + 2 * j], v[16 * i + 2 * j + 1]);
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(yp + 16 * i), "r"(o[0]), "r"(o[1]), "r"(o[2]),
                 "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7]) : "memory");
  }
}

// dbg bit 8: CTA 0 records clock64 stamps of its pipeline (tools/exp_conv.py reads them back)
__device__ long long g_igemm_stamps[16];

template <int MT>
__global__ void __launch_bounds__(kIgemmThreads, 1)
conv3x3_igemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                     const IgemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int NB = MT == 4 ? 3 : kBSlots;     // weight ring slots (MT == 4: two 512-row windows leave room for three)
  // carve: [A stages][B slots][barriers]
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_smem = smem;
  uint8_t* b_smem = a_smem + kAStages * p.a_stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_smem + NB * kBSlotBytes);
  uint64_t* a_full = bars;                      // [kAStages]
  uint64_t* a_empty = a_full + kAStages;        // [kAStages]
  uint64_t* b_full = a_empty + kAStages;        // [kBSlots]
  uint64_t* b_empty = b_full + kBSlots;         // [kBSlots]
  uint64_t* t_full = b_empty + kBSlots;         // [2]
  uint64_t* t_empty = t_full + 2;               // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const bool stamp = (p.dbg & 8) && blockIdx.x == 0;
  if (stamp && threadIdx.x == 0) g_igemm_stamps[0] = clock64();
  // four accumulators (one per issuer warp) of 64 columns, double buffered: all 512 TMEM columns
  constexpr uint32_t kTmemCols = 512;
  constexpr int NACC = kIssuers / MT;           // accumulators that sum to one 128-row sub-tile

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
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
  if (stamp && threadIdx.x == 0) g_igemm_stamps[1] = clock64();
  pdl_trigger();
  pdl_wait();          // everything above overlapped the previous kernel; its results are visible from here on

  const int rows_per_tile = 128 * MT;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {   // elect.sync: ptxas then keeps descriptors and barrier addresses in uniform registers
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0;
      for (int work = blockIdx.x; work < p.num_work; work += gridDim.x) {
        const int nt = work % p.n_ntiles;
        const int mt_idx = work / p.n_ntiles;
        const int q0 = mt_idx * rows_per_tile;
        for (int kb = 0; kb < p.KB; ++kb) {
          mbar_wait(&a_empty[sa], pa ^ 1, 1);
          if ((p.dbg & 2) && (work != blockIdx.x || kb >= kAStages)) mbar_arrive(&a_full[sa]);
          else {
          mbar_expect_tx(&a_full[sa], p.a_stage_bytes);
          for (int bx = 0; bx < p.a_nbox; ++bx)
            tma_load_2d(a_smem + sa * p.a_stage_bytes + bx * p.a_box_rows * 128, &tmap_x, p.x_coff + kb * 64,
                        q0 - p.halo + bx * p.a_box_rows, &a_full[sa]);
          }
          if (++sa == kAStages) { sa = 0; pa ^= 1; }
          for (int dy = 0; dy < 3; ++dy) {
            mbar_wait(&b_empty[sb], pb ^ 1, 2);
            if ((p.dbg & 1) && (work != blockIdx.x || kb > 0)) { mbar_arrive(&b_full[sb]); }
            else {
              mbar_expect_tx(&b_full[sb], kBSlotBytes);
              for (int dx = 0; dx < 3; ++dx)
                tma_load_2d(b_smem + sb * kBSlotBytes + dx * kBTileBytes, &tmap_w, 0,
                            ((dy * 3 + dx) * p.KB + kb) * p.cout + nt * 64, &b_full[sb]);
            }
            if (++sb == NB) { sb = 0; pb ^= 1; }
          }
        }
      }
    }
  } else if (warp <= kIssuers) {
    // ------------------------------------------------------------------ MMA issuers (warps 1..4, one thread each)
    // A single thread sustains roughly one tcgen05.mma per 100 clk through the descriptor arithmetic and the
    // barrier traffic, but a 128x64x16 MMA only occupies the tensor pipe for 32 clk (tools/probe1.py rate).  So the
    // issue work is split over four warps on four different schedulers.  MMAs from different threads are unordered,
    // hence every issuer owns a private 64-column accumulator (no "first MMA clears" hazard); the epilogue adds them.
    //   MT == 4: issuer i -> sub-tile i, all four k-steps (one accumulator per sub-tile, nothing to add);
    //   MT == 2: issuer i -> sub-tile i >> 1, k-steps {2(i&1), 2(i&1)+1};   MT == 1: issuer i -> k-step i.
    if (elect_one()) {   // elect.sync: ptxas then keeps descriptors and barrier addresses in uniform registers
      const int iss = warp - 1;
      const int mt = MT == 4 ? iss : (MT == 2 ? (iss >> 1) : 0);
      const int k0 = MT == 4 ? 0 : (MT == 2 ? 2 * (iss & 1) : iss);
      constexpr int NK = MT == 4 ? 4 : (MT == 2 ? 2 : 1);
      constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
      constexpr uint64_t desc_hi = (uint64_t(16 >> 4) << 16) | (uint64_t(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      const uint32_t a0 = ((smem_u32(a_smem) & 0x3FFFFu) >> 4) + mt * 1024 + k0 * 2;
      const uint32_t b0 = ((smem_u32(b_smem) & 0x3FFFFu) >> 4) + k0 * 2;
      const uint32_t a_stage16 = p.a_stage_bytes >> 4;
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0, as = 0, pt = 0;
      for (int work = blockIdx.x; work < p.num_work; work += gridDim.x) {
        mbar_wait(&t_empty[as], pt ^ 1, 3);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * 256 + iss * 64;
        for (int kb = 0; kb < p.KB; ++kb) {
          mbar_wait(&a_full[sa], pa, 4);
          if (stamp && iss == 0 && kb == 0) g_igemm_stamps[work == (int)blockIdx.x ? 2 : 5] = clock64();
          const uint32_t a_lo = a0 + sa * a_stage16;
          uint32_t row16 = static_cast<uint32_t>(p.halo - p.pitch - 1) * 8u;     // tap (-1,-1), in 16-byte units
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
            row16 += static_cast<uint32_t>(p.pitch) * 8u;
            if (++sb == NB) { sb = 0; pb ^= 1; }
          }
          umma_commit(&a_empty[sa]);
          if (kb == p.KB - 1) umma_commit(&t_full[as]);
          if (stamp && iss == 0 && kb == p.KB - 1) g_igemm_stamps[work == (int)blockIdx.x ? 3 : 6] = clock64();
          if (++sa == kAStages) { sa = 0; pa ^= 1; }
        }
        if (++as == 2) { as = 0; pt ^= 1; }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps 5..12
    const int lg = warp & 3;                    // TMEM lane group this warp may access
    const int half = (warp - 5) >> 2;           // which 32 of the 64 columns
    uint32_t as = 0, pt = 0;
    float vs[32];                               // p.colsum: per-column sums over the rows this lane stored
#pragma unroll
    for (int i = 0; i < 32; ++i) vs[i] = 0.f;
    for (int work = blockIdx.x; work < p.num_work; work += gridDim.x) {
      const int nt = work % p.n_ntiles;
      const int mt_idx = work / p.n_ntiles;
      igemm_prefetch_mask<MT>(p, mt_idx, nt, lg, lane, half);
      mbar_wait(&t_full[as], pt, 6);
      tc_fence_after();
#pragma unroll 1
      for (int mt = 0; mt < MT; ++mt) {
        const int q = mt_idx * rows_per_tile + mt * 128 + lg * 32 + lane;
        const int r = q % p.P;
        const int py = r / p.pitch, px = r - py * p.pitch;
        const bool valid = (q < p.R) && (px < p.W) && (py < p.H) && !(p.dbg & 4);
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
          if (p.ext) igemm_store_row<false, true>(p, v, q, py, px, nt, half);
          else igemm_store_row<false, false>(p, v, q, py, px, nt, half);
        }
        if (p.colsum && valid) {
#pragma unroll
          for (int i = 0; i < 32; ++i) vs[i] += v[i];
        }
      }
      tc_fence_before();
      __syncwarp();
      if (stamp && warp == 5 && lane == 0) g_igemm_stamps[work == (int)blockIdx.x ? 4 : 7] = clock64();
      if (lane == 0) mbar_arrive(&t_empty[as]);
      if (++as == 2) { as = 0; pt ^= 1; }
    }
    if (p.colsum) atomicAdd(p.colsum + half * 32 + lane, warp_colsum32(vs, lane));   // one butterfly per launch
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
  if (stamp && threadIdx.x == 0) g_igemm_stamps[8] = clock64();
}

// ---------------------------------------------------------------------------------- host side
int conv3x3_row_launch(const srb_conv3x3_args* a, cudaStream_t stream);     // conv_row.cu
// Kernel choice: the one-tap-per-MMA kernel below is the default; SRB_CONV_ROW=1 (read once per process) or flags bit
// 16 selects the filter-row kernel, flags bit 17 forces the one-tap kernel (tests run both on the same inputs).
static bool conv_row_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("SRB_CONV_ROW"); v = (e && e[0] == '1') ? 1 : 0; }
  return v != 0;
}
static size_t igemm_smem_bytes(uint32_t a_stage_bytes, int MT) {
  return 1024 + size_t(kAStages) * a_stage_bytes + size_t(MT == 4 ? 3 : kBSlots) * kBSlotBytes + 512;
}
// SRB_CONV_MT4=0 (read once per process) keeps the 512-row tiles off.
static bool conv_mt4_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("SRB_CONV_MT4"); v = (e && e[0] == '0') ? 0 : 1; }
  return v != 0;
}

// Geometry-dependent part of a launch (shared by every layer of a chained launch, conv_chain.cu).
struct IgemmPlan {
  int MT, rows_tile, box_rows, nbox, n_mtiles;
  uint32_t a_stage_bytes;
  size_t smem;
  long long R;
};

int igemm_check_args(const srb_conv3x3_args* a) {
  if (a->cin % 64 || a->cout % 64 || a->cin <= 0 || a->cout <= 0) return set_error("conv3x3: channels must be multiples of 64");
  if ((a->x_cstride % 8) || (a->x_coff % 64))
    return set_error("conv3x3: input channel stride / offset must keep 16-byte alignment (offset: a multiple of 64)");
  // the epilogue moves its per-row streams (output, bf16 / fp32 residuals, fp32 copy) with 256-bit accesses
  auto al32 = [](const void* ptr, int cstride, int coff, int elem) {
    return ptr == nullptr || (((reinterpret_cast<uintptr_t>(ptr) & 31) == 0) && ((cstride * elem) % 32 == 0) && ((coff * elem) % 32 == 0));
  };
  if (!al32(a->y, a->y_cstride, a->y_coff, 2) || !al32(a->res1, a->res1_cstride, a->res1_coff, (a->res_f32 & 1) ? 4 : 2) ||
      !al32(a->res2, a->res2_cstride, a->res2_coff, (a->res_f32 & 2) ? 4 : 2) ||
      !al32(a->res3, a->res3_cstride, a->res3_coff, 2) || !al32(a->y32, a->y32_cstride, a->y32_coff, 4))
    return set_error("conv3x3: output / residual pointers, channel strides and offsets must keep 32-byte alignment");
  if (a->store_mode == SRB_STORE_SHUFFLE && a->cout % 256) return set_error("conv3x3: shuffle store needs cout % 256 == 0");
  const long long R = 1LL * a->batch * (a->H + 1) * (a->W + 1);
  if (R <= 0 || R > 0x7fffffffLL) return set_error("conv3x3: bad geometry");
  return 0;
}

// Tile height.  Every (tile, 64-channel k-block) step re-reads the k-block's nine weight tiles (72 KB) from L2 next to
// the halo window of the tile (37 KB per 256 rows): at 256 rows per tile the weights are two thirds of the L2 -> SM
// traffic, and that traffic (7.6 TB/s at the marginal rate) bounds the kernel.  512-row tiles (MT = 4: one issuer and
// one accumulator per 128-row sub-tile) halve the weight traffic per pixel and drop the partial-accumulator adds of the
// epilogue; they are used when the two windows fit next to a three-slot weight ring (LR geometries) and the coarser
// tiles still fill the machine about as evenly.  Otherwise two 128-row sub-tiles per CTA when that fills the machine,
// else one.  flags bit 18 forbids, bit 19 forces (if it fits) the 512-row tiles.
static bool igemm_geom(const srb_conv3x3_args* a, int MT, IgemmPlan* pl) {
  const int pitch = a->W + 1, P = (a->H + 1) * (a->W + 1);
  const long long R = 1LL * a->batch * P;
  const int halo = pitch + 1;
  const int rows_tile = 128 * MT;
  const int win = rows_tile + 2 * halo;
  const int win8 = (win + 7) & ~7;
  const int nbox = (win8 + 255) / 256;
  const int box_rows = (((win8 + nbox - 1) / nbox) + 7) & ~7;
  pl->MT = MT; pl->rows_tile = rows_tile; pl->box_rows = box_rows; pl->nbox = nbox;
  pl->a_stage_bytes = uint32_t(nbox) * box_rows * 128;
  pl->smem = igemm_smem_bytes(pl->a_stage_bytes, MT);
  pl->R = R;
  pl->n_mtiles = int((R + rows_tile - 1) / rows_tile);
  return pl->smem <= 227 * 1024;
}

int igemm_make_plan(const srb_conv3x3_args* a, IgemmPlan* pl) {
  const int sms = sm_count();
  const int n_ntiles = a->cout / 64;
  auto fill = [&](const IgemmPlan& q) {       // fraction of CTA-slots doing work over the launch's waves
    const long long work = 1LL * q.n_mtiles * n_ntiles;
    return double(work) / double(((work + sms - 1) / sms) * sms);
  };
  IgemmPlan p4, p2;
  const bool fit4 = igemm_geom(a, 4, &p4), fit2 = igemm_geom(a, 2, &p2);
  const bool allow4 = fit4 && !(a->flags & 0x40000) && (conv_mt4_enabled() || (a->flags & 0x80000));
  if (allow4 && ((a->flags & 0x80000) ||
                 (1LL * p4.n_mtiles * n_ntiles >= sms && fill(p4) >= fill(p2) - 0.04))) { *pl = p4; return 0; }
  if (fit2 && 1LL * p2.n_mtiles * n_ntiles >= sms) { *pl = p2; return 0; }
  if (!igemm_geom(a, 1, pl)) return set_error("conv3x3: image too wide for the tcgen05 kernel (W <= ~400)");
  return 0;
}

// Per-layer part: tensor maps and kernel parameters of one convolution under plan `pl`.
int igemm_fill(const srb_conv3x3_args* a, const IgemmPlan& pl, IgemmParams* pp, CUtensorMap* tx, CUtensorMap* tw) {
  const int pitch = a->W + 1, P = (a->H + 1) * (a->W + 1);
  if (make_tmap_2d_bf16(tx, a->x, (uint64_t)a->x_cstride, (uint64_t)pl.R, (uint64_t)a->x_cstride * 2, 64, pl.box_rows))
    return 1;
  if (make_tmap_2d_bf16(tw, a->w, 64, (uint64_t)9 * (a->cin / 64) * a->cout, 128, 64, 64)) return 1;
  IgemmParams p{};
  p.R = (int)pl.R; p.P = P; p.pitch = pitch; p.H = a->H; p.W = a->W;
  p.KB = a->cin / 64; p.cout = a->cout; p.n_ntiles = a->cout / 64;
  p.num_work = pl.n_mtiles * p.n_ntiles;
  p.x_coff = a->x_coff; p.halo = pitch + 1; p.a_box_rows = pl.box_rows; p.a_nbox = pl.nbox;
  p.a_stage_bytes = pl.a_stage_bytes;
  p.bias = a->bias;
  p.y = (__nv_bfloat16*)a->y; p.y_cstride = a->y_cstride; p.y_coff = a->y_coff; p.store_mode = a->store_mode;
  if (a->store_mode == SRB_STORE_SHUFFLE) { p.y_pitch = 2 * a->W + 1; p.y_P = (2 * a->H + 1) * (2 * a->W + 1); }
  else { p.y_pitch = pitch; p.y_P = P; }
  p.alpha = a->alpha; p.act = a->act; p.slope = a->slope;
  p.mask = (const __nv_bfloat16*)a->mask; p.mask_cstride = a->mask_cstride; p.mask_coff = a->mask_coff;
  p.res1 = (const __nv_bfloat16*)a->res1; p.res1_cstride = a->res1_cstride; p.res1_coff = a->res1_coff; p.res1_scale = a->res1_scale;
  p.res2 = (const __nv_bfloat16*)a->res2; p.res2_cstride = a->res2_cstride; p.res2_coff = a->res2_coff; p.res2_scale = a->res2_scale;
  p.base_offset_mode = a->flags & 1;
  p.dbg = (a->flags >> 1) & 0x7fff;
  p.sign_out = (uint32_t*)a->sign_out; p.sign_cstride = a->sign_cstride; p.sign_coff = a->sign_coff;
  p.mask_bits = (const uint32_t*)a->mask_bits; p.maskb_cstride = a->maskb_cstride; p.maskb_coff = a->maskb_coff;
  if (p.sign_out && a->store_mode != SRB_STORE_PLAIN) return set_error("conv3x3: sign_out needs a plain store");
  p.colsum = a->colsum;
  if (p.colsum && (a->cout != 64 || a->store_mode != SRB_STORE_PLAIN))
    return set_error("conv3x3: colsum needs cout == 64 and a plain store");
  p.res_f32 = a->res_f32;
  p.y32 = a->y32; p.y32_cstride = a->y32_cstride; p.y32_coff = a->y32_coff;
  p.res3 = (const __nv_bfloat16*)a->res3; p.res3_cstride = a->res3_cstride; p.res3_coff = a->res3_coff; p.res3_scale = a->res3_scale;
  p.chan_scale = a->chan_scale;
  p.ext = (p.res_f32 || p.y32 || p.res3 || p.chan_scale) ? 1 : 0;
  if ((p.y32 || p.chan_scale) && a->store_mode != SRB_STORE_PLAIN)
    return set_error("conv3x3: y32 / chan_scale need a plain store");
  if (p.y32 && ((a->y32_cstride | a->y32_coff) % 4)) return set_error("conv3x3: y32 stride / offset must keep 16-byte alignment");
  if (((p.res_f32 & 1) && ((a->res1_cstride | a->res1_coff) % 4)) || ((p.res_f32 & 2) && ((a->res2_cstride | a->res2_coff) % 4)))
    return set_error("conv3x3: fp32 residual stride / offset must keep 16-byte alignment");
  if (p.act == SRB_ACT_LRELU_MASK && !p.mask && !p.mask_bits)
    return set_error("conv3x3: mask activation without mask tensor");
  *pp = p;
  return 0;
}

bool conv_t21_enabled();                                                                          // conv_chain21.cu
bool conv_t21_plan(const srb_conv3x3_args* a, struct IgemmPlan* pl);
bool conv_pair_enabled();                                                                         // conv_chain2.cu
bool conv_pair_eligible(const srb_conv3x3_args* a, int n, const IgemmPlan& pl);
int conv3x3_chain_launch(const srb_conv3x3_args* a, int n, int* flags, size_t flag_bytes, cudaStream_t stream);

int conv3x3_igemm_launch(const srb_conv3x3_args* a, cudaStream_t stream) {
  if (int rc = igemm_check_args(a)) return rc;
  // 2+1-tap formulation (conv_chain21.cu): a chain of one layer
  if (((a->flags & 0x100000) || (conv_t21_enabled() && !(a->flags & 0x200000))) && ((a->flags >> 1) & 0x7fff) == 0) {
    IgemmPlan pl21;
    if (conv_t21_plan(a, &pl21)) return conv3x3_chain_launch(a, 1, nullptr, 0, stream);    // else: image too wide for it
  }
  // CTA-pair kernel (conv_chain2.cu): a chain of one layer
  if (((a->flags & 0x400000) || (conv_pair_enabled() && !(a->flags & 0x800000))) && a->impl == SRB_IMPL_TCGEN05 &&
      ((a->flags >> 1) & 0x7fff) == 0 && !(a->flags & 0x30000)) {
    IgemmPlan plp;
    if (igemm_make_plan(a, &plp) == 0 && conv_pair_eligible(a, 1, plp)) return conv3x3_chain_launch(a, 1, nullptr, 0, stream);
  }
  const bool want_row = (a->flags & 0x10000) || (conv_row_enabled() && !(a->flags & 0x20000));
  if (want_row && ((a->flags >> 1) & 0x7fff) == 0 && !a->colsum && !a->sign_out && !a->mask_bits && !a->res_f32 &&
      !a->y32 && !a->res3 && !a->chan_scale) {      // filter-row kernel (conv_row.cu); -1: geometry does not fit
    const int rc = conv3x3_row_launch(a, stream);
    if (rc >= 0) return rc;
  }
  IgemmPlan pl;
  if (int rc = igemm_make_plan(a, &pl)) return rc;
  CUtensorMap tx, tw;
  IgemmParams p;
  if (int rc = igemm_fill(a, pl, &p, &tx, &tw)) return rc;
  const int sms = sm_count();
  const int grid = p.num_work < sms ? p.num_work : sms;
  const size_t smem = pl.smem;
  cudaError_t e;
  if (pl.MT == 4) {
    ensure_max_smem(reinterpret_cast<const void*>(conv3x3_igemm_kernel<4>), 227 * 1024);
    e = launch_pdl(conv3x3_igemm_kernel<4>, dim3(grid), dim3(kIgemmThreads), smem, stream, tx, tw, p);
  } else if (pl.MT == 2) {
    ensure_max_smem(reinterpret_cast<const void*>(conv3x3_igemm_kernel<2>), 227 * 1024);
    e = launch_pdl(conv3x3_igemm_kernel<2>, dim3(grid), dim3(kIgemmThreads), smem, stream, tx, tw, p);
  } else {
    ensure_max_smem(reinterpret_cast<const void*>(conv3x3_igemm_kernel<1>), 227 * 1024);
    e = launch_pdl(conv3x3_igemm_kernel<1>, dim3(grid), dim3(kIgemmThreads), smem, stream, tx, tw, p);
  }
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("conv3x3_igemm launch", e);
  return 0;
}

}  // namespace srb

extern "C" int srb_igemm_read_stamps(long long* out16) {
  cudaError_t e = cudaMemcpyFromSymbol(out16, srb::g_igemm_stamps, sizeof(long long) * 16);
  return e == cudaSuccess ? 0 : srb::set_cuda_error("read stamps", e);
}
