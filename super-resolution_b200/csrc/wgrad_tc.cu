// K3: weight gradient of the 3x3 convolution on tcgen05 tensor cores (sm_100a).
// Replaces the filter half of aten::convolution_backward for every generator conv (autograd of
// models.py:19,67,87,97).
//
//   dW[co][ci][t] = alpha * sum_q  X[q][ci] * GZ[q - s_t][co],      s_t = dy*(W+1) + dx   (PNHWC flat rows)
//
// GEMM view: the reduction (K) dimension is the flat pixel index, so both operands are "MN-major" in shared
// memory exactly as TMA delivers PNHWC rows (128 B = 64 channels per pixel, 128B swizzle):
//   A = X tile    [K = 64 pixels][M = 128 input channels]  (two 64-channel TMA boxes, LBO = box size)
//   B = GZ window [K = 64 pixels][N = 192 = 3 x 64]        the three 64-wide N blocks are the SAME smem
//       window advanced by one pixel row each (LBO = 128 B): N block j is tap dx = 1 - j.  One MMA therefore
//       produces three taps of one filter row; dy selects the window's row offset.
// The M = 128 rows of an MMA are two independent 64-row halves that only have to share the B window.  A unit fills them
// in one of two ways (round 2; before, a unit was a pair of ADJACENT input-channel blocks for one dy, which left half
// of the M tile empty on every odd 64-channel block: 17 % of the MMAs of a dense block):
//   row-pair unit (one per 64-channel input block c):  half 0 = X[q][c], half 1 = X[q + pitch][c] against the dy = -1
//       window - half 1 then IS the dy = 0 tap row (sum_q X[q+pitch] GZ[q+pitch-1+j] = sum_q' X[q'] GZ[q'-1+j]);
//   block-pair unit (one per two input blocks c, c'):  half 0 = X[q][c], half 1 = X[q][c'] against the dy = +1 window.
// A dense block (cin = 64..320) takes 24 units instead of 27, with three idle halves instead of nine; 148 / 24 gives
// 6 K splits = 144 CTAs (was 5 x 27 = 135).  The K range starts ceil(pitch / 64) chunks before row 0 so that the
// shifted half sees the first image row (rows < 0 are TMA out-of-bounds zeros).
// Work item = (unit, 64-wide output-channel block, K split).  Each CTA accumulates its whole K range in 192 TMEM
// columns per issuer, then writes partial [3 taps][64 ci][64 co] fp32 tiles (one per half) to the workspace; a second
// kernel reduces the K splits, scales, and scatters to OIHW (state_dict layout).
#include <mutex>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "sm100.cuh"

namespace srb {

constexpr int kWgThreads = 224;                 // warp 0 TMA, warps 1-2 MMA issuers, warps 3-6 epilogue
constexpr int kWgIssuers = 2;
constexpr int kWgKc = 64;                       // pixels per pipeline stage
constexpr int kWgStages = 6;
constexpr int kWgXBytes = 2 * kWgKc * 128;      // two 64-channel boxes
constexpr int kWgGRows = kWgKc + 8;
constexpr int kWgGBytes = kWgGRows * 128;
constexpr int kWgStageBytes = kWgXBytes + kWgGBytes;

struct WgradParams {
  int R, pitch;
  int cin, cout;
  int nk;            // cin / 64
  int n_cb;          // cout / 64
  int lead;          // chunks in front of row 0 (ceil(pitch / 64))
  int splits;        // K splits
  int chunks_per_split;
  int x_coff, gy_coff;
  float* ws;         // [splits][9][cin][cout]
  int base_offset_mode;
  float alpha;
  int unshuffle;
};

// Several weight gradients that read the same activation buffer and the same gradient buffer (the five convolutions of
// a dense block) run as ONE launch: more work per CTA (fewer K splits, fewer partial tiles to reduce) and one
// prologue / tail instead of five.
constexpr int kWgMaxGroup = 8;
constexpr int kWgColsumFloats = 2048;           // widest gradient buffer the bias pass handles (channels)
struct WgradGroup {
  int n, splits;
  int dbg;                           // timing experiments (SRB_WGRAD_DBG): 1 = operands loaded for the first ring pass only
  int unit_start[kWgMaxGroup + 1];   // first (unit, cb) work item of every problem
  WgradParams prob[kWgMaxGroup];
};

__global__ void __launch_bounds__(kWgThreads, 1)
wgrad3x3_tc_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_g,
                   const __grid_constant__ WgradGroup g) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kWgStages * kWgStageBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + kWgStages;
  uint64_t* done = empty + kWgStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // decode work item
  const int split = blockIdx.x % g.splits;
  const int unit = blockIdx.x / g.splits;
  int pi = 0;
  while (pi + 1 < g.n && unit >= g.unit_start[pi + 1]) ++pi;
  const WgradParams& p = g.prob[pi];
  int w = unit - g.unit_start[pi];
  const int cb = w % p.n_cb; w /= p.n_cb;
  // halves of the M tile: input block, row shift of its X rows, tap row (dy + 1) it produces; dyB = dy of the B window
  int c0, c1, shift1, dyi0, dyi1, dyB;
  if (w < p.nk) { c0 = w; c1 = w; shift1 = p.pitch; dyi0 = 0; dyi1 = 1; dyB = -1; }
  else { const int v = w - p.nk; c0 = 2 * v; c1 = 2 * v + 1; shift1 = 0; dyi0 = 2; dyi1 = 2; dyB = 1; }
  const bool has1 = c1 < p.nk;
  const int n_chunks_total = p.lead + (p.R + kWgKc - 1) / kWgKc;
  const int c_begin = split * p.chunks_per_split;
  int c_end = c_begin + p.chunks_per_split;
  if (c_end > n_chunks_total) c_end = n_chunks_total;
  const int n_chunks = c_end > c_begin ? c_end - c_begin : 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_g);
    for (int i = 0; i < kWgStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], kWgIssuers); }
    mbar_init(done, kWgIssuers);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();

  if (warp == 0) {
    if (elect_one()) {   // elect.sync: ptxas then keeps descriptors and barrier addresses in uniform registers
      uint32_t s = 0, ph = 0;
      for (int c = 0; c < n_chunks; ++c) {
        const int k0 = (c_begin + c - p.lead) * kWgKc;
        mbar_wait(&empty[s], ph ^ 1, 11);
        if ((g.dbg & 1) && c >= kWgStages) { mbar_arrive(&full[s]); if (++s == kWgStages) { s = 0; ph ^= 1; } continue; }
        mbar_expect_tx(&full[s], has1 ? kWgStageBytes : kWgStageBytes - kWgKc * 128);
        uint8_t* st = smem + s * kWgStageBytes;
        tma_load_2d(st, &tmap_x, p.x_coff + c0 * 64, k0, &full[s]);
        if (has1) tma_load_2d(st + kWgKc * 128, &tmap_x, p.x_coff + c1 * 64, k0 + shift1, &full[s]);
        tma_load_2d(st + kWgXBytes, &tmap_g, p.gy_coff + cb * 64, k0 - dyB * p.pitch - 1, &full[s]);
        if (++s == kWgStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp <= kWgIssuers) {
    // two issuer threads (see conv_igemm.cu): issuer i takes k-steps {2i, 2i+1} of every 64-pixel chunk and owns a
    // private 192-column accumulator; the epilogue adds the two.
    if (elect_one()) {   // elect.sync: ptxas then keeps descriptors and barrier addresses in uniform registers
      constexpr uint32_t idesc = make_idesc_bf16(128, 192, 1, 1);
      constexpr uint64_t a_hi = (uint64_t((kWgKc * 128) >> 4) << 16) | (uint64_t(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      constexpr uint64_t b_hi = (uint64_t(128 >> 4) << 16) | (uint64_t(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      const int iss = warp - 1;
      const uint32_t d_tmem = tmem_base + iss * 192;
      const uint32_t x0 = ((smem_u32(smem) & 0x3FFFFu) >> 4) + iss * 2 * (2048 >> 4);
      uint32_t s = 0, ph = 0;
      for (int c = 0; c < n_chunks; ++c) {
        mbar_wait(&full[s], ph, 12);
        tc_fence_after();
        const uint32_t xa = x0 + s * (kWgStageBytes >> 4);
        const uint32_t ga = xa + (kWgXBytes >> 4);
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const uint64_t ad = a_hi | static_cast<uint64_t>(xa + k * (2048 >> 4));
          const uint64_t bd = b_hi | static_cast<uint64_t>(ga + k * (2048 >> 4));
          umma_bf16(d_tmem, ad, bd, idesc, k != 0 ? 1u : static_cast<uint32_t>(c != 0));
        }
        umma_commit(&empty[s]);
        if (c == n_chunks - 1) umma_commit(done);
        if (++s == kWgStages) { s = 0; ph ^= 1; }
      }
    }
  } else {
    const int lg = warp & 3;
    const int row = lg * 32 + lane;               // M index: half = row / 64, input channel within the block = row % 64
    const int hf = row >> 6;
    const int ci = (hf ? c1 : c0) * 64 + (row & 63);
    const int dyi = hf ? dyi1 : dyi0;
    if (n_chunks > 0) {
      mbar_wait(done, 0, 13);
      tc_fence_after();
    }
    const bool live = !hf || has1;                // warp-uniform (a warp lies inside one half)
#pragma unroll 1
    for (int j = 0; j < 3; ++j) {
      if (!live) break;
      const int t = dyi * 3 + (2 - j);            // dx = 1 - j  ->  kx = dx + 1 = 2 - j
      float* dst = p.ws + ((static_cast<size_t>(split) * 9 + t) * p.cin + ci) * p.cout + cb * 64;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t acc[32], acc2[32];
        if (n_chunks > 0) {
          tmem_ld32(tmem_base + (static_cast<uint32_t>(lg * 32) << 16) + j * 64 + half * 32, acc);
          tmem_ld32(tmem_base + (static_cast<uint32_t>(lg * 32) << 16) + 192 + j * 64 + half * 32, acc2);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) { acc[i] = 0u; acc2[i] = 0u; }
        }
        float4* d4 = reinterpret_cast<float4*>(dst + half * 32);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          d4[i] = make_float4(__uint_as_float(acc[4 * i]) + __uint_as_float(acc2[4 * i]),
                              __uint_as_float(acc[4 * i + 1]) + __uint_as_float(acc2[4 * i + 1]),
                              __uint_as_float(acc[4 * i + 2]) + __uint_as_float(acc2[4 * i + 2]),
                              __uint_as_float(acc[4 * i + 3]) + __uint_as_float(acc2[4 * i + 3]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// Sum the K splits, scale, and scatter to OIHW.  One thread per (t, ci, co); co fastest so workspace reads coalesce.
// blockIdx.y selects the problem of the group.
// Bias gradients: column sums of the gradient buffer over all flat rows (pad rows are zero), every channel of the
// buffer in one pass (the five convolutions of a dense block read adjacent 64-channel slices of it).  A thread owns
// one 16-byte channel group and strides over rows; block partials go to `out` with atomics (zeroed by the caller).
__global__ void __launch_bounds__(256)
colsum_kernel(const __nv_bfloat16* __restrict__ g, long long R, int cstride, float* __restrict__ out) {
  __shared__ float red[8 * 256];
  pdl_trigger();
  pdl_wait();
  const int c8n = cstride >> 3;                       // 16-byte groups per row
  const int rpb = 256 / c8n;                          // rows per block per step
  const int ck = threadIdx.x % c8n, rr = threadIdx.x / c8n;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  if (rr < rpb) {
    for (long long r = blockIdx.x * (long long)rpb + rr; r < R; r += (long long)gridDim.x * rpb) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(g + r * cstride) + ck);
      acc[0] += bf16lo(v.x); acc[1] += bf16hi(v.x); acc[2] += bf16lo(v.y); acc[3] += bf16hi(v.y);
      acc[4] += bf16lo(v.z); acc[5] += bf16hi(v.z); acc[6] += bf16lo(v.w); acc[7] += bf16hi(v.w);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) red[i * 256 + threadIdx.x] = acc[i];
  __syncthreads();
  for (int i = threadIdx.x; i < cstride; i += 256) {
    const int k = i >> 3, e = i & 7;
    float sum = 0.f;
    for (int j = 0; j < rpb; ++j) sum += red[e * 256 + j * c8n + k];
    atomicAdd(out + i, sum);
  }
}

// Sum the K splits, scale, and scatter to OIHW.  One thread per (t, ci, co); co fastest so workspace reads coalesce.
// blockIdx.y selects the problem of the group.  The first `cout` threads of blockIdx.x == 0 also finish the bias
// gradient from the column sums.
struct ReduceItem { const float* ws; float* dw; int cin, cin_pad, cout; float alpha; float* colsum; float* dbias; };
struct ReduceGroup { int n, splits, accumulate, unshuffle, colsum_clear; ReduceItem it[kWgMaxGroup]; };

__global__ void wgrad_reduce_kernel(const __grid_constant__ ReduceGroup rg) {
  const ReduceItem& r_ = rg.it[blockIdx.y];
  const int cin = r_.cin, cout = r_.cout, cin_pad = r_.cin_pad;
  const long long total = 9LL * cin * cout;
  pdl_trigger();
  pdl_wait();
  if (blockIdx.x == 0 && r_.dbias) {
    for (int co = threadIdx.x; co < cout; co += blockDim.x) {
      const int cq = cout / 4;
      const int d = rg.unshuffle ? (4 * (co % cq) + co / cq) : co;
      const float v = r_.alpha * r_.colsum[co];
      r_.dbias[d] = rg.accumulate ? r_.dbias[d] + v : v;
      if (rg.colsum_clear) r_.colsum[co] = 0.f;     // caller-owned sums (srb_wgrad3x3_args.gy_colsum): consumed, reset
    }
  }
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int co = int(idx % cout);
    long long r = idx / cout;
    const int ci = int(r % cin);
    const int t = int(r / cin);
    float s = 0.f;
    for (int sp = 0; sp < rg.splits; ++sp) s += r_.ws[((static_cast<size_t>(sp) * 9 + t) * cin_pad + ci) * cout + co];
    s *= r_.alpha;
    const int cq = cout / 4;
    const int co_dst = rg.unshuffle ? (4 * (co % cq) + co / cq) : co;
    float* d = r_.dw + (static_cast<size_t>(co_dst) * cin + ci) * 9 + t;
    *d = rg.accumulate ? *d + s : s;
  }
}

static void wgrad_group_plan(long long R, int pitch, int n, const int* cin, const int* cout, int* splits,
                             int* chunks_per_split, int* unit_start) {
  int units = 0;
  for (int i = 0; i < n; ++i) {
    unit_start[i] = units;
    const int nk = cin[i] / 64;
    units += (nk + (nk + 1) / 2) * (cout[i] / 64);     // row-pair units + block-pair units, per 64-wide output block
  }
  unit_start[n] = units;
  const int chunks = (pitch + kWgKc - 1) / kWgKc + int((R + kWgKc - 1) / kWgKc);
  int s = sm_count() / units;                         // one CTA per SM (150 KB of smem each), one wave, no tail wave
  if (s < 1) s = 1;
  const int max_s = (chunks + 7) / 8;                 // at least 8 chunks per split
  if (s > max_s) s = max_s;
  if (s < 1) s = 1;
  int cps = (chunks + s - 1) / s;
  s = (chunks + cps - 1) / cps;
  *splits = s; *chunks_per_split = cps;
}

static size_t wgrad_group_ws(int n, const int* cin, const int* cout, int splits, size_t* offs) {
  size_t total = 0;
  for (int i = 0; i < n; ++i) {
    if (offs) offs[i] = total;
    total += static_cast<size_t>(splits) * 9 * cin[i] * cout[i];
  }
  return (total + kWgColsumFloats) * sizeof(float);   // + column sums of the gradient buffer (bias gradients)
}

int wgrad3x3_tc_launch_group(const srb_wgrad3x3_args* a, int n, cudaStream_t stream) {
  if (n < 1 || n > kWgMaxGroup) return set_error("wgrad3x3 group: 1..8 problems");
  const srb_wgrad3x3_args& a0 = a[0];
  const int pitch = a0.W + 1, P = (a0.H + 1) * (a0.W + 1);
  const long long R = 1LL * a0.batch * P;
  if (R <= 0 || R > 0x7fffffffLL) return set_error("wgrad3x3: bad geometry");
  int cin[kWgMaxGroup], cout[kWgMaxGroup];
  for (int i = 0; i < n; ++i) {
    const srb_wgrad3x3_args& ai = a[i];
    if (ai.batch != a0.batch || ai.H != a0.H || ai.W != a0.W || ai.x != a0.x || ai.gy != a0.gy ||
        ai.x_cstride != a0.x_cstride || ai.gy_cstride != a0.gy_cstride || ai.accumulate != a0.accumulate ||
        ai.unshuffle != a0.unshuffle)
      return set_error("wgrad3x3 group: problems must share buffers, geometry and flags");
    if (ai.cin % 64 || ai.cout % 64 || !ai.dw) return set_error("wgrad3x3: channels must be multiples of 64");
    if ((ai.x_cstride % 8) || (ai.gy_cstride % 8) || (ai.x_coff % 8) || (ai.gy_coff % 8))
      return set_error("wgrad3x3: channel strides/offsets must keep 16-byte alignment");
    cin[i] = ai.cin; cout[i] = ai.cout;
  }
  WgradGroup g{};
  int splits, cps;
  wgrad_group_plan(R, pitch, n, cin, cout, &splits, &cps, g.unit_start);
  size_t offs[kWgMaxGroup];
  const size_t need = wgrad_group_ws(n, cin, cout, splits, offs);
  if (!a0.workspace || a0.workspace_bytes < need) return set_error("wgrad3x3: workspace too small");

  CUtensorMap tx, tg;
  if (make_tmap_2d_bf16(&tx, a0.x, (uint64_t)a0.x_cstride, (uint64_t)R, (uint64_t)a0.x_cstride * 2, 64, kWgKc)) return 1;
  if (make_tmap_2d_bf16(&tg, a0.gy, (uint64_t)a0.gy_cstride, (uint64_t)R, (uint64_t)a0.gy_cstride * 2, 64, kWgGRows)) return 1;

  g.n = n; g.splits = splits;
  { static int dbg = -1; if (dbg < 0) { const char* e = getenv("SRB_WGRAD_DBG"); dbg = e ? atoi(e) : 0; } g.dbg = dbg; }
  for (int i = 0; i < n; ++i)
    if (a[i].gy_colsum != a0.gy_colsum) return set_error("wgrad3x3 group: problems must share gy_colsum");
  const bool ext_colsum = a0.gy_colsum != nullptr;
  float* colsum = ext_colsum ? a0.gy_colsum : (float*)a0.workspace + (need / sizeof(float) - kWgColsumFloats);
  bool any_bias = false;
  ReduceGroup rg{};
  rg.n = n; rg.splits = splits; rg.accumulate = a0.accumulate; rg.unshuffle = a0.unshuffle;
  rg.colsum_clear = ext_colsum ? 1 : 0;
  long long max_total = 0;
  for (int i = 0; i < n; ++i) {
    const srb_wgrad3x3_args& ai = a[i];
    WgradParams& p = g.prob[i];
    p.R = (int)R; p.pitch = pitch; p.cin = ai.cin; p.cout = ai.cout;
    p.nk = ai.cin / 64; p.n_cb = ai.cout / 64; p.lead = (pitch + kWgKc - 1) / kWgKc;
    p.splits = splits; p.chunks_per_split = cps; p.x_coff = ai.x_coff; p.gy_coff = ai.gy_coff;
    p.ws = (float*)a0.workspace + offs[i]; p.base_offset_mode = 0;
    p.alpha = ai.alpha; p.unshuffle = ai.unshuffle;
    rg.it[i] = ReduceItem{p.ws, ai.dw, ai.cin, ai.cin, ai.cout, ai.alpha,
                          ai.dbias ? colsum + ai.gy_coff : nullptr, ai.dbias};
    any_bias = any_bias || ai.dbias != nullptr;
    const long long tot = 9LL * ai.cin * ai.cout;
    if (tot > max_total) max_total = tot;
  }
  if (any_bias && !ext_colsum) {
    if (a0.gy_cstride > kWgColsumFloats) return set_error("wgrad3x3: gradient buffer too wide for the bias pass");
    cudaMemsetAsync(colsum, 0, sizeof(float) * a0.gy_cstride, stream);
    const int rpb = 256 / (a0.gy_cstride / 8);
    long long blocks = (R + rpb - 1) / rpb;
    if (blocks > 8LL * sm_count()) blocks = 8LL * sm_count();
    colsum_kernel<<<(unsigned)blocks, 256, 0, stream>>>((const __nv_bfloat16*)a0.gy, R, a0.gy_cstride, colsum);
    cudaError_t ce = cudaGetLastError();
    if (ce != cudaSuccess) return set_cuda_error("colsum launch", ce);
  }
  const size_t smem = 1024 + size_t(kWgStages) * kWgStageBytes + 256;
  ensure_max_smem(reinterpret_cast<const void*>(wgrad3x3_tc_kernel), (int)smem);
  const int grid = g.unit_start[n] * splits;
  cudaError_t e = launch_pdl(wgrad3x3_tc_kernel, dim3(grid), dim3(kWgThreads), smem, stream, tx, tg, g);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("wgrad3x3_tc launch", e);
  int rblocks = ceil_div(max_total, 256);
  const int cap = 2 * sm_count() / n > 1 ? 2 * sm_count() / n : 1;
  if (rblocks > cap * 4) rblocks = cap * 4;
  e = launch_pdl(wgrad_reduce_kernel, dim3(rblocks, n), dim3(256), 0, stream, rg);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("wgrad_reduce launch", e);
  return 0;
}

int wgrad3x3_tc_launch(const srb_wgrad3x3_args* a, cudaStream_t stream) { return wgrad3x3_tc_launch_group(a, 1, stream); }

}  // namespace srb

extern "C" size_t srb_wgrad3x3_workspace_bytes(int batch, int H, int W, int cin, int cout) {
  int splits, cps, us[srb::kWgMaxGroup + 1];
  srb::wgrad_group_plan(1LL * batch * (H + 1) * (W + 1), W + 1, 1, &cin, &cout, &splits, &cps, us);
  return srb::wgrad_group_ws(1, &cin, &cout, splits, nullptr);
}

extern "C" size_t srb_wgrad3x3_group_workspace_bytes(int batch, int H, int W, int n, const int* cin, const int* cout) {
  if (n < 1 || n > srb::kWgMaxGroup) return 0;
  int splits, cps, us[srb::kWgMaxGroup + 1];
  srb::wgrad_group_plan(1LL * batch * (H + 1) * (W + 1), W + 1, n, cin, cout, &splits, &cps, us);
  return srb::wgrad_group_ws(n, cin, cout, splits, nullptr);
}

extern "C" int srb_wgrad3x3_group(const srb_wgrad3x3_args* items, int n, void* stream) {
  if (!items) return srb::set_error("wgrad3x3 group: null pointer");
  for (int i = 0; i < n; ++i)
    if (!items[i].x || !items[i].gy || !items[i].dw) return srb::set_error("wgrad3x3 group: null pointer");
  return srb::wgrad3x3_tc_launch_group(items, n, (cudaStream_t)stream);
}

// ---- allocation-free variant: split-K scratch in a static device array ------------------------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).  No counterpart
// in the reference.  The array is part of the library image: every GPU
// that loads libsrb200 holds kWgStaticWsBytes of it until the process exits.  All calls share it and it keeps the last
// call's contents: the partial tiles need no clearing (every tile the reduction reads is overwritten by the main kernel
// of the same call), the bias column sums at its tail are cleared by the call itself.  Two calls that overlap on the
// GPU (different streams, or two host threads) race on it.
constexpr size_t kWgStaticWsBytes = size_t(16) << 20;      // 16 MiB: any single 3x3 problem up to cin * cout = 320 * 64
static __device__ __align__(256) unsigned char g_wgrad_static_ws[kWgStaticWsBytes];

namespace srb {
// In host code g_wgrad_static_ws is only a handle of the symbol, not a device address: the address is looked up once per
// device, with that device current, and that pointer is what the kernels get.
static int wgrad_static_ws_ptr(void** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static void* ptr[kMaxDev] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return set_cuda_error("wgrad3x3_static: cudaGetDevice", e);
  if (dev < 0 || dev >= kMaxDev) return set_error("wgrad3x3_static: device ordinal out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_wgrad_static_ws);
    if (e != cudaSuccess || !p) return set_cuda_error("wgrad3x3_static: cudaGetSymbolAddress", e);
    if (reinterpret_cast<uintptr_t>(p) & 255) return set_error("wgrad3x3_static: scratch array is not 256-byte aligned");
    ptr[dev] = p;
  }
  *out = ptr[dev];
  return 0;
}
}  // namespace srb

extern "C" size_t srb_wgrad3x3_static_max_bytes(void) { return kWgStaticWsBytes; }

extern "C" int srb_wgrad3x3_static(const srb_wgrad3x3_args* a, void* stream) {
  if (!a || !a->x || !a->gy || !a->dw) return srb::set_error("wgrad3x3_static: null pointer");
  if (a->cin <= 0 || a->cout <= 0 || a->cin % 64 || a->cout % 64)
    return srb::set_error("wgrad3x3_static: channels must be multiples of 64");
  if (a->batch <= 0 || a->H <= 0 || a->W <= 0) return srb::set_error("wgrad3x3_static: bad geometry");
  const size_t need = srb_wgrad3x3_workspace_bytes(a->batch, a->H, a->W, a->cin, a->cout);
  if (need > kWgStaticWsBytes) {
    char msg[256];
    snprintf(msg, sizeof(msg), "wgrad3x3_static: this problem needs %zu bytes of split-K scratch, the built-in array "
             "holds %zu (use srb_wgrad3x3 with a caller-owned workspace)", need, (size_t)kWgStaticWsBytes);
    return srb::set_error(msg);
  }
  void* ws = nullptr;
  if (srb::wgrad_static_ws_ptr(&ws)) return 1;
  srb_wgrad3x3_args b = *a;
  b.workspace = ws;
  b.workspace_bytes = kWgStaticWsBytes;
  b.impl = SRB_IMPL_TCGEN05;
  return srb::wgrad3x3_tc_launch(&b, (cudaStream_t)stream);
}
