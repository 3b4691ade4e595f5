// Data path: constituent lists -> dense HR / LR image batches on the device.
//
// Replaces, for a whole batch in one launch, the per-item Python loop of the reference's sparse datasets:
//   datasets.py:136-145  extract(data, etaBins, phiBins): for every (flat index, energy) pair in list order
//                        phi = idx // etaBins, eta = idx % etaBins, img[0, eta, phi] += energy, stopping after the first
//                        pair whose energy is 0 (lists are zero padded);
//   datasets.py:247-249  img_lr = SumPool2d(factor)(img) (models.py:297-305), img_hr = img.
// One CTA per image: the image is built in shared memory by ONE thread walking the list in order (fp32 adds in the
// reference's order, so repeated cells give the same bits), then all threads stream it out (HR, coalesced) and pool
// it (LR, the summation order of srb_sumpool_fwd).  HBM-bound on the image write: H*W*4 (1 + 1/factor^2) bytes/image.
#include "common.cuh"

namespace srb {

__global__ void __launch_bounds__(256)
extract_batch_kernel(const float2* __restrict__ pairs, int N, int etaBins, int phiBins, int factor,
                     float* __restrict__ hr, float* __restrict__ lr, int* __restrict__ n_bad) {
  extern __shared__ float sm[];
  const int HW = etaBins * phiBins;
  float* img = sm;
  float2* lst = reinterpret_cast<float2*>(sm + ((HW + 1) & ~1));
  const int b = blockIdx.x;
  for (int i = threadIdx.x; i < HW; i += blockDim.x) img[i] = 0.f;
  for (int i = threadIdx.x; i < N; i += blockDim.x) lst[i] = __ldg(pairs + static_cast<size_t>(b) * N + i);
  __syncthreads();
  if (threadIdx.x == 0) {
    int bad = 0;
    for (int i = 0; i < N; ++i) {
      const float2 e = lst[i];
      if (e.y == 0.f) break;                       // the zero entry itself adds nothing (datasets.py:143-144)
      const int ii = __float2int_rz(e.x);
      const int phi = ii / etaBins, eta = ii - phi * etaBins;
      if (e.x < 0.f || phi >= phiBins) { ++bad; continue; }      // the reference raises IndexError here
      img[eta * phiBins + phi] += e.y;
    }
    if (bad && n_bad) atomicAdd(n_bad, bad);
  }
  __syncthreads();
  float* out = hr + static_cast<size_t>(b) * HW;
  for (int i = threadIdx.x; i < HW; i += blockDim.x) out[i] = img[i];
  if (lr) {
    const int Ho = etaBins / factor, Wo = phiBins / factor;
    float* lo = lr + static_cast<size_t>(b) * Ho * Wo;
    for (int i = threadIdx.x; i < Ho * Wo; i += blockDim.x) {
      const int yo = i / Wo, xo = i - yo * Wo;
      const float* p = img + (yo * factor) * phiBins + xo * factor;
      float s = 0.f;
      for (int a = 0; a < factor; ++a)
        for (int c = 0; c < factor; ++c) s += p[a * phiBins + c];
      lo[i] = s;
    }
  }
}

}  // namespace srb

extern "C" int srb_extract_batch(const float* pairs, int B, int N, int etaBins, int phiBins, int factor, float* hr,
                                 float* lr, int* n_bad, void* stream) {
  using namespace srb;
  if (!pairs || !hr) return set_error("extract_batch: null pointer");
  if (B <= 0 || N <= 0 || etaBins <= 0 || phiBins <= 0) return set_error("extract_batch: bad sizes");
  if (lr && (factor <= 0 || etaBins % factor || phiBins % factor))
    return set_error("extract_batch: image size must be a multiple of the pooling factor");
  const size_t smem = (static_cast<size_t>((etaBins * phiBins + 1) & ~1)) * 4 + static_cast<size_t>(N) * 8;
  if (smem > 227 * 1024) return set_error("extract_batch: image + list do not fit in shared memory");
  if (smem > 48 * 1024) ensure_max_smem(reinterpret_cast<const void*>(extract_batch_kernel), 227 * 1024);
  extract_batch_kernel<<<B, 256, smem, (cudaStream_t)stream>>>(reinterpret_cast<const float2*>(pairs), N, etaBins,
                                                               phiBins, factor, hr, lr, n_bad);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_cuda_error("extract_batch launch", e);
  return 0;
}
