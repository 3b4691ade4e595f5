// extern "C" entry points of libsrb200 (see include/srb200.h).
#include "common.cuh"
#include "sm100.cuh"

namespace srb {
int conv3x3_igemm_launch(const srb_conv3x3_args* a, cudaStream_t stream);
int conv3x3_naive_launch(const srb_conv3x3_args* a, cudaStream_t stream);
int conv3x3_chain_launch(const srb_conv3x3_args* a, int n, int* flags, size_t flag_bytes, cudaStream_t stream);
size_t conv3x3_chain_flag_bytes(int batch, int H, int W, int n);
int wgrad3x3_naive_launch(const srb_wgrad3x3_args* a, cudaStream_t stream);
int wgrad3x3_tc_launch(const srb_wgrad3x3_args* a, cudaStream_t stream);
}  // namespace srb

extern "C" int srb_read_fault(unsigned int* fault_out) {
  unsigned int v = 0, z = 0;
  cudaError_t e = cudaMemcpyFromSymbol(&v, srb::g_srb_fault, sizeof(v));
  if (e != cudaSuccess) return srb::set_cuda_error("read fault word", e);
  if (v) cudaMemcpyToSymbol(srb::g_srb_fault, &z, sizeof(z));
  *fault_out = v;
  return 0;
}

extern "C" int srb_conv3x3(const srb_conv3x3_args* a, void* stream) {
  if (!a || !a->x || !a->w || !a->y) return srb::set_error("conv3x3: null pointer");
  if (a->impl == SRB_IMPL_NAIVE) {
    if (a->colsum || a->sign_out || a->mask_bits)
      return srb::set_error("conv3x3: colsum / sign_out / mask_bits are tcgen05-kernel options");
    return srb::conv3x3_naive_launch(a, (cudaStream_t)stream);
  }
  return srb::conv3x3_igemm_launch(a, (cudaStream_t)stream);
}

extern "C" int srb_conv3x3_chain(const srb_conv3x3_args* layers, int n, void* flags, size_t flag_bytes, void* stream) {
  if (!layers) return srb::set_error("conv3x3_chain: null pointer");
  return srb::conv3x3_chain_launch(layers, n, static_cast<int*>(flags), flag_bytes, (cudaStream_t)stream);
}
extern "C" size_t srb_conv3x3_chain_flag_bytes(int batch, int H, int W, int n) {
  return srb::conv3x3_chain_flag_bytes(batch, H, W, n);
}

extern "C" int srb_wgrad3x3(const srb_wgrad3x3_args* a, void* stream) {
  if (!a || !a->x || !a->gy || !a->dw) return srb::set_error("wgrad3x3: null pointer");
  if (a->cin % 64 || a->cout % 64) return srb::set_error("wgrad3x3: channels must be multiples of 64");
  if (a->impl == SRB_IMPL_NAIVE) return srb::wgrad3x3_naive_launch(a, (cudaStream_t)stream);
  return srb::wgrad3x3_tc_launch(a, (cudaStream_t)stream);
}
