// Unity translation unit: one device link, one copy of the device-side fault word.
//
// SRB_EXPERIMENTAL (nvcc -DSRB_EXPERIMENTAL, i.e. SRB_EXPERIMENTAL=1 in the environment of the build) also compiles the
// opt-in kernels that lost their measurements in round 1 - the filter-row convolution (conv_row.cu: 10 % slower in the
// step), the 2+1-tap convolution (conv_chain21.cu: 70.7 vs 70.0 us) - and the tcgen05 issue-rate probe (probe.cu).  The
// write-ups are profiles/r01_s27_conv_row_experiments.txt, r01_s29_kernel_experiments.txt, r01_s21_mma_rate_probe.txt.
// The product library is built without them.
#include "common.cu"
#include "pack.cu"
#include "conv_naive.cu"
#include "conv_igemm.cu"
#ifdef SRB_EXPERIMENTAL
#include "conv_row.cu"
#endif
#include "conv_chain.cu"
#ifdef SRB_EXPERIMENTAL
#include "conv_chain21.cu"
#endif
#include "conv_chain2.cu"
#include "wgrad_tc.cu"
#include "gen_edge.cu"
#include "disc_fast.cu"
#include "disc.cu"
#include "disc_mma.cu"
#include "conv_out_tc.cu"
#include "optim_loss.cu"
#include "opt_losses.cu"
#include "sinkhorn.cu"
#include "data_path.cu"
#include "eval_stats.cu"
#include "api.cu"
#ifdef SRB_EXPERIMENTAL
#include "probe.cu"
extern "C" int srb_has_experimental(void) { return 1; }
#else
namespace srb {
int conv3x3_row_launch(const srb_conv3x3_args*, cudaStream_t) { return -1; }          // "does not fit": default kernel
bool conv_t21_enabled() { return false; }
bool conv_t21_plan(const srb_conv3x3_args*, IgemmPlan*) { return false; }
int conv3x3_chain21_launch_kernel(const ChainParams&, const IgemmPlan&, int, cudaStream_t) {
  return set_error("the 2+1-tap kernel is not part of this build (SRB_EXPERIMENTAL)");
}
}  // namespace srb
extern "C" int srb_probe_mma_rate(int, int, int, int, int, int, int, long long*, void*) {
  return srb::set_error("the issue-rate probe is not part of this build (SRB_EXPERIMENTAL)");
}
extern "C" int srb_has_experimental(void) { return 0; }
#endif
