// Unity translation unit: one device link, one copy of the device-side fault word.
#include "common.cu"
#include "pack.cu"
#include "conv_naive.cu"
#include "conv_igemm.cu"
#include "wgrad_tc.cu"
#include "gen_edge.cu"
#include "api.cu"
#include "probe.cu"
