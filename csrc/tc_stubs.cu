// Tensor-core entry points that are not built yet report "unsupported" (removed as they land).
#include "common.cuh"
#ifndef B200_HAVE_TC_CONV2
extern "C" int b200_conv2_fwd_tc(const float*, const float*, const float*, const void*, float*, uint8_t*, int, void*) {
  b200_set_error("conv2_fwd_tc: not built"); return B200_ERR_UNSUPPORTED; }
extern "C" int b200_conv2_dgrad_tc(const float*, const float*, const float*, float*, int, void*) {
  b200_set_error("conv2_dgrad_tc: not built"); return B200_ERR_UNSUPPORTED; }
extern "C" int b200_conv2_wgrad_tc(const float*, const float*, float*, float*, float*, int, void*) {
  b200_set_error("conv2_wgrad_tc: not built"); return B200_ERR_UNSUPPORTED; }
#endif
