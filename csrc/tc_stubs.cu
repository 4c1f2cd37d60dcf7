// Temporary: tensor-core entry points not built yet report "unsupported".
#include "common.cuh"
#ifndef B200_HAVE_TC
extern "C" int b200_tc_gemm(const float*, const float*, float*, int, int, int, int, int, int, int, const float*, float, void*) {
  b200_set_error("tc_gemm: not built"); return B200_ERR_UNSUPPORTED; }
extern "C" int b200_conv2_fwd_tc(const float*, const float*, const float*, const void*, float*, uint8_t*, int, void*) {
  b200_set_error("conv2_fwd_tc: not built"); return B200_ERR_UNSUPPORTED; }
extern "C" int b200_conv2_dgrad_tc(const float*, const float*, const float*, float*, int, void*) {
  b200_set_error("conv2_dgrad_tc: not built"); return B200_ERR_UNSUPPORTED; }
extern "C" int b200_conv2_wgrad_tc(const float*, const float*, float*, float*, float*, int, void*) {
  b200_set_error("conv2_wgrad_tc: not built"); return B200_ERR_UNSUPPORTED; }
#endif
#ifndef B200_HAVE_TC
extern "C" int b200_has_tc(void) { return 0; }
#endif
