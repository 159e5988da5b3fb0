#include "common.cuh"
int clsm_conv_pool_fwd_tc(const void* X, int64_t S, int L, int Dpad, const void* Wt, int k, int F, int Fpad,
                          uint64_t* keys, int ldk, int col0, cudaStream_t st) {
  clsm_set_error("conv_pool_fwd: tcgen05 path not built");
  return CLSM_EUNSUPPORTED;
}
