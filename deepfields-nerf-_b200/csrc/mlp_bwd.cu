// placeholder until the dgrad/wgrad kernels land
#include "mlp_common.cuh"
extern "C" long long dfn_mlp_dz_bytes(long long P) {
  const long long nt = (P + mlp::kTile - 1) / mlp::kTile;
  return nt * (10LL * 65536);
}
extern "C" int dfn_mlp_backward(const dfn_point_source*, long long, const void*, const float*, const float*,
                                const float*, const float*, const float*, const void*, void*, float*, void*) {
  dfn_set_error("mlp_backward: not built yet");
  return DFN_ERR_ARG;
}
