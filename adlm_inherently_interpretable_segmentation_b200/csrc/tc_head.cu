#include "tc_head.cuh"

namespace adlm {

int tc_head_fwd(const HeadShape&, int, const float*, const float*, const float*, float*, float*, float*,
                int32_t*, const StepExtras&, cudaStream_t) {
  set_error("tc_head_fwd: tcgen05 path not built yet");
  return ADLM_E_UNSUPPORTED;
}
int tc_head_bwd(const HeadShape&, int, const float*, const float*, const float*, const float*, const float*,
                const float*, float*, float*, float*, const StepExtras&, cudaStream_t) {
  set_error("tc_head_bwd: tcgen05 path not built yet");
  return ADLM_E_UNSUPPORTED;
}
int tc_push_scan(const float*, int, int, int, const uint32_t*, const float*, int, const int32_t*, long long,
                 int, unsigned long long*, cudaStream_t) {
  set_error("tc_push_scan: tcgen05 path not built yet");
  return ADLM_E_UNSUPPORTED;
}

}  // namespace adlm
