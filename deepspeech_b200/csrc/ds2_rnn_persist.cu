// Persistent weights-resident recurrent kernel (placeholder until the tcgen05 version lands).
#include "common.cuh"
namespace ds2 {
int rnn_persist_supported(int, int, int) { return 0; }
size_t rnn_persist_workspace_bytes(int, int, int, int) { return 0; }
int rnn_relu_fwd_persist(const float*, float, const void*, const float*, const int*, void*, int, int, int, int, int, float, void*,
                         size_t, cudaStream_t) {
  set_error("persistent recurrent kernel not built");
  return DS2_ERR_UNSUPPORTED;
}
int rnn_relu_bwd_persist(const float*, const void*, const int*, const void*, void*, int, int, int, int, int, float, void*, size_t,
                         cudaStream_t) {
  set_error("persistent recurrent kernel not built");
  return DS2_ERR_UNSUPPORTED;
}
}  // namespace ds2
