// Association core of the backward on the tensor pipe (placeholder until the kernels land in this file).
#include "hop_common.cuh"

namespace hop {

bool tc_backward_supported(int, int, int, int) { return false; }
size_t tc_backward_workspace(int, int, int, int) { return 0; }
int launch_tc_backward(const TcBwdArgs&, cudaStream_t) {
  HOP_FAIL(HOP_ERR_UNSUPPORTED, "tensor-core backward is not built");
}

}  // namespace hop
