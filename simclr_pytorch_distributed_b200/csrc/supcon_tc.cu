// Tensor-core engine of the loss (placeholder until the tcgen05 kernels land).
#include "supcon_tc.cuh"

namespace scb {

int64_t tc_workspace_bytes(const scb_supcon_problem*, int) { return 0; }

int tc_forward(const scb_supcon_problem*, int, int, float*, int*, cudaStream_t) {
    set_error("tensor-core engine not built yet");
    return SCB_ERR_UNSUPPORTED;
}

int tc_backward(const scb_supcon_problem*, int, int, const float*, const float*, const float*, int,
                float*, int, void*, cudaStream_t) {
    set_error("tensor-core engine not built yet");
    return SCB_ERR_UNSUPPORTED;
}

}  // namespace scb
