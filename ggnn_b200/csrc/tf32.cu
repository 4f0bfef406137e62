// tf32.cu -- tcgen05 (kind::tf32) timestep.  (placeholder until the kernels land)
#include "tf32.cuh"

int ggnn_tf32_pack_weights(ggnn_ctx*, int, bool, cudaStream_t) { return GGNN_OK; }
bool ggnn_tf32_supported(const ggnn_ctx*) { return false; }
int ggnn_step_tf32(ggnn_ctx* ctx, int, const ggnn_plan&, const ggnn_step_bufs&, float*, cudaStream_t) {
  GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "TF32 path not built yet");
}
