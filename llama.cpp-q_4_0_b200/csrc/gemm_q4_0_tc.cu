// gemm_q4_0_tc.cu -- kernel 3: prefill GEMM (TMA + tcgen05 kind::i8 + TMEM).  [stub until built]
#include "state.h"

namespace b200 {
int  gemm_init() { return GGML_B200_OK; }
void gemm_shutdown() {}
int  gemm_set_tuning(const char *, int) { return GGML_B200_ERR_ARG; }
bool gemm_available(const ggml_b200_weight *) { return false; }
int  gemm_prepare_weight(ggml_b200_weight *) { return GGML_B200_OK; }
void gemm_release_weight(ggml_b200_weight *) {}
int  launch_gemm(const ggml_b200_weight *, float *, long long, int, int *) {
    set_error("tensor-core GEMM path not built");
    return GGML_B200_ERR_STATE;
}
}  // namespace b200
