// gemv_mma.cu -- variant 2 of the quantised matmul (placeholder until the TMA-staged tensor-core kernel lands).
#include "pib_common.cuh"

bool pib_gemv_mma_supported(const pib_weight * w, int b) { (void) w; (void) b; return false; }

int pib_gemv_mma(const pib_weight * w, const void * act_dev, int b, float * y_dev, const int * cancel, cudaStream_t st) {
    (void) cancel; (void) w; (void) act_dev; (void) b; (void) y_dev; (void) st;
    pib_set_error("pib_gemv: variant 2 not built");
    return PIB_ERR_UNSUPP;
}
