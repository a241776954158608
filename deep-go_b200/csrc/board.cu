// placeholder: GPU board analysis lands with the c5 inference path
#include "common.cuh"
#include "../../include/deepgo.h"
extern "C" int dg_summarize_boards(dg_ctx *ctx, const uint8_t *stones, uint8_t *out, int n) {
    dg_set_error("dg_summarize_boards: not built yet");
    return 3;
}
extern "C" int dg_predict_sym8(dg_model *m, float *probs_out, int32_t *argmax_out) {
    dg_set_error("dg_predict_sym8: not built yet");
    return 3;
}
