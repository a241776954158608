// placeholder until the tcgen05 path lands
#include "model.cuh"
int dg_bf16_create(dg_model *m) { dg_set_error("DG_PREC_BF16 path not built yet"); return 3; }
void dg_bf16_destroy(dg_model *m) {}
int dg_bf16_params_changed(dg_model *m) { return 0; }
int dg_bf16_stage_planes(dg_model *m) { return 3; }
int dg_bf16_stage_raw(dg_model *m) { return 3; }
int dg_bf16_forward(dg_model *m) { return 3; }
int dg_bf16_backward(dg_model *m, float scale) { return 3; }
