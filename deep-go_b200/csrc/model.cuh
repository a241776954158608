// dg_model: parameters in Torch7 flat order + per-precision activation storage.
#pragma once
#include "common.cuh"

struct Bf16Path;   // conv_umma.cu
struct Tf32Path;   // conv_tf32.cu

// one layer of the launch that folds weight-gradient accumulation blocks [tap][cin/4][cout][4] into the Torch7-order gradient
struct FinalizeLayer {
    const float *gacc; float *gW; int cin_true, cin, cout, ksize;
    const float *gab; float *ga, *gb; int pitch, pad, first;      // bias gradients accumulated by the weight-gradient kernel (gab may be NULL)
};
int dg_bf16_wgrad_multi(dg_ctx *c, int num_sms, int nl, const void *const *dz_bf16, const void *const *x_bf16, float *const *gacc, int n, int cin_pad, int cout, int ksize);   // conv_umma.cu
int dg_launch_wgrad_finalize(dg_ctx *ctx, const FinalizeLayer *table_dev, int nlayers, cudaStream_t st);    // conv_umma.cu

struct dg_model {
    dg_ctx *ctx = nullptr;
    int L = 0, C = 0, in_planes = 37, k_first = 5, prec = 0, max_batch = 0;
    size_t P = 0;
    std::vector<LayerInfo> layers;
    float *params = nullptr, *grads = nullptr;
    // staged batch
    int n = 0;
    int global_batch = 0;
    bool have_targets = false, have_raw = false, forward_done = false;
    bool fused_step = false;           // inside dg_train_step: forward and backward run back to back
    uint8_t *d_raw = nullptr, *d_player = nullptr, *d_rank = nullptr, *d_sym = nullptr, *d_x = nullptr, *d_y = nullptr;
    int32_t *d_targets = nullptr, *d_argmax = nullptr;
    float *d_logp = nullptr, *d_loss_n = nullptr, *d_loss = nullptr;
    int *d_errors = nullptr;
    uint8_t *h_meta = nullptr;         // pinned staging ring for the small per-position arrays of dg_model_set_batch_raw
    cudaEvent_t ev_meta[4] = {nullptr, nullptr, nullptr, nullptr};
    unsigned meta_slot = 0;
    bool loss_reduced = false;         // d_loss / d_errors already hold the reduction of the last pass (done inside the backward pass)
    // captured CUDA graph of one fused training step (bf16 path, single GPU)
    cudaGraphExec_t step_graph = nullptr;
    int step_graph_n = 0;
    int step_graph_mode = 0;           // input mode the graph was captured in (raw records vs blocked image)
    long long step_graph_launches = 0;
    // rate for the graph's SGD node: device scalar, fed from a ring of pinned slots (a slot is rewritten only after the copy that
    // reads it has completed: steps that return without a host sync would otherwise see the next step's rate)
    // Record staging is double-buffered and copied on a stream of its own: while step k runs on the context's stream, the records
    // of minibatch k+1 travel into the other buffer (dg_train_step_begin / dg_model_set_batch_raw / dg_train_step_result).
    // d_raw, d_player .. d_y above always point into the buffer of the minibatch staged last.
    uint8_t *d_raw_buf[2] = {nullptr, nullptr}, *d_meta_buf[2] = {nullptr, nullptr};
    int stage_cur = 0;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copy[2] = {nullptr, nullptr}, ev_enc[2] = {nullptr, nullptr};    // copy into buffer b landed / last encode that read buffer b
    bool enc_recorded[2] = {false, false};
    float *h_loss_ring = nullptr;      // pinned, 4 slots: the loss of a step begun with dg_train_step_begin
    cudaEvent_t ev_loss[4] = {nullptr, nullptr, nullptr, nullptr};
    unsigned loss_slot = 0;
    int loss_pending = -1;
    float *h_res = nullptr;            // pinned {loss, error count}: where the step's scalars land before they go to the caller
    float *d_rate = nullptr, *h_rate = nullptr;
    cudaEvent_t ev_rate[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    unsigned rate_slot = 0;
    uint8_t *d_sym_keep = nullptr;     // dg_predict_sym8: the caller's symmetry ids while the eight passes overwrite d_sym
    // fp32 path (NCHW)
    std::vector<float *> act;          // act[0] = input planes, act[l] = h_l
    float *dz[2] = {nullptr, nullptr};
    // bf16 tcgen05 path
    Bf16Path *bf = nullptr;
    // tf32 tcgen05 path
    Tf32Path *t32 = nullptr;
    // model on a multi-device context: one replica per device, the staged batch split on dim 0 (multi.cu)
    std::vector<dg_model *> kids;
    std::vector<int> kid_lo;           // first position of the staged batch on replica k
    std::vector<float *> kid_gsum;     // per replica: the summed gradient vector of the separate-call path
};

// api.cu internals shared with multi.cu
int dg_model_layout(dg_model *m, int num_layers, int channel_size, int in_planes, int k_first);
int dg_forward_impl(dg_model *m);
int dg_backward_impl(dg_model *m);
int dg_read_loss(dg_model *m, float *loss_out, int *errors_out);
int dg_model_mark_encoded(dg_model *m);     // an encode that reads the current record buffer has been queued on the context's stream
int dg_bf16_dp_prepare_local(dg_model *const *kids, int n);   // conv_umma.cu

// conv_tf32.cu
int dg_tf32_create(dg_model *m);
void dg_tf32_destroy(dg_model *m);
int dg_tf32_params_changed(dg_model *m);                 // refresh the fp32 operand copies from the master
int dg_tf32_forward(dg_model *m);                        // conv layers 0 .. L-2; leaves h_{L-1} (NCHW) in act[L-1]
int dg_tf32_backward(dg_model *m, const float *dz_top);  // dz_top: dZ of conv layer L-2, NCHW; gradients of the conv layers into m->grads

// conv_umma.cu
int dg_bf16_create(dg_model *m);
void dg_bf16_destroy(dg_model *m);
int dg_bf16_params_changed(dg_model *m);
int dg_bf16_sgd_step(dg_model *m, float rate, const float *rate_dev);   // fused SGD update + operand repack                 // refresh bf16 operand copies from the fp32 master
int dg_bf16_stage_planes(dg_model *m);                   // act[0] (fp32 NCHW) -> blocked bf16 input image
int dg_bf16_stage_raw(dg_model *m);                      // raw records -> blocked bf16 input image (deferred into the first conv kernel)
bool dg_bf16_input_mode_fused();                         // DG_FUSED_ENCODE=1: no separate encode kernel (labels need their own launch then)
int dg_bf16_input_mode(const dg_model *m);               // 1: the next forward pass encodes raw records itself
int dg_bf16_forward(dg_model *m);                        // fills m->d_logp-ready logits via head
int dg_bf16_backward(dg_model *m, float scale);
int dg_bf16_dp_prepare(dg_model *m);                     // data parallel: map the peers' gradient buffers before the first fused step
int dg_bf16_sync_grads(dg_model *m);                     // fused step leaves the conv gradients in accumulation buffers: fold them into m->grads
bool dg_bf16_defers_grads(const dg_model *m);            // fused steps keep the conv gradients in the accumulation buffers
void dg_bf16_after_fused_step(dg_model *m);              // host-side state after a (replayed) fused step
void dg_bf16_drop_pending_grads(dg_model *m);            // ... or forget them (grads:zero())
