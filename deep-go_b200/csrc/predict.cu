// c5 inference: move probabilities averaged over the 8 board symmetries (BASELINE config 5).
//   p_avg[n][q] = 1/8 * sum_s exp(logp_s[n][T_s(q)])      logp_s = forward(encode(record, sym = s))
// The prediction the net makes at the transformed point T_s(q) belongs to the original point q (DESIGN.md symmetry spec).
// Records must have been staged with dg_model_set_batch_raw; encode runs on the device once per symmetry.
#include "model.cuh"
#include "../../include/deepgo.h"

__global__ void __launch_bounds__(384) sym_accumulate_kernel(const float *__restrict__ logp, float *__restrict__ acc, int sym, int first) {
    const int n = blockIdx.x, q = threadIdx.x;
    if (q >= DG_POINTS) return;
    int x = q / DG_BOARD, y = q - x * DG_BOARD;                  // forward action T_s on the original point
    if (sym & 1) { int t = x; x = y; y = t; }
    if (sym & 2) x = DG_BOARD - 1 - x;
    if (sym & 4) y = DG_BOARD - 1 - y;
    const float p = 0.125f * expf(logp[(size_t)n * DG_POINTS + x * DG_BOARD + y]);
    float *o = acc + (size_t)n * DG_POINTS + q;
    *o = first ? p : *o + p;
}

// first index of the maximum per row (preds:max(2), train.lua:29, on the averaged distribution)
__global__ void __launch_bounds__(128) row_argmax_kernel(const float *__restrict__ acc, int32_t *__restrict__ out, int n) {
    const int row = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    float mx = -1.f;
    int best = 0;
    for (int p = lane; p < DG_POINTS; p += 32) {
        const float v = acc[(size_t)row * DG_POINTS + p];
        if (v > mx) { mx = v; best = p; }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, mx, s);
        const int ob = __shfl_xor_sync(0xffffffffu, best, s);
        if (om > mx || (om == mx && ob < best)) { mx = om; best = ob; }
    }
    if (lane == 0) out[row] = best;
}

__global__ void fill_u8_kernel(uint8_t *p, int n, uint8_t v) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

int dgm_predict_sym8(dg_model *m, float *probs_out, int32_t *argmax_out);      // multi.cu

extern "C" int dg_predict_sym8(dg_model *m, float *probs_out, int32_t *argmax_out) {
    DG_REQUIRE(m, "dg_predict_sym8: model is NULL");
    if (!m->kids.empty()) return dgm_predict_sym8(m, probs_out, argmax_out);
    DG_REQUIRE(m->have_raw && m->n > 0, "dg_predict_sym8: stage records with dg_model_set_batch_raw first");
    dg_ctx *c = m->ctx;
    DG_CUDA(cudaSetDevice(c->dev));
    float *acc = nullptr;
    if (int rc = dg_ctx_scratch(c, (size_t)m->n * DG_POINTS * 4, (void **)&acc)) return rc;
    const bool had_targets = m->have_targets;
    m->have_targets = false;                                   // pure inference: no NLL / target lookups
    // the eight passes overwrite the staged symmetry ids: keep the caller's and put them back afterwards
    DG_CUDA(cudaMemcpyAsync(m->d_sym_keep, m->d_sym, m->n, cudaMemcpyDeviceToDevice, c->stream));
    auto restore = [&]() { m->have_targets = had_targets; cudaMemcpyAsync(m->d_sym, m->d_sym_keep, m->n, cudaMemcpyDeviceToDevice, c->stream); };
    for (int s = 0; s < 8; ++s) {
        fill_u8_kernel<<<(m->n + 255) / 256, 256, 0, c->stream>>>(m->d_sym, m->n, (uint8_t)s);
        c->launches++;
        if (int rc = dg_model_encode_resident(m)) { restore(); return rc; }
        if (int rc = dg_forward(m, nullptr)) { restore(); return rc; }
        sym_accumulate_kernel<<<m->n, 384, 0, c->stream>>>(m->d_logp, acc, s, s == 0);
        c->launches++;
    }
    restore();
    if (int rc = dg_model_encode_resident(m)) return rc;       // staged planes and labels are the caller's again (they were those of symmetry 7)
    row_argmax_kernel<<<(m->n + 3) / 4, 128, 0, c->stream>>>(acc, m->d_argmax, m->n);
    c->launches++;
    DG_CUDA(cudaGetLastError());
    if (probs_out) DG_CUDA(cudaMemcpyAsync(probs_out, acc, (size_t)m->n * DG_POINTS * 4, cudaMemcpyDeviceToHost, c->stream));
    if (argmax_out) DG_CUDA(cudaMemcpyAsync(argmax_out, m->d_argmax, (size_t)m->n * 4, cudaMemcpyDeviceToHost, c->stream));
    DG_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
