// C ABI of libdeepgo_b200.so (include/deepgo.h): context, model container, staging, and the
// step orchestration (eval() of train.lua:4-12 + SGD:step of optimizer.lua:24-27).
#include "model.cuh"
#include "../../include/deepgo.h"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>

static thread_local char g_err[1024] = "";

// multi.cu: one process, several devices
int dgm_ctx_destroy_kids(dg_ctx *c);
int dgm_ctx_synchronize(dg_ctx *c);
int dgm_model_create(dg_ctx *c, int num_layers, int channel_size, int in_planes, int k_first, int precision, int max_batch, dg_model **out);
int dgm_model_destroy(dg_model *m);
int dgm_set_params(dg_model *m, const float *host, size_t n);
int dgm_zero_grads(dg_model *m);
int dgm_set_batch_raw(dg_model *m, const uint8_t *raw, const uint8_t *player, const uint8_t *rank, const uint8_t *sym, const uint8_t *x, const uint8_t *y, int n);
int dgm_set_batch_planes(dg_model *m, const float *planes, const int32_t *targets0, int n);
int dgm_encode_resident(dg_model *m);
int dgm_forward(dg_model *m, float *logprobs_out);
int dgm_argmax(dg_model *m, int32_t *out);
int dgm_eval_batch(dg_model *m, int *errors_out, float *cost_out);
int dgm_loss_backward(dg_model *m, float *loss_out);
int dgm_sgd_step(dg_model *m, float rate);
int dgm_train_step(dg_model *m, float rate, float *loss_out);
#define DG_MULTI(m) (!(m)->kids.empty())

void dg_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

extern "C" const char *dg_last_error(void) { return g_err; }
extern "C" int dg_version(void) { return 100; }

extern "C" int dg_device_count(int *out) {
    DG_REQUIRE(out, "dg_device_count: out is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { n = 0; cudaGetLastError(); }
    *out = n;
    return 0;
}

extern "C" int dg_ctx_create(int device, dg_ctx **out) {
    DG_REQUIRE(out, "dg_ctx_create: out is NULL");
    int ndev = 0;
    DG_CUDA(cudaGetDeviceCount(&ndev));
    DG_REQUIRE(device >= 0 && device < ndev, "dg_ctx_create: device %d out of range (%d visible); there is no CPU fallback", device, ndev);
    cudaDeviceProp prop;
    DG_CUDA(cudaGetDeviceProperties(&prop, device));
    DG_REQUIRE(prop.major == 10, "dg_ctx_create: device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor);
    DG_CUDA(cudaSetDevice(device));
    dg_ctx *c = new dg_ctx();
    c->dev = device;
    DG_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    if (getenv("DG_TL")) {
        unsigned long long h[3 * DG_TL_COUNT];
        for (int i = 0; i < 3 * DG_TL_COUNT; ++i) h[i] = (i % 3 == 2) ? 0ull : ~0ull;
        DG_CUDA(cudaMalloc((void **)&c->tl, sizeof(h)));
        DG_CUDA(cudaMemcpy(c->tl, h, sizeof(h), cudaMemcpyHostToDevice));
    }
    DG_CUDA(cudaEventCreate(&c->t0));
    DG_CUDA(cudaEventCreate(&c->t1));
    *out = c;
    return 0;
}

extern "C" int dg_ctx_destroy(dg_ctx *c) {
    if (!c) return 0;
    if (c->live_models > 0) { c->zombie = true; return 0; }   // a model still uses this context: free it with the last model
    if (!c->kids.empty()) dgm_ctx_destroy_kids(c);
    cudaSetDevice(c->dev);
    cudaStreamSynchronize(c->stream);
    if (c->scratch) cudaFree(c->scratch);
    if (c->flush_buf) cudaFree(c->flush_buf);
    cudaEventDestroy(c->t0);
    cudaEventDestroy(c->t1);
    cudaStreamDestroy(c->stream);
    delete c;
    return 0;
}

extern "C" int dg_ctx_synchronize(dg_ctx *c) {
    DG_REQUIRE(c, "dg_ctx_synchronize: ctx is NULL");
    if (!c->kids.empty()) return dgm_ctx_synchronize(c);
    DG_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int dg_ctx_launch_count(dg_ctx *c, long long *out) {
    DG_REQUIRE(c && out, "dg_ctx_launch_count: NULL argument");
    *out = c->launches;
    for (dg_ctx *k : c->kids) *out += k->launches;
    return 0;
}

extern "C" int dg_timer_start(dg_ctx *c) {
    DG_REQUIRE(c, "dg_timer_start: ctx is NULL");
    if (c->kids.empty()) DG_CUDA(cudaSetDevice(c->dev));
    if (!c->kids.empty()) { for (dg_ctx *k : c->kids) if (int rc = dg_timer_start(k)) return rc; return 0; }
    DG_CUDA(cudaEventRecord(c->t0, c->stream));
    return 0;
}
extern "C" int dg_timer_stop(dg_ctx *c, float *ms) {
    DG_REQUIRE(c && ms, "dg_timer_stop: NULL argument");
    if (!c->kids.empty()) {                          // the slowest device
        *ms = 0.f;
        for (dg_ctx *k : c->kids) { float t = 0.f; if (int rc = dg_timer_stop(k, &t)) return rc; if (t > *ms) *ms = t; }
        return 0;
    }
    DG_CUDA(cudaSetDevice(c->dev));
    DG_CUDA(cudaEventRecord(c->t1, c->stream));
    DG_CUDA(cudaEventSynchronize(c->t1));
    DG_CUDA(cudaEventElapsedTime(ms, c->t0, c->t1));
    return 0;
}

extern "C" int dg_prof_begin(dg_ctx *c) {
    DG_REQUIRE(c, "dg_prof_begin: ctx is NULL");
    for (int k = 0; k < DG_K_COUNT; ++k) { c->prof_ms[k] = 0; c->prof_n[k] = 0; }
    c->prof_on = true;
    return 0;
}
extern "C" int dg_prof_end(dg_ctx *c) {
    DG_REQUIRE(c, "dg_prof_end: ctx is NULL");
    c->prof_on = false;
    DG_CUDA(cudaStreamSynchronize(c->stream));
    for (dg_prof_rec &r : c->prof_recs) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) { c->prof_ms[r.kind] += ms; c->prof_n[r.kind]++; }
        cudaEventDestroy(r.e0); cudaEventDestroy(r.e1);
    }
    c->prof_recs.clear();
    return 0;
}
extern "C" int dg_prof_get(dg_ctx *c, int kind, float *ms_total, long long *launches) {
    DG_REQUIRE(c && kind >= 0 && kind < DG_K_COUNT, "dg_prof_get: bad argument");
    if (ms_total) *ms_total = (float)c->prof_ms[kind];
    if (launches) *launches = c->prof_n[kind];
    return 0;
}

extern "C" int dg_prof_timeline(dg_ctx *c, unsigned long long *out, int n) {
    DG_REQUIRE(c && out && n >= 0, "dg_prof_timeline: bad argument");
    DG_REQUIRE(c->tl, "dg_prof_timeline: start the process with DG_TL=1");
    DG_CUDA(cudaSetDevice(c->dev));
    DG_CUDA(cudaStreamSynchronize(c->stream));
    unsigned long long h[3 * DG_TL_COUNT];
    DG_CUDA(cudaMemcpy(h, c->tl, sizeof(h), cudaMemcpyDeviceToHost));
    for (int i = 0; i < n && i < 3 * DG_TL_COUNT; ++i) out[i] = h[i];
    for (int i = 0; i < 3 * DG_TL_COUNT; ++i) h[i] = (i % 3 == 2) ? 0ull : ~0ull;
    DG_CUDA(cudaMemcpy(c->tl, h, sizeof(h), cudaMemcpyHostToDevice));
    return 0;
}

extern "C" int dg_ctx_flush_l2(dg_ctx *c) {
    DG_REQUIRE(c, "dg_ctx_flush_l2: ctx is NULL");
    if (!c->kids.empty()) { for (dg_ctx *k : c->kids) if (int rc = dg_ctx_flush_l2(k)) return rc; return 0; }
    DG_CUDA(cudaSetDevice(c->dev));
    if (!c->flush_buf) {
        c->flush_bytes = (size_t)256 << 20;       // > 126 MB L2
        DG_CUDA(cudaMalloc(&c->flush_buf, c->flush_bytes));
    }
    return dg_launch_fill(c, (float *)c->flush_buf, c->flush_bytes / 4, 0.f);
}

int dg_ctx_scratch(dg_ctx *c, size_t bytes, void **out) {
    if (bytes > c->scratch_bytes) {
        DG_CUDA(cudaStreamSynchronize(c->stream));
        if (c->scratch) DG_CUDA(cudaFree(c->scratch));
        c->scratch = nullptr;
        c->scratch_bytes = 0;
        DG_CUDA(cudaMalloc(&c->scratch, bytes));
        c->scratch_bytes = bytes;
    }
    *out = c->scratch;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// encode (stand-alone, host buffers)
// ---------------------------------------------------------------------------------------------
static int check_records(const uint8_t *player, const uint8_t *rank, const uint8_t *sym, int n) {
    for (int i = 0; i < n; ++i) {
        DG_REQUIRE(player[i] == 1 || player[i] == 2, "record %d: player %d not in {1,2}", i, player[i]);
        // dataloader.lua:87 indexes input[RANK+rank]: rank outside 1..9 is out of range there
        DG_REQUIRE(rank[i] >= 1 && rank[i] <= 9, "record %d: rank %d outside 1..9 (dataloader.lua:87 would index out of range)", i, rank[i]);
        DG_REQUIRE(!sym || sym[i] < 8, "record %d: symmetry id %d outside 0..7", i, sym[i]);
    }
    return 0;
}

extern "C" int dg_encode(dg_ctx *c, const uint8_t *raw, const uint8_t *player, const uint8_t *rank, const uint8_t *sym,
                         int n, void *out, int outfmt) {
    DG_REQUIRE(c, "dg_encode: ctx is NULL");
    DG_REQUIRE(n >= 0, "dg_encode: n < 0");
    if (n == 0) return 0;
    DG_REQUIRE(raw && player && rank && out, "dg_encode: NULL buffer");
    DG_REQUIRE(outfmt == DG_FMT_F32_NCHW || outfmt == DG_FMT_BITPLANES, "dg_encode: unknown output format %d", outfmt);
    if (int rc = check_records(player, rank, sym, n)) return rc;
    DG_CUDA(cudaSetDevice(c->dev));
    const size_t raw_b = (size_t)n * DG_RAW_PLANES * DG_POINTS;
    const size_t meta_off = (raw_b + 255) / 256 * 256;
    const size_t out_off = (meta_off + 3 * (size_t)n + 255) / 256 * 256;
    const size_t out_b = outfmt == DG_FMT_F32_NCHW ? (size_t)n * DG_PLANES * DG_POINTS * 4 : (size_t)n * DG_PLANES * 12 * 4;
    char *d = nullptr;
    if (int rc = dg_ctx_scratch(c, out_off + out_b, (void **)&d)) return rc;
    uint8_t *d_raw = (uint8_t *)d, *d_pl = (uint8_t *)d + meta_off, *d_rk = d_pl + n, *d_sym = d_rk + n;
    DG_CUDA(cudaMemcpyAsync(d_raw, raw, raw_b, cudaMemcpyHostToDevice, c->stream));
    DG_CUDA(cudaMemcpyAsync(d_pl, player, n, cudaMemcpyHostToDevice, c->stream));
    DG_CUDA(cudaMemcpyAsync(d_rk, rank, n, cudaMemcpyHostToDevice, c->stream));
    if (sym) DG_CUDA(cudaMemcpyAsync(d_sym, sym, n, cudaMemcpyHostToDevice, c->stream));
    int rc = outfmt == DG_FMT_F32_NCHW ? dg_launch_encode_f32(c, d_raw, d_pl, d_rk, sym ? d_sym : nullptr, n, (float *)(d + out_off))
                                       : dg_launch_encode_bits(c, d_raw, d_pl, d_rk, sym ? d_sym : nullptr, n, (uint32_t *)(d + out_off));
    if (rc) return rc;
    DG_CUDA(cudaMemcpyAsync(out, d + out_off, out_b, cudaMemcpyDeviceToHost, c->stream));
    DG_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int dg_encode_labels(const uint8_t *x, const uint8_t *y, const uint8_t *sym, int n, int32_t *out) {
    DG_REQUIRE(n >= 0, "dg_encode_labels: n < 0");
    if (n == 0) return 0;
    DG_REQUIRE(x && y && out, "dg_encode_labels: NULL buffer");
    for (int i = 0; i < n; ++i) {
        DG_REQUIRE(x[i] >= 1 && x[i] <= 19 && y[i] >= 1 && y[i] <= 19, "move %d: (%d,%d) outside the board", i, x[i], y[i]);
        DG_REQUIRE(!sym || sym[i] < 8, "move %d: symmetry id %d outside 0..7", i, sym[i]);
        int a = x[i] - 1, b = y[i] - 1;
        const int s = sym ? sym[i] : 0;
        if (s & 1) { int t = a; a = b; b = t; }
        if (s & 2) a = 18 - a;
        if (s & 4) b = 18 - b;
        out[i] = 19 * a + b;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// model
// ---------------------------------------------------------------------------------------------
// layer table and flat-vector offsets (experiments.lua:88-93,105-107)
int dg_model_layout(dg_model *m, int num_layers, int channel_size, int in_planes, int k_first) {
    m->L = num_layers; m->C = channel_size; m->in_planes = in_planes; m->k_first = k_first;
    m->layers.clear();
    size_t off = 0;
    for (int l = 0; l < num_layers; ++l) {
        LayerInfo li;
        li.cin = l == 0 ? in_planes : channel_size;
        li.cout = l == num_layers - 1 ? 1 : channel_size;
        li.k = l == 0 ? k_first : 3;
        li.nW = (size_t)li.cout * li.cin * li.k * li.k;
        li.offW = off; off += li.nW;
        li.offB = off; off += li.cout;
        li.offA = off; off += (size_t)li.cout * DG_POINTS;
        m->layers.push_back(li);
    }
    m->P = off;
    return 0;
}

// d_raw, d_player .. point into record buffer b
static void select_record_buffer(dg_model *m, int b) {
    const size_t B = (size_t)m->max_batch;
    m->stage_cur = b;
    m->d_raw = m->d_raw_buf[b];
    m->d_player = m->d_meta_buf[b];
    m->d_rank = m->d_player + B; m->d_sym = m->d_rank + B; m->d_x = m->d_sym + B; m->d_y = m->d_x + B; m->d_sym_keep = m->d_y + B;
}
// call after queueing an encode that reads the current record buffer: the next copy into that buffer waits for it
int dg_model_mark_encoded(dg_model *m) {
    DG_CUDA(cudaEventRecord(m->ev_enc[m->stage_cur], m->ctx->stream));
    m->enc_recorded[m->stage_cur] = true;
    return 0;
}

extern "C" int dg_model_create(dg_ctx *c, int num_layers, int channel_size, int in_planes, int k_first, int precision,
                               int max_batch, dg_model **out) {
    DG_REQUIRE(c && out, "dg_model_create: NULL argument");
    DG_REQUIRE(num_layers >= 2, "dg_model_create: numLayers must be >= 2");
    DG_REQUIRE(channel_size >= 1 && max_batch >= 1, "dg_model_create: bad channelSize/max_batch");
    DG_REQUIRE(k_first == 3 || k_first == 5, "dg_model_create: first kernel must be 3 or 5");
    DG_REQUIRE(in_planes == DG_PLANES, "dg_model_create: in_planes must be %d (dataloader.lua TOTAL)", DG_PLANES);
    DG_REQUIRE(precision == DG_PREC_FP32 || precision == DG_PREC_BF16 || precision == DG_PREC_TF32, "dg_model_create: unknown precision %d", precision);
    if (!c->kids.empty()) return dgm_model_create(c, num_layers, channel_size, in_planes, k_first, precision, max_batch, out);
    DG_CUDA(cudaSetDevice(c->dev));
    dg_model *m = new dg_model();
    m->ctx = c; m->L = num_layers; m->C = channel_size; m->in_planes = in_planes; m->k_first = k_first;
    m->prec = precision; m->max_batch = max_batch;
    if (int rc = dg_model_layout(m, num_layers, channel_size, in_planes, k_first)) return rc;
    c->live_models++;
    const size_t B = max_batch;
    DG_CUDA(cudaMalloc(&m->params, m->P * 4));
    DG_CUDA(cudaMalloc(&m->grads, m->P * 4));
    DG_CUDA(cudaMemsetAsync(m->params, 0, m->P * 4, c->stream));
    DG_CUDA(cudaMemsetAsync(m->grads, 0, m->P * 4, c->stream));
    for (int i = 0; i < 2; ++i) {
        DG_CUDA(cudaMalloc(&m->d_raw_buf[i], B * DG_RAW_PLANES * DG_POINTS));
        DG_CUDA(cudaMalloc(&m->d_meta_buf[i], B * 6));
        DG_CUDA(cudaEventCreateWithFlags(&m->ev_copy[i], cudaEventDisableTiming));
        DG_CUDA(cudaEventCreateWithFlags(&m->ev_enc[i], cudaEventDisableTiming));
    }
    DG_CUDA(cudaStreamCreateWithFlags(&m->copy_stream, cudaStreamNonBlocking));
    DG_CUDA(cudaMallocHost(&m->h_loss_ring, 4 * 4));
    for (int i = 0; i < 4; ++i) DG_CUDA(cudaEventCreateWithFlags(&m->ev_loss[i], cudaEventDisableTiming));
    select_record_buffer(m, 0);
    DG_CUDA(cudaMalloc(&m->d_targets, B * 4));
    DG_CUDA(cudaMalloc(&m->d_argmax, B * 4));
    DG_CUDA(cudaMalloc(&m->d_logp, B * DG_POINTS * 4));
    DG_CUDA(cudaMalloc(&m->d_loss_n, B * 4));
    DG_CUDA(cudaMalloc(&m->d_loss, 4));
    DG_CUDA(cudaMalloc(&m->d_errors, 4));
    DG_CUDA(cudaMalloc(&m->d_rate, 4));
    DG_CUDA(cudaMallocHost(&m->h_rate, 8 * 4));
    DG_CUDA(cudaMallocHost(&m->h_res, 2 * 4));
    for (int i = 0; i < 8; ++i) DG_CUDA(cudaEventCreateWithFlags(&m->ev_rate[i], cudaEventDisableTiming));
    m->act.assign(num_layers + 1, nullptr);
    DG_CUDA(cudaMalloc(&m->act[0], B * in_planes * DG_POINTS * 4));
    if (precision == DG_PREC_FP32) {
        for (int l = 0; l < num_layers; ++l) DG_CUDA(cudaMalloc(&m->act[l + 1], B * m->layers[l].cout * DG_POINTS * 4));
        for (int i = 0; i < 2; ++i) DG_CUDA(cudaMalloc(&m->dz[i], B * channel_size * DG_POINTS * 4));
    } else if (precision == DG_PREC_TF32) {
        // the fp32 head works on NCHW maps: h_{L-1}, the 1-channel output map, and the two gradient maps around the head layer
        for (int l = num_layers - 1; l <= num_layers; ++l) DG_CUDA(cudaMalloc(&m->act[l], B * m->layers[l - 1].cout * DG_POINTS * 4));
        for (int i = 0; i < 2; ++i) DG_CUDA(cudaMalloc(&m->dz[i], B * channel_size * DG_POINTS * 4));
        if (int rc = dg_tf32_create(m)) { return rc; }
    } else {
        if (int rc = dg_bf16_create(m)) { return rc; }
    }
    *out = m;
    return 0;
}

extern "C" int dg_model_destroy(dg_model *m) {
    if (!m) return 0;
    if (DG_MULTI(m)) return dgm_model_destroy(m);
    cudaSetDevice(m->ctx->dev);
    cudaStreamSynchronize(m->ctx->stream);
    if (m->bf) dg_bf16_destroy(m);
    if (m->t32) dg_tf32_destroy(m);
    for (float *p : m->act) if (p) cudaFree(p);
    for (int i = 0; i < 2; ++i) if (m->dz[i]) cudaFree(m->dz[i]);
    cudaFree(m->params); cudaFree(m->grads);
    for (int i = 0; i < 2; ++i) { cudaFree(m->d_raw_buf[i]); cudaFree(m->d_meta_buf[i]); if (m->ev_copy[i]) cudaEventDestroy(m->ev_copy[i]); if (m->ev_enc[i]) cudaEventDestroy(m->ev_enc[i]); }
    if (m->copy_stream) { cudaStreamSynchronize(m->copy_stream); cudaStreamDestroy(m->copy_stream); }
    if (m->h_loss_ring) cudaFreeHost(m->h_loss_ring);
    for (int i = 0; i < 4; ++i) if (m->ev_loss[i]) cudaEventDestroy(m->ev_loss[i]);
    if (m->h_meta) { cudaFreeHost(m->h_meta); for (int i = 0; i < 4; ++i) cudaEventDestroy(m->ev_meta[i]); }
    cudaFree(m->d_targets); cudaFree(m->d_argmax); cudaFree(m->d_logp); cudaFree(m->d_loss_n);
    cudaFree(m->d_loss); cudaFree(m->d_errors); cudaFree(m->d_rate); cudaFreeHost(m->h_rate); cudaFreeHost(m->h_res);
    if (m->step_graph) cudaGraphExecDestroy(m->step_graph);
    for (int i = 0; i < 8; ++i) if (m->ev_rate[i]) cudaEventDestroy(m->ev_rate[i]);
    dg_ctx *c = m->ctx;
    delete m;
    if (--c->live_models == 0 && c->zombie) { c->zombie = false; return dg_ctx_destroy(c); }
    return 0;
}

extern "C" int dg_model_param_count(dg_model *m, size_t *out) {
    DG_REQUIRE(m && out, "dg_model_param_count: NULL argument");
    *out = m->P;
    return 0;
}

extern "C" int dg_model_set_params(dg_model *m, const float *host, size_t n) {
    DG_REQUIRE(m && host, "dg_model_set_params: NULL argument");
    DG_REQUIRE(n == m->P, "dg_model_set_params: got %zu values, the model has %zu parameters", n, m->P);
    if (DG_MULTI(m)) return dgm_set_params(m, host, n);
    DG_CUDA(cudaSetDevice(m->ctx->dev));
    DG_CUDA(cudaMemcpyAsync(m->params, host, n * 4, cudaMemcpyHostToDevice, m->ctx->stream));
    if (m->bf) if (int rc = dg_bf16_params_changed(m)) return rc;
    if (m->t32) if (int rc = dg_tf32_params_changed(m)) return rc;
    DG_CUDA(cudaStreamSynchronize(m->ctx->stream));
    return 0;
}
extern "C" int dg_model_get_params(dg_model *m, float *host, size_t n) {
    DG_REQUIRE(m && host, "dg_model_get_params: NULL argument");
    DG_REQUIRE(n == m->P, "dg_model_get_params: buffer holds %zu values, the model has %zu parameters", n, m->P);
    if (DG_MULTI(m)) return dg_model_get_params(m->kids[0], host, n);        // the replicas are identical
    DG_CUDA(cudaSetDevice(m->ctx->dev));
    DG_CUDA(cudaMemcpyAsync(host, m->params, n * 4, cudaMemcpyDeviceToHost, m->ctx->stream));
    DG_CUDA(cudaStreamSynchronize(m->ctx->stream));
    return 0;
}
extern "C" int dg_model_get_grads(dg_model *m, float *host, size_t n) {
    DG_REQUIRE(m && host, "dg_model_get_grads: NULL argument");
    DG_REQUIRE(n == m->P, "dg_model_get_grads: buffer holds %zu values, the model has %zu parameters", n, m->P);
    if (DG_MULTI(m)) return dg_model_get_grads(m->kids[0], host, n);         // every replica holds the summed gradient
    DG_CUDA(cudaSetDevice(m->ctx->dev));
    if (m->bf) if (int rc = dg_bf16_sync_grads(m)) return rc;
    DG_CUDA(cudaMemcpyAsync(host, m->grads, n * 4, cudaMemcpyDeviceToHost, m->ctx->stream));
    DG_CUDA(cudaStreamSynchronize(m->ctx->stream));
    return 0;
}
extern "C" int dg_zero_grads(dg_model *m) {
    DG_REQUIRE(m, "dg_zero_grads: model is NULL");
    if (DG_MULTI(m)) return dgm_zero_grads(m);
    DG_CUDA(cudaSetDevice(m->ctx->dev));
    if (m->bf) dg_bf16_drop_pending_grads(m);
    DG_CUDA(cudaMemsetAsync(m->grads, 0, m->P * 4, m->ctx->stream));
    return 0;
}
extern "C" int dg_model_set_global_batch(dg_model *m, int gb) {
    DG_REQUIRE(m && gb >= 0, "dg_model_set_global_batch: bad argument");
    if (DG_MULTI(m)) { m->global_batch = gb; return 0; }                      // handed to the replicas when a batch is staged
    if (gb != m->global_batch && m->step_graph) {
        // the captured step holds 1/global_batch as a by-value kernel argument: capture again
        DG_CUDA(cudaSetDevice(m->ctx->dev));
        DG_CUDA(cudaStreamSynchronize(m->ctx->stream));
        cudaGraphExecDestroy(m->step_graph); m->step_graph = nullptr;
    }
    m->global_batch = gb;
    return 0;
}

extern "C" int dg_model_set_batch_raw(dg_model *m, const uint8_t *raw, const uint8_t *player, const uint8_t *rank,
                                      const uint8_t *sym, const uint8_t *x, const uint8_t *y, int n) {
    DG_REQUIRE(m, "dg_model_set_batch_raw: model is NULL");
    if (DG_MULTI(m)) {
        DG_REQUIRE(raw && player && rank, "dg_model_set_batch_raw: NULL buffer");
        DG_REQUIRE((x == nullptr) == (y == nullptr), "dg_model_set_batch_raw: x and y must both be given or both NULL");
        return dgm_set_batch_raw(m, raw, player, rank, sym, x, y, n);
    }
    DG_REQUIRE(n >= 1 && n <= m->max_batch, "dg_model_set_batch_raw: batch %d outside 1..%d", n, m->max_batch);
    DG_REQUIRE(raw && player && rank, "dg_model_set_batch_raw: NULL buffer");
    DG_REQUIRE((x == nullptr) == (y == nullptr), "dg_model_set_batch_raw: x and y must both be given or both NULL");
    if (int rc = check_records(player, rank, sym, n)) return rc;
    dg_ctx *c = m->ctx;
    DG_CUDA(cudaSetDevice(c->dev));
    // The records go into the buffer the previous minibatch did NOT use, on the copy stream: a step that is still running on the
    // context's stream (dg_train_step_begin) keeps its records, and the copy runs under it.  The copy waits for the last encode
    // that read this buffer; the context's stream waits for the copy.  (With DG_FUSED_ENCODE the step graph itself reads the
    // records through pointers fixed at capture: one buffer, the context's stream.)
    const bool two = !dg_bf16_input_mode_fused();
    cudaStream_t cs = two ? m->copy_stream : c->stream;
    if (two) {
        select_record_buffer(m, m->stage_cur ^ 1);
        if (m->enc_recorded[m->stage_cur]) DG_CUDA(cudaStreamWaitEvent(cs, m->ev_enc[m->stage_cur], 0));
    }
    DG_CUDA(cudaMemcpyAsync(m->d_raw, raw, (size_t)n * DG_RAW_PLANES * DG_POINTS, cudaMemcpyHostToDevice, cs));
    m->n = n; m->have_raw = true; m->forward_done = false;
    m->have_targets = x != nullptr;
    if (x)
        for (int i = 0; i < n; ++i)
            DG_REQUIRE(x[i] >= 1 && x[i] <= 19 && y[i] >= 1 && y[i] <= 19, "move %d: (%d,%d) outside the board", i, x[i], y[i]);
    {
        // player / rank / sym / x / y are rows of one device block: gather them in a pinned staging slot and send ONE copy instead of
        // five (each small copy costs a few microseconds of stream time).  Four slots, an event each: a slot is reused only after its
        // copy has completed.
        const size_t B = (size_t)m->max_batch;
        if (!m->h_meta) {
            DG_CUDA(cudaMallocHost((void **)&m->h_meta, 4 * 5 * B));
            for (int i = 0; i < 4; ++i) DG_CUDA(cudaEventCreateWithFlags(&m->ev_meta[i], cudaEventDisableTiming));
        }
        const int slot = m->meta_slot++ & 3;
        DG_CUDA(cudaEventSynchronize(m->ev_meta[slot]));
        uint8_t *h = m->h_meta + (size_t)slot * 5 * B;
        memcpy(h, player, n); memcpy(h + B, rank, n);
        if (sym) memcpy(h + 2 * B, sym, n); else memset(h + 2 * B, 0, n);
        if (x) { memcpy(h + 3 * B, x, n); memcpy(h + 4 * B, y, n); }
        DG_CUDA(cudaMemcpyAsync(m->d_player, h, x ? 4 * B + n : 2 * B + n, cudaMemcpyHostToDevice, cs));
        DG_CUDA(cudaEventRecord(m->ev_meta[slot], cs));
    }
    if (two) {
        DG_CUDA(cudaEventRecord(m->ev_copy[m->stage_cur], cs));
        DG_CUDA(cudaStreamWaitEvent(c->stream, m->ev_copy[m->stage_cur], 0));
    }
    if (x) {
        // the bf16 path's encode kernel computes the labels as well (one launch less per batch)
        if (m->prec != DG_PREC_BF16 || dg_bf16_input_mode_fused()) if (int rc = dg_launch_labels(c, m->d_x, m->d_y, m->d_sym, n, m->d_targets)) return rc;
    }
    if (int rc = m->prec != DG_PREC_BF16 ? dg_launch_encode_f32(c, m->d_raw, m->d_player, m->d_rank, m->d_sym, n, m->act[0]) : dg_bf16_stage_raw(m)) return rc;
    return dg_model_mark_encoded(m);
}

extern "C" int dg_model_encode_resident(dg_model *m) {
    DG_REQUIRE(m, "dg_model_encode_resident: model is NULL");
    if (DG_MULTI(m)) return dgm_encode_resident(m);
    DG_REQUIRE(m->have_raw && m->n > 0, "dg_model_encode_resident: no raw records staged (dg_model_set_batch_raw)");
    dg_ctx *c = m->ctx;
    DG_CUDA(cudaSetDevice(c->dev));
    m->forward_done = false;
    if (m->have_targets && (m->prec != DG_PREC_BF16 || dg_bf16_input_mode_fused())) if (int rc = dg_launch_labels(c, m->d_x, m->d_y, m->d_sym, m->n, m->d_targets)) return rc;
    if (int rc = m->prec != DG_PREC_BF16 ? dg_launch_encode_f32(c, m->d_raw, m->d_player, m->d_rank, m->d_sym, m->n, m->act[0]) : dg_bf16_stage_raw(m)) return rc;
    return dg_model_mark_encoded(m);
}

extern "C" int dg_model_set_batch_planes(dg_model *m, const float *planes, const int32_t *targets0, int n) {
    DG_REQUIRE(m && planes, "dg_model_set_batch_planes: NULL argument");
    if (DG_MULTI(m)) return dgm_set_batch_planes(m, planes, targets0, n);
    DG_REQUIRE(n >= 1 && n <= m->max_batch, "dg_model_set_batch_planes: batch %d outside 1..%d", n, m->max_batch);
    dg_ctx *c = m->ctx;
    DG_CUDA(cudaSetDevice(c->dev));
    if (targets0) {
        for (int i = 0; i < n; ++i)
            DG_REQUIRE(targets0[i] >= 0 && targets0[i] < DG_POINTS, "target %d: class %d outside 0..360", i, targets0[i]);
        DG_CUDA(cudaMemcpyAsync(m->d_targets, targets0, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    }
    DG_CUDA(cudaMemcpyAsync(m->act[0], planes, (size_t)n * m->in_planes * DG_POINTS * 4, cudaMemcpyHostToDevice, c->stream));
    m->n = n; m->have_raw = false; m->have_targets = targets0 != nullptr; m->forward_done = false;
    if (m->prec == DG_PREC_BF16) return dg_bf16_stage_planes(m);
    return 0;
}

static float grad_scale(dg_model *m) { return 1.0f / (float)(m->global_batch > 0 ? m->global_batch : m->n); }

// forward up to the ReLU'd last-layer map, then log-softmax (+ per-sample NLL, argmax)
int dg_forward_impl(dg_model *m) {
    dg_ctx *c = m->ctx;
    DG_REQUIRE(m->n > 0, "forward: no batch staged (call dg_model_set_batch_* first)");
    if (m->prec == DG_PREC_FP32) {
        for (int l = 0; l < m->L; ++l) {
            const LayerInfo &li = m->layers[l];
            if (int rc = dg_simt_conv(c, m->act[l], m->params + li.offW, m->params + li.offB, m->params + li.offA, nullptr,
                                      m->act[l + 1], m->n, li.cin, li.cout, li.k, 0, 1)) return rc;
        }
        if (int rc = dg_launch_logsoftmax_nll(c, m->act[m->L], m->have_targets ? m->d_targets : nullptr, m->d_logp, m->d_loss_n,
                                              nullptr, m->d_argmax, m->n, 0.f)) return rc;
    } else if (m->prec == DG_PREC_TF32) {
        if (int rc = dg_tf32_forward(m)) return rc;
        const LayerInfo &lh = m->layers[m->L - 1];
        if (int rc = dg_simt_conv(c, m->act[m->L - 1], m->params + lh.offW, m->params + lh.offB, m->params + lh.offA, nullptr,
                                  m->act[m->L], m->n, lh.cin, lh.cout, lh.k, 0, 1)) return rc;
        if (int rc = dg_launch_logsoftmax_nll(c, m->act[m->L], m->have_targets ? m->d_targets : nullptr, m->d_logp, m->d_loss_n,
                                              nullptr, m->d_argmax, m->n, 0.f)) return rc;
    } else {
        if (int rc = dg_bf16_forward(m)) return rc;
    }
    m->forward_done = true;
    m->loss_reduced = false;           // a new forward pass: the reduction of the previous one is stale
    return 0;
}

int dg_backward_impl(dg_model *m) {
    dg_ctx *c = m->ctx;
    DG_REQUIRE(m->forward_done, "backward: call dg_forward first");
    DG_REQUIRE(m->have_targets, "backward: the staged batch has no targets");
    const float scale = grad_scale(m);
    if (m->prec == DG_PREC_BF16) return dg_bf16_backward(m, scale);
    if (m->prec == DG_PREC_TF32) {
        // the 1-channel head layer on the fp32 kernels, then the conv stack on the tensor cores
        const LayerInfo &lh = m->layers[m->L - 1];
        if (int rc = dg_launch_logsoftmax_nll(c, m->act[m->L], m->d_targets, m->d_logp, m->d_loss_n, m->dz[1], m->d_argmax, m->n, scale)) return rc;
        if (int rc = dg_simt_bias_grad(c, m->dz[1], m->grads + lh.offA, m->grads + lh.offB, m->n, lh.cout)) return rc;
        if (int rc = dg_simt_wgrad(c, m->act[m->L - 1], m->dz[1], m->grads + lh.offW, m->n, lh.cin, lh.cout, lh.k)) return rc;
        if (int rc = dg_simt_conv(c, m->dz[1], m->params + lh.offW, nullptr, nullptr, m->act[m->L - 1], m->dz[0], m->n, lh.cout, lh.cin, lh.k, 1, 0)) return rc;
        if (int rc = dg_tf32_backward(m, m->dz[0])) return rc;
        if (m->fused_step) if (int rc = dg_dp_bucket(c, m->grads, m->P)) return rc;
        return 0;
    }
    int cur = 0;
    if (int rc = dg_launch_logsoftmax_nll(c, m->act[m->L], m->d_targets, m->d_logp, m->d_loss_n, m->dz[cur], m->d_argmax, m->n, scale)) return rc;
    for (int l = m->L - 1; l >= 0; --l) {
        const LayerInfo &li = m->layers[l];
        if (int rc = dg_simt_bias_grad(c, m->dz[cur], m->grads + li.offA, m->grads + li.offB, m->n, li.cout)) return rc;
        if (int rc = dg_simt_wgrad(c, m->act[l], m->dz[cur], m->grads + li.offW, m->n, li.cin, li.cout, li.k)) return rc;
        if (m->fused_step) if (int rc = dg_dp_bucket(c, m->grads + li.offW, li.nW + li.cout + (size_t)li.cout * DG_POINTS)) return rc;
        if (l > 0) {
            // input gradient through the conv, then the ReLU mask of the layer below (h > 0 <=> z > 0)
            if (int rc = dg_simt_conv(c, m->dz[cur], m->params + li.offW, nullptr, nullptr, m->act[l], m->dz[cur ^ 1], m->n,
                                      li.cout, li.cin, li.k, 1, 0)) return rc;
            cur ^= 1;
        }
    }
    return 0;
}

int dg_read_loss(dg_model *m, float *loss_out, int *errors_out) {
    dg_ctx *c = m->ctx;
    if (!m->loss_reduced) if (int rc = dg_launch_loss_reduce(c, m->d_loss_n, m->d_argmax, m->d_targets, m->n, m->d_loss, m->d_errors)) return rc;
    // through a pinned word pair: a copy into the caller's pageable memory goes through the driver's own staging and blocks twice
    if (loss_out) DG_CUDA(cudaMemcpyAsync(m->h_res, m->d_loss, 4, cudaMemcpyDeviceToHost, c->stream));
    if (errors_out) DG_CUDA(cudaMemcpyAsync(m->h_res + 1, m->d_errors, 4, cudaMemcpyDeviceToHost, c->stream));
    DG_CUDA(cudaStreamSynchronize(c->stream));
    if (loss_out) *loss_out = m->h_res[0];
    if (errors_out) memcpy(errors_out, m->h_res + 1, 4);
    return 0;
}

extern "C" int dg_forward(dg_model *m, float *logprobs_out) {
    DG_REQUIRE(m, "dg_forward: model is NULL");
    if (DG_MULTI(m)) return dgm_forward(m, logprobs_out);
    DG_CUDA(cudaSetDevice(m->ctx->dev));
    if (int rc = dg_forward_impl(m)) return rc;
    if (logprobs_out) {
        DG_CUDA(cudaMemcpyAsync(logprobs_out, m->d_logp, (size_t)m->n * DG_POINTS * 4, cudaMemcpyDeviceToHost, m->ctx->stream));
        DG_CUDA(cudaStreamSynchronize(m->ctx->stream));
    }
    return 0;
}

extern "C" int dg_loss_backward(dg_model *m, float *loss_out) {
    DG_REQUIRE(m, "dg_loss_backward: model is NULL");
    if (DG_MULTI(m)) return dgm_loss_backward(m, loss_out);
    DG_CUDA(cudaSetDevice(m->ctx->dev));
    if (m->bf) if (int rc = dg_bf16_sync_grads(m)) return rc;      // this call accumulates into the gradient vector
    if (int rc = dg_backward_impl(m)) return rc;
    if (loss_out) return dg_read_loss(m, loss_out, nullptr);
    return 0;
}

extern "C" int dg_sgd_step(dg_model *m, float rate) {
    DG_REQUIRE(m, "dg_sgd_step: model is NULL");
    if (DG_MULTI(m)) return dgm_sgd_step(m, rate);
    DG_CUDA(cudaSetDevice(m->ctx->dev));
    if (m->bf) { if (int rc = dg_bf16_sync_grads(m)) return rc; return dg_bf16_sgd_step(m, rate, nullptr); }
    if (int rc = dg_launch_sgd(m->ctx, m->params, m->grads, m->P, rate)) return rc;
    if (m->t32) return dg_tf32_params_changed(m);
    return 0;
}

// one fused step on the context's stream: eval() of train.lua:4-12 followed by SGD:step (train.lua:131)
static int train_step_body(dg_model *m, float rate, bool rate_from_device) {
    dg_ctx *c = m->ctx;
    // a deferring bf16 step keeps the whole gradient in the accumulation buffers (cleared inside the backward pass)
    if (!(m->bf && dg_bf16_defers_grads(m))) DG_CUDA(cudaMemsetAsync(m->grads, 0, m->P * 4, c->stream));      // train.lua:5
    m->fused_step = true;
    int rc = dg_forward_impl(m);                                              // train.lua:6
    if (!rc) rc = dg_backward_impl(m);                                        // train.lua:8-10
    m->fused_step = false;
    if (rc) return rc;
    if (int rc2 = dg_dp_finish(c)) return rc2;                             // buckets were issued layer by layer in backward
    if (m->bf) return dg_bf16_sgd_step(m, rate, rate_from_device ? m->d_rate : nullptr);   // reads the accumulation buffers directly
    return dg_sgd_step(m, rate);                                           // train.lua:131
}

extern "C" int dg_train_step(dg_model *m, float rate, float *loss_out) {
    DG_REQUIRE(m, "dg_train_step: model is NULL");
    if (DG_MULTI(m)) return dgm_train_step(m, rate, loss_out);
    dg_ctx *c = m->ctx;
    DG_CUDA(cudaSetDevice(c->dev));
    DG_REQUIRE(m->have_targets, "dg_train_step: the staged batch has no targets");
    DG_REQUIRE(m->n > 0, "dg_train_step: no batch staged");
    if (m->bf) if (int rc = dg_bf16_dp_prepare(m)) return rc;
    static const bool no_graph = getenv("DG_NO_GRAPH") != nullptr || getenv("DG_TIMELINE") != nullptr;
    // with data parallelism the gradient exchange is captured too: the peer-memory kernel is just another node; the NCCL fallback's
    // communication stream joins the capture through events
    static const bool no_dp_graph = getenv("DG_NO_DP_GRAPH") != nullptr;
    const bool use_graph = m->prec == DG_PREC_BF16 && (c->world == 1 || !no_dp_graph) && !c->prof_on && !no_graph;
    if (use_graph) {
        // the launches and the cross-stream event edges of a step (10 nodes for the chained bf16 net) are captured once per batch size
        // and replayed
        const int rs = m->rate_slot++ & 7;
        DG_CUDA(cudaEventSynchronize(m->ev_rate[rs]));      // the copy that last read this slot is done (8 steps ago)
        m->h_rate[rs] = rate;
        DG_CUDA(cudaMemcpyAsync(m->d_rate, m->h_rate + rs, 4, cudaMemcpyHostToDevice, c->stream));
        DG_CUDA(cudaEventRecord(m->ev_rate[rs], c->stream));
        const int in_mode = dg_bf16_input_mode(m);
        if (!m->step_graph || m->step_graph_n != m->n || m->step_graph_mode != in_mode) {
            if (m->step_graph) { cudaGraphExecDestroy(m->step_graph); m->step_graph = nullptr; }
            const long long l0 = c->launches;
            cudaGraph_t graph = nullptr;
            DG_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
            const int rc = train_step_body(m, rate, true);
            cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
            if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
            DG_REQUIRE(e == cudaSuccess && graph, "dg_train_step: graph capture failed: %s", cudaGetErrorString(e));
            if (const char *dot = getenv("DG_GRAPH_DOT")) cudaGraphDebugDotPrint(graph, dot, 0);     // dev aid: the captured step as graphviz
            e = cudaGraphInstantiate(&m->step_graph, graph, 0);
            cudaGraphDestroy(graph);
            DG_REQUIRE(e == cudaSuccess, "dg_train_step: cudaGraphInstantiate failed: %s", cudaGetErrorString(e));
            m->step_graph_n = m->n;
            m->step_graph_mode = in_mode;
            m->step_graph_launches = c->launches - l0;
            c->launches = l0;
        }
        DG_CUDA(cudaGraphLaunch(m->step_graph, c->stream));
        c->launches += m->step_graph_launches;
        dg_bf16_after_fused_step(m);
        m->forward_done = true;
        m->loss_reduced = m->bf != nullptr;       // the captured backward pass reduces the loss
    } else {
        if (int rc = train_step_body(m, rate, false)) return rc;
    }
    if (loss_out) return dg_read_loss(m, loss_out, nullptr);
    return 0;
}

extern "C" int dg_dp_allreduce_grads(dg_model *m) {
    DG_REQUIRE(m, "dg_dp_allreduce_grads: model is NULL");
    if (DG_MULTI(m)) return 0;                 // dg_loss_backward on a multi-device model has summed the replicas' gradients already
    DG_CUDA(cudaSetDevice(m->ctx->dev));
    if (m->bf) if (int rc = dg_bf16_sync_grads(m)) return rc;
    if (int rc = dg_dp_bucket(m->ctx, m->grads, m->P)) return rc;
    return dg_dp_finish(m->ctx);
}

extern "C" int dg_argmax(dg_model *m, int32_t *out) {
    DG_REQUIRE(m && out, "dg_argmax: NULL argument");
    if (DG_MULTI(m)) return dgm_argmax(m, out);
    DG_REQUIRE(m->forward_done, "dg_argmax: call dg_forward first");
    DG_CUDA(cudaSetDevice(m->ctx->dev));
    DG_CUDA(cudaMemcpyAsync(out, m->d_argmax, (size_t)m->n * 4, cudaMemcpyDeviceToHost, m->ctx->stream));
    DG_CUDA(cudaStreamSynchronize(m->ctx->stream));
    return 0;
}

extern "C" int dg_eval_batch(dg_model *m, int *errors_out, float *cost_out) {
    DG_REQUIRE(m, "dg_eval_batch: model is NULL");
    if (DG_MULTI(m)) return dgm_eval_batch(m, errors_out, cost_out);
    DG_REQUIRE(m->have_targets, "dg_eval_batch: the staged batch has no targets");
    DG_CUDA(cudaSetDevice(m->ctx->dev));
    if (int rc = dg_forward_impl(m)) return rc;
    return dg_read_loss(m, cost_out, errors_out);
}

// ---------------------------------------------------------------------------------------------
// a step without the wait: begin queues eval() + optimizer:step and the read-back of the cost, result waits for THAT step
// ---------------------------------------------------------------------------------------------
int dgm_train_step_begin(dg_model *m, float rate);
int dgm_train_step_result(dg_model *m, float *loss_out);

extern "C" int dg_train_step_begin(dg_model *m, float rate) {
    DG_REQUIRE(m, "dg_train_step_begin: model is NULL");
    if (DG_MULTI(m)) return dgm_train_step_begin(m, rate);
    DG_REQUIRE(m->loss_pending < 0, "dg_train_step_begin: the previous step's result has not been collected (dg_train_step_result)");
    if (int rc = dg_train_step(m, rate, nullptr)) return rc;
    dg_ctx *c = m->ctx;
    if (!m->loss_reduced) {
        if (int rc = dg_launch_loss_reduce(c, m->d_loss_n, m->d_argmax, m->d_targets, m->n, m->d_loss, m->d_errors)) return rc;
        m->loss_reduced = true;
    }
    const int slot = m->loss_slot++ & 3;
    DG_CUDA(cudaMemcpyAsync(m->h_loss_ring + slot, m->d_loss, 4, cudaMemcpyDeviceToHost, c->stream));
    DG_CUDA(cudaEventRecord(m->ev_loss[slot], c->stream));
    m->loss_pending = slot;
    return 0;
}

extern "C" int dg_train_step_result(dg_model *m, float *loss_out) {
    DG_REQUIRE(m && loss_out, "dg_train_step_result: NULL argument");
    if (DG_MULTI(m)) return dgm_train_step_result(m, loss_out);
    DG_REQUIRE(m->loss_pending >= 0, "dg_train_step_result: no step has been begun (dg_train_step_begin)");
    DG_CUDA(cudaEventSynchronize(m->ev_loss[m->loss_pending]));
    *loss_out = m->h_loss_ring[m->loss_pending];
    m->loss_pending = -1;
    return 0;
}
