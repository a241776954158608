// Several GPUs inside ONE process: the reference's own data-parallel option (experiments.lua:10,97-103,155-168:
// numGPUs = n wraps the model in nn.DataParallelTable(1), which splits every batch on dim 1 (the batch dimension) over the
// GPUs, runs the replicas and sums their gradients).  dg_ctx_create_multi returns a context that stands for n devices; a
// model created on it is n replicas, one per device, each with its own stream.  Every call is asynchronous on the
// replicas' streams, so one caller thread drives all devices: no worker threads, no NCCL, no IPC -- the replicas reach
// each other's gradient buffers through peer access (dp.cu: dg_dp_p2p_setup_local) and dg_train_step runs the same
// captured step graph with the same in-graph exchange kernel as the process-per-GPU mode.
#include "model.cuh"
#include "../../include/deepgo.h"
#include <cstring>

// DataParallelTable's split of n positions over `world` replicas: near-equal contiguous chunks, the first n % world one longer
static void shard(int n, int r, int world, int *lo, int *cnt) {
    const int base = n / world, rem = n % world;
    *lo = r * base + (r < rem ? r : rem);
    *cnt = base + (r < rem ? 1 : 0);
}

extern "C" int dg_ctx_create_multi(int ndev, const int *devs, dg_ctx **out) {
    DG_REQUIRE(out && devs, "dg_ctx_create_multi: NULL argument");
    DG_REQUIRE(ndev >= 1 && ndev <= DG_MAX_RANKS, "dg_ctx_create_multi: %d devices (1..%d)", ndev, DG_MAX_RANKS);
    for (int i = 0; i < ndev; ++i)
        for (int j = 0; j < i; ++j) DG_REQUIRE(devs[i] != devs[j], "dg_ctx_create_multi: device %d listed twice", devs[i]);
    if (ndev == 1) return dg_ctx_create(devs[0], out);
    dg_ctx *parent = nullptr;
    if (int rc = dg_ctx_create(devs[0], &parent)) return rc;
    for (int i = 0; i < ndev; ++i) {
        dg_ctx *k = nullptr;
        if (int rc = dg_ctx_create(devs[i], &k)) {
            for (dg_ctx *q : parent->kids) dg_ctx_destroy(q);
            parent->kids.clear();
            dg_ctx_destroy(parent);
            return rc;
        }
        k->rank = i; k->world = ndev;
        parent->kids.push_back(k);
    }
    *out = parent;
    return 0;
}

int dgm_ctx_destroy_kids(dg_ctx *c) {
    for (dg_ctx *k : c->kids) dg_ctx_destroy(k);
    c->kids.clear();
    return 0;
}

int dgm_ctx_synchronize(dg_ctx *c) {
    for (dg_ctx *k : c->kids) if (int rc = dg_ctx_synchronize(k)) return rc;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// model
// ---------------------------------------------------------------------------------------------
int dgm_model_create(dg_ctx *c, int num_layers, int channel_size, int in_planes, int k_first, int precision, int max_batch, dg_model **out) {
    const int world = (int)c->kids.size();
    dg_model *m = new dg_model();
    m->ctx = c; m->prec = precision; m->max_batch = max_batch;
    if (int rc = dg_model_layout(m, num_layers, channel_size, in_planes, k_first)) { delete m; return rc; }
    const int per = (max_batch + world - 1) / world;
    for (int k = 0; k < world; ++k) {
        dg_model *r = nullptr;
        if (int rc = dg_model_create(c->kids[k], num_layers, channel_size, in_planes, k_first, precision, per, &r)) {
            for (dg_model *q : m->kids) dg_model_destroy(q);
            delete m;
            return rc;
        }
        m->kids.push_back(r);
    }
    m->kid_lo.assign(world, 0);
    m->kid_gsum.assign(world, nullptr);
    c->live_models++;
    *out = m;
    return 0;
}

int dgm_model_destroy(dg_model *m) {
    for (size_t k = 0; k < m->kids.size(); ++k) {
        if (m->kid_gsum[k]) { cudaSetDevice(m->kids[k]->ctx->dev); cudaFree(m->kid_gsum[k]); }
        dg_model_destroy(m->kids[k]);
    }
    dg_ctx *c = m->ctx;
    delete m;
    if (--c->live_models == 0 && c->zombie) { c->zombie = false; return dg_ctx_destroy(c); }
    return 0;
}

int dgm_set_params(dg_model *m, const float *host, size_t n) {
    for (dg_model *k : m->kids) if (int rc = dg_model_set_params(k, host, n)) return rc;      // every replica starts from the same vector
    return 0;
}

int dgm_zero_grads(dg_model *m) {
    for (dg_model *k : m->kids) if (int rc = dg_zero_grads(k)) return rc;
    return 0;
}

// stage a batch: replica k gets positions [lo_k, lo_k + n_k) and scales its loss gradient by 1/n (the sum over the replicas is
// then the mean gradient of the whole batch, experiments.lua:160 + ClassNLLCriterion's sizeAverage)
static int stage_check(dg_model *m, int n, const char *who) {
    const int world = (int)m->kids.size();
    DG_REQUIRE(n >= world && n <= m->max_batch, "%s: batch %d outside %d..%d (at least one position per device)", who, n, world, m->max_batch);
    m->n = n;
    return 0;
}

int dgm_set_batch_raw(dg_model *m, const uint8_t *raw, const uint8_t *player, const uint8_t *rank, const uint8_t *sym, const uint8_t *x,
                      const uint8_t *y, int n) {
    if (int rc = stage_check(m, n, "dg_model_set_batch_raw")) return rc;
    const int world = (int)m->kids.size();
    const int gb = m->global_batch > 0 ? m->global_batch : n;
    for (int k = 0; k < world; ++k) {
        int lo, cnt;
        shard(n, k, world, &lo, &cnt);
        m->kid_lo[k] = lo;
        if (int rc = dg_model_set_global_batch(m->kids[k], gb)) return rc;
        if (int rc = dg_model_set_batch_raw(m->kids[k], raw + (size_t)lo * DG_RAW_PLANES * DG_POINTS, player + lo, rank + lo, sym ? sym + lo : nullptr,
                                            x ? x + lo : nullptr, y ? y + lo : nullptr, cnt)) return rc;
    }
    m->have_raw = true; m->have_targets = x != nullptr; m->forward_done = false;
    return 0;
}

int dgm_set_batch_planes(dg_model *m, const float *planes, const int32_t *targets0, int n) {
    if (int rc = stage_check(m, n, "dg_model_set_batch_planes")) return rc;
    const int world = (int)m->kids.size();
    const int gb = m->global_batch > 0 ? m->global_batch : n;
    for (int k = 0; k < world; ++k) {
        int lo, cnt;
        shard(n, k, world, &lo, &cnt);
        m->kid_lo[k] = lo;
        if (int rc = dg_model_set_global_batch(m->kids[k], gb)) return rc;
        if (int rc = dg_model_set_batch_planes(m->kids[k], planes + (size_t)lo * m->in_planes * DG_POINTS, targets0 ? targets0 + lo : nullptr, cnt)) return rc;
    }
    m->have_raw = false; m->have_targets = targets0 != nullptr; m->forward_done = false;
    return 0;
}

int dgm_encode_resident(dg_model *m) {
    DG_REQUIRE(m->have_raw && m->n > 0, "dg_model_encode_resident: no raw records staged (dg_model_set_batch_raw)");
    for (dg_model *k : m->kids) if (int rc = dg_model_encode_resident(k)) return rc;
    m->forward_done = false;
    return 0;
}

int dgm_forward(dg_model *m, float *logprobs_out) {
    DG_REQUIRE(m->n > 0, "forward: no batch staged (call dg_model_set_batch_* first)");
    for (dg_model *k : m->kids) {                                  // all replicas first: they run side by side
        DG_CUDA(cudaSetDevice(k->ctx->dev));
        if (int rc = dg_forward_impl(k)) return rc;
    }
    m->forward_done = true;
    if (!logprobs_out) return 0;
    for (size_t i = 0; i < m->kids.size(); ++i) {
        dg_model *k = m->kids[i];
        DG_CUDA(cudaSetDevice(k->ctx->dev));
        DG_CUDA(cudaMemcpyAsync(logprobs_out + (size_t)m->kid_lo[i] * DG_POINTS, k->d_logp, (size_t)k->n * DG_POINTS * 4, cudaMemcpyDeviceToHost, k->ctx->stream));
    }
    return dgm_ctx_synchronize(m->ctx);
}

int dgm_argmax(dg_model *m, int32_t *out) {
    DG_REQUIRE(m->forward_done, "dg_argmax: call dg_forward first");
    for (size_t i = 0; i < m->kids.size(); ++i) {
        dg_model *k = m->kids[i];
        DG_CUDA(cudaSetDevice(k->ctx->dev));
        DG_CUDA(cudaMemcpyAsync(out + m->kid_lo[i], k->d_argmax, (size_t)k->n * 4, cudaMemcpyDeviceToHost, k->ctx->stream));
    }
    return dgm_ctx_synchronize(m->ctx);
}

// mean NLL over the whole batch and the error count from the replicas' own reductions
static int gather_loss(dg_model *m, float *loss_out, int *errors_out) {
    double s = 0.0;
    int e = 0;
    for (dg_model *k : m->kids) {
        float l = 0.f;
        int ek = 0;
        DG_CUDA(cudaSetDevice(k->ctx->dev));
        if (int rc = dg_read_loss(k, &l, errors_out ? &ek : nullptr)) return rc;
        s += (double)l * k->n;
        e += ek;
    }
    if (loss_out) *loss_out = (float)(s / m->n);
    if (errors_out) *errors_out = e;
    return 0;
}

int dgm_eval_batch(dg_model *m, int *errors_out, float *cost_out) {
    DG_REQUIRE(m->have_targets, "dg_eval_batch: the staged batch has no targets");
    if (int rc = dgm_forward(m, nullptr)) return rc;
    return gather_loss(m, cost_out, errors_out);
}

// ---------------------------------------------------------------------------------------------
// gradient sum of the separate-call path (eval() then SGD:step, train.lua:111,131) and of the fp32 path: every replica
// reads all replicas' gradient vectors through peer access and adds them in replica order (the same bits everywhere)
// ---------------------------------------------------------------------------------------------
struct SumArgs { const float *src[DG_MAX_RANKS]; float *dst; size_t n4; int world; };
__global__ void __launch_bounds__(256) multi_sum_kernel(const __grid_constant__ SumArgs a) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 s = __ldcg(reinterpret_cast<const float4 *>(a.src[0]) + i);
        for (int r = 1; r < a.world; ++r) {
            const float4 v = __ldcg(reinterpret_cast<const float4 *>(a.src[r]) + i);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        reinterpret_cast<float4 *>(a.dst)[i] = s;
    }
}

static int enable_peers(dg_model *m) {
    const int world = (int)m->kids.size();
    for (int r = 0; r < world; ++r) {
        DG_CUDA(cudaSetDevice(m->kids[r]->ctx->dev));
        for (int q = 0; q < world; ++q) {
            if (q == r) continue;
            int can = 0;
            DG_CUDA(cudaDeviceCanAccessPeer(&can, m->kids[r]->ctx->dev, m->kids[q]->ctx->dev));
            DG_REQUIRE(can, "multi-device model: device %d cannot access device %d", m->kids[r]->ctx->dev, m->kids[q]->ctx->dev);
            cudaError_t pe = cudaDeviceEnablePeerAccess(m->kids[q]->ctx->dev, 0);
            DG_REQUIRE(pe == cudaSuccess || pe == cudaErrorPeerAccessAlreadyEnabled, "cudaDeviceEnablePeerAccess failed: %s", cudaGetErrorString(pe));
            cudaGetLastError();
        }
    }
    return 0;
}

// m->grads of every replica <- sum over the replicas.  Stream order across devices through events: nobody reads a vector before
// its backward pass is done, nobody overwrites its vector before everyone has read it.
static int sum_grads(dg_model *m) {
    const int world = (int)m->kids.size();
    const size_t P4 = (m->P + 3) / 4;                // the vectors are cudaMalloc'ed: padding floats exist
    if (!m->kid_gsum[0]) {
        if (int rc = enable_peers(m)) return rc;
        for (int k = 0; k < world; ++k) {
            DG_CUDA(cudaSetDevice(m->kids[k]->ctx->dev));
            DG_CUDA(cudaMalloc((void **)&m->kid_gsum[k], P4 * 16));
        }
    }
    std::vector<cudaEvent_t> ea(world), eb(world);
    for (int k = 0; k < world; ++k) {
        dg_ctx *c = m->kids[k]->ctx;
        DG_CUDA(cudaSetDevice(c->dev));
        DG_CUDA(cudaEventCreateWithFlags(&ea[k], cudaEventDisableTiming));
        DG_CUDA(cudaEventCreateWithFlags(&eb[k], cudaEventDisableTiming));
        DG_CUDA(cudaEventRecord(ea[k], c->stream));
    }
    for (int k = 0; k < world; ++k) {
        dg_ctx *c = m->kids[k]->ctx;
        DG_CUDA(cudaSetDevice(c->dev));
        for (int j = 0; j < world; ++j) if (j != k) DG_CUDA(cudaStreamWaitEvent(c->stream, ea[j], 0));
        SumArgs a{};
        for (int j = 0; j < world; ++j) a.src[j] = m->kids[j]->grads;
        a.dst = m->kid_gsum[k]; a.n4 = P4; a.world = world;
        multi_sum_kernel<<<296, 256, 0, c->stream>>>(a);
        c->launches++;
        DG_CUDA(cudaGetLastError());
        DG_CUDA(cudaEventRecord(eb[k], c->stream));
    }
    for (int k = 0; k < world; ++k) {
        dg_ctx *c = m->kids[k]->ctx;
        DG_CUDA(cudaSetDevice(c->dev));
        for (int j = 0; j < world; ++j) if (j != k) DG_CUDA(cudaStreamWaitEvent(c->stream, eb[j], 0));
        DG_CUDA(cudaMemcpyAsync(m->kids[k]->grads, m->kid_gsum[k], m->P * 4, cudaMemcpyDeviceToDevice, c->stream));
    }
    for (int k = 0; k < world; ++k) { cudaEventDestroy(ea[k]); cudaEventDestroy(eb[k]); }    // released once the work that uses them has run
    return 0;
}

// criterion + backward on every replica, then the gradient sum.  dg_loss_backward ACCUMULATES into the gradient vector: what the
// vector held before the call is kept once (on replica 0), not once per replica.
int dgm_loss_backward(dg_model *m, float *loss_out) {
    DG_REQUIRE(m->forward_done, "backward: call dg_forward first");
    DG_REQUIRE(m->have_targets, "backward: the staged batch has no targets");
    for (size_t i = 0; i < m->kids.size(); ++i) {
        dg_model *k = m->kids[i];
        DG_CUDA(cudaSetDevice(k->ctx->dev));
        if (k->bf) if (int rc = dg_bf16_sync_grads(k)) return rc;
        if (i > 0) DG_CUDA(cudaMemsetAsync(k->grads, 0, k->P * 4, k->ctx->stream));
        if (int rc = dg_backward_impl(k)) return rc;
    }
    if (int rc = sum_grads(m)) return rc;
    if (loss_out) return gather_loss(m, loss_out, nullptr);
    return 0;
}

int dgm_sgd_step(dg_model *m, float rate) {
    for (dg_model *k : m->kids) if (int rc = dg_sgd_step(k, rate)) return rc;
    return 0;
}

// eval() + optimizer:step on all devices.  bf16: every replica replays its captured step graph, whose exchange kernel sums the
// gradient accumulation buffers over the replicas (peer memory) before the SGD kernel; fp32: separate calls + the sum above.
int dgm_train_step(dg_model *m, float rate, float *loss_out) {
    DG_REQUIRE(m->have_targets, "dg_train_step: the staged batch has no targets");
    DG_REQUIRE(m->n > 0, "dg_train_step: no batch staged");
    const bool fused = m->kids[0]->bf && dg_bf16_defers_grads(m->kids[0]);
    if (fused) {
        if (int rc = dg_bf16_dp_prepare_local(m->kids.data(), (int)m->kids.size())) return rc;
        for (dg_model *k : m->kids) if (int rc = dg_train_step(k, rate, nullptr)) return rc;
    } else {
        for (dg_model *k : m->kids) {
            DG_CUDA(cudaSetDevice(k->ctx->dev));
            if (int rc = dg_zero_grads(k)) return rc;
            if (int rc = dg_forward_impl(k)) return rc;
            if (int rc = dg_backward_impl(k)) return rc;
        }
        if (int rc = sum_grads(m)) return rc;
        if (int rc = dgm_sgd_step(m, rate)) return rc;
    }
    m->forward_done = true;
    if (loss_out) return gather_loss(m, loss_out, nullptr);
    return 0;
}

// c5 inference on all devices: the replicas work on their shards one after the other from the host's point of view only -- each
// call below queues work and returns; the copies at the end wait
int dgm_predict_sym8(dg_model *m, float *probs_out, int32_t *argmax_out) {
    DG_REQUIRE(m->have_raw && m->n > 0, "dg_predict_sym8: stage records with dg_model_set_batch_raw first");
    for (size_t i = 0; i < m->kids.size(); ++i) {
        dg_model *k = m->kids[i];
        if (int rc = dg_predict_sym8(k, probs_out ? probs_out + (size_t)m->kid_lo[i] * DG_POINTS : nullptr, argmax_out ? argmax_out + m->kid_lo[i] : nullptr)) return rc;
    }
    return 0;
}

int dgm_train_step_begin(dg_model *m, float rate) {
    DG_REQUIRE(m->have_targets && m->n > 0, "dg_train_step_begin: no batch with targets staged");
    const bool fused = m->kids[0]->bf && dg_bf16_defers_grads(m->kids[0]);
    DG_REQUIRE(fused, "dg_train_step_begin on a multi-device model needs the bf16 path (the other precisions sum gradients between separate calls)");
    if (int rc = dg_bf16_dp_prepare_local(m->kids.data(), (int)m->kids.size())) return rc;
    for (dg_model *k : m->kids) if (int rc = dg_train_step_begin(k, rate)) return rc;
    m->forward_done = true;
    return 0;
}

int dgm_train_step_result(dg_model *m, float *loss_out) {
    double s = 0.0;
    for (dg_model *k : m->kids) {
        float l = 0.f;
        if (int rc = dg_train_step_result(k, &l)) return rc;
        s += (double)l * k->n;
    }
    *loss_out = (float)(s / m->n);
    return 0;
}
