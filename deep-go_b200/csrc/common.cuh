// Shared declarations for libdeepgo_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stddef.h>
#include <vector>
#include <string>

#define DG_BOARD 19
#define DG_POINTS 361
#define DG_RAW_PLANES 9
#define DG_PLANES 37

void dg_set_error(const char *fmt, ...);

#define DG_CUDA(expr)                                                                              \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            dg_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return 1;                                                                              \
        }                                                                                          \
    } while (0)

#define DG_REQUIRE(cond, ...)                                                                      \
    do {                                                                                           \
        if (!(cond)) {                                                                             \
            dg_set_error(__VA_ARGS__);                                                             \
            return 2;                                                                              \
        }                                                                                          \
    } while (0)

enum {
    DG_K_ENCODE = 0, DG_K_CONV_FIRST, DG_K_CONV, DG_K_DGRAD, DG_K_WGRAD, DG_K_WGRAD_REDUCE, DG_K_BIAS_GRAD,
    DG_K_HEAD, DG_K_SGD, DG_K_PACK, DG_K_ALLREDUCE, DG_K_OTHER, DG_K_COUNT
};

struct dg_prof_rec { int kind; cudaEvent_t e0, e1; };

struct dg_ctx {
    int dev = 0;
    bool prof_on = false;
    int live_models = 0;               // dg_ctx_destroy is deferred until the last model is gone
    bool zombie = false;
    std::vector<dg_prof_rec> prof_recs;
    double prof_ms[DG_K_COUNT] = {0};
    long long prof_n[DG_K_COUNT] = {0};
    cudaStream_t stream = nullptr;
    long long launches = 0;
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    // grow-only device scratch for the stand-alone entry points (dg_encode, dg_summarize_boards)
    void *scratch = nullptr;
    size_t scratch_bytes = 0;
    void *flush_buf = nullptr;
    size_t flush_bytes = 0;
    // data parallel (dp.cu)
    void *nccl_lib = nullptr;
    void *comm = nullptr;
    int rank = 0, world = 1;
    cudaStream_t comm_stream = nullptr;
    cudaEvent_t ev_grad_ready = nullptr, ev_comm_done = nullptr;
    bool dp_pending = false;           // buckets issued on the communication stream since the last dg_dp_finish
    // DG_TL=1: device buffer of globaltimer stamps {first CTA in, first CTA past the dependency wait, last CTA out} per kernel of the
    // step -- the only clock that sees inside a replayed CUDA graph (dg_prof_timeline)
    unsigned long long *tl = nullptr;
    // dg_ctx_create_multi: this context stands for several devices of ONE process; kids[i] is the context of device i (multi.cu)
    std::vector<dg_ctx *> kids;
};
enum { DG_TL_CONV0 = 0, DG_TL_CONV, DG_TL_HEAD, DG_TL_DGRAD, DG_TL_WGRAD, DG_TL_WGRAD0, DG_TL_FINALIZE, DG_TL_SGD, DG_TL_HEADRED, DG_TL_P2P, DG_TL_P2P_B, DG_TL_COUNT };

struct LayerInfo {
    int cin, cout, k;
    size_t offW, offB, offA;   // offsets into the Torch7-order flat vector
    size_t nW;
};

int dg_ctx_scratch(dg_ctx *ctx, size_t bytes, void **out);

// Programmatic dependent launch: the kernel may be scheduled while its predecessor in the stream is still running;
// it must execute dg_pdl_wait() before touching global memory (all kernels of the step do, so completion is transitive).
#ifdef __CUDACC__
__device__ __forceinline__ void dg_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void dg_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ unsigned long long dg_gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define DG_TL_IN(tl, id) do { if ((tl) && threadIdx.x == 0) atomicMin((tl) + 3 * (id), dg_gtimer()); } while (0)
#define DG_TL_GO(tl, id) do { if ((tl) && threadIdx.x == 0) atomicMin((tl) + 3 * (id) + 1, dg_gtimer()); } while (0)
#define DG_TL_OUT(tl, id) do { if ((tl) && threadIdx.x == 0) atomicMax((tl) + 3 * (id) + 2, dg_gtimer()); } while (0)
template <typename... KArgs, typename... Args>
static inline cudaError_t dg_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    static const bool no_pdl = getenv("DG_NO_PDL") != nullptr;
    cfg.attrs = at; cfg.numAttrs = no_pdl ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
// same, as thread-block clusters of `cluster` CTAs along x (grid.x must be a multiple of it)
template <typename KArg, typename Arg>
static inline cudaError_t dg_launch_pdl(void (*kernel)(KArg), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Arg arg, int cluster) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = cluster; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 2;
    return cudaLaunchKernelEx(&cfg, kernel, KArg(arg));
}
#endif

// per-kernel CUDA-event timing (dg_prof_*): brackets one launch on the context's stream when enabled
struct dg_prof_scope {
    dg_ctx *c; int idx; cudaStream_t st;
    dg_prof_scope(dg_ctx *ctx, int kind, cudaStream_t stream = nullptr) : c(ctx), idx(-1), st(stream ? stream : ctx->stream) {
        if (!c->prof_on) return;
        dg_prof_rec r; r.kind = kind;
        cudaEventCreate(&r.e0); cudaEventCreate(&r.e1);
        cudaEventRecord(r.e0, st);
        c->prof_recs.push_back(r);
        idx = (int)c->prof_recs.size() - 1;
    }
    ~dg_prof_scope() { if (idx >= 0) cudaEventRecord(c->prof_recs[idx].e1, st); }
};

// ---- encode.cu
int dg_launch_encode_f32(dg_ctx *ctx, const uint8_t *raw, const uint8_t *player, const uint8_t *rank,
                         const uint8_t *sym, int n, float *out);
int dg_launch_encode_bits(dg_ctx *ctx, const uint8_t *raw, const uint8_t *player, const uint8_t *rank,
                          const uint8_t *sym, int n, uint32_t *out);
int dg_launch_labels(dg_ctx *ctx, const uint8_t *x, const uint8_t *y, const uint8_t *sym, int n, int32_t *out);

// ---- conv_simt.cu (fp32 reference-precision path, NCHW)
int dg_simt_conv(dg_ctx *ctx, const float *in, const float *W, const float *b, const float *a, const float *mask,
                 float *out, int n, int cin, int cout, int k, int transposed, int relu);
int dg_simt_wgrad(dg_ctx *ctx, const float *x, const float *dz, float *gW, int n, int cin, int cout, int k);
int dg_simt_bias_grad(dg_ctx *ctx, const float *dz, float *ga, float *gb, int n, int cout);

// ---- head.cu (log-softmax + NLL + dlogits, argmax, loss reduction)
int dg_launch_logsoftmax_nll(dg_ctx *ctx, const float *h_last, const int32_t *targets, float *logp, float *loss_n,
                             float *dz_last, int32_t *argmax, int n, float scale);
int dg_launch_loss_reduce(dg_ctx *ctx, const float *loss_n, const int32_t *argmax, const int32_t *targets, int n,
                          float *loss_out, int *errors_out);

// ---- sgd.cu
int dg_launch_sgd(dg_ctx *ctx, float *p, const float *g, size_t n, float rate);
int dg_launch_sgd_dev_rate(dg_ctx *ctx, float *p, const float *g, size_t n, const float *rate_dev);
int dg_launch_fill(dg_ctx *ctx, float *p, size_t n, float v);

// ---- dp.cu
int dg_dp_bucket(dg_ctx *ctx, float *buf, size_t n, cudaStream_t producer = nullptr);
int dg_dp_finish(dg_ctx *ctx);
// gradient exchange over NVLink peer memory (dp.cu): collective setup, then one kernel per step
#define DG_MAX_RANKS 8
struct dg_p2p;
int dg_dp_p2p_setup(dg_ctx *ctx, float *buf, size_t nfloats, dg_p2p **out);
// the same mapping when all ranks are contexts of this process: peer access and plain pointers, no IPC, no NCCL
int dg_dp_p2p_setup_local(dg_ctx *const *ctxs, float *const *bufs, size_t nfloats, int world, dg_p2p **out);
int dg_dp_p2p_allreduce(dg_ctx *ctx, dg_p2p *p, cudaStream_t st, size_t off_floats, size_t n_floats, int chan, int scatter_only);
int dg_dp_p2p_barrier(dg_ctx *ctx, dg_p2p *p, cudaStream_t st);
void dg_dp_p2p_peers(const dg_p2p *p, const float **buf, int *world, size_t *per4);
void dg_dp_p2p_destroy(dg_p2p *p);
