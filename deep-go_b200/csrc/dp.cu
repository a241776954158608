// Data-parallel gradient exchange (replaces nn.DataParallelTable(1), experiments.lua:155-168).
// One process per GPU; NCCL is bound at run time (dlopen) so the library loads without it.
#include "common.cuh"
#include "../../include/deepgo.h"
#include <dlfcn.h>
#include <cstring>

typedef struct { char internal[128]; } nccl_uid_t;
typedef int (*fn_get_uid)(nccl_uid_t *);
typedef int (*fn_comm_init)(void **, int, nccl_uid_t, int);
typedef int (*fn_allreduce)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef const char *(*fn_errstr)(int);
typedef int (*fn_comm_destroy)(void *);

static void *g_nccl = nullptr;
static fn_get_uid p_get_uid;
static fn_comm_init p_comm_init;
static fn_allreduce p_allreduce;
static fn_errstr p_errstr;

static int load_nccl() {
    if (g_nccl) return 0;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
        g_nccl = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl) break;
    }
    DG_REQUIRE(g_nccl, "data parallel: cannot dlopen libnccl.so.2 (%s)", dlerror());
    p_get_uid = (fn_get_uid)dlsym(g_nccl, "ncclGetUniqueId");
    p_comm_init = (fn_comm_init)dlsym(g_nccl, "ncclCommInitRank");
    p_allreduce = (fn_allreduce)dlsym(g_nccl, "ncclAllReduce");
    p_errstr = (fn_errstr)dlsym(g_nccl, "ncclGetErrorString");
    DG_REQUIRE(p_get_uid && p_comm_init && p_allreduce && p_errstr, "data parallel: libnccl lacks a required symbol");
    return 0;
}

extern "C" int dg_dp_unique_id(char id_out[128]) {
    DG_REQUIRE(id_out, "dg_dp_unique_id: NULL buffer");
    if (int rc = load_nccl()) return rc;
    nccl_uid_t uid;
    int e = p_get_uid(&uid);
    DG_REQUIRE(e == 0, "ncclGetUniqueId failed: %s", p_errstr(e));
    memcpy(id_out, uid.internal, 128);
    return 0;
}

extern "C" int dg_ctx_init_dp(dg_ctx *c, int rank, int world, const char id[128]) {
    DG_REQUIRE(c && id, "dg_ctx_init_dp: NULL argument");
    DG_REQUIRE(world >= 1 && rank >= 0 && rank < world, "dg_ctx_init_dp: rank %d / world %d invalid", rank, world);
    c->rank = rank;
    c->world = world;
    if (world == 1) return 0;
    if (int rc = load_nccl()) return rc;
    DG_CUDA(cudaSetDevice(c->dev));
    nccl_uid_t uid;
    memcpy(uid.internal, id, 128);
    int e = p_comm_init(&c->comm, world, uid, rank);
    DG_REQUIRE(e == 0, "ncclCommInitRank failed: %s", p_errstr(e));
    DG_CUDA(cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking));
    DG_CUDA(cudaEventCreateWithFlags(&c->ev_grad_ready, cudaEventDisableTiming));
    DG_CUDA(cudaEventCreateWithFlags(&c->ev_comm_done, cudaEventDisableTiming));
    return 0;
}

// Gradient bucket exchange, overlapped with the rest of the backward pass: the bucket (one layer's W, b, a slice of
// the flat gradient vector, contiguous in Torch7 order) is summed over ranks on the communication stream as soon as
// the kernels that produce it have finished on the compute stream (ncclFloat32 = 7, ncclSum = 0).
int dg_dp_bucket(dg_ctx *c, float *buf, size_t n, cudaStream_t producer) {
    if (c->world <= 1 || n == 0) return 0;
    DG_REQUIRE(c->comm, "data parallel: dg_ctx_init_dp was not called");
    DG_CUDA(cudaEventRecord(c->ev_grad_ready, producer ? producer : c->stream));
    DG_CUDA(cudaStreamWaitEvent(c->comm_stream, c->ev_grad_ready, 0));
    int e = p_allreduce(buf, buf, n, 7, 0, c->comm, c->comm_stream);
    DG_REQUIRE(e == 0, "ncclAllReduce failed: %s", p_errstr(e));
    c->launches++;
    return 0;
}
// the compute stream continues (SGD) only after every bucket has been reduced
int dg_dp_finish(dg_ctx *c) {
    if (c->world <= 1) return 0;
    DG_CUDA(cudaEventRecord(c->ev_comm_done, c->comm_stream));
    DG_CUDA(cudaStreamWaitEvent(c->stream, c->ev_comm_done, 0));
    return 0;
}
