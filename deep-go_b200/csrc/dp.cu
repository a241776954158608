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
typedef int (*fn_allgather)(const void *, void *, size_t, int, void *, cudaStream_t);

static void *g_nccl = nullptr;
static fn_get_uid p_get_uid;
static fn_comm_init p_comm_init;
static fn_allreduce p_allreduce;
static fn_errstr p_errstr;
static fn_allgather p_allgather;

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
    p_allgather = (fn_allgather)dlsym(g_nccl, "ncclAllGather");
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
    DG_REQUIRE(c->kids.empty(), "dg_ctx_init_dp: a multi-device context is data parallel already (one process drives its devices)");
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
    c->dp_pending = true;
    return 0;
}
// the compute stream continues (SGD) only after every bucket has been reduced
int dg_dp_finish(dg_ctx *c) {
    if (c->world <= 1 || !c->dp_pending) return 0;     // nothing went to the communication stream (peer-memory exchange)
    c->dp_pending = false;
    DG_CUDA(cudaEventRecord(c->ev_comm_done, c->comm_stream));
    DG_CUDA(cudaStreamWaitEvent(c->stream, c->ev_comm_done, 0));
    return 0;
}


// ---------------------------------------------------------------------------------------------
// Gradient exchange over NVLink peer memory (the default inside dg_train_step when every peer can be mapped).
// Every rank maps every other rank's accumulation buffer (CUDA IPC, handles exchanged once through the NCCL communicator)
// and ONE kernel does a two-shot allreduce: barrier, rank r sums slice r over all ranks in a fixed order and writes the sum
// into every rank's buffer, barrier.  No proxy thread, no protocol staging buffers, no second stream: 8.9 MB take a
// fraction of the NCCL call's latency, and the sums are bit-identical on all ranks (replicas stay identical).
// ---------------------------------------------------------------------------------------------
#define DG_P2P_BLOCKS 144                   // one block per SM (a few SMs stay free for whatever else is resident)
struct P2PSignals {                          // lives in every rank's IPC-shared flag buffer
    unsigned int begin[DG_P2P_BLOCKS][DG_MAX_RANKS];   // begin[b][r]: rank r's block b has arrived at the first barrier of call #value
    unsigned int end[DG_P2P_BLOCKS][DG_MAX_RANKS];
    unsigned int calls[DG_P2P_BLOCKS];       // this rank's call counter, per block (graph replays carry no host-side counter)
};
struct P2PArgs {
    float *buf[DG_MAX_RANKS];                // every rank's accumulation buffer as mapped HERE (buf[rank] is the local one)
    P2PSignals *sig[DG_MAX_RANKS];           // [2] signal sets per rank
    float *stage[DG_MAX_RANKS];              // every rank's staging area [world][per4] float4: contributions pushed by the peers
    int rank, world;
    size_t off4, n4;                         // float4 range of the buffer this launch exchanges
    int chan;                                // signal set: launches that may overlap on a GPU use different ones
    int scatter_only;                        // 1: the sum of slice r stays on rank r (the consumer pulls it), 0: written to every rank
    unsigned long long *tl;                  // optional step timeline (DG_TL=1)
};

__device__ __forceinline__ void p2p_st_release_sys(unsigned int *p, unsigned int v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned int p2p_ld_acquire_sys(const unsigned int *p) { unsigned int v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
// data moves with plain weak accesses (L2-level: nothing of these buffers can be stale in L1, each address is touched once per
// launch, after the barrier); ordering comes from the barriers' fence + release/acquire at system scope
__device__ __forceinline__ float4 p2p_ld_v4(const float4 *p) { return __ldcg(p); }
__device__ __forceinline__ void p2p_st_v4(float4 *p, float4 v) { __stcg(p, v); }

// all ranks' block b meet: thread r tells rank r, then waits for rank r's word in the local buffer (bounded: a dead peer traps)
__device__ __forceinline__ void p2p_barrier(const P2PArgs &a, unsigned int (P2PSignals::*which)[DG_P2P_BLOCKS][DG_MAX_RANKS], unsigned int call) {
    __syncthreads();
    if (threadIdx.x < (unsigned)a.world) {
        __threadfence_system();
        p2p_st_release_sys(&((a.sig[threadIdx.x] + a.chan)->*which)[blockIdx.x][a.rank], call);
        const unsigned int *mine = &((a.sig[a.rank] + a.chan)->*which)[blockIdx.x][threadIdx.x];
        unsigned long long spins = 0;
        while (p2p_ld_acquire_sys(mine) != call)
            if (++spins > (1ull << 27)) __trap();      // ~2 minutes of polling: a peer that never arrives kills the step instead of hanging the GPU
    }
    __syncthreads();
}

__global__ void __launch_bounds__(512) p2p_allreduce_kernel(const __grid_constant__ P2PArgs a) {
    __shared__ unsigned int call_s;
    dg_pdl_launch_dependents();              // the SGD kernel may take its SMs early; it waits for this grid before it reads
    DG_TL_IN(a.tl, DG_TL_P2P + a.chan);
    dg_pdl_wait();                           // launched under the tail of the last weight-gradient kernel
    if (threadIdx.x == 0) call_s = ++(a.sig[a.rank] + a.chan)->calls[blockIdx.x];
    __syncthreads();
    const unsigned int call = call_s;
    // every rank has finished its weight-gradient kernels (stream order on each GPU) before anybody reads its buffer
    p2p_barrier(a, &P2PSignals::begin, call);
    DG_TL_GO(a.tl, DG_TL_P2P + a.chan);
    const size_t per = (a.n4 + a.world - 1) / a.world;
    const size_t lo = a.off4 + (per * a.rank < a.n4 ? per * a.rank : a.n4), hi = a.off4 + (per * (a.rank + 1) < a.n4 ? per * (a.rank + 1) : a.n4);
    // every thread first issues the loads of two elements from ALL ranks (they overlap on the wire), then adds in rank order.
    // (Measured on 2 GPUs: more loads in flight per thread, or more threads, do not shorten this phase -- 24 us for 8.7 MB.)
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += 2 * stride) {
        const size_t i2 = i + stride;
        const bool two = i2 < hi;
        float4 v[DG_MAX_RANKS], w[DG_MAX_RANKS];
#pragma unroll
        for (int r = 0; r < DG_MAX_RANKS; ++r)
            if (r < a.world) {
                v[r] = p2p_ld_v4(reinterpret_cast<const float4 *>(a.buf[r]) + i);
                if (two) w[r] = p2p_ld_v4(reinterpret_cast<const float4 *>(a.buf[r]) + i2);
            }
        float4 s = v[0], t = w[0];
#pragma unroll
        for (int r = 1; r < DG_MAX_RANKS; ++r)
            if (r < a.world) {                                        // fixed order: the same bits on every rank
                s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w;
                if (two) { t.x += w[r].x; t.y += w[r].y; t.z += w[r].z; t.w += w[r].w; }
            }
        if (a.scatter_only) {
            p2p_st_v4(reinterpret_cast<float4 *>(a.buf[a.rank]) + i, s);
            if (two) p2p_st_v4(reinterpret_cast<float4 *>(a.buf[a.rank]) + i2, t);
        } else {
#pragma unroll
            for (int r = 0; r < DG_MAX_RANKS; ++r)
                if (r < a.world) {
                    p2p_st_v4(reinterpret_cast<float4 *>(a.buf[r]) + i, s);
                    if (two) p2p_st_v4(reinterpret_cast<float4 *>(a.buf[r]) + i2, t);
                }
        }
    }
    // nobody leaves (and lets its next kernel read, or its next step clear, the buffer) before every rank has written every slice
    p2p_barrier(a, &P2PSignals::end, call);
    DG_TL_OUT(a.tl, DG_TL_P2P + a.chan);
}

// Push variant (DG_DP_WRITE=1; measured SLOWER on 2 GPUs: 50 us against 31 us -- 4.35 MB of remote writes take 20 us, no faster
// than the remote reads of the kernel above, and there are two data phases instead of one): remote writes only.
//   1. every rank pushes its contribution to slice o into rank o's staging area, for all o != rank.  No waiting: the staging areas
//      were consumed before the previous call's closing barrier, and a skewed peer's weight-gradient kernels are not disturbed;
//   2. barrier (all pushes have landed);  3. rank r sums slice r -- its own values and the staged ones, in rank order, all local
//      reads -- and writes the sum into every rank's buffer;  4. barrier.
__global__ void __launch_bounds__(512) p2p_push_allreduce_kernel(const __grid_constant__ P2PArgs a) {
    __shared__ unsigned int call_s;
    dg_pdl_launch_dependents();
    DG_TL_IN(a.tl, DG_TL_P2P + a.chan);
    dg_pdl_wait();
    if (threadIdx.x == 0) call_s = ++(a.sig[a.rank] + a.chan)->calls[blockIdx.x];
    __syncthreads();
    const unsigned int call = call_s;
    const size_t per = (a.n4 + a.world - 1) / a.world;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    const float4 *mine = reinterpret_cast<const float4 *>(a.buf[a.rank]);
    for (int o = 0; o < a.world; ++o) {
        if (o == a.rank) continue;
        const size_t lo = per * o < a.n4 ? per * o : a.n4, hi = per * (o + 1) < a.n4 ? per * (o + 1) : a.n4;
        float4 *dst = reinterpret_cast<float4 *>(a.stage[o]) + per * a.rank;
        for (size_t i = lo + tid; i < hi; i += 2 * stride) {
            const size_t i2 = i + stride;
            const float4 v = p2p_ld_v4(mine + i);
            float4 w = v;
            if (i2 < hi) w = p2p_ld_v4(mine + i2);
            p2p_st_v4(dst + (i - lo), v);
            if (i2 < hi) p2p_st_v4(dst + (i2 - lo), w);
        }
    }
    p2p_barrier(a, &P2PSignals::begin, call);
    DG_TL_GO(a.tl, DG_TL_P2P + a.chan);
    {
        const size_t lo = per * a.rank < a.n4 ? per * a.rank : a.n4, hi = per * (a.rank + 1) < a.n4 ? per * (a.rank + 1) : a.n4;
        const float4 *st = reinterpret_cast<const float4 *>(a.stage[a.rank]);
        for (size_t i = lo + tid; i < hi; i += stride) {
            float4 v[DG_MAX_RANKS];
#pragma unroll
            for (int r = 0; r < DG_MAX_RANKS; ++r)
                if (r < a.world) v[r] = r == a.rank ? p2p_ld_v4(mine + i) : p2p_ld_v4(st + per * r + (i - lo));
            float4 s = v[0];
#pragma unroll
            for (int r = 1; r < DG_MAX_RANKS; ++r)
                if (r < a.world) { s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w; }      // rank order: the same bits every run
#pragma unroll
            for (int r = 0; r < DG_MAX_RANKS; ++r)
                if (r < a.world) p2p_st_v4(reinterpret_cast<float4 *>(a.buf[r]) + i, s);
        }
    }
    p2p_barrier(a, &P2PSignals::end, call);
    DG_TL_OUT(a.tl, DG_TL_P2P + a.chan);
}

struct dg_p2p {
    P2PArgs args;
    void *opened[3 * DG_MAX_RANKS];
    int n_opened = 0;
    P2PSignals *sig_local = nullptr;
    float *stage_local = nullptr;
};

// collective over the ranks of the context: map every rank's `buf` (a cudaMalloc base pointer of `nfloats` floats); 0 = ready.
// Any failure leaves *out NULL with a message in dg_last_error -- the caller keeps using NCCL.
int dg_dp_p2p_setup(dg_ctx *c, float *buf, size_t nfloats, dg_p2p **out) {
    *out = nullptr;
    DG_REQUIRE(c->world > 1 && c->world <= DG_MAX_RANKS && c->comm && p_allgather, "peer exchange: needs 2..%d ranks and ncclAllGather", DG_MAX_RANKS);
    DG_REQUIRE(nfloats % 4 == 0, "peer exchange: buffer length must be a multiple of 4 floats");
    dg_p2p *p = new dg_p2p();
    DG_CUDA(cudaMalloc((void **)&p->sig_local, 2 * sizeof(P2PSignals)));
    DG_CUDA(cudaMemset(p->sig_local, 0, 2 * sizeof(P2PSignals)));
    const size_t per4 = (nfloats / 4 + c->world - 1) / c->world;
    DG_CUDA(cudaMalloc((void **)&p->stage_local, per4 * c->world * 16));
    struct Rec { cudaIpcMemHandle_t buf, sig, stage; int dev; int ok; };
    Rec mine{}; mine.dev = c->dev; mine.ok = 1;
    if (cudaIpcGetMemHandle(&mine.buf, buf) != cudaSuccess || cudaIpcGetMemHandle(&mine.sig, p->sig_local) != cudaSuccess ||
        cudaIpcGetMemHandle(&mine.stage, p->stage_local) != cudaSuccess) { cudaGetLastError(); mine.ok = 0; }
    Rec *d_all = nullptr;
    DG_CUDA(cudaMalloc((void **)&d_all, sizeof(Rec) * (c->world + 1)));
    DG_CUDA(cudaMemcpy(d_all + c->world, &mine, sizeof(Rec), cudaMemcpyHostToDevice));
    int e = p_allgather(d_all + c->world, d_all, sizeof(Rec), 0 /* ncclChar */, c->comm, c->comm_stream);
    DG_REQUIRE(e == 0, "peer exchange: ncclAllGather failed: %s", p_errstr(e));
    DG_CUDA(cudaStreamSynchronize(c->comm_stream));
    std::vector<Rec> all(c->world);
    DG_CUDA(cudaMemcpy(all.data(), d_all, sizeof(Rec) * c->world, cudaMemcpyDeviceToHost));
    cudaFree(d_all);
    bool ok = true;
    for (int r = 0; r < c->world; ++r) ok = ok && all[r].ok;
    for (int r = 0; r < c->world && ok; ++r) {
        if (r == c->rank) { p->args.buf[r] = buf; p->args.sig[r] = p->sig_local; p->args.stage[r] = p->stage_local; continue; }
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, c->dev, all[r].dev) != cudaSuccess || !can) { ok = false; break; }
        cudaError_t pe = cudaDeviceEnablePeerAccess(all[r].dev, 0);
        if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) { ok = false; break; }
        cudaGetLastError();
        void *pb = nullptr, *ps = nullptr;
        if (cudaIpcOpenMemHandle(&pb, all[r].buf, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = false; break; }
        p->opened[p->n_opened++] = pb;
        if (cudaIpcOpenMemHandle(&ps, all[r].sig, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = false; break; }
        p->opened[p->n_opened++] = ps;
        void *pg = nullptr;
        if (cudaIpcOpenMemHandle(&pg, all[r].stage, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = false; break; }
        p->opened[p->n_opened++] = pg;
        p->args.buf[r] = (float *)pb; p->args.sig[r] = (P2PSignals *)ps; p->args.stage[r] = (float *)pg;
    }
    cudaGetLastError();
    // every rank must take the same path: agree on success through one more tiny allgather
    int *d_ok = nullptr;
    DG_CUDA(cudaMalloc((void **)&d_ok, sizeof(int) * (c->world + 1)));
    const int my_ok = ok ? 1 : 0;
    DG_CUDA(cudaMemcpy(d_ok + c->world, &my_ok, sizeof(int), cudaMemcpyHostToDevice));
    e = p_allgather(d_ok + c->world, d_ok, sizeof(int), 0, c->comm, c->comm_stream);
    DG_REQUIRE(e == 0, "peer exchange: ncclAllGather failed: %s", p_errstr(e));
    DG_CUDA(cudaStreamSynchronize(c->comm_stream));
    std::vector<int> oks(c->world);
    DG_CUDA(cudaMemcpy(oks.data(), d_ok, sizeof(int) * c->world, cudaMemcpyDeviceToHost));
    cudaFree(d_ok);
    for (int r = 0; r < c->world; ++r) ok = ok && oks[r];
    if (!ok) {
        dg_dp_p2p_destroy(p);
        DG_REQUIRE(false, "peer exchange: a rank could not map its peers' memory (CUDA IPC / P2P unavailable): using NCCL");
    }
    p->args.rank = c->rank; p->args.world = c->world; p->args.off4 = 0; p->args.n4 = nfloats / 4; p->args.chan = 0; p->args.scatter_only = 0;
    *out = p;
    return 0;
}

// Single process, one context per device (dg_ctx_create_multi): every context enables peer access to the others and uses their
// buffers and flag blocks through plain device pointers.  out[r] receives rank r's mapping; all or nothing.
int dg_dp_p2p_setup_local(dg_ctx *const *ctxs, float *const *bufs, size_t nfloats, int world, dg_p2p **out) {
    DG_REQUIRE(world > 1 && world <= DG_MAX_RANKS, "peer exchange: needs 2..%d devices", DG_MAX_RANKS);
    DG_REQUIRE(nfloats % 4 == 0, "peer exchange: buffer length must be a multiple of 4 floats");
    for (int r = 0; r < world; ++r) out[r] = nullptr;
    const size_t per4 = (nfloats / 4 + world - 1) / world;
    for (int r = 0; r < world; ++r) {
        DG_CUDA(cudaSetDevice(ctxs[r]->dev));
        for (int q = 0; q < world; ++q) {
            if (q == r) continue;
            int can = 0;
            DG_CUDA(cudaDeviceCanAccessPeer(&can, ctxs[r]->dev, ctxs[q]->dev));
            DG_REQUIRE(can, "peer exchange: device %d cannot access device %d (numGPUs > 1 in one process needs peer access)", ctxs[r]->dev, ctxs[q]->dev);
            cudaError_t pe = cudaDeviceEnablePeerAccess(ctxs[q]->dev, 0);
            DG_REQUIRE(pe == cudaSuccess || pe == cudaErrorPeerAccessAlreadyEnabled, "cudaDeviceEnablePeerAccess(%d -> %d) failed: %s", ctxs[r]->dev, ctxs[q]->dev, cudaGetErrorString(pe));
            cudaGetLastError();
        }
        dg_p2p *p = new dg_p2p();
        out[r] = p;
        DG_CUDA(cudaMalloc((void **)&p->sig_local, 2 * sizeof(P2PSignals)));
        DG_CUDA(cudaMemset(p->sig_local, 0, 2 * sizeof(P2PSignals)));
        DG_CUDA(cudaMalloc((void **)&p->stage_local, per4 * world * 16));
    }
    for (int r = 0; r < world; ++r) {
        dg_p2p *p = out[r];
        for (int q = 0; q < world; ++q) { p->args.buf[q] = bufs[q]; p->args.sig[q] = out[q]->sig_local; p->args.stage[q] = out[q]->stage_local; }
        p->args.rank = r; p->args.world = world; p->args.off4 = 0; p->args.n4 = nfloats / 4; p->args.chan = 0; p->args.scatter_only = 0;
    }
    return 0;
}

void dg_dp_p2p_destroy(dg_p2p *p) {
    if (!p) return;
    for (int i = 0; i < p->n_opened; ++i) cudaIpcCloseMemHandle(p->opened[i]);
    cudaFree(p->sig_local);
    cudaFree(p->stage_local);
    delete p;
}

// sum floats [off, off + n) of the mapped buffers over ranks, in place on every rank, as one kernel on `st` (captured into the
// step graph like any kernel).  Two launches that may be in flight together on a GPU must use different signal sets (chan 0 / 1).
int dg_dp_p2p_allreduce(dg_ctx *c, dg_p2p *p, cudaStream_t st, size_t off, size_t n, int chan, int scatter_only) {
    DG_REQUIRE(off % 4 == 0 && n % 4 == 0 && chan >= 0 && chan < 2, "peer exchange: bad range");
    P2PArgs a = p->args;
    a.off4 = off / 4; a.n4 = n / 4; a.chan = chan; a.tl = c->tl; a.scatter_only = scatter_only;
    static const bool push_rs = getenv("DG_DP_WRITE") != nullptr;      // the write-only variant, for comparison
    if (!scatter_only && push_rs && off == 0 && n / 4 == p->args.n4) DG_CUDA(dg_launch_pdl(p2p_push_allreduce_kernel, dim3(DG_P2P_BLOCKS), dim3(512), 0, st, a));
    else DG_CUDA(dg_launch_pdl(p2p_allreduce_kernel, dim3(DG_P2P_BLOCKS), dim3(512), 0, st, a));
    c->launches++;
    return 0;
}

// every rank has reached this point of `st` (one block, signal set 1): used before a rank clears a buffer its peers may still be reading
__global__ void __launch_bounds__(32) p2p_barrier_kernel(const __grid_constant__ P2PArgs a) {
    __shared__ unsigned int call_s;
    if (threadIdx.x == 0) call_s = ++(a.sig[a.rank] + a.chan)->calls[0];
    __syncthreads();
    p2p_barrier(a, &P2PSignals::begin, call_s);
}
int dg_dp_p2p_barrier(dg_ctx *c, dg_p2p *p, cudaStream_t st) {
    P2PArgs a = p->args;
    a.chan = 1;
    p2p_barrier_kernel<<<1, 32, 0, st>>>(a);
    c->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
// the mapping, for kernels that pull a summed slice from its owner: buf[r] = rank r's buffer as mapped here, per4 = float4 per slice
void dg_dp_p2p_peers(const dg_p2p *p, const float **buf, int *world, size_t *per4) {
    for (int r = 0; r < DG_MAX_RANKS; ++r) buf[r] = r < p->args.world ? p->args.buf[r] : nullptr;
    *world = p->args.world;
    *per4 = (p->args.n4 + p->args.world - 1) / p->args.world;
}
