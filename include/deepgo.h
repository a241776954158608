/*
 * deepgo.h -- C ABI of libdeepgo_b200.so: the B200-native replacement for the hot path of
 * vipmath/deep-go (Torch7 move-prediction net).  Plain C, no macros in the declared part, so the
 * same text can be fed to LuaJIT `ffi.cdef` and to Python `cffi`.
 *
 * Conventions
 *   - every function returns int: 0 = ok, non-zero = error; dg_last_error() gives the text
 *     (thread-local).  No C++ exception or longjmp crosses this boundary, so the reference's
 *     pcall() guard (train.lua:106) keeps working through the Lua shim.
 *   - indices are 0-based on this side (labels, argmax, board x/y); the Lua shim adds/subtracts 1
 *     (dataloader.lua:89, train.lua:29-36 are 1-based).
 *   - caller owns every host buffer; the library owns all device memory.  Handles are opaque.
 *   - one caller thread per dg_ctx.  GPU work is asynchronous on the context's stream; a call
 *     synchronises only when it returns a scalar or fills a host buffer.
 *   - there is NO CPU fallback: every compute entry point fails with an error if no B200 is present.
 *
 * References are to files of the reference repository (file:line).
 */
#ifndef DEEPGO_H
#define DEEPGO_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dg_ctx dg_ctx;
typedef struct dg_model dg_model;

/* precision of the conv stack */
/* FP32: CUDA cores, the reference's CUDA arithmetic (train.lua:99-100); BF16: tcgen05 kind::f16 with bf16 operands and stored
 * activations, fp32 accumulation -- the fast path; TF32: tcgen05 kind::tf32, fp32 storage, 10-bit-mantissa products, fp32
 * accumulation -- the tensor-core type closest to the reference's */
enum { DG_PREC_FP32 = 0, DG_PREC_BF16 = 1, DG_PREC_TF32 = 2 };
/* output formats of dg_encode */
enum { DG_FMT_F32_NCHW = 0, DG_FMT_BITPLANES = 1 };

/* ---- errors / info ----------------------------------------------------------------------- */
/* replaces Lua error() text seen by pcall (train.lua:106-109) */
const char *dg_last_error(void);
int dg_version(void);
/* number of visible CUDA devices (cutorch.getDeviceCount, experiments.lua:157) */
int dg_device_count(int *out);

/* ---- context: one device, one stream (cutorch.setDevice, default-experiment.lua:6) --------- */
int dg_ctx_create(int device, dg_ctx **out);
/* numGPUs > 1 inside ONE process (experiments.lua:10,97-103,155-168: makeDataParallel wraps the model in
 * nn.DataParallelTable(1)): a context that stands for ndev devices.  A model created on it is one replica per device;
 * dg_model_set_batch_* split the batch on dim 0 into near-equal contiguous chunks as DataParallelTable does (n >= ndev),
 * every replica scales its loss gradient by 1/n, and dg_train_step / dg_loss_backward sum the replicas' gradients over
 * NVLink peer memory (cudaDeviceEnablePeerAccess; no NCCL, no second process) before the update, so the replicas stay
 * bit-identical.  Results (log-probs, argmax, loss, errors) come back for the whole batch; parameters and gradients are
 * read from replica 0.  One caller thread drives all devices.  ndev == 1 is dg_ctx_create.  The stand-alone entry points
 * (dg_encode, dg_summarize_boards, dg_update_boards, timers, profiling) use the first device. */
int dg_ctx_create_multi(int ndev, const int *devs, dg_ctx **out);
int dg_ctx_destroy(dg_ctx *ctx);
int dg_ctx_synchronize(dg_ctx *ctx);
/* kernels this library has launched on ctx so far (bench.py "gpu_launches") */
int dg_ctx_launch_count(dg_ctx *ctx, long long *out);
/* CUDA-event timer on the context's stream (sys.clock() pairs, train.lua:91,94,113,137) */
int dg_timer_start(dg_ctx *ctx);
int dg_timer_stop(dg_ctx *ctx, float *ms_out);
/* overwrite a 256 MB device buffer (L2 flush between timed iterations) */
int dg_ctx_flush_l2(dg_ctx *ctx);
/* per-kernel device timing for the roofline report: between dg_prof_begin and dg_prof_end every launch
 * is bracketed by CUDA events on the context's stream; dg_prof_get returns the summed duration and the
 * launch count of one kernel kind (0 encode, 1 first conv, 2 conv, 3 input-gradient conv, 4 weight gradient,
 * 5 weight-gradient reduction, 6 bias gradient, 7 head, 8 SGD, 9 operand pack, 10 allreduce, 11 other). */
int dg_prof_begin(dg_ctx *ctx);
int dg_prof_end(dg_ctx *ctx);
int dg_prof_get(dg_ctx *ctx, int kind, float *ms_total, long long *launches);
/* Step timeline that also sees inside a replayed CUDA graph (process started with DG_TL=1): per kernel of the training step
 * {first CTA in, first CTA past its dependency wait, last CTA out} in globaltimer nanoseconds, in the order first conv, conv
 * chain, head, input-gradient chain, weight gradient (3x3 layers), weight gradient (first layer), gradient finalize, SGD, head
 * reduction, peer-memory gradient exchange (two launches); copies min(n, 33) values, then resets the stamps.  Profiling aid, no reference counterpart. */
int dg_prof_timeline(dg_ctx *ctx, unsigned long long *out, int n);

/* ---- data parallel (nn.DataParallelTable(1), experiments.lua:155-168) ---------------------- */
/* One process per GPU.  Rank 0 calls dg_dp_unique_id and ships the 128 bytes to the other ranks
 * (any side channel); every rank then calls dg_ctx_init_dp.  After that dg_train_step sums the
 * gradient over ranks before the SGD update: on the bf16 path ONE kernel after the last
 * weight-gradient kernel, a two-shot allreduce over NVLink peer memory (CUDA IPC mapping of the
 * gradient accumulation buffers; bit-identical sums on every rank), captured in the step graph; NCCL
 * allreduce as the
 * fallback when peer mapping is unavailable, and on the fp32 path).  Each rank must scale its loss
 * gradient by 1/global_batch (dg_model_set_global_batch) so the sum is the global-batch mean
 * gradient -- what DataParallelTable + ClassNLLCriterion on the gathered output produce. */
int dg_dp_unique_id(char id_out[128]);
int dg_ctx_init_dp(dg_ctx *ctx, int rank, int world, const char id[128]);
/* sum the whole gradient vector over ranks with one NCCL allreduce (for callers that run eval() and SGD:step as
 * separate calls, train.lua:111 then :131); dg_train_step does the exchange itself, see above */
int dg_dp_allreduce_grads(dg_model *m);

/* ---- feature planes (dataloader.lua) ------------------------------------------------------ */
/* preprocess() over a batch (dataloader.lua:50-92,113-121) with the D4 symmetry hook
 * (transform(), dataloader.lua:41-44; sym 0 = reference behaviour).
 *   raw    u8[n][9][19][19]  flatten_data() planes (dataloader.lua:30-39)
 *   player u8[n] in {1,2}    data.move.player
 *   rank   u8[n] in 1..9     data.ranks[player]
 *   sym    u8[n] in 0..7 or NULL (all 0)
 *   out    DG_FMT_F32_NCHW:  float[n][37][19][19]
 *          DG_FMT_BITPLANES: uint32[n][37][12]  (bit p%32 of word p/32 = plane value at p=19*x+y)
 * Host buffers in and out. */
int dg_encode(dg_ctx *ctx, const uint8_t *raw, const uint8_t *player, const uint8_t *rank,
              const uint8_t *sym, int n, void *out, int outfmt);
/* dg_encode for callers that cannot allocate device memory.  No counterpart in the reference: this entry point exists
 * only because of how the library is embedded.  Loader threads that must not stall a running training step
 * (cudaMalloc/cudaFree synchronise the device) and processes whose HBM is already owned by another allocator
 * cannot let dg_encode grow its staging buffer; this variant stages through a fixed ~3.7 MB array that is part of
 * the library image on every GPU.  Like dg_encode it synchronises the context's stream before it returns.  Same arguments and results as dg_encode, but
 * n <= 64 (larger n is an error).  The first call on a device resolves the array's device address -- make it before
 * any stream capture.  Calls on one GPU share the array: run them one at a time (overlapping calls race); the array
 * keeps the last call's bytes and needs no clearing.  Calls no cudaMalloc/cudaFree. */
int dg_encode_noalloc(dg_ctx *ctx, const uint8_t *raw, const uint8_t *player, const uint8_t *rank,
                      const uint8_t *sym, int n, void *out, int outfmt);
/* label = 19*(x-1)+y of the transformed move (dataloader.lua:89); x,y 1-based as stored in the
 * records, result 0-based.  Pure host helper, reentrant. */
int dg_encode_labels(const uint8_t *x, const uint8_t *y, const uint8_t *sym, int n, int32_t *out);

/* torch.load of one position record (dataloader.lua:123-125) without Torch7: fills raw u8[9*19*19] and the
 * 1-based move / player / ranks exactly as stored.  Host-only, reentrant (loader threads, data.lua:14-19). */
int dg_t7_read_record(const char *path, uint8_t *raw, int *x, int *y, int *player, int ranks[2]);

/* ---- checkpoints (experiment:save / Experiment:load, experiments.lua:55-72,124-131) -------------------------------- */
/* One format for every host of this library: "DGCKPT02" + key=value lines (numLayers, channelSize, kFirst, inPlanes, P written here,
 * then the caller's own lines `meta`: rate, rateDecay, iterations, ..) + "end" + P float32 parameters in Torch7 getParameters
 * order.  dg_checkpoint_read returns the header lines in meta (may be NULL), the count in *n_out and, when params is not NULL, the
 * parameters.  Host only apart from the parameter read-back. */
int dg_checkpoint_save(dg_model *m, const char *path, const char *meta);
int dg_checkpoint_read(const char *path, char *meta, size_t meta_cap, float *params, size_t cap, size_t *n_out);
/* A checkpoint written by the REFERENCE (torch.save of the pickled experiment, experiments.lua:124-127): extracts the flat
 * parameter vector (field modelParameters, experiments.lua:107; Float, Cuda or Double tensor) and, where present, optimizer.rate,
 * optimizer.rate_decay (optimizer.lua:18-21) and iterations.  params may be NULL to query the count.  Host only, reentrant. */
int dg_t7_read_model(const char *path, float *params, size_t cap, size_t *n_out, double *rate, double *rate_decay, long long *iterations);

/* ---- packed corpus + native minibatch loader -------------------------------- */
/* The reference keeps one Torch7 file per position and loads a minibatch with n torch.load calls in a pool of Lua threads
 * (dataloader.lua:113-125, data.lua:14-19,29-37,82-92).  dg_corpus_pack reads n such files (nthreads host threads) and
 * writes ONE file of fixed 3,264-byte records; dg_corpus_read_batch then assembles a minibatch -- the arguments of
 * dg_model_set_batch_raw -- with n memcpys out of a read-only mapping.  Host only, reentrant per handle. */
typedef struct dg_corpus dg_corpus;
int dg_corpus_pack(const char *const *paths, long long n, const char *out_path, int nthreads);
int dg_corpus_open(const char *path, dg_corpus **out);
int dg_corpus_close(dg_corpus *c);
int dg_corpus_count(dg_corpus *c, long long *out);
/* records idx[0..n) -> raw u8[n][9][19][19], player u8[n], rank u8[n] (= ranks[player], dataloader.lua:87), and the
 * 1-based move x, y u8[n] (both NULL: not wanted) */
int dg_corpus_read_batch(dg_corpus *c, const long long *idx, int n, uint8_t *raw, uint8_t *player, uint8_t *rank,
                         uint8_t *x, uint8_t *y, int nthreads);

/* ---- SGF front end (makedata.lua:24-120) ------------------------------------- */
/* One game's SGF text -> dan ranks (get_ranks; *has_ranks = 0: the reference writes no record for such a game), handicap
 * stones (handicaps) and the move list without passes (all_moves), coordinates 1-based, player 1 = black.  Follows the
 * reference's tokenisation to the letter (lines end at CR LF only, a piece is a token only with exactly one '[' ...: see
 * csrc/sgf.cpp).  The output arrays may be NULL to count only.  Host only, reentrant.  Feed the result to dg_update_boards /
 * dg_summarize_boards (all_boards, makedata.lua:156-186: summarise BEFORE each move, then play it). */
int dg_sgf_parse(const char *sgf, size_t len, int ranks[2], int *has_ranks, uint8_t *hx, uint8_t *hy, uint8_t *hplayer,
                 int max_handicaps, int *n_handicaps, uint8_t *mx, uint8_t *my, uint8_t *mplayer, int max_moves, int *n_moves);

/* ---- board analysis (makedata.lua) --------------------------------------------------------- */
/* summarize_board (makedata.lua:143-153): stones u8[n][19][19] in {0,1,2} ->
 * out u8[n][7][19][19]: liberties, liberties_after[p=1,2], kills[p=1,2], ladders[p=1,2]. */
int dg_summarize_boards(dg_ctx *ctx, const uint8_t *stones, uint8_t *out, int n);
/* update_board (makedata.lua:329-354) for n independent boards, in place: ages 1..254 grow by one, the stone
 * (x,y 1-based, player in {1,2}) is placed, captured groups and, on suicide, the own group are removed, and every
 * point the move set gets age 1.  An occupied target is an error (makedata.lua:352); that board is left unchanged. */
int dg_update_boards(dg_ctx *ctx, uint8_t *stones, uint8_t *age, const uint8_t *x, const uint8_t *y,
                     const uint8_t *player, int n);

/* ---- model (experiments.lua:133-153 getBasicModel + :105-107 getParameters) ---------------- */
/* numLayers x [pad, conv k, +b[c], +a[c,x,y], ReLU] + log-softmax(361); kernels {kFirst,3,3,..},
 * channels {inPlanes, C,..,C, 1} (experiments.lua:88-93).  max_batch sizes the device buffers. */
int dg_model_create(dg_ctx *ctx, int num_layers, int channel_size, int in_planes, int k_first,
                    int precision, int max_batch, dg_model **out);
int dg_model_destroy(dg_model *m);
int dg_model_param_count(dg_model *m, size_t *out);
/* flat parameter / gradient vectors in Torch7 getParameters order: per layer W[cout][cin][k][k],
 * b[cout], a[cout][19][19] (experiments.lua:105-107) */
int dg_model_set_params(dg_model *m, const float *host, size_t n);
int dg_model_get_params(dg_model *m, float *host, size_t n);
int dg_model_get_grads(dg_model *m, float *host, size_t n);
/* grads:zero() (train.lua:5) */
int dg_zero_grads(dg_model *m);
/* loss-gradient scale 1/global_batch for data-parallel shards; 0 = use the local batch size */
int dg_model_set_global_batch(dg_model *m, int global_batch);

/* batch staging (cudaInput:resize():copy(), train.lua:98-104)
 *   _raw:    records + on-device encode; x,y 1-based move (label computed on device)
 *   _planes: the reference's own staging: float[n][37][19][19] planes + 0-based targets */
int dg_model_set_batch_raw(dg_model *m, const uint8_t *raw, const uint8_t *player, const uint8_t *rank,
                           const uint8_t *sym, const uint8_t *x, const uint8_t *y, int n);
int dg_model_set_batch_planes(dg_model *m, const float *planes, const int32_t *targets0, int n);
/* re-run the on-device encode of the records staged by the last dg_model_set_batch_raw (inputs already
 * resident in HBM; no host copy) */
int dg_model_encode_resident(dg_model *m);

/* model:forward (train.lua:6,28): log-probs float[n][361] to host if logprobs_out != NULL */
int dg_forward(dg_model *m, float *logprobs_out);
/* criterion:forward + criterion:backward + model:backward (train.lua:8-10); needs dg_forward first.
 * Accumulates into the gradient vector (call dg_zero_grads first, as eval() does). */
int dg_loss_backward(dg_model *m, float *loss_out);
/* SGD:step (optimizer.lua:24-25): p += -rate*g.  The rate decay (:26) stays on the host side. */
int dg_sgd_step(dg_model *m, float rate);
/* eval() + optimizer:step fused (train.lua:111,131): zero grads, forward, NLL, backward,
 * [allreduce], SGD.  loss_out may be NULL (no host sync). */
int dg_train_step(dg_model *m, float rate, float *loss_out);
/* The same step without the wait (no counterpart in the reference, whose loop is synchronous, train.lua:106-131): _begin queues
 * eval() + optimizer:step and the read-back of the cost and returns; _result waits for THAT step and returns its cost.  Between the
 * two the caller may stage the next minibatch: dg_model_set_batch_raw copies on a second stream into the other of two record
 * buffers, so the host->device copy of minibatch k+1 runs under step k (its encode and its step queue up behind step k).  One
 * step may be outstanding. */
int dg_train_step_begin(dg_model *m, float rate);
int dg_train_step_result(dg_model *m, float *loss_out);
/* preds:max(2) (train.lua:29): 0-based argmax of the last forward, int32[n] to host */
int dg_argmax(dg_model *m, int32_t *out);
/* eval_validation inner loop (train.lua:28-37): forward + top-1 errors + mean NLL of the staged batch */
int dg_eval_batch(dg_model *m, int *errors_out, float *cost_out);

/* c5 inference: 8-fold symmetry-averaged move probabilities for n records (already staged with
 * dg_model_set_batch_raw, sym ignored): probs float[n][361] (may be NULL), argmax int32[n] (may be NULL).  The staged batch
 * (symmetry ids, planes, labels) is left as the caller staged it. */
int dg_predict_sym8(dg_model *m, float *probs_out, int32_t *argmax_out);

#ifdef __cplusplus
}
#endif
#endif /* DEEPGO_H */
