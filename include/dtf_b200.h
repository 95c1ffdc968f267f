/* C ABI of libdtf_b200.so — the hand-written sm_100a kernels and the native host runtime
 * of distributedtf_b200.  All device pointers are raw CUDA pointers; `stream` is a
 * cudaStream_t.  Every function returns DTF_OK (0) or a negative DTF_ERR_* code and never
 * allocates device memory unless its comment says so.  Python binds these with ctypes
 * (distributedtf_b200/ops/native.py); the struct layouts below are mirrored there.
 *
 * Reference parity is cited per entry point (paths relative to the reference repository).
 */
#ifndef DTF_B200_H
#define DTF_B200_H

#include <stdint.h>

#ifdef __cplusplus
#include <cuda_bf16.h>
#include <cuda_runtime.h>
typedef __nv_bfloat16 dtf_bf16;
extern "C" {
#else
typedef uint16_t dtf_bf16;
typedef void* cudaStream_t;
#endif

enum { DTF_OK = 0, DTF_ERR_SHAPE = -1, DTF_ERR_SMEM = -2, DTF_ERR_CUDA = -3, DTF_ERR_ARG = -4,
       DTF_ERR_LIMIT = -5, DTF_ERR_IO = -6 };

/* operand prologues (applied while gathering the im2col tile) */
enum { DTF_PRO_NONE = 0, DTF_PRO_BN_RELU = 1, DTF_PRO_BN_BWD = 3 };
/* epilogues */
enum { DTF_EPI_STORE = 0, DTF_EPI_RESIDUAL = 1, DTF_EPI_BIAS_RELU = 2, DTF_EPI_RELU_MASK = 3, DTF_EPI_BIAS = 4 };

/* A BatchNorm layer as the kernels see it.  Training statistics are raw accumulators
 * stats[0..C) = sum(x), stats[C..2C) = sum(x^2) filled by the producer's epilogue.
 * (reference: resnet/resnet_model.py:45-52, decay 0.997, eps 1e-5) */
typedef struct {
    const float* stats;
    const float* gamma;
    const float* beta;
    float* run_mean;
    float* run_var;
    float inv_count;     /* 1 / (N*H*W) of the normalised tensor */
    float eps;
    float decay;
    int channels;
    int use_running;     /* 1 = inference: normalise with run_mean / run_var */
} DtfBnRef;

/* Element type of the activation / gradient tensors and packed weights of a launch.  DTF_BF16: bf16 operands,
 * kind::f16 MMA.  DTF_F32: fp32 storage, kind::tf32 MMA (10-bit mantissa products, fp32 accumulation) -- the
 * reference's default dtype=fp32 mode (official/utils/flags/_performance.py:107) on tensor cores.  Pointers stay
 * typed dtf_bf16* in the structs; with DTF_F32 they address floats and channel counts must be multiples of 4. */
enum { DTF_BF16 = 0, DTF_F32 = 1 };

/* One gathered GEMM operand: an NHWC tensor read through a (KH x KW, stride, pad) window. */
typedef struct {
    const dtf_bf16* src;     /* [NB, Hs, Ws, C] */
    const dtf_bf16* src2;    /* DTF_PRO_BN_BWD: the BN input x (src is then t = dL/d(bn out) masked) */
    const float* bwd_stats;  /* DTF_PRO_BN_BWD: [0..C) sum(t), [C..2C) sum(t * xhat)  (= dbeta, dgamma) */
    DtfBnRef bn;
    int mode;                /* DTF_PRO_* */
    int update_running;      /* CTA 0 also updates the moving averages (training, once per BN) */
    int Hs, Ws, C;
    int KH, KW, stride, pad;
    int transposed;          /* 1 = gradient gather of a strided conv (dgrad) */
    int dtype;               /* DTF_BF16 | DTF_F32: element type of src / src2 (and of every tensor of the launch) */
    int cpt_shift;           /* filled by the launcher */
    int kchunks_real;        /* filled by the launcher */
} DtfGather;

/* out[M][N] = epilogue( gather(a)[M][K] x wpk[N][K]^T ),  M = NB*Ho*Wo.
 * Forward conv (K1/K2/K5 of SURVEY §2.5; resnet_model.py:55-92) and dgrad. */
typedef struct {
    DtfGather a;
    const dtf_bf16* wpk;     /* packed weights [K/8][N][8] (see dtf_pack_weights) */
    dtf_bf16* out;           /* [M][N] */
    const dtf_bf16* epi_res; /* DTF_EPI_RESIDUAL: tensor added to the result */
    const float* epi_bias;   /* DTF_EPI_BIAS(_RELU) */
    const dtf_bf16* epi_x;   /* DTF_EPI_RELU_MASK: pre-BN input at the output pixels */
    DtfBnRef epi_bn;         /* DTF_EPI_RELU_MASK: BN whose ReLU mask / xhat is applied */
    float* epi_stats_out;    /* optional [2][N] accumulators (sum,sumsq) or (sum t, sum t*xhat) */
    int epi_mode;
    int epi_accumulate;      /* RELU_MASK: add to what `out` already holds */
    int NB, Ho, Wo, N;
    int kchunks, kb_chunks, smem_b_off, smem_a_off;   /* filled by the launcher */
} DtfConvGemm;

/* dw[Cout][KH*KW*Cin] += sum_pixels gy'[pixel][co] * gather(x)[pixel][k]   (fp32 atomics) */
typedef struct {
    DtfGather x;             /* conv input side (window = the conv's) */
    DtfGather gy;            /* output-gradient side; only src/src2/bwd_stats/bn/mode are read */
    float* dw;
    int NB, Ho, Wo, Cout;
    int kchunks, kchunks_real, chunks_per_cta, smem_a_off;   /* filled by the launcher */
} DtfConvWgrad;

int dtf_conv_gemm(DtfConvGemm* p, cudaStream_t stream);
int dtf_conv_wgrad(DtfConvWgrad* p, cudaStream_t stream);
/* Backward of one convolution: input gradient + weight gradient.  One fused launch when the conv is
 * 3x3 / stride 1 (the two GEMMs share the SMs of a single grid), otherwise wgrad then dgrad. */
int dtf_conv_bwd(DtfConvGemm* dgrad, DtfConvWgrad* wgrad, cudaStream_t stream);

/* ---- classifier head: final BN + ReLU + global average pool + dense + softmax-CE ----------
 * forward and backward in one launch, one CTA per image (resnet_model.py:537-552,
 * resnet_run_loop.py:230-246,303; mnist_model.py:97-126 with has_bn = 0 and HW = 1). */
typedef struct {
    const dtf_bf16* x;       /* [batch, HW, C] pre-BN activations */
    const int* labels;       /* [batch] */
    const float* w;          /* dense kernel [classes, C] fp32 */
    const float* bias;       /* [classes] */
    DtfBnRef bn;
    float* loss_out;         /* += mean cross-entropy */
    int* correct_out;        /* += number of argmax hits */
    float* logits_out;       /* optional [batch, classes] */
    float* dw;               /* += d loss / d w   (training) */
    float* db;
    dtf_bf16* t_out;         /* [batch, HW, C]: dL/d(bn out) masked by the ReLU (training) */
    float* bwd_stats_out;    /* [2][C] += (sum t, sum t*xhat)  (= dbeta, dgamma) */
    float loss_scale;
    int batch, HW, C, classes;
    int has_bn, training, update_running;
    unsigned int* board_hits;  /* optional (evaluation): this member's DtfScoreSlot on the master-visible score board, a
                                  peer pointer when the master sits on another GPU.  Every hit is added to slot->hits and
                                  the batch size to slot->seen with system-scope reductions, so the eval score reaches the
                                  master without a host round trip (pbt_cluster.py:114-128 gathered it over MPI). */
    int dtype;               /* DTF_BF16 | DTF_F32: element type of x and t_out */
} DtfHead;

/* ---- optimizer ------------------------------------------------------------------------ */
enum { DTF_OPT_GD = 0, DTF_OPT_MOMENTUM = 1, DTF_OPT_ADAGRAD = 2, DTF_OPT_ADADELTA = 3, DTF_OPT_ADAM = 4,
       DTF_OPT_RMSPROP = 5 };
enum { DTF_REG_NONE = 0, DTF_REG_L1 = 1, DTF_REG_L2 = 2, DTF_REG_L1_L2 = 3 };
#define DTF_OPT_CHUNK 1024
/* worst case of the CIFAR schedule: decay_steps = 1 gives ceil(100/1) - 1 = 99 boundaries (cifar10_main.py:192-208) */
#define DTF_MAX_LR_BOUNDARIES 100

/* Device-resident hyper-parameter block of one member: explore rewrites it in place, so a
 * captured CUDA graph of the step never needs re-capture (constants.py:14-43 ranges). */
typedef struct {
    int optimizer;           /* DTF_OPT_* */
    int regularizer;         /* DTF_REG_* */
    float momentum;
    float grad_decay;
    float weight_decay;
    float inv_loss_scale;
    int n_boundaries;
    int batch_size;
    long long boundaries[DTF_MAX_LR_BOUNDARIES];
    float lr_values[DTF_MAX_LR_BOUNDARIES + 1];
    float reserved[3];
} DtfHparams;

/* One parameter tensor of the flat arena. */
typedef struct {
    int offset, count;       /* elements, in params / grads / slot0 / slot1 */
    int regularized;         /* conv kernels only (resnet_model.py:91) */
    int pack;                /* 1 = conv kernel: emit GEMM-ready bf16 copies [K/8][N][8]; 2 = dense: plain bf16 copy;
                                3 = conv kernel, fp32 operands (DTF_F32): [K/4][N][4] floats, pack offsets in floats */
    int Cout, taps, Cin;     /* OHWI dims of a conv kernel */
    int fwd_pack_off;        /* element offsets in the bf16 pack arena, -1 = none */
    int dgrad_pack_off;
    int reserved[3];
} DtfOptSegment;

int dtf_bn_bwd_apply(const dtf_bf16* t, const dtf_bf16* x, const dtf_bf16* res, dtf_bf16* out,
                     const DtfBnRef* bn, const float* bwd_stats, long long pixels, int C, cudaStream_t stream);
/* Same for tensors of element type `dtype` (DTF_BF16 | DTF_F32). */
int dtf_bn_bwd_apply_dt(const dtf_bf16* t, const dtf_bf16* x, const dtf_bf16* res, dtf_bf16* out,
                        const DtfBnRef* bn, const float* bwd_stats, long long pixels, int C, int dtype,
                        cudaStream_t stream);
int dtf_head(const DtfHead* p, cudaStream_t stream);
/* ResNet v1 (post-activation) block tail (resnet_model.py:127-168, 215-264): out = relu(bn(x) + shortcut) with
 * shortcut = none (sc NULL) | sc (bn_sc NULL: identity) | bn_sc(sc) (projection conv followed by its own BN);
 * update_running: the launch also moves the moving averages of the BN layer(s) it applies. */
int dtf_bn_act_apply(const dtf_bf16* x, const DtfBnRef* bn, const dtf_bf16* sc, const DtfBnRef* bn_sc, dtf_bf16* out,
                     long long pixels, int C, int update_running, cudaStream_t stream);
/* Backward of that tail: t_out = dy * (y > 0); stats[0..C) += sum t, stats[C..2C) += sum t * xhat(x; bn)
 * (= dbeta, dgamma of bn); the same into stats2 against (x2, bn2) when the block has a projection. */
int dtf_relu_bwd_stats(const dtf_bf16* dy, const dtf_bf16* y, dtf_bf16* t_out, const dtf_bf16* x, const DtfBnRef* bn,
                       float* stats, const dtf_bf16* x2, const DtfBnRef* bn2, float* stats2, long long pixels, int C,
                       cudaStream_t stream);
/* The same two for tensors of element type `dtype` (DTF_BF16 | DTF_F32). */
int dtf_bn_act_apply_dt(const dtf_bf16* x, const DtfBnRef* bn, const dtf_bf16* sc, const DtfBnRef* bn_sc, dtf_bf16* out,
                        long long pixels, int C, int update_running, int dtype, cudaStream_t stream);
int dtf_relu_bwd_stats_dt(const dtf_bf16* dy, const dtf_bf16* y, dtf_bf16* t_out, const dtf_bf16* x, const DtfBnRef* bn,
                          float* stats, const dtf_bf16* x2, const DtfBnRef* bn2, float* stats2, long long pixels, int C,
                          int dtype, cudaStream_t stream);
/* chunks: n_chunks x {segment index, first element} pairs; each CTA handles DTF_OPT_CHUNK elements */
int dtf_pack_weights(const float* params, dtf_bf16* pack, const DtfOptSegment* segs, const int* chunks,
                     int n_chunks, cudaStream_t stream);
/* TF-1.x Apply<Optimizer> (resnet_run_loop.py:552-586) + regulariser gradient + LR schedule +
 * bf16 re-pack in one launch.  `step` holds the 1-based index of this update. */
int dtf_optimizer_step(float* params, const float* grads, float* slot0, float* slot1, dtf_bf16* pack,
                       const DtfOptSegment* segs, const int* chunks, int n_chunks, const DtfHparams* hp,
                       const long long* step, float grad_scale, cudaStream_t stream);
int dtf_step_begin(float* zero_base, long long zero_floats, long long* step, int bump, cudaStream_t stream);
int dtf_fill(float* p, long long n, float v, cudaStream_t stream);
/* Parameter initialisation on the device (reference: tf.layers initialisers chosen per member,
 * cifar10_main.py:167-199).  One segment per tensor of the flat arena; Philox4x32-10 keyed by `seed`, counter =
 * (element / 4, segment), so a member's weights depend on (seed, layout) only. */
enum { DTF_INIT_CONST = 0, DTF_INIT_UNIFORM = 1, DTF_INIT_TRUNC_NORMAL = 2 };
typedef struct {
    int offset, count;       /* elements in params */
    int kind;                /* DTF_INIT_*: scale = the constant | half-width a of U(-a, a) | std of N(0, std) cut at 2 std */
    float scale;
    int c_pad, c_real;       /* innermost (input-channel) dim padded from c_real to c_pad: the padding is zeroed; 0 = none */
    int reserved[2];
} DtfInitSegment;
int dtf_init_params(float* params, const DtfInitSegment* segs_dev, int n_segs, int max_count, unsigned long long seed,
                    cudaStream_t stream);
/* uint8 NHWC -> bf16 NHWC (channels padded to 8); mode 0 cast, 1 standardise, 2 crop+flip+standardise
 * (cifar10_main.py:96-109).  crop_flip: optional [n][3] = (oy, ox, flip) overriding the RNG.
 * mode | DTF_PP_F32 writes fp32 instead of bf16 (same 8 channels per pixel). */
#define DTF_PP_F32 (DTF_F32 << 8)
int dtf_preprocess(const uint8_t* images, const int* indices, dtf_bf16* out, const int* crop_flip, int n,
                   int H, int W, int C, int mode, unsigned long long seed, unsigned long long counter,
                   const long long* step, cudaStream_t stream);
/* `step` (optional device step counter) is added to `counter` x batch so that the random crop / flip draw stays
 * fresh when the launch is replayed from a CUDA graph.
 * Same, with the batch gathered from a resident dataset inside the kernel: image n = images[indices[n]],
 * labels_out[n] = labels_all[indices[n]]. */
int dtf_preprocess_indexed(const uint8_t* images, const int* indices, const int* labels_all, int* labels_out,
                           dtf_bf16* out, const int* crop_flip, int n, int H, int W, int C, int mode,
                           unsigned long long seed, unsigned long long counter, const long long* step,
                           cudaStream_t stream);

/* ---- communication-bound hot paths over NVLink peer memory (csrc/comm_kernels.cu) ------------- */
#define DTF_MAX_DP 8

/* One member's cell of the score board.  hits / seen are accumulated by the evaluation kernels (dtf_head with
 * board_hits set); `epoch` is release-stored once the member's evaluation of a round is complete.  hits ==
 * DTF_SCORE_DROPPED marks a member whose loss went non-finite (the worker drops it, training_worker.py:63-80). */
typedef struct {
    unsigned int hits;
    unsigned int seen;
    unsigned long long epoch;
} DtfScoreSlot;
#define DTF_SCORE_DROPPED 0xFFFFFFFFu

/* Exploit transfer (pbt_cluster.py:145-147,168-181 did this with `cp`): copy `bytes` of member state
 * from `src` to `dst` — either may be a peer-mapped pointer — skip the hyper-parameter block at
 * `hp_offset` and write `*new_hp` (the explored values) there instead, zero `zero_region`, then
 * release-store `epoch` to `*done_flag` once every CTA has finished (arrive_counter: 4 zeroed bytes). */
int dtf_exploit_copy(void* dst, const void* src, long long bytes, long long hp_offset, const DtfHparams* new_hp,
                     void* zero_region, long long zero_bytes, unsigned long long* done_flag,
                     unsigned long long epoch, unsigned int* arrive_counter, int max_ctas, cudaStream_t stream);
/* Same, device-ordered against the producer of `src`: every CTA first waits (ld.acquire.sys, watchdog-protected)
 * until *wait_flag >= wait_value.  The winner publishes its round snapshot with dtf_exploit_copy(done_flag =
 * &snapshot_epoch[parity], epoch = round + 1); the loser passes that flag's (peer) address here, so the pull may
 * be launched before the winner's publish has finished and no host barrier sits between the two. */
int dtf_exploit_copy_ordered(void* dst, const void* src, long long bytes, long long hp_offset,
                             const DtfHparams* new_hp, void* zero_region, long long zero_bytes,
                             unsigned long long* done_flag, unsigned long long epoch, unsigned int* arrive_counter,
                             const unsigned long long* wait_flag, unsigned long long wait_value, int max_ctas,
                             cudaStream_t stream);

/* Eval-score reduction (pbt_cluster.py:114-128).  dtf_score_publish stores (*correct, total) into the
 * master-visible board slot (peer pointer allowed) followed by a release-store of `epoch` — the path of a
 * data-parallel member, whose replicas first sum their counts.  A single-GPU member lets its evaluation
 * kernels add straight into the slot (DtfHead.board_hits) and brackets them with dtf_score_reset (hits = seen
 * = 0 at the start of the evaluation) and dtf_score_release (epoch release once the last batch is in; writes
 * DTF_SCORE_DROPPED instead of the hits when *loss is not finite; loss may be NULL). */
int dtf_score_publish(DtfScoreSlot* board_slot, const int* correct, int total, unsigned long long epoch,
                      int reset_counter, int* correct_rw, cudaStream_t stream);
int dtf_score_reset(DtfScoreSlot* board_slot, cudaStream_t stream);
int dtf_score_release(DtfScoreSlot* board_slot, unsigned long long epoch, const float* loss, int dropped,
                      cudaStream_t stream);

/* Fused one-shot gradient all-reduce + optimizer update for a data-parallel member
 * (distribution_utils.py:41-47 + resnet_run_loop.py:552-586). */
typedef struct {
    float* params; float* slot0; float* slot1; dtf_bf16* pack;
    const float* peer_grads[DTF_MAX_DP];             /* gradient arena of every replica (incl. self) */
    unsigned long long* peer_ready[DTF_MAX_DP];      /* ready[] array of every replica */
    unsigned long long* peer_done[DTF_MAX_DP];       /* done[] array of every replica */
    const unsigned long long* local_ready;           /* my ready[world] */
    const DtfOptSegment* segs; const int* chunks;
    const DtfHparams* hp; const long long* step;
    unsigned int* arrive_counter;
    const unsigned long long* epoch_ctr;             /* exchange counter of the group (bumped by dtf_dp_step_begin);
                                                        lives OUTSIDE the member state, so exploit / restore never touch it */
    unsigned long long epoch;                        /* offset added to *epoch_ctr to form the exchange epoch */
    float grad_scale;
    int world, rank, n_chunks;
    unsigned long long* timing;                      /* optional [4] device cells: ns blocked on the peers' ready flags,
                                                        ns of the remote loads, ns of the kernel (all CTA 0), calls */
} DtfAllreduceUpdate;
int dtf_allreduce_update(const DtfAllreduceUpdate* p, cudaStream_t stream);
/* Step begin of a data-parallel member: waits until every peer has consumed this replica's gradients
 * of the previous exchange (done[w] >= *epoch_ctr + epoch_offset), zeroes the accumulators, bumps *step and
 * *epoch_ctr.  The exchange counter is separate from the step counter because exploit and restore overwrite
 * the step (members run different batch sizes, hence different step counts) while the flags keep counting. */
int dtf_dp_step_begin(float* zero_base, long long zero_floats, long long* step, unsigned long long* epoch_ctr,
                      const unsigned long long* local_done, int world, unsigned long long epoch_offset,
                      unsigned int* arrive_counter, cudaStream_t stream);

/* Symmetric-memory plumbing.  These allocate / map device memory (cudaMalloc, CUDA IPC). */
int dtf_symm_alloc(long long bytes, void** out);
int dtf_symm_free(void* p);
int dtf_ipc_get_handle(void* p, unsigned char* handle64);
int dtf_ipc_open_handle(const unsigned char* handle64, void** out);
int dtf_ipc_close(void* p);
int dtf_enable_peer(int peer_device);
int dtf_last_cuda_error(char* buf, int n);

/* ---- exploit transfer without any allocation ------------------------------------------------
 * (No counterpart in the reference: its exploit is a file copy run by the master.)
 * Same copy as dtf_exploit_copy for callers that cannot provide device scratch and must not
 * trigger an allocation while other members train on the same GPU: the arrival counter lives in
 * a 256-byte static __device__ array owned by the library (one per GPU context, kept until exit).
 * Limits and rules: at most DTF_NOALLOC_MAX_BYTES per call (DTF_ERR_LIMIT beyond, text from
 * dtf_noalloc_last_error()); calls on one GPU share the array, so they must not overlap; the
 * counter is cleared on `stream` at the start of each call; the first call on a device looks up
 * the array's address and should be made outside stream capture. */
#define DTF_NOALLOC_MAX_BYTES (1ll << 30)
int dtf_exploit_copy_noalloc(void* dst, const void* src, long long bytes, long long hp_offset,
                             const DtfHparams* new_hp, void* zero_region, long long zero_bytes,
                             unsigned long long* done_flag, unsigned long long epoch, cudaStream_t stream);
const char* dtf_noalloc_last_error(void);

/* ---- dense layers and pooling (MNIST member: mnist_model.py:62-126) ------------------------------
 * D[M][N] = A x B^T on tcgen05.  An operand is row-major bf16, either K-major X[i][k] (ld = row pitch,
 * K multiple of 8) or MN-major X[k][i] (the M/N extent multiple of 8).  Epilogue: + bias[n], ReLU,
 * inverted dropout with a counter-based mask (seed, *step, element index); output bf16 or fp32. */
typedef struct {
    const dtf_bf16* a; const dtf_bf16* b;
    void* out; const float* bias; const long long* step;
    long long lda, ldb, ldo;
    int M, N, K;
    int a_mn_major, b_mn_major, out_fp32, relu;
    float dropout_keep;      /* 1.0 = no dropout */
    unsigned int seed;
    /* optional split-K: `ws` = fp32 [M][N] workspace and `tile_counters` = one uint per 128x128 output tile
     * (max_tiles of them), both zero before the first call; the kernel leaves them zero again. */
    float* ws;
    unsigned int* tile_counters;
    int max_tiles;
    int k_split;             /* filled by the launcher */
} DtfGemm;
int dtf_gemm(const DtfGemm* p, cudaStream_t stream);
int dtf_maxpool_fwd(const dtf_bf16* x, dtf_bf16* y, int NB, int H, int W, int C, cudaStream_t stream);
/* gx = gy at the first maximum of each 2x2 window, masked by x > 0 (x is post-ReLU); db[c] += sum gx */
int dtf_maxpool_bwd(const dtf_bf16* x, const dtf_bf16* y, const dtf_bf16* gy, dtf_bf16* gx, float* db,
                    int NB, int H, int W, int C, cudaStream_t stream);
/* g_out = g_in * (act > 0) * scale; db[c] += column sums   (ReLU + dropout backward, bias gradient) */
int dtf_act_mask(const dtf_bf16* g_in, const dtf_bf16* act, dtf_bf16* g_out, float* db, float scale,
                 long long rows, int C, cudaStream_t stream);

/* SMs the persistent kernels occupy (default: all).  Lower it when several members train at once on one GPU
 * so that their kernels run side by side; the value in force at CUDA-graph capture is baked into the graph.
 * Returns the effective count.  (No reference counterpart: the reference trains a worker's members one by one,
 * training_worker.py:60-75.) */
int dtf_set_sm_limit(int n);
int dtf_sm_count(void);

/* A CUDA stream of its own (cudaStreamNonBlocking) on `device`.  torch.cuda.Stream() hands out streams from a pool of
 * 32 per device round-robin, so the 33rd "new" stream of a process is the first one again: with dozens of members per
 * worker, one member's graph-capture stream was another member's compute stream (or the master's board-polling
 * stream), and work of the one ended up in -- or was rejected by -- the capture of the other.  Every long-lived
 * stream of the framework comes from here instead (utils/streams.py wraps it as torch.cuda.ExternalStream). */
int dtf_stream_create(int device, cudaStream_t* out);
int dtf_stream_destroy(int device, cudaStream_t stream);

/* ---- host runtime: background checkpoint writer (csrc/host_io.cpp) -------------------------------------
 * Replaces the synchronous `saver.save(sess, ...)` the reference runs after every train call
 * (mnist_model.py / cifar10 model, via TensorFlow's C++ Saver).  A job = wait until everything queued on
 * `cuda_stream` of `device` at submit time has run (the caller's device->host copy into `host_ptr`; pass
 * device = -1 for a host-only job), write `data_path`, then `index_path`, then
 * `pointer_path`, each through a temp file + rename.  Strings are copied at submit; `host_ptr` must stay valid
 * and unmodified until the ticket has been waited for.  Returns a ticket > 0 or a DTF_ERR_* code. */
void* dtf_io_create(int threads);
long long dtf_io_submit(void* writer, const char* data_path, const void* host_ptr, unsigned long long nbytes,
                        void* cuda_stream, int device, const char* index_path, const char* index_text,
                        const char* pointer_path, const char* pointer_text);
int dtf_io_wait(void* writer, long long ticket);
int dtf_io_wait_all(void* writer);
int dtf_io_pending(void* writer);
long long dtf_io_bytes_written(void* writer);
void dtf_io_destroy(void* writer);
/* ---- host runtime: dataset readers (csrc/host_data.cpp) -----------------------------------------------------
 * CIFAR-10 binary records (label + CHW planes; cifar10_main.py:54-109) -> HWC uint8 images + int32 labels with a
 * few threads; raw file read; idx (MNIST, mnist_model.py:128-150) header parse.  Host memory only. */
int dtf_cifar_decode(const uint8_t* records, long long n, uint8_t* images, int* labels, int threads);
long long dtf_read_file(const char* path, long long offset, void* dst, long long nbytes);
int dtf_idx_header(const uint8_t* buf, long long nbytes, int* ndim, int* dims, int max_dims);
/* debugging aid: native backtrace of the faulting thread on SIGSEGV, then the previous handler */
void dtf_install_crash_handler(void);

#ifdef __cplusplus
}
#endif
#endif /* DTF_B200_H */
