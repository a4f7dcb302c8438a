/*
 * erasurehead.h -- C ABI of liberasurehead.so (sm_100a).
 *
 * Every entry point returns 0 on success, a positive cudaError_t, or a negative EH_E* code;
 * eh_last_error() gives a human-readable message for the calling thread's last failure.
 * Pointers named d_* are device pointers (possibly peer-mapped), h_* are host pointers.
 * All launches go to the cudaStream_t passed as `stream` (a void* here so the header needs no
 * CUDA include); nothing synchronises unless documented.
 *
 * Reference citations (`file:line`) point into Distributed-Deep-Learning/ErasureHead.
 */
#ifndef ERASUREHEAD_H_
#define ERASUREHEAD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EH_OK            0
#define EH_EINVAL       -1   /* bad argument / unsupported shape */
#define EH_ETIMEOUT     -2   /* device-side watchdog fired */
#define EH_ENOTSUP      -3
#define EH_ESTATE       -4

/* model ids (erasurehead_b200/models/base.py) */
#define EH_MODEL_LOGISTIC 0
#define EH_MODEL_LINEAR   1

/* decode kinds (erasurehead_b200/schemes.py) */
#define EH_DECODE_SUM             0
#define EH_DECODE_FIRST_PER_GROUP 1
#define EH_DECODE_LSTSQ           2

/* message kinds */
#define EH_MSG_CODED   0
#define EH_MSG_PRIVATE 1

#define EH_MAX_WORKERS 64
#define EH_MAX_MSGS     2
#define EH_MAX_PEERS   16

const char* eh_last_error(void);
int  eh_version(void);
/* SM count / co-resident CTA capacity of the current device for the persistent kernels. */
int  eh_device_info(int* sm_count, int* grad_dense_ctas, int* decoder_ctas);

/* ------------------------------------------------------------------------------------------
 * Workspace: one cudaMalloc'd block per GPU that peers map through CUDA IPC.  Holds the
 * double-buffered parameters, the gradient-message slots and all flags.
 * Replaces the reference's mpi4py Isend/Irecv buffers (naive.py:68-79).
 * ---------------------------------------------------------------------------------------- */
typedef struct eh_ws_layout {
    uint64_t bytes;          /* total size */
    uint64_t off_ctl;        /* eh control words (see csrc/workspace.cuh) */
    uint64_t off_beta;       /* float  [2][d_pad]                      */
    uint64_t off_flags;      /* uint32 [n_workers][EH_MAX_MSGS], padded to 128 B each */
    uint64_t off_slots;      /* float  [n_workers][n_msgs][2][d_pad]   */
    uint32_t n_workers, n_msgs, d_pad, reserved;
} eh_ws_layout;

int eh_ws_layout_make(int n_workers, int n_msgs, int d, eh_ws_layout* out);
int eh_ws_create(const eh_ws_layout* lay, void** d_ws, unsigned char ipc_handle[64]);
int eh_ws_open(const unsigned char ipc_handle[64], void** d_ws_peer);
int eh_ws_close(void* d_ws_peer);
int eh_ws_destroy(void* d_ws);
int eh_ws_reset(const eh_ws_layout* lay, void* d_ws, void* stream);
/* Device addresses inside a workspace block (for transfers issued by the host program). */
void* eh_ws_slot_ptr(const eh_ws_layout* lay, void* d_ws, int worker, int msg, int parity);   /* float[d_pad] */
void* eh_ws_beta_ptr(const eh_ws_layout* lay, void* d_ws, int parity);                        /* float[d_pad] */
/* Bridge to GPUs outside this NVLink domain (other nodes): messages and parameters arrive through
 * stream-ordered NCCL point-to-point receives into the slots above; these calls, enqueued behind the
 * receive on the same stream, publish what a peer store would have published -- the arrival record of
 * (worker, msg) for round stamp-1, stamped with THIS GPU's clock (decoder clock offset of such a worker
 * is 0), and the parameter flag of the receiving GPU. */
int eh_bridge_arrival(const eh_ws_layout* lay, void* d_ws, int worker, int msg, uint32_t stamp, void* stream);
int eh_bridge_beta_flag(const eh_ws_layout* lay, void* d_ws, uint32_t stamp, void* stream);
/* Read the sticky device error word (0 = none) -- set by any watchdog. */
int eh_ws_error(const eh_ws_layout* lay, void* d_ws, int* err, void* stream);

/* ------------------------------------------------------------------------------------------
 * K1+K2'+K3 fused: encoded partial gradient of one worker message, dense fp32 features.
 *   g = sum_segs coeff[seg] * sum_rows  X_seg[r,:] * resid(X_seg[r,:].beta, y_seg[r])
 * logistic: resid = -y/(exp(p*y)+1) (naive.py:137-139, encode fold coded.py:92-95,184);
 * linear:   resid = -2 (y-p)        (naive.py:345-346).
 * X is streamed once from HBM by TMA bulk copies into a shared-memory ring; the row tile
 * stays on chip between the forward dot and the rank-1 accumulate.  Per-CTA partial sums are
 * combined through `d_scratch` (>= eh_grad_dense_scratch_bytes) by a device-wide barrier and
 * the result is written to `d_out` (local or peer memory).  If `d_flag` is non-NULL the kernel
 * first waits for the parameter flag (`d_beta_flag`, value >= stamp) and, once `d_out` is
 * complete, publishes `stamp` to `d_flag` with release/system semantics.  `d_flag` addresses a
 * 16-byte ARRIVAL RECORD (8-byte aligned; 128 bytes apart in the workspace): word 0 receives the
 * stamp, bytes 8..15 the publishing GPU's %globaltimer at completion -- every gradient entry point
 * below follows this contract.
 * ---------------------------------------------------------------------------------------- */
#define EH_MAX_SEGS 16

typedef struct eh_grad_job {
    const void*     d_X[EH_MAX_SEGS];   /* per segment: [rows_per_seg, ld] row-major fp32 (or bf16), 16B aligned */
    const float*    d_y[EH_MAX_SEGS];   /* per segment: [rows_per_seg] labels / targets */
    float           coeff[EH_MAX_SEGS]; /* per segment encode weight (B[w, partition] or 1) */
    const float*    d_beta;         /* [d] parameters (buffer of this round) */
    float*          d_out;          /* [d] destination (may be peer memory) */
    uint32_t*       d_flag;         /* arrival flag to publish (may be peer memory) or NULL */
    const uint32_t* d_beta_flag;    /* local parameter flag to wait on, or NULL */
    uint32_t*       d_err;          /* sticky error word or NULL */
    int32_t         n_segs;         /* partitions streamed for this message (<= EH_MAX_SEGS) */
    int32_t         rows_per_seg;
    int32_t         d, ld;          /* ld % 4 == 0, ld >= d, columns d..ld must be zero */
    int32_t         model;
    uint32_t        stamp;          /* round + 1 */
    uint64_t        timeout_ns;     /* watchdog for the beta wait (0 = 30 s) */
    int32_t         x_bf16;         /* 0: d_X are fp32 rows; 1: bf16 rows (ld in elements, ld % 8 == 0) */
    int32_t         abandon;        /* 1: while streaming, watch d_beta_flag; once it passes `stamp` (the round this
                                       message belongs to has been decided without it) stop reading X and publish
                                       nothing -- a straggler's compute is skipped, not just its message.  The
                                       reference can only cancel the send (coded.py:178-180, README.md:122) */
} eh_grad_job;

size_t eh_grad_dense_scratch_bytes(int d);
int eh_grad_dense(const eh_grad_job* job, void* d_scratch, size_t scratch_bytes, void* stream);
/* Up to 8 messages of identical shape (same d, ld, rows_per_seg, model, parameters, stamp; at most 32
 * partitions in total) in ONE launch: the CTAs are split between the messages, each message is
 * reduced, shipped and flagged separately.  Used by the round loop for GPUs hosting several workers. */
int eh_grad_dense_multi(const eh_grad_job* jobs, int n_jobs, void* d_scratch, size_t scratch_bytes, void* stream);

/* Same kernel for callers that cannot allocate device memory (fixed-latency loops that forbid
 * cudaMalloc after start-up, processes whose HBM is fully committed, plain C hosts without an
 * allocator): the scratch lives in a file-scope device array inside the library (2.5 MiB per
 * GPU, held until exit).  No counterpart in the reference.  Limits: ld <= 4096 floats
 * (eh_grad_dense_noalloc_max_ld()), devices with <= 160 SMs; larger shapes return EH_EINVAL.
 * The first call on each device resolves the array's address and triggers the runtime's one-time
 * module load for that device (make it at start-up, before memory is fully committed and before
 * any stream capture); calls sharing a device must be stream-ordered -- overlapping calls race on
 * the array.  The entry point itself calls no cudaMalloc/cudaFree and no allocating library routine. */
size_t eh_grad_dense_noalloc_max_ld(void);
int eh_grad_dense_noalloc(const eh_grad_job* job, void* stream);

/* One-hot ELL features (every value 1.0; arrange_real_data.py:41-64): gather-sum margins,
 * scatter-add gradient with the always-on bias column reduced in registers. */
typedef struct eh_grad_ell_job {
    const int32_t*  d_idx[EH_MAX_SEGS]; /* per segment: [rows_per_seg, nnz] column ids */
    const float*    d_y[EH_MAX_SEGS];
    float           coeff[EH_MAX_SEGS];
    const float*    d_beta;
    float*          d_out;
    uint32_t*       d_flag;
    const uint32_t* d_beta_flag;
    uint32_t*       d_err;
    int32_t         n_segs, rows_per_seg;
    int32_t         d, nnz;
    int32_t         model;
    int32_t         bias_col;       /* -1 if none */
    uint32_t        stamp;
    uint32_t        reserved;
    uint64_t        timeout_ns;
} eh_grad_ell_job;

size_t eh_grad_ell_scratch_bytes(int d);
int eh_grad_ell(const eh_grad_ell_job* job, void* d_scratch, size_t scratch_bytes, void* stream);
/* Up to 16 one-hot messages of identical shape (same d, nnz, rows_per_seg, model, bias_col, parameters, stamp;
 * at most 64 partitions in total) in ONE launch: CTA ranges per message, per-CTA partial gradients in shared
 * memory when the gradient fits, else one accumulation vector per message in scratch. */
int eh_grad_ell_multi(const eh_grad_ell_job* jobs, int n_jobs, void* d_scratch, size_t scratch_bytes, void* stream);

/* Multi-column model on one-hot ELL features: softmax regression with C <= 8 classes (model 2, labels
 * 0..C-1) or C independent logistic columns (model 3, labels +-1) -- "Covertype-shaped multi-class" at
 * the one-hot shape 396112 x 15509 (BASELINE.json config 5; the reference binarises Covertype,
 * arrange_real_data.py:154-156, so the model has no reference counterpart).  Jobs are eh_grad_ell_job
 * with d_beta / d_out in COLUMN-major layout W[d][8] (eh_grad_ell_mc_padded_classes() floats per
 * feature: one 32-byte sector) -- a flat vector of 8 d floats for the decoder; bias_col is ignored.
 * Up to 16 messages of one shape per launch (CTA ranges per message, one accumulation vector each in
 * scratch >= eh_grad_ell_mc_scratch_bytes(d)); scatter by 16-byte L2 reductions with a per-lane
 * run-length cache (csrc/grad_ell_mc.cu). */
int eh_grad_ell_mc_padded_classes(void);
size_t eh_grad_ell_mc_scratch_bytes(int d);
int eh_grad_ell_mc_multi(const eh_grad_ell_job* jobs, int n_jobs, int classes, void* d_scratch, size_t scratch_bytes,
                         void* stream);

/* Multi-column models on the tensor cores (tcgen05.mma kind::tf32 for both products): softmax
 * regression with C classes (model 2, labels 0..C-1) and C independent logistic columns on the
 * same data (model 3, labels +-1).  Parameters / gradients use the class-major layout
 * W[CN][ldw] (CN = 16 for C <= 16, else 32; ldw = d padded to 32) so a C-column model is one flat
 * vector of CN*ldw floats for the decoder.  d_out (may be peer memory) receives the encoded
 * gradient, d_loss (optional, zeroed by the caller) the summed loss.  No counterpart in the
 * reference (single-vector models only); BASELINE.json config "Covertype-shaped multi-class".
 *
 * Kernel selection for a gradient call (d_out set, d_loss NULL): d <= 256 -> one pass, one launch
 * (warp-level MMA forward + tcgen05 backward from one shared-memory tile, grad_mc_fused.cu);
 * d <= 1024 -> the same on thread-block clusters (grad_mc_cluster.cu); wider, or with d_loss ->
 * two tcgen05 passes (grad_mc_tc.cu).  bf16 feature rows (x_bf16, d <= 640): both products on
 * tcgen05 from one shared-memory tile (grad_mc_bf16.cu).  The scratch buffer must be zero-filled when it is
 * allocated (the one-pass kernels keep their barrier words in it and leave them zero). */
typedef struct eh_mc_job {
    const float*    d_X[EH_MAX_SEGS];
    const float*    d_y[EH_MAX_SEGS];
    float           coeff[EH_MAX_SEGS];
    const float*    d_W;            /* [CN, ldw] parameters of this round */
    float*          d_out;          /* [CN, ldw] destination or NULL (loss only) */
    double*         d_loss;         /* [1] or NULL */
    uint32_t*       d_flag;
    const uint32_t* d_beta_flag;
    uint32_t*       d_err;
    int32_t         n_segs, rows_per_seg;
    int32_t         d, ld, ldw, C;
    int32_t         model;
    uint32_t        stamp;
    uint64_t        timeout_ns;
    int32_t         x_bf16;         /* d_X are bf16 rows (ld in elements, ld % 8 == 0): all-tcgen05 one-pass kernel, d <= 640 */
    int32_t         reserved;
} eh_mc_job;

size_t eh_grad_mc_scratch_bytes(int64_t rows_per_seg, int d, int C);
int eh_grad_mc(const eh_mc_job* job, void* d_scratch, size_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Decoder round (rank 0): poll arrival flags, apply the injected per-message delay, evaluate
 * the scheme stop rule after every arrival, derive decode weights (sum / first-per-group /
 * least-squares row by table lookup or an in-kernel solve), combine the arrived slots, apply
 * the GD or AGD update and push the new parameters + flag to every peer -- one kernel.
 * Reference: master loop approximate_coding.py:136-180, decode coded.py:147-149,
 * update naive.py:112-124.
 * ---------------------------------------------------------------------------------------- */
typedef struct eh_decoder_desc {
    /* scheme */
    int32_t  n_workers, n_msgs, n_groups, s;
    int32_t  need_private, need_count, need_groups, decode;
    int32_t  d, d_pad;
    int32_t  update_agd;            /* 0 = GD, 1 = AGD */
    int32_t  n_peers;               /* GPUs (including self) receiving the parameter push */
    const int32_t* d_group_of;      /* [n_workers] */
    const double*  d_B;             /* [n_workers, n_workers] row-major or NULL */
    const double*  d_A;             /* decode table [C(n,s), n] or NULL (-> in-kernel solve) */
    const int64_t* d_binom;         /* [(n+1), (n+2)] binomials for the combinadic rank or NULL */
    /* per-round tables, indexed by local round r = round - first_round */
    const uint64_t* d_delay_ns;     /* [rounds, n_workers]; UINT64_MAX = never */
    const double*   d_eta;          /* [rounds] learning rates */
    /* state (rank 0 memory) */
    double*  d_beta64;              /* [d] master copy */
    double*  d_u64;                 /* [d] AGD state */
    void*    d_ws_local;            /* rank 0 workspace */
    void*    d_ws_peer[EH_MAX_PEERS]; /* every rank's workspace as mapped on rank 0 ([0] = local) */
    /* outputs */
    double*  d_betaset;             /* [rounds, d] or NULL */
    double*  d_timeset;             /* [rounds] seconds */
    double*  d_worker_timeset;      /* [rounds, n_workers] seconds, -1 if not collected */
    double*  d_weights;             /* [rounds, n_workers] decode weights applied */
    int32_t* d_ncollected;          /* [rounds] */
    uint64_t* d_tstart;             /* [1] round-start timestamp carried between launches */
    void*    d_scratch;             /* >= eh_decoder_scratch_bytes() */
    float*   d_beta_mc;             /* optional NVLS multicast address of a [2][d_pad] parameter buffer that is
                                       mapped on every GPU: the push becomes ONE multimem.st per 16 bytes instead
                                       of n_peers stores (NULL = peer stores into each workspace) */
    double   alpha, n_rows, lr_scale;
    uint64_t timeout_ns;            /* watchdog for one round (0 = 60 s) */
    eh_ws_layout lay;
    const int64_t* d_clock_off;     /* [n_workers] clock of the GPU hosting worker w minus this GPU's clock, ns
                                       (eh_clock_sync_root); NULL = one clock */
    double*  d_done;                /* optional output [rounds, n_workers]: seconds from round start to the
                                       completion stamp of w's coded message (before the injected delay) */
    int32_t  dec_ctas;              /* grid size cap (0 = one CTA per SM); eh_decoder_ctas(d_pad) when the decoder
                                       runs next to gradient kernels that leave that many SMs free */
    int32_t  reserved;
} eh_decoder_desc;

size_t eh_decoder_scratch_bytes(void);
int eh_decoder_ctas(int d_pad);
/* SMs of the current device that gradient kernels leave free (for a concurrent decoder). */
int eh_set_reserved_sms(int n);
int eh_worker_sms(void);
/* Clock alignment between the root GPU and one peer (run both sides concurrently, root on the
 * root's stream, peer on the peer's; `base` distinguishes exchanges).  Results (device scalars on
 * the root): peer clock minus root clock and the best round trip, in ns. */
int eh_clock_sync_root(const eh_ws_layout* lay, void* d_ws_local, void* d_ws_peer, uint32_t base, int iters,
                       long long* d_offset_ns, long long* d_rtt_ns, void* stream);
int eh_clock_sync_peer(const eh_ws_layout* lay, void* d_ws_local, void* d_ws_root, uint32_t base, int iters,
                       void* stream);
/* round = global round index (AGD's theta uses it); r = index into the per-round tables. */
int eh_decoder_round(const eh_decoder_desc* desc, int round, int r, void* stream);
/* Publish beta_0 (from d_beta64) to every peer and open round `round` (flag = round+1). */
int eh_decoder_publish(const eh_decoder_desc* desc, int round, void* stream);

/* ------------------------------------------------------------------------------------------
 * Native round loop: enqueue `n_rounds` complete rounds on `stream` without returning to the
 * host in between.  Per round and per local worker message: one fused gradient kernel that
 * waits for this round's parameters, computes the encoded gradient and ships it into the
 * root's slot; on the root rank additionally one decoder kernel.  Plays the role of the
 * `for i in range(rounds)` loop of every reference driver (approximate_coding.py:123-207).
 * ---------------------------------------------------------------------------------------- */
typedef struct eh_worker_job {
    int32_t  worker;                 /* global worker id (slot row on the root) */
    int32_t  msg;                    /* EH_MSG_CODED / EH_MSG_PRIVATE */
    int32_t  sparse;                 /* 0: dense fp32 rows, 1: one-hot ELL, 2: dense multi-column (tensor cores), 3: dense bf16 rows,
                                        4: dense multi-column with bf16 rows (d <= 640), 5: one-hot ELL multi-column (<= 8 classes) */
    int32_t  model;
    int32_t  n_segs, rows_per_seg;
    int32_t  d, ld_or_nnz;           /* dense: row stride; ELL: ones per row */
    int32_t  bias_col, classes;      /* classes: columns of a multi-column model (kind 2) */
    const void*  d_X[EH_MAX_SEGS];   /* float* rows or int32* indices */
    const float* d_y[EH_MAX_SEGS];
    float    coeff[EH_MAX_SEGS];
} eh_worker_job;

typedef struct eh_round_plan {
    int32_t  n_jobs;
    int32_t  abandon;                /* dense single-column messages: see eh_grad_job.abandon */
    const eh_worker_job* jobs;       /* host array */
    void*    d_ws_local;             /* this GPU's workspace (parameters + parameter flag) */
    void*    d_ws_root;              /* the root's workspace as mapped here (slots + arrival flags) */
    void*    d_scratch;              /* gradient-kernel scratch (max of the per-job needs) */
    const float* d_beta_local;       /* optional: this GPU's view of the multicast parameter buffer [2][d_pad]
                                        (workers read parameters here instead of the workspace) */
    size_t   scratch_bytes;
    const eh_decoder_desc* decoder;  /* non-NULL on the root rank only */
    uint64_t timeout_ns;
    eh_ws_layout lay;
    void*    dec_stream;             /* optional (root): stream of the decoder kernels.  When set and different
                                        from `stream`, the decoder of round i runs CONCURRENTLY with the gradient
                                        kernels of round i of this GPU (fork/join by events inside eh_run_rounds);
                                        call eh_set_reserved_sms(desc.dec_ctas) first so that it always finds SMs */
    void*    ev_fork;                /* cudaEvent_t pair owned by the caller (cudaEventDisableTiming) */
    void*    ev_join;
} eh_round_plan;

/* r0 = index of `first_round` in the decoder's per-round tables. */
int eh_run_rounds(const eh_round_plan* plan, int first_round, int n_rounds, int r0, void* stream);
/* Same work captured into ONE CUDA graph (stream capture of the enqueue above) and launched once:
 * the host pays one graph launch instead of n_rounds * kernels-per-round launches and the kernels of
 * consecutive rounds are dependency-linked on the device.  *graph_exec receives the instantiated
 * graph; release it with eh_graph_destroy after the stream has been synchronised. */
int eh_run_rounds_graph(const eh_round_plan* plan, int first_round, int n_rounds, int r0, void* stream,
                        void** graph_exec);
int eh_graph_destroy(void* graph_exec);

/* ------------------------------------------------------------------------------------------
 * Evaluation (reference epilogue naive.py:154-209, util.py:136-141, sklearn roc_curve+auc).
 * ---------------------------------------------------------------------------------------- */
/* margins P[rows, R] = X[rows, d] . Bt[R, d]^T for all stored iterates at once. */
int eh_margins_dense(const float* d_X, int64_t rows, int d, int ld, const float* d_Bt, int R,
                     int ldb, float* d_P, void* stream);
int eh_margins_ell(const int32_t* d_idx, int64_t rows, int nnz, const float* d_Bt, int R, int ldb,
                   float* d_P, void* stream);
/* Tensor-core version (tcgen05.mma kind::tf32, TMEM accumulators, TMA tensor-map loads): one pass
 * over X evaluates all R iterates; d_loss_sums[R] (zeroed by the caller) receives the per-column
 * SUM of log(1+exp(-y p)) (kind 0) or (y-p)^2 (kind 1); d_P (optional, [rows, ldp]) receives the
 * margins.  X and Bt rows must be zero-padded to a multiple of 32 floats (ld, ldb >= that).
 * Reference: the per-iterate evaluation loop naive.py:186-198 + util.py:136-141. */
int eh_eval_tc(const float* d_X, int64_t rows, int d, int ld, const float* d_Bt, int R, int ldb,
               const float* d_y, int kind, double* d_loss_sums, float* d_P, int ldp, void* stream);
/* column-wise loss of P[rows, R] against y: kind 0 = logistic, 1 = squared error; out[R] = sum.
 * d_out must be zeroed by the caller; accumulation is in double. */
int eh_loss_columns(const float* d_P, const float* d_y, int64_t rows, int R, int kind,
                    double* d_out, void* stream);
/* tie-aware ROC-AUC of one score column (labels +1/-1): device radix sort (descending) + scans,
 * trapezoid over distinct thresholds == sklearn auc(roc_curve(y, score)) (naive.py:196-197). */
int eh_auc(const float* d_scores, const float* d_labels, int64_t n, double* d_auc, void* d_scratch,
           size_t scratch_bytes, void* stream);
size_t eh_auc_scratch_bytes(int64_t n);
/* The same for R score columns at once (d_scores is [R][n], one column per stored iterate): ONE device-wide
 * sort of (column, score) keys and two scans instead of R sort + scan sequences; d_auc receives R values. */
int eh_auc_batch(const float* d_scores, const float* d_labels, int64_t n, int R, double* d_auc, void* d_scratch,
                 size_t scratch_bytes, void* stream);
size_t eh_auc_batch_scratch_bytes(int64_t n, int R);

/* ------------------------------------------------------------------------------------------
 * Data / code generation on the device.
 * ---------------------------------------------------------------------------------------- */
/* GMM features + logistic labels, counter-based (Philox) so that any (partition,row,col) is
 * reproducible on any GPU (generate_data.py:23-45, util.py:39-47). */
/* task 0: labels +-1 ~ Bernoulli(sigmoid(x.beta_star)); 1: real targets (x.beta_star) / sqrt(d) + noise; 2: class ids
 * 0..classes-1 = argmax_k(0.3 x.w_k + Gumbel), w_k in {-1,+1}^d (classes <= 8; multi-class extension). */
int eh_synth_gmm(float* d_X, float* d_y, int64_t rows, int d, int ld, uint64_t seed,
                 uint64_t partition, int task, int classes, void* stream);
/* One-hot ELL data of a named dataset's shape (arrange_real_data.py:41-64,151-178,211-221 builds this from
 * the real files): idx[rows, nnz] int32 with one feature per categorical block [d_edges[f], d_edges[f+1])
 * for f < nnz-1 (heavy-tailed level popularity) and the always-on bias feature d-1 last; labels as above
 * from hidden N(0, 0.5^2) parameters per feature (and class). */
int eh_synth_onehot(int32_t* d_idx, float* d_y, int64_t rows, int d, int nnz, const int32_t* d_edges,
                    uint64_t seed, uint64_t partition, int task, int classes, void* stream);
/* Cyclic-MDS code B from the parity matrix H (util.py:64-83): one warp per row. */
int eh_code_cyclic(const double* d_H, int n, int s, double* d_B, void* stream);
/* Decode table A (util.py:85-103): one warp per straggler pattern. */
int eh_code_decode_table(const double* d_B, int n, int s, const int32_t* d_missing,
                         int64_t n_patterns, double* d_A, void* stream);

/* Native reader of the reference's dense text partitions (util.py:13-15 np.loadtxt of util.py:26-30 files):
 * memory-mapped, parsed on all cores, float32 or float64 output with a caller-chosen row stride (host
 * memory, e.g. the pinned staging buffer of a partition).  eh_text_shape counts rows (non-blank lines)
 * and columns (numbers on the first line). */
int eh_text_shape(const char* path, int64_t* rows, int64_t* cols);
int eh_load_dense_text(const char* path, int64_t rows, int64_t cols, void* h_out, int64_t ld, int elem_bytes);
/* Native writer of the same format (util.py:26-30: " ".join(str(x)) per row): every value as Python's repr of a
 * double (shortest round-trip digits; positional for 1e-4 <= |x| < 1e16, else d[.ddd]e+XX), formatted on all cores,
 * byte-identical to the reference's files.  eh_format_repr writes one value (NUL-terminated, < 32 bytes) and
 * returns its length. */
int eh_save_dense_text(const char* path, const double* h_data, int64_t rows, int64_t cols, int64_t ld);
int eh_format_repr(double v, char* out32);

/* One-hot feature construction (K15): idx[r,c] = offsets[c] + codes[r,c]; pair features as the
 * mixed-radix id of two label-encoded columns (arrange_real_data.py:41-64, util.py:49-55). */
int eh_onehot_assemble(const long long* d_codes, const long long* d_offsets, int64_t rows, int cols,
                       int32_t* d_idx, void* stream);
int eh_pair_codes(const long long* d_codes, int64_t rows, int cols, const int32_t* d_pairs, int n_pairs,
                  const long long* d_radix, long long* d_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ERASUREHEAD_H_ */
