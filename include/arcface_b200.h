/*
 * arcface_b200 -- C ABI of the B200-native ArcFace head (libarcface_b200.so).
 *
 * Drop-in boundary for ONE hot path of XiaoyuWant/DDP_Classification_pytorch: the ArcFace head
 * `ArcMarginProduct` (ARCFACE/arc_main.py:130-176) plus the `nn.CrossEntropyLoss` that consumes it
 * (ARCFACE/arc_main.py:245, applied at :327 and :369) and their autograd backward (:329).
 * The reference has no FFI of its own (it is eager PyTorch); these entry points are what a binding
 * for this path would call -- see INTEGRATION.md for the ctypes stub that the Python module
 * `ddp_classification_pytorch_b200.ArcMarginProduct` uses.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless stated otherwise; the caller owns every buffer,
 *     including workspaces (sizes from the *_bytes queries); nothing is allocated or freed here;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, no host sync;
 *   - return value 0 = success, negative = error; arcb_last_error() returns a thread-local message;
 *   - matrices are row-major fp32 with an explicit leading dimension where one is given;
 *   - labels are int64 GLOBAL class ids; a shard owns classes [class_offset, class_offset + Cs);
 *   - there is no process-wide mode state: whatever selects a code path is the `flags` argument of the call
 *     (ARCB_* bits below, 0 = the library's defaults), so two heads in one process may differ and every *_bytes
 *     query is a pure function of its arguments.  The library may be called from several host threads (forward on
 *     the main thread, backward on autograd's); its one-time per-device caches are thread-safe.
 */
#ifndef ARCFACE_B200_H_
#define ARCFACE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ARCB_OK 0
#define ARCB_ERR_INVALID (-1)   /* bad argument (null pointer, size, alignment) */
#define ARCB_ERR_CUDA (-2)      /* a CUDA call / launch failed */
#define ARCB_ERR_WORKSPACE (-3) /* workspace / scratch too small */
#define ARCB_ERR_TIMEOUT (-4)   /* a kernel was aborted by the device-side pipeline watchdog */

/* ---- per-call flags (OR them; 0 = defaults) ---- */
/* arcb_shard_forward / _prenorm: where F.normalize(self.weight) (arc_main.py:159) runs.  Default: inside the forward
 * GEMM kernel for the shapes where that is faster on B200 (512 <= B <= 1024, >= 65536 classes; extra warps normalise the
 * rows the coming tiles need, so the HBM-bound pass overlaps the tensor-bound one), else in its own kernel first.
 * Same w_hat / inv_norm bits either way. */
#define ARCB_FWD_NORM_SEPARATE 0x1 /* always its own kernel before the GEMM */
#define ARCB_FWD_NORM_FUSED 0x2    /* inside the GEMM kernel whenever the shape allows (D % 4 == 0, D <= 2048) */
#define ARCB_FWD_PAIR 0x4          /* forward GEMM on CTA pairs (cta_group::2); default: one CTA per 128 x 256 tile */
#define ARCB_FWD_STREAMING 0x8     /* never use the x_hat-resident schedule (default for Dp <= 512 and many classes: a
                                      CTA keeps its 128 samples in shared memory and only w_hat streams) */
#define ARCB_FWD_BLOCK_MAXIMA 0x200 /* keep the per-128-class block maxima: REQUIRED by a following arcb_shard_top3 */
/* arcb_shard_backward: schedule.  Default: from stored probabilities where that applies (below), else the recompute
 * schedule: ONE persistent cooperative kernel whose CTA pairs (cta_group::2, 256 x 256 tile per pair of SMs) claim G / dX /
 * dW tiles from the global tile list as they go.  The next three flags pick other variants of the recompute schedule (and
 * imply ARCB_BWD_RECOMPUTE).  All schedules give the same gradients within the parity tolerance. */
#define ARCB_BWD_PER_CHUNK 0x10    /* recompute, one launch per GEMM and class chunk (no co-residency needed) */
#define ARCB_BWD_SINGLE_CTA 0x20   /* recompute, persistent kernel with one CTA per 128 x 256 tile */
#define ARCB_BWD_STATIC_SHARE 0x40 /* recompute, persistent kernel, fixed share of the tile list per CTA pair (A/B runs) */
#define ARCB_BWD_NO_COOP 0x80      /* PROFILING ONLY: default kernels launched without the cooperative attribute (ncu cannot
                                      replay cooperative + cluster launches).  Safe only on an otherwise idle GPU, where
                                      one CTA per SM is co-resident anyway; never set it in production. */
/* Stored probabilities (the default where it applies).  The forward keeps the softmax numerators
 * P' = bf16(2^(s log2e (cos - 1))) of the whole shard in the workspace, class-major ([Cs, B] bf16: 2 GB at B = 1024,
 * Cs = 1M), and the backward runs as two plain GEMMs on them (dX, dW) instead of recomputing the logits tile by tile:
 * 6 B C D executed flops instead of 8 B C D, paid for with 2 B C bytes of HBM per pass.  Applies for 256 < D <= 512,
 * s <= 41, B >= 128 and B * Cs >= 2^24 logits per shard (arcb_keep_probs_active tells); elsewhere, and whenever one of the
 * recompute-schedule flags above or ARCB_FWD_PAIR is given, the backward recomputes.  Same gradients within the parity
 * tolerance either way.  arcb_ws_bytes, arcb_shard_forward* and arcb_shard_backward must be given the SAME flags: the
 * workspace layout and what the forward leaves for the backward depend on them (a backward that expects probabilities the
 * forward did not store is caught by a device-side check: its first kernel traps, arcb_stream_check reports it). */
#define ARCB_BWD_RECOMPUTE 0x100 /* never keep probabilities (forward-only callers: saves the 2 B C bytes of stores) */
/* arcb_gemm_bf16 (unit tests of the GEMM core) */
#define ARCB_GEMM_PAIR 0x4         /* CTA-pair core (same bit as ARCB_FWD_PAIR) */
#define ARCB_GEMM_PAIR_WIDE 0x8    /* with ARCB_GEMM_PAIR, a_mn = b_mn = 1: a pair computes two 256-column tiles from one A stage */

/* Library version (major * 10000 + minor * 100 + patch). */
int arcb_version(void);

/* Message for the last error raised on the calling thread ("" if none). */
const char* arcb_last_error(void);

/* Synchronises `stream` and reports whether a kernel was aborted (ARCB_ERR_TIMEOUT).  The pipeline kernels carry a
 * watchdog: a barrier / counter wait that outlives ~20 s of SM clock (no legitimate wait comes near it) records where
 * it was and TRAPS, so the launch fails loudly -- this and every later CUDA call on the context report the error --
 * instead of producing results from a pipeline that never completed.  Tests call it after every case; a training
 * loop needs nothing: its next synchronisation raises. */
int arcb_check_watchdog(void* stream);

/* ---- row L2 normalisation: F.normalize(input) / F.normalize(self.weight), arc_main.py:159 ----
 * x [rows, D] fp32 (leading dim ldx)  ->  out [rows, Dp] bf16 (zero padded, Dp % 64 == 0, Dp >= D)
 * and inv_norm[rows] = 1 / max(||x||_2, 1e-12). */
int arcb_l2norm_rows(const float* x, int64_t rows, int64_t D, int64_t ldx, void* out_bf16, int64_t Dp,
                     float* inv_norm, void* stream);

/* ---- F.normalize backward for the embeddings (autograd behind arc_main.py:329) ----
 * dx = (dxhat - xhat (xhat . dxhat)) / ||x||. */
int arcb_l2norm_rows_bwd(const float* x, int64_t ldx, const float* dxhat, int64_t ldg, int64_t rows, int64_t D,
                         float* dx, int64_t ldo, void* stream);

/* ---- tcgen05 GEMM core, exposed for unit tests: out[M,N] (+)= A * B^T with bf16 operands ----
 * a_mn / b_mn = 0: operand stored [M|N, K] (K contiguous); 1: stored [K, M|N] (M|N contiguous).
 * lda / ldb in elements (multiples of 8).  split_k > 1 accumulates with atomics into `out`
 * (which must then be zeroed by the caller). */
int arcb_gemm_bf16(const void* A, int a_mn, const void* B, int b_mn, int64_t M, int64_t N, int64_t K, int64_t lda,
                   int64_t ldb, float* out, int64_t ldo, int split_k, int flags, void* stream);

/* Bytes of persistent workspace a shard needs between forward and backward, for calls made with `flags`. */
int64_t arcb_ws_bytes(int64_t B, int64_t Cs, int64_t D, int flags);

/* 1 when calls with these sizes, scale and flags use the stored-probabilities schedule (see ARCB_BWD_RECOMPUTE), else 0.
 * In that case arcb_shard_backward needs no scratch (scratch may be NULL / scratch_bytes 0). */
int arcb_keep_probs_active(int64_t B, int64_t Cs, int64_t D, float s, int flags);

/* Bytes of backward scratch that hold `chunk_classes` classes of bf16 dLoss/dcos for batch B.
 * arcb_shard_backward accepts any scratch >= arcb_scratch_bytes(B, 128); its default persistent
 * kernel splits the scratch into four chunk buffers (fewer when the scratch is small; below
 * arcb_scratch_bytes(B, 2 * 256) it falls back to one kernel launch per GEMM and chunk). */
int64_t arcb_scratch_bytes(int64_t B, int64_t chunk_classes);

/* Chunk length (classes, a multiple of 256) that suits the persistent backward for this problem on the current
 * device: size the scratch as arcb_scratch_bytes(B, 4 * arcb_suggest_chunk_classes(B, Cs, D)). */
int64_t arcb_suggest_chunk_classes(int64_t B, int64_t Cs, int64_t D);

/* ---- fused forward on one class shard: arc_main.py:159-172 + the softmax statistics of :245 ----
 * x [B, D] fp32, w [Cs, D] fp32 (the shard's rows of ArcMarginProduct.weight), labels [B] int64.
 * Writes stats [B, 4] fp32 = (row max of logits * log2e, sum 2^(logit*log2e - max), target logit
 * or -inf if another shard owns the label, 0).  The B x Cs logits are never written. */
int arcb_shard_forward(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels, int64_t B,
                       int64_t Cs, int64_t D, int64_t class_offset, float s, float m, int easy_margin, void* ws,
                       int64_t ws_bytes, float* stats, int flags, void* stream);

/* ---- fused forward whose shard record is exchanged by the kernels themselves (class-sharded mode) ----
 * Replaces "arcb_shard_forward -> all_gather of stats -> arcb_merge_loss" (the DDP all-reduce of arc_main.py:243 is what
 * class sharding removed; this is the exchange that took its place).  The kernel that merges the shard's
 * statistics stores its [B, 4] record straight into slot `rank` of every rank's gather buffer -- `gather_bufs[p]` is rank
 * p's buffer [world, B, 4] fp32 as mapped into THIS process (peer / symmetric memory; plain stores, no TMA) -- and then
 * raises `flags[p][rank]` to `epoch`; arcb_merge_loss_peer on each rank waits until all `world` flags of its own array have
 * reached the epoch and reduces its local buffer.  No collective launch, no host synchronisation.
 *   - epoch must grow by one per call and start above the flags' initial value (zero them once, then use 1, 2, ...);
 *   - alternate between TWO gather buffers (epoch & 1): a rank may be one step ahead of its peers, never two;
 *   - world <= ARCB_MAX_PEERS; all pointers must stay valid until every rank has consumed the step. */
#define ARCB_MAX_PEERS 16
typedef struct arcb_peer_exchange {
  int rank, world;
  unsigned int epoch;
  void* gather_bufs[ARCB_MAX_PEERS]; /* [p]: rank p's [world, B, 4] fp32 buffer of this step */
  void* flags[ARCB_MAX_PEERS];       /* [p]: rank p's uint32 flags [world] */
} arcb_peer_exchange;
int arcb_shard_forward_peer(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels, int64_t B,
                            int64_t Cs, int64_t D, int64_t class_offset, float s, float m, int easy_margin, void* ws,
                            int64_t ws_bytes, float* stats, int flags, const arcb_peer_exchange* px, void* stream);
/* gathered: THIS rank's gather buffer [world, B, 4]; own_flags: THIS rank's flags [world]. */
int arcb_merge_loss_peer(const float* gathered, const void* own_flags, unsigned int epoch, int64_t P, int64_t B, float* lse,
                         float* target_logit, float* loss, void* stream);

/* ---- sum of the class shards' dX_hat partials + F.normalize backward, exchanged by the kernel itself ----
 * Replaces "all_reduce(dxhat) -> arcb_l2norm_rows_bwd" of the class-sharded backward (autograd behind arc_main.py:329).
 * Every rank's partial (the `dxhat` it passed to arcb_shard_backward) must live in peer-mapped (symmetric) memory:
 * partials[p] = rank p's [B, arcb_padded_dim(D)] fp32 of this step as mapped into THIS process.  Rank r reduces the rows
 * [r * ceil(B / world), ...) over all partials (peer loads), applies dx = (dxh - xh (xh . dxh)) / ||x|| and keeps the
 * finished rows in slices[r] ([ceil(B / world), D] fp32, symmetric); then every rank copies all slices into its local
 * dx [B, ldo].  own_only = 1 (data-parallel trainer, reduce_scatter semantics): dx receives only this rank's rows
 * ([rows, ldo]) and nothing is collected.  ready / done: per rank `world` uint32 flags each, zeroed once; counter: one
 * zeroed uint32 in local memory; epoch grows by one per call from 1; alternate between two sets of partial / slice
 * buffers (epoch & 1).  D % 4 == 0, D <= 2048, 16-byte aligned rows.  No collective launch, no host synchronisation. */
typedef struct arcb_peer_reduce {
  int rank, world;
  unsigned int epoch;
  int own_only;
  void* partials[ARCB_MAX_PEERS];
  void* slices[ARCB_MAX_PEERS];
  void* ready[ARCB_MAX_PEERS];
  void* done[ARCB_MAX_PEERS];
  void* counter;
} arcb_peer_reduce;
int arcb_dx_peer_reduce(const float* x, int64_t ldx, int64_t B, int64_t D, const arcb_peer_reduce* pr, float* dx,
                        int64_t ldo, void* stream);

/* ---- fused forward without a caller workspace (no counterpart in the reference) ----
 * For hosts that are not allowed to obtain device memory when the call happens: a plugin living inside another
 * runtime whose allocator owns all of HBM, a serving loop that has frozen its memory plan, or a foreign-language
 * binding that can pass pointers it was given but has no allocator of its own.  Those callers cannot provide the
 * `ws` buffer arcb_shard_forward needs, and neither this function nor anything it calls allocates or frees
 * (no cudaMalloc / cudaFree, no allocating library routine): the workspace is a fixed ARCB_STATIC_WS_BYTES block
 * that is part of the library image on every GPU that loads it (held until the process exits).
 *   - Arguments and the stats output are exactly those of arcb_shard_forward, minus ws / ws_bytes.
 *   - Sizes whose workspace (arcb_ws_bytes(B, Cs, D, ARCB_BWD_RECOMPUTE)) exceeds ARCB_STATIC_WS_BYTES are rejected with
 *     ARCB_ERR_WORKSPACE and a message naming both numbers; nothing is launched.  (32 MiB covers e.g.
 *     B=256, Cs=16384, D=512 or B=1024, Cs=8192, D=1024.)
 *   - The first call on each device looks up the block's device address (with that device current) and caches
 *     it; make that first call before any stream capture begins.  The block itself is materialised by the driver
 *     when the library's device code is loaded on that GPU (with lazy loading: at the first call into the library
 *     there), so a host that freezes its memory plan should make one library call per device during start-up, before
 *     the freeze; the cached address does not survive cudaDeviceReset().
 *   - Arguments are those of arcb_shard_forward with flags = ARCB_BWD_RECOMPUTE (forward only: nothing is kept for
 *     a backward; otherwise the library defaults).
 *   - All calls on a device share the one block: two calls whose kernels overlap on the GPU (different streams,
 *     or different host threads) race on it -- serialise them.  The block keeps the previous call's contents;
 *     the forward rewrites every region before reading it, so nothing needs clearing between calls.
 *   - The block is a forward workspace only: arcb_shard_backward still needs the caller-owned `ws` variant. */
#define ARCB_STATIC_WS_BYTES (32ll << 20)
int arcb_shard_forward_static(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels,
                              int64_t B, int64_t Cs, int64_t D, int64_t class_offset, float s, float m,
                              int easy_margin, float* stats, void* stream);

/* ---- forward on a weight whose normalised copy is already in the workspace ----
 * Same arguments, outputs and arithmetic as arcb_shard_forward, but the F.normalize(self.weight) pass
 * (arc_main.py:159, weight half) is skipped: `ws` must already hold w_hat / inv_norm of exactly this `w` in its
 * weight slots -- put there by arcb_sgd_step / arcb_adam_step (below), or left by an earlier forward on the same
 * workspace while `w` has not changed since (validation loop, arc_main.py:357-404: the weight is constant). */
int arcb_shard_forward_prenorm(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels,
                               int64_t B, int64_t Cs, int64_t D, int64_t class_offset, float s, float m,
                               int easy_margin, void* ws, int64_t ws_bytes, float* stats, int flags, void* stream);

/* Where a workspace keeps the normalised weight: w_hat [Cs, arcb_padded_dim(D)] bf16 and inv_norm_w [Cs] fp32.
 * The two regions lead the workspace layout, so the addresses do not depend on the batch size. */
int arcb_ws_weight_slots(void* ws, int64_t Cs, int64_t D, void** w_hat_bf16, float** inv_norm_w);

/* ---- optimizer step on the weight shard, fused with the next step's F.normalize(weight) ----
 * Replaces optimizer.step() (arc_main.py:330) for the head's parameter group: torch.optim.SGD(lr, weight_decay=5e-4,
 * momentum=0.9) (arc_main.py:252-253) and torch.optim.Adam(lr) (arc_main.py:249-250), with the update rules of
 * torch/optim/sgd.py and torch/optim/adam.py (single-tensor form, amsgrad / maximize off).  w, dw and the state
 * tensors are [rows, D] fp32 with their own leading dimensions; w and the state are updated IN PLACE.
 * If w_hat_bf16 / inv_norm_w are given (both or neither; e.g. from arcb_ws_weight_slots of the workspace the next
 * forward will use) the unit rows of the UPDATED weight are written there in the same pass, so that
 * arcb_shard_forward_prenorm can skip its normalisation kernel: one read of w per training step instead of two.
 *   SGD : g = dw + weight_decay * w;  buf = g on the first step, else momentum * buf + (1 - dampening) * g;
 *         g = g + momentum * buf if nesterov else buf;  w -= lr * g.   momentum == 0: momentum_buf may be NULL.
 *   Adam: g = dw + weight_decay * w;  m += (1 - beta1)(g - m);  v = beta2 v + (1 - beta2) g^2;
 *         w -= lr / (1 - beta1^step) * m / (sqrt(v) / sqrt(1 - beta2^step) + eps);   step counts from 1.
 * Hyper-parameters are doubles (as Python / torch hold them); 1 - beta, 1 - dampening and the bias corrections are
 * formed in double and rounded to fp32 once; the per-element arithmetic is fp32. */
int arcb_sgd_step(float* w, int64_t ldw, const float* dw, int64_t ldg, float* momentum_buf, int64_t ldb, int64_t rows,
                  int64_t D, double lr, double momentum, double dampening, double weight_decay, int nesterov,
                  int first_step, void* w_hat_bf16, int64_t Dp, float* inv_norm_w, void* stream);
int arcb_adam_step(float* w, int64_t ldw, const float* dw, int64_t ldg, float* exp_avg, float* exp_avg_sq, int64_t lds,
                   int64_t rows, int64_t D, double lr, double beta1, double beta2, double eps, double weight_decay,
                   int64_t step, void* w_hat_bf16, int64_t Dp, float* inv_norm_w, void* stream);

/* ---- materialising forward: the literal ArcMarginProduct.forward output, arc_main.py:157-176 ----
 * logits [B, ldl] fp32 <- s * (phi at the label column, cos elsewhere) for this shard's columns. */
int arcb_shard_logits(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels, int64_t B,
                      int64_t Cs, int64_t D, int64_t class_offset, float s, float m, int easy_margin, void* ws,
                      int64_t ws_bytes, float* logits, int64_t ldl, void* stream);

/* ---- exact top-3 / argmax of the logits: torch.max(out, 1) / out.topk(3), arc_main.py:338-345, :377-388 ----
 * Needs the workspace left by arcb_shard_forward (called with ARCB_FWD_BLOCK_MAXIMA and otherwise the same flags as
 * passed here) for the same inputs; after a
 * forward without that flag the kernel traps rather than rank stale data.  Only the class blocks whose bf16
 * block maximum can still reach the top 3 are re-scored, in fp32 from the original x / w (margin value at
 * the label column).  top_val [B,3] = logits (descending), top_idx [B,3] = GLOBAL class ids (-1 if the shard
 * has fewer than 3 classes).  Class-sharded: all_gather the [B,3] pairs and keep the best 3. */
int arcb_shard_top3(const float* x, int64_t ldx, const float* w, int64_t ldw, int64_t B, int64_t Cs, int64_t D,
                    int64_t class_offset, float s, void* ws, int64_t ws_bytes, float* top_val, int64_t* top_idx,
                    int flags, void* stream);

/* ---- merge shard records into the loss: nn.CrossEntropyLoss() mean, arc_main.py:245,:327 ----
 * stats [P, B, 4] (P shard records, e.g. after an all_gather; P = 1 unsharded)
 * -> lse [B], target_logit [B] (may be NULL), loss [1] = mean(lse - target_logit). */
int arcb_merge_loss(const float* stats, int64_t P, int64_t B, float* lse, float* target_logit, float* loss,
                    void* stream);

/* ---- fused backward on one class shard (autograd of arc_main.py:326-329) ----
 * Needs the workspace left by arcb_shard_forward / arcb_shard_logits for the same inputs.
 * lse [B] and target_logit [B] are the GLOBAL values from arcb_merge_loss; grad_loss is a device
 * scalar (NULL = 1.0).  If grad_logits != NULL (materialising path) it is dLoss/dlogits [B, ldgl]
 * for this shard's columns and lse / target_logit / grad_loss are ignored.
 * Writes dw [Cs, D] (final gradient of the weight shard) and ACCUMULATES this shard's partial of
 * d loss / d xhat into dxhat [B, Dp] fp32 after zeroing it (sum over shards, then
 * arcb_l2norm_rows_bwd gives dx). */
int arcb_shard_backward(int64_t B, int64_t Cs, int64_t D, float s, void* ws, int64_t ws_bytes, const float* lse,
                        const float* target_logit, const float* grad_loss, const float* grad_logits, int64_t ldgl,
                        void* scratch, int64_t scratch_bytes, float* dw, float* dxhat, int flags, void* stream);

/* Leading dimension (elements) of the dxhat buffer for feature size D. */
int64_t arcb_padded_dim(int64_t D);

/* ---- measurement aids (bench.py / tests; no counterpart in the reference) ----
 * arcb_launch_count: kernels this library has launched so far in the process.
 * arcb_profile_enable(1): from now on every launch is bracketed by CUDA events on its own stream
 *   (the event pool is created here, so call it outside any timed region); (0) switches it off.
 * arcb_profile_collect: synchronises the device, sums the elapsed milliseconds and launch counts
 *   per kernel kind (ARCB_K_*) and clears the records. */
enum {
  ARCB_K_L2NORM_X = 0,    /* F.normalize(input)                      */
  ARCB_K_L2NORM_W = 1,    /* F.normalize(weight)                     */
  ARCB_K_TARGET_PREP = 2, /* F.normalize(input) + fp32 target cosine + margin branch */
  ARCB_K_FWD_GEMM = 3,    /* cos GEMM + online softmax statistics    */
  ARCB_K_MERGE = 4,       /* merge of per-tile statistics            */
  ARCB_K_LOSS = 5,        /* cross-entropy from merged statistics    */
  ARCB_K_BWD_PREP = 6,    /* backward set-up (tile list, zeroing)    */
  ARCB_K_BWD_G_GEMM = 7,  /* recompute cos, write bf16 dLoss/dcos    */
  ARCB_K_DX_GEMM = 8,     /* dxhat += G * what                       */
  ARCB_K_DW_GEMM = 9,     /* dW = G^T * xhat with normalise-backward */
  ARCB_K_L2NORM_BWD = 10, /* F.normalize backward for the embeddings */
  ARCB_K_LOGITS_GEMM = 11,/* materialising forward                   */
  ARCB_K_GEMM_TEST = 12,  /* arcb_gemm_bf16                          */
  ARCB_K_BWD_PERSISTENT = 13, /* whole backward of a shard, one launch */
  ARCB_K_TOP3 = 14,       /* exact top-3 re-score                    */
  ARCB_K_OPTIM = 15,      /* optimizer step + next step's normalise  */
  ARCB_K_COUNT = 16
};
int64_t arcb_launch_count(void);
int arcb_profile_enable(int on);
int arcb_profile_collect(double* ms_by_kind, int64_t* launches_by_kind, int nkinds);

#ifdef __cplusplus
}
#endif
#endif /* ARCFACE_B200_H_ */
