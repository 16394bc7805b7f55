// C ABI of the B200-native ArcFace head.  See include/arcface_b200.h for the contract and the
// reference lines (ARCFACE/arc_main.py) each entry point replaces.
//
// Build (in-tree, see __graft_entry__.build()):
//   nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC \
//        -Iinclude -o ddp_classification_pytorch_b200/libarcface_b200.so csrc/arcface_b200.cu
#include "../../include/arcface_b200.h"

#include <cuda_bf16.h>
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <atomic>
#include <mutex>
#include <stdio.h>
#include <string.h>

#include "bwd_persistent.cuh"
#include "bwd_probs.cuh"
#include "epilogues.cuh"
#include "gemm2_sm100.cuh"
#include "gemm_sm100.cuh"
#include "fwd_resident.cuh"
#include "rowwise.cuh"
#include "optim_rows.cuh"

using namespace arcb;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define ARCB_CUDA(expr)                                                                        \
  do {                                                                                         \
    cudaError_t e_ = (expr);                                                                   \
    if (e_ != cudaSuccess) return fail(ARCB_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)

inline int64_t round_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

// ---- launch accounting and optional per-kernel CUDA-event timing (bench / tests only) ----
// The counter is always on (one relaxed atomic per launch).  Event timing is off unless
// arcb_profile_enable(1) was called; the event pool is created there, never on the hot path.
std::atomic<int64_t> g_launches{0};
constexpr int kProfMax = 8192;
struct Profiler {
  std::atomic<bool> on{false};
  std::mutex mu;  // guards n / kind / created: forward (main thread) and backward (autograd thread) both launch
  int n = 0;
  cudaEvent_t beg[kProfMax];
  cudaEvent_t end[kProfMax];
  int kind[kProfMax];
  bool created = false;
} g_prof;

struct LaunchScope {
  cudaStream_t st;
  int slot;
  LaunchScope(int kind, cudaStream_t s, int nkernels = 1) : st(s), slot(-1) {
    g_launches.fetch_add(nkernels, std::memory_order_relaxed);
    if (g_prof.on.load(std::memory_order_relaxed)) {
      {
        std::lock_guard<std::mutex> lock(g_prof.mu);
        if (g_prof.n < kProfMax) {
          slot = g_prof.n++;
          g_prof.kind[slot] = kind;
        }
      }
      if (slot >= 0) cudaEventRecord(g_prof.beg[slot], st);
    }
  }
  ~LaunchScope() {
    if (slot >= 0) cudaEventRecord(g_prof.end[slot], st);
  }
};

constexpr int kBNTile = 256;
constexpr int kMergeSlices = 64;

// Tuning / debugging knobs from the environment.  Read ONCE (std::call_once) and immutable afterwards, so there is
// nothing for concurrent callers to race on; everything that selects a code path per call is the `flags` argument of
// the entry points (include/arcface_b200.h), never process state.
struct EnvKnobs {
  int lag = 2;            // ARCB_LAG: phases between a chunk's G tiles and its X / W tiles
  int x_splits = 0;       // ARCB_X_SPLITS: class slices per dX tile (0 = heuristic)
  int kind_mask = 7;      // ARCB_DEBUG_KINDS: tile kinds that run (wrong results unless 7)
  int debug_epi = 0;      // ARCB_DEBUG_EPI: G epilogue without math (1) / without TMEM loads (2)
  int scratch_bufs = 4;   // ARCB_SCRATCH_BUFS: chunk buffers of the persistent backward
  int trace = 0;          // ARCB_TRACE: per-tile clock stamps into the tail of the scratch
  int norm_look = 0;      // ARCB_NORM_LOOK: look-ahead (B tiles) of the in-GEMM normaliser
  int default_flags = 0;  // ARCB_DEFAULT_FLAGS: OR-ed into every call's flags (A/B runs of the tools)
};
const EnvKnobs& env_knobs() {
  static EnvKnobs k;
  static std::once_flag once;
  std::call_once(once, [] {
    auto geti = [](const char* name, int dflt) {
      const char* e = getenv(name);
      return e ? static_cast<int>(strtol(e, nullptr, 0)) : dflt;
    };
    int v = geti("ARCB_LAG", 2);
    k.lag = (v >= 1 && v <= 4) ? v : 2;
    k.x_splits = geti("ARCB_X_SPLITS", 0);
    v = geti("ARCB_DEBUG_KINDS", 7) & 7;
    k.kind_mask = v ? v : 7;
    k.debug_epi = geti("ARCB_DEBUG_EPI", 0);
    v = geti("ARCB_SCRATCH_BUFS", 4);
    k.scratch_bufs = (v >= 2 && v <= 6) ? v : 4;
    k.trace = geti("ARCB_TRACE", 0) == 1;
    k.norm_look = geti("ARCB_NORM_LOOK", 0);
    k.default_flags = geti("ARCB_DEFAULT_FLAGS", 0);
  });
  return k;
}
inline int effective_flags(int flags) { return flags | env_knobs().default_flags; }

// Persistent per-shard workspace, carved deterministically from one caller-owned buffer.
struct Workspace {
  int64_t B, Cs, D, Bp, Dp;
  __nv_bfloat16* xhat;   // [B, Dp]
  __nv_bfloat16* what;   // [Cs, Dp]
  float* inv_nx;         // [Bp]
  float* inv_nw;         // [Cs]
  int* lbl_local;        // [Bp]
  float* sub;            // [Bp]
  float* zt;             // [Bp]
  float* dphi;           // [Bp]
  float* cos_t;          // [Bp]
  float2* part;          // [2 * tiles_n, Bp] one partial per (column block, epilogue warpgroup)
  float2* rec;           // [Bp]
  float2* rec2;          // [kMergeSlices, Bp] first-level merge
  float* lse2;           // [Bp]
  float* zt2;            // [Bp]
  float* gscale;         // [1]
  float* tproj;          // [Cs]
  unsigned int* counters;  // [2 * ceil(Cs / 128)] chunk completion counters of the persistent backward
  int4* items;             // [items_cap][2] tile list of the persistent backward (bwd_items_kernel)
  int64_t items_cap;
  __nv_bfloat16* probs;    // stored probabilities: P'^T [Cs, Bp] class-major, else nullptr
  __nv_bfloat16* xc;       // stored probabilities: c (.) x_hat [B, Dp], the dW operand
  float* cvec;             // stored probabilities: c_i [Bp]
  float* bmax_p;           // stored probabilities: block maxima [2 * tiles_n][Bp] (the partials carry a fixed reference there)
  float* bmax;             // resident forward: block maxima [2 * tiles_n][Bp], second half of the `part` region
  int* fwd_mode;           // which forward filled the workspace (kFwd*): tells arcb_shard_top3 where the block maxima are
  int64_t total;
};

// Stored-probabilities schedule (bwd_probs.cuh): when it applies.  The feature row must span exactly two 256-column
// slices (EpiDWT), and with the fixed reference s2 = s log2(e) every P' must stay a normal bf16 number:
// 2^(-2 s2) >= 2^-126.  Otherwise, or when a recompute schedule is asked for by flag, the backward recomputes.
constexpr int kRecomputeFlags =
    ARCB_BWD_RECOMPUTE | ARCB_BWD_PER_CHUNK | ARCB_BWD_SINGLE_CTA | ARCB_BWD_STATIC_SHARE | ARCB_FWD_PAIR;
constexpr int kFwdKeptProbs = 0x100;  // OR-ed into the workspace's forward-schedule word: P'^T was stored
bool keep_probs_applies(int64_t B, int64_t Cs, int64_t D, float s, int flags) {
  if (flags & kRecomputeFlags) return false;
  if (D <= 256 || D > 512 || !(s > 0.f) || 2.f * s * 1.4426950408889634f > 120.f) return false;
  // small shards are bound by the host's launch rate: the recompute schedule needs two launches fewer (measured at
  // B = 512, D = 512: ahead from 50k classes per shard on, behind at 25k)
  return B >= 128 && Cs >= 4096 && B * Cs >= (int64_t{1} << 24);
}

Workspace carve(void* base, int64_t B, int64_t Cs, int64_t D, bool keep_probs = false) {
  Workspace w;
  w.B = B; w.Cs = Cs; w.D = D;
  w.Bp = round_up(B, kBNTile);
  w.Dp = round_up(D, 64);
  const int64_t tiles_n = (Cs + kBNTile - 1) / kBNTile;
  char* p = reinterpret_cast<char*>(base);
  int64_t off = 0;
  auto take = [&](int64_t bytes) {
    char* r = p ? p + off : nullptr;
    off += round_up(bytes, 256);
    return r;
  };
  // the two weight-derived regions come first so that their offsets do not depend on the batch size
  // (arcb_ws_weight_slots: the fused optimizer step deposits them for the next forward)
  w.what = reinterpret_cast<__nv_bfloat16*>(take(Cs * w.Dp * 2));
  w.inv_nw = reinterpret_cast<float*>(take(Cs * 4));
  w.xhat = reinterpret_cast<__nv_bfloat16*>(take(B * w.Dp * 2));
  w.inv_nx = reinterpret_cast<float*>(take(w.Bp * 4));
  w.lbl_local = reinterpret_cast<int*>(take(w.Bp * 4));
  w.sub = reinterpret_cast<float*>(take(w.Bp * 4));
  w.zt = reinterpret_cast<float*>(take(w.Bp * 4));
  w.dphi = reinterpret_cast<float*>(take(w.Bp * 4));
  w.cos_t = reinterpret_cast<float*>(take(w.Bp * 4));
  w.part = reinterpret_cast<float2*>(take(2 * tiles_n * w.Bp * 8));
  w.rec = reinterpret_cast<float2*>(take(w.Bp * 8));
  w.rec2 = reinterpret_cast<float2*>(take(kMergeSlices * w.Bp * 8));
  w.lse2 = reinterpret_cast<float*>(take(w.Bp * 4));
  w.zt2 = reinterpret_cast<float*>(take(w.Bp * 4));
  w.gscale = reinterpret_cast<float*>(take(4));
  w.tproj = reinterpret_cast<float*>(take(Cs * 4));
  w.counters = reinterpret_cast<unsigned int*>(take((2 * ((Cs + 127) / 128) + 4) * 4));  // + the feed's claim counter
  {
    // upper bound of the backward's tile count for any chunking (chunks are multiples of 128 classes, a dX tile
    // covers at least 4 blocks of 64 classes): recompute + dW tiles, then dX tiles incl. per-chunk rounding
    const int64_t T = (Cs + 127) / 128, dn = (D + 255) / 256;
    w.items_cap = T * (w.Bp / kBNTile + dn) + ((B + 127) / 128) * dn * (2 * T + 1);
    w.items = reinterpret_cast<int4*>(take(w.items_cap * 32));
  }
  // the resident forward leaves only 2 * lanes partials per row (lanes <= tiles_n / 4) in the first half of `part`
  w.bmax = w.part ? reinterpret_cast<float*>(w.part + tiles_n * w.Bp) : nullptr;
  w.fwd_mode = reinterpret_cast<int*>(take(8));  // [0] forward schedule, [1] block counter of the publishing merge
  w.probs = nullptr; w.xc = nullptr; w.cvec = nullptr;
  w.bmax_p = nullptr;
  if (keep_probs) {  // at the end: every other offset is the same with and without
    w.xc = reinterpret_cast<__nv_bfloat16*>(take(B * w.Dp * 2));
    w.cvec = reinterpret_cast<float*>(take(w.Bp * 4));
    w.bmax_p = reinterpret_cast<float*>(take(2 * tiles_n * w.Bp * 4));
    w.probs = reinterpret_cast<__nv_bfloat16*>(take(round_up(Cs, 128) * w.Bp * 2));  // P'^T [Cs, Bp]: class-major
  }
  w.total = off;
  return w;
}

int check_dims(int64_t B, int64_t Cs, int64_t D) {
  if (B <= 0 || Cs <= 0 || D <= 0) return fail(ARCB_ERR_INVALID, "B, Cs, D must be positive (got %lld, %lld, %lld)",
                                                (long long)B, (long long)Cs, (long long)D);
  if (B > (1ll << 30) || Cs > (1ll << 30) || D > (1 << 20))
    return fail(ARCB_ERR_INVALID, "size out of range (B=%lld, Cs=%lld, D=%lld)", (long long)B, (long long)Cs,
                (long long)D);
  return ARCB_OK;
}

int launch_l2norm(const float* x, int64_t rows, int64_t D, int64_t ldx, __nv_bfloat16* out, int64_t Dp, float* inv,
                  cudaStream_t st, int kind) {
  if (rows == 0) return ARCB_OK;
  LaunchScope ls(kind, st);
  const int warps_per_block = 8;
  int64_t blocks = (rows + warps_per_block - 1) / warps_per_block;
  if (blocks > 148 * 64) blocks = 148 * 64;  // grid-stride beyond that
  const bool vec = (D % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
  const int g = static_cast<int>(blocks);
#define ARCB_L2N(NV) l2norm_rows_kernel<NV><<<g, 256, 0, st>>>(x, rows, (int)D, ldx, out, (int)Dp, inv)
  if (vec && Dp <= 128) ARCB_L2N(1);
  else if (vec && Dp <= 256) ARCB_L2N(2);
  else if (vec && Dp <= 512) ARCB_L2N(4);
  else if (vec && Dp <= 1024) ARCB_L2N(8);
  else if (vec && Dp <= 2048) ARCB_L2N(16);
  else ARCB_L2N(0);
#undef ARCB_L2N
  ARCB_CUDA(cudaGetLastError());
  return ARCB_OK;
}

// Steps shared by the fused and the materialising forward: normalise, target preparation.
int forward_common(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels, int64_t B,
                   int64_t Cs, int64_t D, int64_t class_offset, float s, float m, int easy, void* ws_ptr,
                   int64_t ws_bytes, Workspace* out_ws, cudaStream_t st, int w_mode = 0, int fwd_mode = kFwdStreaming,
                   bool keep_probs = false) {
  // w_mode: 0 = normalise the weight here, 1 = the workspace already holds w_hat / inv_norm of this weight,
  //         2 = the forward GEMM kernel normalises it on the fly (launch_gemm_norm)
  int rc = check_dims(B, Cs, D);
  if (rc) return rc;
  if (!x || !w || !labels || !ws_ptr) return fail(ARCB_ERR_INVALID, "null pointer argument");
  if (ldx < D || ldw < D) return fail(ARCB_ERR_INVALID, "leading dimension smaller than D");
  if (!(s > 0.f)) return fail(ARCB_ERR_INVALID, "scale s must be positive (got %g)", (double)s);
  if (reinterpret_cast<uintptr_t>(ws_ptr) & 255) return fail(ARCB_ERR_INVALID, "workspace must be 256-byte aligned");
  Workspace W = carve(ws_ptr, B, Cs, D, keep_probs);
  if (W.total > ws_bytes)
    return fail(ARCB_ERR_WORKSPACE, "workspace too small: need %lld bytes, got %lld", (long long)W.total,
                (long long)ws_bytes);
  if (w_mode == 0) {
    rc = launch_l2norm(w, Cs, D, ldw, W.what, W.Dp, W.inv_nw, st, ARCB_K_L2NORM_W);
    if (rc) return rc;
  }
  // w_mode 2: the per-B-tile "rows normalised" counters of the fused kernel are cleared by the kernel below
  const int n_zero = w_mode == 2 ? static_cast<int>((Cs + kBNTile - 1) / kBNTile) : 0;
  const double md = static_cast<double>(m);
  const float cos_m = static_cast<float>(cos(md)), sin_m = static_cast<float>(sin(md));
  const float th = static_cast<float>(cos(M_PI - md)), mm = static_cast<float>(sin(M_PI - md) * md);
  const int blocks = static_cast<int>((W.Bp + 7) / 8);
  LaunchScope ls(ARCB_K_TARGET_PREP, st);
  // per-sample work in one launch: F.normalize(input) and the exact fp32 target cosine / margin branch
  const int xvec = (D % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
  target_prep_kernel<<<blocks, 256, 0, st>>>(x, ldx, w, ldw, labels, W.xhat, (int)W.Dp, W.inv_nx, xvec, (int)B,
                                             (int)W.Bp, (int)D, Cs, class_offset, s, cos_m, sin_m, th, mm, easy,
                                             W.lbl_local, W.sub, W.zt, W.dphi, W.cos_t, W.counters, n_zero, W.fwd_mode,
                                             fwd_mode | (keep_probs ? kFwdKeptProbs : 0));
  ARCB_CUDA(cudaGetLastError());
  *out_ws = W;
  return ARCB_OK;
}

constexpr int kLaunchRefused = 1;  // internal: the cooperative launch was refused, try the next schedule

// ONE cooperative launch for the whole backward of a shard (bwd_persistent.cuh).
template <bool EXT, bool PAIR, bool TRACE = false, bool DYN = false>
int launch_bwd_persistent(const Workspace& W, float s, const float* grad_logits, int64_t ldgl, __nv_bfloat16* gt,
                          int64_t chunk_rows, int nbuf, float* dw, float* dxhat, cudaStream_t st, BwdSetup su,
                          unsigned long long* trace = nullptr, bool cooperative = true) {
  const int64_t B = W.B, Cs = W.Cs, D = W.D;
  BwdPlan pl;
  pl.chunk_rows = static_cast<int>(chunk_rows);
  pl.nbuf = nbuf;
  {
    const int lag = env_knobs().lag;  // nbuf >= lag + 2 keeps buffer-reuse waits a phase away
    pl.lag = lag < nbuf ? lag : (nbuf > 1 ? nbuf - 1 : 1);
  }
  pl.n_chunks = static_cast<int>((Cs + chunk_rows - 1) / chunk_rows);
  pl.Cs = (int)Cs; pl.B = (int)B; pl.Bp = (int)W.Bp; pl.D = (int)D; pl.Dp = (int)W.Dp;
  pl.g_tiles_n = (int)(W.Bp / kPBN);
  pl.g_kb = (int)(W.Dp / kBK);
  pl.tile_rows = PAIR ? 2 * kBM : kBM;
  pl.x_tiles_m = (int)((B + pl.tile_rows - 1) / pl.tile_rows);
  pl.x_tiles_n = (int)((D + kPBN - 1) / kPBN);
  pl.w_tiles_n = pl.x_tiles_n;
  pl.w_kb = (int)((B + kBK - 1) / kBK);
  // dX tiles are cut along the chunk's classes so that one costs about as much as a dW tile and a
  // chunk yields at least ~one tile per SM
  {
    const int kbt = (int)((chunk_rows + kBK - 1) / kBK);
    const int tiles_mn = pl.x_tiles_m * pl.x_tiles_n;
    const int groups = device_sm_count() / (PAIR ? 2 : 1);  // CTAs or CTA pairs working in parallel
    // enough class slices that a chunk yields a few dozen dX tiles, but not more than 4: every extra slice is
    // another 128 KB of fp32 reduce-add traffic per tile column (measured best on B200: 4, see profiles/)
    int want_splits = (groups + tiles_mn - 1) / tiles_mn;
    if (want_splits > 4) want_splits = 4;
    if (env_knobs().x_splits > 0) want_splits = env_knobs().x_splits;
    int per = (kbt + want_splits - 1) / want_splits;
    if (per < 4) per = 4;
    pl.x_kb_per_split = per;
  }
  pl.g_done = W.counters;
  pl.kind_mask = env_knobs().kind_mask;
  pl.debug_epi = env_knobs().debug_epi;
  pl.rot = 0;
  pl.xw_done = W.counters + pl.n_chunks;
  pl.claim = W.counters + 2 * pl.n_chunks;
  {
    // the tile list: written once here, then read (one 32-byte record per tile) by every role of every CTA
    const int nph = pl.n_chunks + pl.lag;
    const long long total = phase_begin_of(pl, nph);
    if (total > W.items_cap || total > 0x7fffffffll) {
      fail(ARCB_ERR_WORKSPACE, "persistent backward: %lld tiles exceed the workspace's tile list (%lld)", total,
           (long long)W.items_cap);
      return kLaunchRefused;  // the per-chunk schedule needs no list
    }
    pl.items = W.items;
    pl.total_items = static_cast<int>(total);
    su.zero_c = W.counters;
    su.n_c = 2 * pl.n_chunks + 1;
    su.n_phases = nph;
    LaunchScope ls(ARCB_K_BWD_PREP, st);
    const int blocks = nph > 2 * device_sm_count() ? nph : 2 * device_sm_count();
    bwd_setup_kernel<<<blocks, 256, 0, st>>>(pl, W.items, su);
    ARCB_CUDA(cudaGetLastError());
  }

  CUtensorMap tmWhatK, tmXhatK, tmGtSt, tmGtMN, tmWhatMN, tmGtK, tmXhatMN, tmDx, tmDw;
  const CUtensorMapDataType bf = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, f32 = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  bool ok = true;
  ok = ok && make_tmap_2d(&tmWhatK, W.what, bf, 2, W.Dp, Cs, W.Dp * 2, kBK, kBM);
  constexpr int kBRows = PAIR ? kPBN / 2 : kPBN;  // rows / columns of the B tile one CTA stages
  ok = ok && make_tmap_2d(&tmXhatK, W.xhat, bf, 2, W.Dp, B, W.Dp * 2, kBK, kBRows);
  const void* gsrc = gt;  // G^T as the dX / dW tiles see it: the chunk scratch
  const int64_t grows = nbuf * chunk_rows;
  ok = ok && make_tmap_2d(&tmGtSt, gsrc, bf, 2, W.Bp, grows, W.Bp * 2, 64, kBM);
  ok = ok && make_tmap_mn3(&tmGtMN, gsrc, W.Bp, grows, W.Bp * 2, kBM / 64);
  ok = ok && make_tmap_mn3(&tmWhatMN, W.what, W.Dp, Cs, W.Dp * 2, kBRows / 64);
  ok = ok && make_tmap_2d(&tmGtK, gsrc, bf, 2, W.Bp, grows, W.Bp * 2, kBK, kBM);
  ok = ok && make_tmap_mn3(&tmXhatMN, W.xhat, W.Dp, B, W.Dp * 2, kBRows / 64);
  ok = ok && make_tmap_2d(&tmDx, dxhat, f32, 4, D, B, W.Dp * 4, 32, kBM);
  if (!ok) return fail(ARCB_ERR_CUDA, "tensor maps for the persistent backward could not be encoded");
  const bool dw_tma = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(dw) & 15) == 0) &&
                      make_tmap_2d(&tmDw, dw, f32, 4, D, Cs, D * 4, 32, kBM);
  if (!dw_tma) tmDw = tmDx;

  BwdArgs ar;
  ar.lse2 = W.lse2; ar.gscale = W.gscale; ar.lbl_local = W.lbl_local; ar.zt2 = W.zt2; ar.dphi = W.dphi;
  ar.cos_t = W.cos_t; ar.tproj = W.tproj; ar.s2 = s * kLog2e; ar.grad_out = grad_logits; ar.ldgo = ldgl;
  ar.s_ext = s; ar.dw = dw; ar.what = W.what; ar.inv_nw = W.inv_nw; ar.dw_tma = dw_tma ? 1 : 0;
  ar.trace = trace;

  auto kern = bwd_persistent_kernel<EXT, PAIR, TRACE, DYN>;
  constexpr int kPSmemTotal = PCfg<PAIR>::kSmemTotal;
  static std::atomic<bool> attr_set[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!attr_set[dev].load(std::memory_order_acquire)) {  // idempotent: a race only repeats the call
    ARCB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kPSmemTotal));
    attr_set[dev].store(true, std::memory_order_release);
  }
  const long long total_items = pl.total_items;
  int groups = device_sm_count() / (PAIR ? 2 : 1);
  if (groups > total_items) groups = static_cast<int>(total_items);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(groups * (PAIR ? 2 : 1));
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = kPSmemTotal;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (PAIR) {
    attr[na].id = cudaLaunchAttributeClusterDimension;  // the two CTAs of a tile share a TPC
    attr[na].val.clusterDim.x = 2;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (cooperative) {  // all CTAs co-resident: they wait on one another.  (ARCB_BWD_NO_COOP drops the attribute for
    attr[na].id = cudaLaunchAttributeCooperative;  // profiler runs: one CTA per SM on an idle GPU is co-resident anyway)
    attr[na].val.cooperative = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  LaunchScope ls(ARCB_K_BWD_PERSISTENT, st);
  const cudaError_t le =
      cudaLaunchKernelEx(&cfg, kern, tmWhatK, tmXhatK, tmGtSt, tmGtMN, tmWhatMN, tmGtK, tmXhatMN, tmDx, tmDw, pl, ar);
  if (le != cudaSuccess) {
    // e.g. the device cannot keep the whole grid co-resident right now (shared GPU, MPS limits): nothing has
    // been enqueued, so the caller may fall back to another schedule
    (void)cudaGetLastError();
    fail(ARCB_ERR_CUDA, "persistent backward launch refused: %s", cudaGetErrorString(le));
    return kLaunchRefused;
  }
  return ARCB_OK;
}

// Built-in forward workspace for arcb_shard_forward_static: a fixed block of device memory that is
// part of the loaded library image, so that entry point never has to ask anybody for memory.
// NOTE: in host code the name below is only a handle for the runtime, not a device address; the
// address a kernel can use comes from cudaGetSymbolAddress (static_ws_address()).
constexpr int64_t kStaticWsBytes = ARCB_STATIC_WS_BYTES;
static __device__ __align__(256) unsigned char g_static_ws[kStaticWsBytes];

// Device address of g_static_ws on the CURRENT device; looked up once per device and cached.
int static_ws_address(void** out) {
  static std::mutex mu;
  static void* addr[64] = {nullptr};
  int dev = -1;
  ARCB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(ARCB_ERR_INVALID, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_ws);
    if (e != cudaSuccess || !p)
      return fail(ARCB_ERR_CUDA, "cudaGetSymbolAddress(built-in workspace) on device %d: %s", dev, cudaGetErrorString(e));
    addr[dev] = p;
  }
  *out = addr[dev];
  return ARCB_OK;
}

}  // namespace

namespace {
// Backward from the probabilities the forward kept (bwd_probs.cuh): set-up, dX GEMM, dW kernel.
int backward_from_probs(const Workspace& W, float s, const float* lse, const float* zt, const float* grad_loss, float* dw,
                        float* dxhat, int flags, cudaStream_t st) {
  const int64_t B = W.B, Cs = W.Cs, D = W.D;
  {
    ProbsSetup su;
    su.lse = lse; su.zt = zt; su.grad_loss = grad_loss; su.dphi = W.dphi; su.lbl_local = W.lbl_local; su.xhat = W.xhat;
    su.xc = W.xc; su.cvec = W.cvec; su.probs = W.probs; su.ldB = W.Bp; su.s = s; su.B = (int)B; su.Bp = (int)W.Bp;
    su.Dp = (int)W.Dp; su.zero = dxhat; su.n_zero = B * W.Dp;
    su.zero_t = W.tproj; su.n_t = Cs; su.zero_c = W.counters; su.n_c = (Cs + kBM - 1) / kBM + 1; su.fwd_mode = W.fwd_mode;
    LaunchScope ls(ARCB_K_BWD_PREP, st);
    probs_setup_kernel<<<2 * device_sm_count(), 256, 0, st>>>(su);
    ARCB_CUDA(cudaGetLastError());
  }
  const int sms = device_sm_count();
  CUtensorMap tmDx, tmDw;
  if (!make_tmap_2d(&tmDx, dxhat, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, D, B, W.Dp * 4, 32, kBM))
    return fail(ARCB_ERR_CUDA, "tensor map for dxhat could not be encoded (16-byte alignment?)");
  {
    // dX_hat[B, D] = diag(c) P'[B, Cs] w_hat[Cs, D] on the pair core: every CTA pair keeps ONE output tile in TMEM for
    // its whole slice of the classes and reduce-adds it once at the end
    // A = P' as an MN-major operand: stored [k = classes, m = samples] with the samples contiguous
    Operand a{W.probs, W.Bp, Cs, W.Bp}, b{W.what, W.Dp, Cs, W.Dp};
    EpiRedAdd::Params ep{(int)B, (int)D, W.cvec};
    LaunchScope ls(ARCB_K_DX_GEMM, st);
    // wide tiles: a pair owns 256 samples x the WHOLE feature row (two 256-column accumulators = all of TMEM), so that
    // every element of P'^T is fetched by exactly one pair
    const int ptiles = static_cast<int>(((B + 2 * kBM - 1) / (2 * kBM)) * ((D + 2 * kBNTile - 1) / (2 * kBNTile)));
    int psplit = (sms / 2) / (ptiles > 0 ? ptiles : 1);
    if (psplit < 1) psplit = 1;
    const cudaError_t e = launch_gemm_pair<EpiRedAdd, kBNTile, 3, true, true, 2>(a, b, &tmDx, B, D, Cs, psplit, ep, st);
    if (e != cudaSuccess) return fail(ARCB_ERR_CUDA, "dX GEMM launch: %s", cudaGetErrorString(e));
  }
  {
    // dW[Cs, D] = inv_norm (P'^T (c.x_hat) - t w_hat): A = P'^T [Cs, Bp] K-major, B = c.x_hat [B, Dp] MN-major
    const bool dw_tma = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(dw) & 15) == 0) &&
                        make_tmap_2d(&tmDw, dw, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, D, Cs, D * 4, 32, kBM);
    if (!dw_tma) tmDw = tmDx;
    Operand a{W.probs, Cs, B, W.Bp}, b{W.xc, W.Dp, B, W.Dp};
    EpiDWT<true>::Params ep;
    ep.base = EpiDW::Params{dw, W.what, W.tproj, W.inv_nw, 0, (int)Cs, (int)D, (int)W.Dp, dw_tma ? 1 : 0};
    ep.t_acc = W.tproj;
    ep.arrivals = W.counters;
    ep.slices = static_cast<int>((D + kBNTile - 1) / kBNTile);
    // a CTA's next tile: the pair core walks the class tiles with a stride of (CTAs / slices) 128-row blocks
    const int64_t blocks = (Cs + kBM - 1) / kBM;
    int pairs = sms / 2;
    if (pairs > ((blocks + 1) / 2) * ep.slices) pairs = static_cast<int>(((blocks + 1) / 2) * ep.slices);
    pairs -= pairs % ep.slices;
    ep.m_blocks = static_cast<int>(2 * ((blocks + 1) / 2));
    ep.pair = 1;
    const bool direct = !dw_tma;
    const bool coop = !(flags & ARCB_BWD_NO_COOP);  // profiling only, see the header
    LaunchScope ls(ARCB_K_DW_GEMM, st);
    cudaError_t e;
    for (;;) {
      ep.next_tile_blocks = 2 * pairs / ep.slices;
      if (direct) {
        EpiDWT<false>::Params ed;
        static_assert(sizeof(ed) == sizeof(ep), "same parameter block");
        memcpy(&ed, &ep, sizeof(ed));
        e = launch_gemm_pair_row_slices<EpiDWT<false>, kBNTile, 5, false, true>(a, b, &tmDw, Cs, D, B, ed, st, coop, pairs);
      } else {
        e = launch_gemm_pair_row_slices<EpiDWT<true>, kBNTile, 3, false, true>(a, b, &tmDw, Cs, D, B, ep, st, coop, pairs);
      }
      // part of the GPU is taken (MPS share, another resident kernel): the slices of a class tile still have to run
      // together, so ask for a grid that fits instead of giving up co-residency
      if (e != cudaErrorCooperativeLaunchTooLarge && e != cudaErrorLaunchOutOfResources) break;
      int fewer = pairs / 2;
      fewer -= fewer % ep.slices;
      if (fewer < ep.slices) break;
      pairs = fewer;
    }
    if (e != cudaSuccess) return fail(ARCB_ERR_CUDA, "dW GEMM launch: %s", cudaGetErrorString(e));
  }
  return ARCB_OK;
}

// ---- optimizer step on the weight shard (+ the next forward's normalised copy) -- optim_rows.cuh ----
template <class Rule>
int launch_optim(float* w, int64_t ldw, const float* dw, int64_t ldg, float* s0, float* s1, int64_t lds,
                        int64_t rows, int64_t D, const Rule& rule, void* w_hat, int64_t Dp, float* inv_norm,
                        cudaStream_t st) {
  if (!w || !dw) return fail(ARCB_ERR_INVALID, "null pointer argument");
  if (rows < 0 || D <= 0 || D > (1 << 20) || ldw < D || ldg < D || ((s0 || s1) && lds < D))
    return fail(ARCB_ERR_INVALID, "bad sizes for the optimizer step");
  if ((w_hat != nullptr) != (inv_norm != nullptr))
    return fail(ARCB_ERR_INVALID, "w_hat_bf16 and inv_norm_w must be given together (or both NULL)");
  if (w_hat && (Dp < D || Dp % 64)) return fail(ARCB_ERR_INVALID, "Dp must be arcb_padded_dim(D)");
  if (rows == 0) return ARCB_OK;
  LaunchScope ls(ARCB_K_OPTIM, st);
  int64_t blocks = (rows + 7) / 8;
  if (blocks > 148 * 64) blocks = 148 * 64;  // grid-stride beyond that
  const int64_t Dpad = w_hat ? Dp : round_up(D, 64);
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool vec = (D % 4 == 0) && (ldw % 4 == 0) && (ldg % 4 == 0) && (lds % 4 == 0 || (!s0 && !s1)) && al16(w) &&
                   al16(dw) && al16(s0) && al16(s1);
  const int g = static_cast<int>(blocks);
  __nv_bfloat16* wh = static_cast<__nv_bfloat16*>(w_hat);
#define ARCB_OPT(NV) \
  optim_rows_kernel<NV, Rule><<<g, 256, 0, st>>>(w, ldw, dw, ldg, s0, s1, lds, rows, (int)D, rule, wh, (int)Dpad, inv_norm)
  if (vec && Dpad <= 128) ARCB_OPT(1);
  else if (vec && Dpad <= 256) ARCB_OPT(2);
  else if (vec && Dpad <= 512) ARCB_OPT(4);
  else if (vec && Dpad <= 1024) ARCB_OPT(8);
  else if (vec && Dpad <= 2048) ARCB_OPT(16);
  else ARCB_OPT(0);
#undef ARCB_OPT
  ARCB_CUDA(cudaGetLastError());
  return ARCB_OK;
}

}  // namespace

extern "C" {

int arcb_version(void) { return 20000; }

const char* arcb_last_error(void) { return g_err; }

int arcb_check_watchdog(void* stream) {
  // A fired watchdog has trapped its kernel: the stream (and the context) report a launch failure from here on.
  const cudaError_t e = cudaStreamSynchronize(static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess)
    return fail(ARCB_ERR_TIMEOUT, "a kernel was aborted (device-side pipeline watchdog or a fault): %s",
                cudaGetErrorString(e));
  unsigned int h[4] = {0, 0, 0, 0};
  ARCB_CUDA(cudaMemcpyFromSymbol(h, g_wait_timeout, sizeof(h)));
  if (h[0]) return fail(ARCB_ERR_TIMEOUT, "pipeline watchdog fired: barrier tag 0x%x, block %u, arg %u", h[1], h[2], h[3]);
  return ARCB_OK;
}

int64_t arcb_padded_dim(int64_t D) { return round_up(D, 64); }

int64_t arcb_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int arcb_profile_enable(int on) {
  std::lock_guard<std::mutex> lock(g_prof.mu);
  if (on && !g_prof.created) {
    for (int i = 0; i < kProfMax; ++i) {
      ARCB_CUDA(cudaEventCreate(&g_prof.beg[i]));
      ARCB_CUDA(cudaEventCreate(&g_prof.end[i]));
    }
    g_prof.created = true;
  }
  g_prof.n = 0;
  g_prof.on.store(on != 0, std::memory_order_relaxed);
  return ARCB_OK;
}

int arcb_profile_collect(double* ms_by_kind, int64_t* launches_by_kind, int nkinds) {
  if (!ms_by_kind || !launches_by_kind || nkinds <= 0) return fail(ARCB_ERR_INVALID, "bad profile buffers");
  for (int k = 0; k < nkinds; ++k) {
    ms_by_kind[k] = 0.0;
    launches_by_kind[k] = 0;
  }
  ARCB_CUDA(cudaDeviceSynchronize());
  std::lock_guard<std::mutex> lock(g_prof.mu);
  for (int i = 0; i < g_prof.n; ++i) {
    float ms = 0.f;
    ARCB_CUDA(cudaEventElapsedTime(&ms, g_prof.beg[i], g_prof.end[i]));
    const int k = g_prof.kind[i];
    if (k >= 0 && k < nkinds) {
      ms_by_kind[k] += ms;
      launches_by_kind[k] += 1;
    }
  }
  const int dropped = g_prof.n >= kProfMax;
  g_prof.n = 0;
  return dropped ? fail(ARCB_ERR_WORKSPACE, "profile record pool overflowed (%d records)", kProfMax) : ARCB_OK;
}

int arcb_l2norm_rows(const float* x, int64_t rows, int64_t D, int64_t ldx, void* out_bf16, int64_t Dp, float* inv_norm,
                     void* stream) {
  if (!x || !out_bf16 || !inv_norm) return fail(ARCB_ERR_INVALID, "null pointer argument");
  if (rows < 0 || D <= 0 || ldx < D || Dp < D || Dp % 64) return fail(ARCB_ERR_INVALID, "bad sizes for l2norm_rows");
  return launch_l2norm(x, rows, D, ldx, static_cast<__nv_bfloat16*>(out_bf16), Dp, inv_norm,
                       static_cast<cudaStream_t>(stream), ARCB_K_L2NORM_W);
}

int arcb_l2norm_rows_bwd(const float* x, int64_t ldx, const float* dxhat, int64_t ldg, int64_t rows, int64_t D,
                         float* dx, int64_t ldo, void* stream) {
  if (!x || !dxhat || !dx) return fail(ARCB_ERR_INVALID, "null pointer argument");
  if (rows <= 0 || D <= 0 || ldx < D || ldg < D || ldo < D) return fail(ARCB_ERR_INVALID, "bad sizes");
  const int blocks = static_cast<int>((rows + 7) / 8);
  LaunchScope ls(ARCB_K_L2NORM_BWD, static_cast<cudaStream_t>(stream));
  l2norm_rows_bwd_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, ldx, dxhat, ldg, (int)rows, (int)D,
                                                                               dx, ldo);
  ARCB_CUDA(cudaGetLastError());
  return ARCB_OK;
}

int arcb_gemm_bf16(const void* A, int a_mn, const void* B, int b_mn, int64_t M, int64_t N, int64_t K, int64_t lda,
                   int64_t ldb, float* out, int64_t ldo, int split_k, int flags, void* stream) {
  if (!A || !B || !out) return fail(ARCB_ERR_INVALID, "null pointer argument");
  const bool pair_gemm = (effective_flags(flags) & ARCB_GEMM_PAIR) != 0;
  if (M <= 0 || N <= 0 || K <= 0 || (lda % 8) || (ldb % 8)) return fail(ARCB_ERR_INVALID, "bad GEMM sizes");
  Operand a{A, M, K, lda}, b{B, N, K, ldb};
  EpiStore::Params ep{out, ldo, (int)M, (int)N, split_k > 1 ? 1 : 0};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t e;
  LaunchScope ls(ARCB_K_GEMM_TEST, st);
  if (pair_gemm) {
    // an MN-major operand is addressed in whole 64-element blocks: its (zero padded) leading dimension is the extent
    if ((a_mn && (lda % 64 || lda < M)) || (b_mn && (ldb % 64 || ldb < N)))
      return fail(ARCB_ERR_INVALID, "pair GEMM: an MN-major operand needs a leading dimension that is a multiple of 64");
    if (a_mn) a.mn_extent = lda;
    if (b_mn) b.mn_extent = ldb;
    if (effective_flags(flags) & ARCB_GEMM_PAIR_WIDE) {  // two column tiles per pair tile (the dX GEMM of the backward)
      if (!(a_mn && b_mn)) return fail(ARCB_ERR_INVALID, "wide pair GEMM: built for MN-major operands only");
      e = launch_gemm_pair<EpiStore, 256, 4, true, true, 2>(a, b, nullptr, M, N, K, split_k, ep, st);
    } else
    if (!a_mn && !b_mn) e = launch_gemm_pair<EpiStore, 256, 6, false, false>(a, b, nullptr, M, N, K, split_k, ep, st);
    else if (a_mn && !b_mn) e = launch_gemm_pair<EpiStore, 256, 6, true, false>(a, b, nullptr, M, N, K, split_k, ep, st);
    else if (!a_mn && b_mn) e = launch_gemm_pair<EpiStore, 256, 6, false, true>(a, b, nullptr, M, N, K, split_k, ep, st);
    else e = launch_gemm_pair<EpiStore, 256, 6, true, true>(a, b, nullptr, M, N, K, split_k, ep, st);
  } else
  if (!a_mn && !b_mn) e = launch_gemm<EpiStore, 256, 4, false, false>(a, b, nullptr, M, N, K, split_k, ep, st);
  else if (a_mn && !b_mn) e = launch_gemm<EpiStore, 256, 4, true, false>(a, b, nullptr, M, N, K, split_k, ep, st);
  else if (!a_mn && b_mn) e = launch_gemm<EpiStore, 256, 4, false, true>(a, b, nullptr, M, N, K, split_k, ep, st);
  else e = launch_gemm<EpiStore, 256, 4, true, true>(a, b, nullptr, M, N, K, split_k, ep, st);
  if (e != cudaSuccess) return fail(ARCB_ERR_CUDA, "gemm launch: %s", cudaGetErrorString(e));
  return ARCB_OK;
}

int64_t arcb_ws_bytes(int64_t B, int64_t Cs, int64_t D, int flags) {
  if (B <= 0 || Cs <= 0 || D <= 0) return -1;
  // the stored probabilities are part of the workspace when the schedule applies to the shape (for ANY scale s: the query
  // has no s, so it sizes for the case that the forward uses them)
  return carve(nullptr, B, Cs, D, keep_probs_applies(B, Cs, D, 1.f, effective_flags(flags))).total;
}

// Classes per backward chunk for this problem.  Four chunks of bf16 G^T must stay well inside L2, short chunks pay
// too much fp32 reduce-add traffic per dX tile; within those bounds the choice is nearly flat now that CTA pairs
// claim their tiles (2.73-2.90 ms over 13-25 tile rows at B=1024, D=512, C=1M), so the rule that served the fixed
// dealing is kept: a phase (one chunk's recompute tiles + an earlier chunk's dX / dW tiles) whose tile count sits
// just below a multiple of the number of pairs, never exactly on one.
int64_t arcb_suggest_chunk_classes(int64_t B, int64_t Cs, int64_t D) {
  if (B <= 0 || Cs <= 0 || D <= 0) return -1;
  const int64_t Bp = round_up(B, kBNTile);
  const int groups = device_sm_count() / 2 > 0 ? device_sm_count() / 2 : 74;
  const int64_t g_tiles_n = Bp / kBNTile, dn = (D + 255) / 256, x_tiles = ((B + 255) / 256) * dn;
  const int64_t r_cap = (Cs + 255) / 256;
  int64_t best_r = 0;
  double best_cost = 1e30;
  for (int64_t r = 16; r <= 64; ++r) {
    if (4 * r * 256 * Bp * 2 > (48ll << 20) && best_r) break;  // four G^T buffers <= 48 MB of the 126 MB L2
    int64_t want = (groups + x_tiles - 1) / x_tiles;
    if (want > 4) want = 4;
    int64_t per = (4 * r + want - 1) / want;
    if (per < 4) per = 4;
    const int64_t items = r * (g_tiles_n + dn) + x_tiles * ((4 * r + per - 1) / per);
    if (items % groups == 0) continue;  // the dealing would never rotate: the same pairs get the dX tiles every phase
    const int64_t rounds = (items + groups - 1) / groups;
    const double waste = static_cast<double>(rounds * groups - items) / static_cast<double>(items);
    const double cost = waste - 1e-4 * static_cast<double>(r);  // ties: the longer chunk
    if (cost < best_cost) {
      best_cost = cost;
      best_r = r;
    }
  }
  if (best_r > r_cap) best_r = r_cap;
  return best_r * 256;
}

int arcb_keep_probs_active(int64_t B, int64_t Cs, int64_t D, float s, int flags) {
  if (B <= 0 || Cs <= 0 || D <= 0) return 0;
  return keep_probs_applies(B, Cs, D, s, effective_flags(flags)) ? 1 : 0;
}

int64_t arcb_scratch_bytes(int64_t B, int64_t chunk_classes) {
  if (B <= 0 || chunk_classes <= 0) return -1;
  return round_up(chunk_classes, 128) * round_up(B, kBNTile) * 2;
}

static int shard_forward_impl(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels,
                              int64_t B, int64_t Cs, int64_t D, int64_t class_offset, float s, float m,
                              int easy_margin, void* ws, int64_t ws_bytes, float* stats, int flags, void* stream,
                              bool w_prenormalised, const arcb_peer_exchange* px = nullptr) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!stats) return fail(ARCB_ERR_INVALID, "null stats pointer");
  flags = effective_flags(flags);
  PeerTable peers;
  if (px) {
    if (px->world < 1 || px->world > ARCB_MAX_PEERS || px->rank < 0 || px->rank >= px->world)
      return fail(ARCB_ERR_INVALID, "peer exchange: bad rank / world (%d / %d)", px->rank, px->world);
    for (int p = 0; p < px->world; ++p) {
      if (!px->gather_bufs[p] || !px->flags[p]) return fail(ARCB_ERR_INVALID, "peer exchange: null pointer for rank %d", p);
      peers.buf[p] = static_cast<float*>(px->gather_bufs[p]);
      peers.flag[p] = static_cast<unsigned int*>(px->flags[p]);
    }
  }
  // last merge level: into `stats` only, or (px) also into every rank's gather buffer + flag
  auto final_merge = [&](const float2* part, int n_part, const Workspace& Wk) -> int {
    LaunchScope ls(ARCB_K_MERGE, st, 1);
    if (px)
      merge_publish_kernel<<<static_cast<int>((B + 127) / 128), 128, 0, st>>>(
          part, n_part, Wk.Bp, Wk.zt, (int)B, stats, peers, px->rank, px->world, px->epoch,
          reinterpret_cast<unsigned int*>(Wk.fwd_mode) + 1);
    else
      merge_pack_kernel<<<static_cast<int>((B + 127) / 128), 128, 0, st>>>(part, n_part, Wk.Bp, Wk.zt, (int)B, stats);
    ARCB_CUDA(cudaGetLastError());
    return ARCB_OK;
  };
  if ((flags & ARCB_FWD_NORM_SEPARATE) && (flags & ARCB_FWD_NORM_FUSED))
    return fail(ARCB_ERR_INVALID, "ARCB_FWD_NORM_SEPARATE and ARCB_FWD_NORM_FUSED exclude each other");
  const bool pair_gemm = (flags & ARCB_FWD_PAIR) != 0;
  Workspace W;
  // F.normalize(weight) inside the GEMM kernel; needs 16-byte aligned rows like the vector path of the stand-alone
  // kernel, D <= 2048 and the one-CTA tile shape.  Measured on B200: the overlap pays (4-9 % of the forward) where the
  // GEMM is long enough to hide the pass but short enough that eight extra warps per SM do not cost it more than the
  // pass itself: 512 <= B <= 1024, many classes.  ARCB_FWD_NORM_FUSED forces it for every shape that allows it.
  const bool shape_ok = (flags & ARCB_FWD_NORM_FUSED) || (B >= 512 && B <= 1024 && Cs >= 65536);
  bool fuse = !(flags & ARCB_FWD_NORM_SEPARATE) && shape_ok && !w_prenormalised && !pair_gemm && (D % 4 == 0) &&
              (ldw % 4 == 0) && D <= 2048 && ((reinterpret_cast<uintptr_t>(w) & 15) == 0);
  // x_hat-resident schedule (fwd_resident.cuh): the sample block stays in shared memory, one partial per row and CTA
  const bool want_bmax = (flags & ARCB_FWD_BLOCK_MAXIMA) != 0;
  // the workspace layout follows the flag and the shape (what arcb_ws_bytes sized); whether the probabilities are USED
  // also depends on the scale: too large an s falls back to the recompute backward on the same layout
  const bool probs_layout = keep_probs_applies(B, Cs, D, 1.f, flags);
  const bool keep_probs = keep_probs_applies(B, Cs, D, s, flags);
  const bool resident = !keep_probs && !(flags & ARCB_FWD_STREAMING) && !pair_gemm &&
                        resident_forward_applies(B, Cs, round_up(D, 64));
  // which array arcb_shard_top3 finds the block maxima in: the partials themselves (streaming), or a separate array
  const int fwd_mode = (resident || keep_probs) ? (want_bmax ? kFwdResidentBmax : kFwdResident) : kFwdStreaming;
  int rc = forward_common(x, ldx, w, ldw, labels, B, Cs, D, class_offset, s, m, easy_margin, ws, ws_bytes, &W, st,
                          w_prenormalised ? 1 : (fuse ? 2 : 0), fwd_mode, probs_layout);
  if (rc) return rc;
  Operand a{W.xhat, B, W.Dp, W.Dp}, b{W.what, Cs, W.Dp, W.Dp};
  EpiFwdStats::Params ep{W.lbl_local, W.sub, W.part, (int)W.Bp, (int)B, (int)Cs, s * kLog2e};
  cudaError_t e = cudaSuccess;
  const int tiles_n = static_cast<int>((Cs + kBNTile - 1) / kBNTile);
  EpiFwdStatsP::Params epp{W.lbl_local, W.sub, W.part, want_bmax ? W.bmax_p : nullptr, (int)W.Bp, (int)B, (int)Cs, s * kLog2e};
  CUtensorMap tmP;
  if (keep_probs && !make_tmap_2d(&tmP, W.probs, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, W.Bp, Cs, W.Bp * 2, 64, 64))
    return fail(ARCB_ERR_CUDA, "tensor map for the stored probabilities could not be encoded");
  NormArgs na_res{};
  if (fuse) {
    na_res.w = w; na_res.ldw = ldw; na_res.rows = Cs; na_res.D = (int)D; na_res.Dp = (int)W.Dp; na_res.what = W.what;
    na_res.inv_nw = W.inv_nw; na_res.done = W.counters;
    const int ahead = resident_lanes(B) + 2;  // a CTA's last tile is at most this far from the end
    na_res.lookahead = ahead > 64 ? ahead : 64;
    const int look = env_knobs().norm_look;
    if (look > 0) na_res.lookahead = look > ahead ? look : ahead;
  }
  if (resident) {
    ResParams rp{W.lbl_local, W.sub, W.part, want_bmax ? W.bmax : nullptr, (int)W.Bp, (int)B, (int)Cs, s * kLog2e};
    if (fuse) {
      {
        LaunchScope ls(ARCB_K_FWD_GEMM, st);
        e = launch_fwd_resident<true>(a, b, B, Cs, W.Dp, rp, na_res, st);
      }
      if (e != cudaSuccess) {  // the grid cannot be made co-resident right now: normalise in a kernel of its own
        fuse = false;
        rc = launch_l2norm(w, Cs, D, ldw, W.what, W.Dp, W.inv_nw, st, ARCB_K_L2NORM_W);
        if (rc) return rc;
      }
    }
    if (!fuse) {
      LaunchScope ls(ARCB_K_FWD_GEMM, st);
      e = launch_fwd_resident<false>(a, b, B, Cs, W.Dp, rp, NormArgs{}, st);
    }
    if (e != cudaSuccess) return fail(ARCB_ERR_CUDA, "forward GEMM launch: %s", cudaGetErrorString(e));
    return final_merge(W.part, 2 * resident_lanes(B), W);
  }
  if (fuse) {
    NormArgs na;
    na.w = w; na.ldw = ldw; na.rows = Cs; na.D = (int)D; na.Dp = (int)W.Dp; na.what = W.what; na.inv_nw = W.inv_nw;
    na.done = W.counters;
    const int tiles_m = (int)((B + kBM - 1) / kBM);
    const int ahead = (device_sm_count() + tiles_m - 1) / tiles_m + 2;  // a CTA's last tile is at most this far from the end
    na.lookahead = ahead > 64 ? ahead : 64;                             // 64 B tiles of w_hat = 16 MB at D = 512
    const int look = env_knobs().norm_look;
    if (look > 0) na.lookahead = look > ahead ? look : ahead;
    {
      LaunchScope ls(ARCB_K_FWD_GEMM, st);
      e = keep_probs ? launch_gemm_norm<EpiFwdStatsP, 4>(a, b, B, Cs, W.Dp, epp, na, st, &tmP)
                     : launch_gemm_norm<EpiFwdStats, 4>(a, b, B, Cs, W.Dp, ep, na, st);
    }
    if (e != cudaSuccess) {  // the grid cannot be made co-resident right now: two kernels instead
      fuse = false;
      rc = launch_l2norm(w, Cs, D, ldw, W.what, W.Dp, W.inv_nw, st, ARCB_K_L2NORM_W);
      if (rc) return rc;
    }
  }
  if (!fuse) {
    LaunchScope ls(ARCB_K_FWD_GEMM, st);
    if (keep_probs) e = launch_gemm<EpiFwdStatsP, kBNTile, 4, false, false>(a, b, &tmP, B, Cs, W.Dp, 1, epp, st);
    else if (pair_gemm) e = launch_gemm_pair<EpiFwdStats, kBNTile, 6, false, false>(a, b, nullptr, B, Cs, W.Dp, 1, ep, st);
    else e = launch_gemm<EpiFwdStats, kBNTile, 4, false, false>(a, b, nullptr, B, Cs, W.Dp, 1, ep, st);
  }
  if (e != cudaSuccess) return fail(ARCB_ERR_CUDA, "forward GEMM launch: %s", cudaGetErrorString(e));
  const int blocks = static_cast<int>((B + 255) / 256);
  const int n_part = 2 * tiles_n;
  if (n_part > 4 * kMergeSlices) {  // two levels: 64 slices in parallel, then the 64 results + the shard record
    {
      LaunchScope ls(ARCB_K_MERGE, st, 1);
      merge_stats_kernel<<<dim3(blocks, kMergeSlices), 256, 0, st>>>(W.part, n_part, W.Bp, (int)B, W.rec2, W.Bp);
      ARCB_CUDA(cudaGetLastError());
    }
    return final_merge(W.rec2, kMergeSlices, W);
  }
  return final_merge(W.part, n_part, W);
}

int arcb_shard_forward(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels, int64_t B,
                       int64_t Cs, int64_t D, int64_t class_offset, float s, float m, int easy_margin, void* ws,
                       int64_t ws_bytes, float* stats, int flags, void* stream) {
  return shard_forward_impl(x, ldx, w, ldw, labels, B, Cs, D, class_offset, s, m, easy_margin, ws, ws_bytes, stats,
                            flags, stream, false);
}

int arcb_shard_forward_peer(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels, int64_t B,
                            int64_t Cs, int64_t D, int64_t class_offset, float s, float m, int easy_margin, void* ws,
                            int64_t ws_bytes, float* stats, int flags, const arcb_peer_exchange* px, void* stream) {
  if (!px) return fail(ARCB_ERR_INVALID, "null peer exchange descriptor");
  return shard_forward_impl(x, ldx, w, ldw, labels, B, Cs, D, class_offset, s, m, easy_margin, ws, ws_bytes, stats,
                            flags, stream, false, px);
}

int arcb_merge_loss_peer(const float* gathered, const void* own_flags, unsigned int epoch, int64_t P, int64_t B, float* lse,
                         float* target_logit, float* loss, void* stream) {
  if (!gathered || !own_flags || !lse || !loss) return fail(ARCB_ERR_INVALID, "null pointer argument");
  if (P <= 0 || P > ARCB_MAX_PEERS || B <= 0) return fail(ARCB_ERR_INVALID, "bad sizes for merge_loss_peer");
  LaunchScope ls(ARCB_K_LOSS, static_cast<cudaStream_t>(stream));
  loss_wait_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(gathered, static_cast<const unsigned int*>(own_flags),
                                                                     epoch, (int)P, (int)B, lse, target_logit, loss);
  ARCB_CUDA(cudaGetLastError());
  return ARCB_OK;
}

int arcb_dx_peer_reduce(const float* x, int64_t ldx, int64_t B, int64_t D, const arcb_peer_reduce* pr, float* dx,
                        int64_t ldo, void* stream) {
  if (!x || !pr || !dx) return fail(ARCB_ERR_INVALID, "null pointer argument");
  if (B <= 0 || D <= 0 || D > 2048 || (D % 4) || ldx < D || ldo < D || (ldx % 4) || (ldo % 4) ||
      (reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(dx) & 15))
    return fail(ARCB_ERR_INVALID, "arcb_dx_peer_reduce needs D %% 4 == 0, D <= 2048 and 16-byte aligned rows");
  if (pr->world < 1 || pr->world > ARCB_MAX_PEERS || pr->rank < 0 || pr->rank >= pr->world || !pr->counter)
    return fail(ARCB_ERR_INVALID, "peer reduce: bad rank / world (%d / %d)", pr->rank, pr->world);
  PeerReduceArgs a;
  for (int p = 0; p < pr->world; ++p) {
    if (!pr->partials[p] || !pr->slices[p] || !pr->ready[p] || !pr->done[p])
      return fail(ARCB_ERR_INVALID, "peer reduce: null pointer for rank %d", p);
    a.partial[p] = static_cast<const float*>(pr->partials[p]);
    a.slice[p] = static_cast<const float*>(pr->slices[p]);
    a.ready[p] = static_cast<unsigned int*>(pr->ready[p]);
    a.done[p] = static_cast<unsigned int*>(pr->done[p]);
  }
  a.my_slice = static_cast<float*>(pr->slices[pr->rank]);
  a.counter = static_cast<unsigned int*>(pr->counter);
  a.rank = pr->rank; a.P = pr->world; a.epoch = pr->epoch;
  a.B = (int)B; a.D = (int)D; a.Dp = (int)round_up(D, 64);
  a.rows_per = (int)((B + pr->world - 1) / pr->world);
  a.own_only = pr->own_only ? 1 : 0;
  // every block waits for other ranks: the whole grid must be resident -> at most two blocks per SM
  int grid = a.rows_per;
  if (grid > 2 * device_sm_count()) grid = 2 * device_sm_count();
  if (grid < 1) grid = 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  LaunchScope ls(ARCB_K_L2NORM_BWD, st);
  dx_peer_reduce_kernel<<<grid, 256, 0, st>>>(x, ldx, dx, ldo, a);
  ARCB_CUDA(cudaGetLastError());
  return ARCB_OK;
}

int arcb_shard_forward_prenorm(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels,
                               int64_t B, int64_t Cs, int64_t D, int64_t class_offset, float s, float m,
                               int easy_margin, void* ws, int64_t ws_bytes, float* stats, int flags, void* stream) {
  return shard_forward_impl(x, ldx, w, ldw, labels, B, Cs, D, class_offset, s, m, easy_margin, ws, ws_bytes, stats,
                            flags, stream, true);
}

int arcb_ws_weight_slots(void* ws, int64_t Cs, int64_t D, void** w_hat_bf16, float** inv_norm_w) {
  if (!ws || !w_hat_bf16 || !inv_norm_w) return fail(ARCB_ERR_INVALID, "null pointer argument");
  if (Cs <= 0 || D <= 0) return fail(ARCB_ERR_INVALID, "Cs and D must be positive");
  if (reinterpret_cast<uintptr_t>(ws) & 255) return fail(ARCB_ERR_INVALID, "workspace must be 256-byte aligned");
  Workspace W = carve(ws, 1, Cs, D);  // the two weight regions lead the layout: same place for every B
  *w_hat_bf16 = W.what;
  *inv_norm_w = W.inv_nw;
  return ARCB_OK;
}

int arcb_sgd_step(float* w, int64_t ldw, const float* dw, int64_t ldg, float* momentum_buf, int64_t ldb, int64_t rows,
                  int64_t D, double lr, double momentum, double dampening, double weight_decay, int nesterov,
                  int first_step, void* w_hat_bf16, int64_t Dp, float* inv_norm_w, void* stream) {
  if (momentum != 0.0 && !momentum_buf) return fail(ARCB_ERR_INVALID, "momentum != 0 needs a momentum buffer");
  if (nesterov && (momentum <= 0.0 || dampening != 0.0))
    return fail(ARCB_ERR_INVALID, "Nesterov momentum requires a momentum and zero dampening");
  // hyper-parameters arrive as doubles (Python floats) and are rounded to fp32 once, after any arithmetic on
  // them -- the way torch hands its scalars to its kernels
  SgdRule r;
  r.lr = static_cast<float>(lr); r.momentum = static_cast<float>(momentum);
  r.one_minus_damp = static_cast<float>(1.0 - dampening); r.wd = static_cast<float>(weight_decay);
  r.nesterov = nesterov ? 1 : 0; r.first = first_step ? 1 : 0; r.has_buf = (momentum != 0.0) ? 1 : 0;
  return launch_optim(w, ldw, dw, ldg, r.has_buf ? momentum_buf : nullptr, nullptr, ldb, rows, D, r, w_hat_bf16, Dp,
                      inv_norm_w, static_cast<cudaStream_t>(stream));
}

int arcb_adam_step(float* w, int64_t ldw, const float* dw, int64_t ldg, float* exp_avg, float* exp_avg_sq, int64_t lds,
                   int64_t rows, int64_t D, double lr, double beta1, double beta2, double eps, double weight_decay,
                   int64_t step, void* w_hat_bf16, int64_t Dp, float* inv_norm_w, void* stream) {
  if (!exp_avg || !exp_avg_sq) return fail(ARCB_ERR_INVALID, "null optimizer state");
  if (step < 1) return fail(ARCB_ERR_INVALID, "step counts from 1 (got %lld)", (long long)step);
  if (!(beta1 >= 0.0 && beta1 < 1.0 && beta2 >= 0.0 && beta2 < 1.0)) return fail(ARCB_ERR_INVALID, "betas out of [0, 1)");
  const double bc1 = 1.0 - pow(beta1, static_cast<double>(step));
  const double bc2 = 1.0 - pow(beta2, static_cast<double>(step));
  AdamRule r;
  r.one_minus_b1 = static_cast<float>(1.0 - beta1); r.b2 = static_cast<float>(beta2);
  r.one_minus_b2 = static_cast<float>(1.0 - beta2); r.eps = static_cast<float>(eps);
  r.wd = static_cast<float>(weight_decay);
  r.step_size = static_cast<float>(lr / bc1);
  r.bc2_sqrt = static_cast<float>(sqrt(bc2));
  return launch_optim(w, ldw, dw, ldg, exp_avg, exp_avg_sq, lds, rows, D, r, w_hat_bf16, Dp, inv_norm_w,
                      static_cast<cudaStream_t>(stream));
}

// Forward for hosts that must not allocate: identical to arcb_shard_forward, except that the
// workspace is the library's own fixed device block instead of a caller buffer.
int arcb_shard_forward_static(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels,
                              int64_t B, int64_t Cs, int64_t D, int64_t class_offset, float s, float m,
                              int easy_margin, float* stats, void* stream) {
  int rc = check_dims(B, Cs, D);
  if (rc) return rc;
  const int64_t need = carve(nullptr, B, Cs, D).total;  // forward only: nothing is kept for a backward
  if (need > kStaticWsBytes)
    return fail(ARCB_ERR_WORKSPACE,
                "arcb_shard_forward_static: B=%lld, Cs=%lld, D=%lld needs %lld bytes of workspace but the built-in "
                "block holds %lld; use arcb_shard_forward with a caller-owned workspace for this size",
                (long long)B, (long long)Cs, (long long)D, (long long)need, (long long)kStaticWsBytes);
  void* ws = nullptr;
  rc = static_ws_address(&ws);  // first call on a device resolves the address (not a kernel launch, no allocation)
  if (rc) return rc;
  // Nothing in the forward workspace has to start from zero: every region is fully rewritten
  // (normalised operands incl. their padding, per-sample target data, per-tile partials) before it is read.
  return arcb_shard_forward(x, ldx, w, ldw, labels, B, Cs, D, class_offset, s, m, easy_margin, ws, kStaticWsBytes,
                            stats, ARCB_BWD_RECOMPUTE, stream);
}

int arcb_shard_logits(const float* x, int64_t ldx, const float* w, int64_t ldw, const int64_t* labels, int64_t B,
                      int64_t Cs, int64_t D, int64_t class_offset, float s, float m, int easy_margin, void* ws,
                      int64_t ws_bytes, float* logits, int64_t ldl, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!logits || ldl < Cs) return fail(ARCB_ERR_INVALID, "bad logits buffer");
  Workspace W;
  int rc = forward_common(x, ldx, w, ldw, labels, B, Cs, D, class_offset, s, m, easy_margin, ws, ws_bytes, &W, st);
  if (rc) return rc;
  Operand a{W.xhat, B, W.Dp, W.Dp}, b{W.what, Cs, W.Dp, W.Dp};
  EpiLogits::Params ep{W.lbl_local, W.sub, logits, ldl, (int)B, (int)Cs, s};
  LaunchScope ls(ARCB_K_LOGITS_GEMM, st);
  cudaError_t e = launch_gemm<EpiLogits, kBNTile, 4, false, false>(a, b, nullptr, B, Cs, W.Dp, 1, ep, st);
  if (e != cudaSuccess) return fail(ARCB_ERR_CUDA, "logits GEMM launch: %s", cudaGetErrorString(e));
  return ARCB_OK;
}

int arcb_shard_top3(const float* x, int64_t ldx, const float* w, int64_t ldw, int64_t B, int64_t Cs, int64_t D,
                    int64_t class_offset, float s, void* ws, int64_t ws_bytes, float* top_val, int64_t* top_idx,
                    int flags, void* stream) {
  int rc = check_dims(B, Cs, D);
  if (rc) return rc;
  if (!x || !w || !ws || !top_val || !top_idx) return fail(ARCB_ERR_INVALID, "null pointer argument");
  if (D * 4 > 96 * 1024) return fail(ARCB_ERR_INVALID, "arcb_shard_top3 supports D <= 24576");
  flags = effective_flags(flags);
  Workspace W = carve(ws, B, Cs, D, keep_probs_applies(B, Cs, D, 1.f, flags));
  const bool probs_used = keep_probs_applies(B, Cs, D, s, flags);
  if (W.total > ws_bytes) return fail(ARCB_ERR_WORKSPACE, "workspace too small");
  const int nparts = 2 * static_cast<int>((Cs + kBNTile - 1) / kBNTile);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (D * 4 > 48 * 1024) {
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev].load(std::memory_order_acquire)) {
      ARCB_CUDA(cudaFuncSetAttribute(topk_rescore_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      attr_set[dev].store(true, std::memory_order_release);
    }
  }
  LaunchScope ls(ARCB_K_TOP3, st);
  topk_rescore_kernel<<<static_cast<unsigned>(B), kTopkThreads, D * 4, st>>>(
      x, ldx, w, ldw, W.inv_nx, W.inv_nw, W.lbl_local, W.sub, W.part, probs_used ? W.bmax_p : W.bmax, W.fwd_mode, nparts, W.Bp,
      (int)B, (int)D,
      Cs, class_offset, s, top_val, top_idx);
  ARCB_CUDA(cudaGetLastError());
  return ARCB_OK;
}

int arcb_merge_loss(const float* stats, int64_t P, int64_t B, float* lse, float* target_logit, float* loss,
                    void* stream) {
  if (!stats || !lse || !loss) return fail(ARCB_ERR_INVALID, "null pointer argument");
  if (P <= 0 || B <= 0) return fail(ARCB_ERR_INVALID, "bad sizes for merge_loss");
  LaunchScope ls(ARCB_K_LOSS, static_cast<cudaStream_t>(stream));
  loss_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(stats, (int)P, (int)B, lse, target_logit, loss);
  ARCB_CUDA(cudaGetLastError());
  return ARCB_OK;
}

int arcb_shard_backward(int64_t B, int64_t Cs, int64_t D, float s, void* ws, int64_t ws_bytes, const float* lse,
                        const float* target_logit, const float* grad_loss, const float* grad_logits, int64_t ldgl,
                        void* scratch, int64_t scratch_bytes, float* dw, float* dxhat, int flags, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = check_dims(B, Cs, D);
  if (rc) return rc;
  flags = effective_flags(flags);
  const bool ext = grad_logits != nullptr;
  const bool use_probs = !ext && keep_probs_applies(B, Cs, D, s, flags);  // the forward kept P': no recompute, no scratch
  if (!ws || (!scratch && !use_probs) || !dw || !dxhat) return fail(ARCB_ERR_INVALID, "null pointer argument");
  const bool persistent = !(flags & ARCB_BWD_PER_CHUNK);
  const bool pair_bwd = !(flags & ARCB_BWD_SINGLE_CTA);
  const bool dyn_feed = !(flags & ARCB_BWD_STATIC_SHARE);
  const bool coop = !(flags & ARCB_BWD_NO_COOP);
  if (!ext && (!lse || !target_logit)) return fail(ARCB_ERR_INVALID, "lse / target_logit required");
  if (ext && ldgl < Cs) return fail(ARCB_ERR_INVALID, "grad_logits leading dimension smaller than Cs");
  if (reinterpret_cast<uintptr_t>(scratch) & 255) return fail(ARCB_ERR_INVALID, "scratch must be 256-byte aligned");
  Workspace W = carve(ws, B, Cs, D, keep_probs_applies(B, Cs, D, 1.f, flags));
  if (W.total > ws_bytes) return fail(ARCB_ERR_WORKSPACE, "workspace too small");
  if (use_probs) return backward_from_probs(W, s, lse, target_logit, grad_loss, dw, dxhat, flags, st);
  int64_t chunk = scratch_bytes / (W.Bp * 2) / 128 * 128;
  if (chunk < 128)
    return fail(ARCB_ERR_WORKSPACE, "scratch too small: need at least %lld bytes", (long long)(128 * W.Bp * 2));
  if (chunk > round_up(Cs, 128)) chunk = round_up(Cs, 128);
  __nv_bfloat16* gt = static_cast<__nv_bfloat16*>(scratch);

  // one set-up launch in front of the GEMMs (bwd_setup_kernel); the persistent schedule adds its tile list to it
  BwdSetup su;
  su.lse = ext ? nullptr : lse; su.zt = target_logit; su.grad_loss = grad_loss; su.s = s;
  su.B = (int)B; su.Bp = (int)W.Bp; su.lse2 = W.lse2; su.zt2 = W.zt2; su.gscale = W.gscale;
  su.zero_a = W.tproj; su.n_a = Cs;
  su.zero_b = dxhat; su.n_b = B * W.Dp;
  su.zero_c = nullptr; su.n_c = 0; su.n_phases = 0;

  if (persistent) {
    // multi-buffered scratch: 4 buffers of `pchunk` classes each (fewer if the scratch is small)
    int nbuf = env_knobs().scratch_bufs;
    int64_t pchunk = scratch_bytes / (W.Bp * 2) / nbuf / 128 * 128;
    while (pchunk < 256 && nbuf > 2) {
      --nbuf;
      pchunk = scratch_bytes / (W.Bp * 2) / nbuf / 128 * 128;
    }
    // debug trace (ARCB_TRACE=1, fused path only): the last 1 MiB of the scratch receives per-tile clock stamps
    if (env_knobs().trace && !ext && scratch_bytes > (4ll << 20)) {
      const int64_t usable = scratch_bytes - (1ll << 20);
      unsigned long long* trace = reinterpret_cast<unsigned long long*>(static_cast<char*>(scratch) + usable);
      ARCB_CUDA(cudaMemsetAsync(trace, 0, 1ll << 20, st));
      int64_t pc = usable / (W.Bp * 2) / nbuf / 256 * 256;
      if (pc > round_up(Cs, 256)) pc = round_up(Cs, 256);
      if (pc >= 256) {
        if (pair_bwd) {
          if (dyn_feed)
            return launch_bwd_persistent<false, true, true, true>(W, s, grad_logits, ldgl, gt, pc, nbuf, dw, dxhat, st,
                                                                  su, trace);
          return launch_bwd_persistent<false, true, true>(W, s, grad_logits, ldgl, gt, pc, nbuf, dw, dxhat, st, su, trace);
        }
        return launch_bwd_persistent<false, false, true>(W, s, grad_logits, ldgl, gt, pc, nbuf, dw, dxhat, st, su, trace);
      }
    }
    if (pair_bwd) {  // a pair's tile has 256 rows: chunks (= scratch buffers) must be whole tiles
      int64_t pc2 = pchunk / 256 * 256;
      if (pc2 > round_up(Cs, 256)) pc2 = round_up(Cs, 256);
      if (pc2 >= 256) {
        if (dyn_feed)
          rc = ext ? launch_bwd_persistent<true, true, false, true>(W, s, grad_logits, ldgl, gt, pc2, nbuf, dw, dxhat, st, su,
                                                                    nullptr, coop)
                   : launch_bwd_persistent<false, true, false, true>(W, s, grad_logits, ldgl, gt, pc2, nbuf, dw, dxhat, st, su,
                                                                     nullptr, coop);
        else
          rc = ext ? launch_bwd_persistent<true, true>(W, s, grad_logits, ldgl, gt, pc2, nbuf, dw, dxhat, st, su)
                   : launch_bwd_persistent<false, true>(W, s, grad_logits, ldgl, gt, pc2, nbuf, dw, dxhat, st, su);
        if (rc != kLaunchRefused) return rc;
      }
    }
    if (pchunk >= 128) {
      if (pchunk > round_up(Cs, 128)) pchunk = round_up(Cs, 128);
      rc = ext ? launch_bwd_persistent<true, false>(W, s, grad_logits, ldgl, gt, pchunk, nbuf, dw, dxhat, st, su)
               : launch_bwd_persistent<false, false>(W, s, grad_logits, ldgl, gt, pchunk, nbuf, dw, dxhat, st, su);
      if (rc != kLaunchRefused) return rc;
    }
    // refused twice: one launch per GEMM and chunk needs no co-residency
  }
  {  // per-chunk schedule: the same set-up without a tile list
    LaunchScope ls(ARCB_K_BWD_PREP, st);
    bwd_setup_kernel<<<2 * device_sm_count(), 256, 0, st>>>(BwdPlan{}, nullptr, su);
    ARCB_CUDA(cudaGetLastError());
  }

  const int sms = device_sm_count();
  // fp32 outputs leave the GEMM epilogues as bulk tensor stores / reduce-adds of [128 x 32] boxes
  CUtensorMap tmDx, tmDw;
  if (!make_tmap_2d(&tmDx, dxhat, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, D, B, W.Dp * 4, 32, kBM))
    return fail(ARCB_ERR_CUDA, "tensor map for dxhat could not be encoded (16-byte alignment?)");
  const bool dw_tma = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(dw) & 15) == 0) &&
                      make_tmap_2d(&tmDw, dw, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, D, Cs, D * 4, 32, kBM);
  for (int64_t c0 = 0; c0 < Cs; c0 += chunk) {
    const int64_t rows = (Cs - c0 < chunk) ? (Cs - c0) : chunk;
    const __nv_bfloat16* wch = W.what + c0 * W.Dp;
    // (1) recompute cos^T for the chunk, write G^T = dLoss/dcos (bf16) and the projections t_j
    {
      CUtensorMap tmG;
      if (!make_tmap_2d(&tmG, gt, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, W.Bp, rows, W.Bp * 2, 64, kBM))
        return fail(ARCB_ERR_CUDA, "tensor map for the G^T scratch could not be encoded");
      Operand a{wch, rows, W.Dp, W.Dp}, b{W.xhat, B, W.Dp, W.Dp};
      cudaError_t e;
      LaunchScope ls(ARCB_K_BWD_G_GEMM, st);
      if (!ext) {
        EpiBwdG<false>::Params ep{W.lse2, W.gscale, W.lbl_local, W.zt2, W.dphi, W.cos_t, W.tproj,
                                  (int)c0, (int)rows, (int)B, (int)W.Bp, s * kLog2e, nullptr, 0, 0.f};
        e = launch_gemm<EpiBwdG<false>, kBNTile, 3, false, false>(a, b, &tmG, rows, B, W.Dp, 1, ep, st);
      } else {
        EpiBwdG<true>::Params ep{W.lse2, W.gscale, W.lbl_local, W.zt2, W.dphi, W.cos_t, W.tproj,
                                 (int)c0, (int)rows, (int)B, (int)W.Bp, s * kLog2e, grad_logits + c0, ldgl, s};
        e = launch_gemm<EpiBwdG<true>, kBNTile, 3, false, false>(a, b, &tmG, rows, B, W.Dp, 1, ep, st);
      }
      if (e != cudaSuccess) return fail(ARCB_ERR_CUDA, "backward recompute GEMM launch: %s", cudaGetErrorString(e));
    }
    // (2) dxhat[B, D] += G[B, rows] * what[rows, D]      (both operands have the class index slowest)
    {
      Operand a{gt, W.Bp, rows, W.Bp}, b{wch, W.Dp, rows, W.Dp};
      const int tiles = static_cast<int>(((B + kBM - 1) / kBM) * ((D + kBNTile - 1) / kBNTile));
      int split = sms / (tiles > 0 ? tiles : 1);
      if (split < 1) split = 1;
      EpiRedAdd::Params ep{(int)B, (int)D, nullptr};
      LaunchScope ls(ARCB_K_DX_GEMM, st);
      cudaError_t e = launch_gemm<EpiRedAdd, kBNTile, 3, true, true>(a, b, &tmDx, B, D, rows, split, ep, st);
      if (e != cudaSuccess) return fail(ARCB_ERR_CUDA, "dX GEMM launch: %s", cudaGetErrorString(e));
    }
    // (3) dW[rows, D] = inv_norm * (G^T[rows, B] * xhat[B, D] - t * what)
    {
      Operand a{gt, rows, W.Bp, W.Bp}, b{W.xhat, W.Dp, B, W.Dp};
      EpiDW::Params ep{dw, W.what, W.tproj, W.inv_nw, (int)c0, (int)rows, (int)D, (int)W.Dp, dw_tma ? 1 : 0};
      LaunchScope ls(ARCB_K_DW_GEMM, st);
      cudaError_t e = launch_gemm<EpiDW, kBNTile, 3, false, true>(a, b, dw_tma ? &tmDw : nullptr, rows, D, B, 1, ep, st);
      if (e != cudaSuccess) return fail(ARCB_ERR_CUDA, "dW GEMM launch: %s", cudaGetErrorString(e));
    }
  }
  return ARCB_OK;
}

}  // extern "C"
