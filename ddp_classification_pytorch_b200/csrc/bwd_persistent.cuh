// Persistent backward kernel of the ArcFace head (autograd behind ARCFACE/arc_main.py:329).
//
// One cooperative launch replaces the per-chunk sequence of three GEMM launches.  The classes of
// the shard are cut into chunks; every chunk k contributes three kinds of output tiles (128 x 256 per CTA):
//
//   G(k)  recompute cos^T = w_hat[chunk] * x_hat^T, epilogue -> bf16 dLoss/dcos tile G^T into an
//         L2-resident scratch of nbuf (default 4) chunk buffers (buffer k % nbuf) + per-class projections t_j
//   X(k)  dX_hat[B, D] += G[B, chunk] * w_hat[chunk, D]     (split along the chunk's classes,
//         fp32 bulk reduce-add)
//   W(k)  dW[chunk, D] = inv_norm * (G^T[chunk, B] * x_hat[B, D] - t * w_hat)
//
// There is ONE global tile list, written by bwd_setup_kernel before the launch: phase p = the G tiles of chunk p
// mixed proportionally with the X / W tiles of chunk p - lag (default 2).  CTA pairs claim the next tile of the
// list as they go (TileFeed<true>); the single-CTA build takes a fixed share (ItemWalker).  Dependencies are two
// arrays of global counters:
//   g_done[k]   half-tiles of G(k) whose stores have completed  -> gates the loads of X(k)/W(k)
//   xw_done[k]  X/W tiles of chunk k that finished reading G^T  -> gates the stores of G(k+nbuf)
// Every wait targets a tile that is EARLIER in the list and all CTAs are co-resident (cooperative
// launch), so the schedule cannot deadlock.
//
// Warp roles: 0 = TMA producer, 1 = tcgen05.mma issuer (pair: leader CTA; the follower's warp 1 claims tiles),
// 2 / 3 = G^T store issuers of epilogue warpgroup 1 / 0 (TMA store + completion + release of g_done, off the
// epilogue's critical path; warp 2 also owns the TMEM allocation), 4..11 = epilogue (two warpgroups, half the
// columns each).
//
// Two builds of the same kernel (template parameter PAIR):
//   PAIR = false  one CTA per 128 x 256 tile (cta_group::1), 4 operand stages of 48 KB, one 16 KB output slot
//                 per epilogue warpgroup (a [128 x 64] bf16 box of G^T, or a [128 x 32] fp32 chunk of dX / dW)
//   PAIR = true   a cluster of two CTAs per 256 x 256 tile (cta_group::2, see gemm2_sm100.cuh): each CTA
//                 stages its own 128 rows of A and half of B, 5 stages of 32 KB, two output slots per
//                 warpgroup; the backward tiles are shared-memory-bandwidth bound (operand traffic + epilogue
//                 staging), and the pair cuts the operand traffic per SM by a third.  Tiles, counters and
//                 epilogues are per CTA as before; only the leader issues MMAs.
#pragma once

#include "epilogues.cuh"

namespace arcb {

constexpr int kPBN = 256;
constexpr int kPABytes = kBM * kBK * 2;
template <bool PAIR>
struct PCfg {
  static constexpr int kStages = PAIR ? 5 : 4;
  static constexpr int kBBytes = (PAIR ? kPBN / 2 : kPBN) * kBK * 2;
  static constexpr int kStageBytes = kPABytes + kBBytes;  // 32 KB / 48 KB
  // epilogue slots per warpgroup: with two, a warpgroup never waits inside a tile for the bulk store of its
  // previous G^T box / dX / dW chunk to finish reading shared memory (the pair's smaller stages pay for it)
  static constexpr int kBufs = PAIR ? 2 : 1;
  static constexpr int kRowsOff = 2 * kBufs * kBM * 128;   // staged lse2 / labels behind the slots
  static constexpr int kRing = 8;  // tile records a CTA pair may hold ahead of its slowest role (dynamic feed)
  static constexpr int kNumBars = 2 * kStages + 4 + 4 * kBufs + 3 * kRing;  // + slot_full / slot_free [2 wg][kBufs], feed
  static constexpr int kBarOff = kStages * kStageBytes + kRowsOff + 256 * 8;
  static constexpr int kRingOff = kBarOff + kNumBars * 8 + 16;  // after the barriers and the TMEM slot; 16 B aligned
  static constexpr int kSmemTotal = kRingOff + kRing * 32;      // no slack: the dynamic smem base is 1024 B aligned
  static constexpr int kCtas = PAIR ? 2 : 1;              // CTAs that work on one tile
};
constexpr int kPSlotBytes = kBM * 128;                   // 16 KB: [128 rows x 128 B], one per epilogue warpgroup
static_assert(PCfg<false>::kSmemTotal <= 232448 && PCfg<true>::kSmemTotal <= 232448,
              "persistent backward kernel exceeds the shared memory budget");

enum : int { kItemG = 0, kItemX = 1, kItemW = 2 };

struct BwdPlan {
  int n_chunks, chunk_rows;  // chunk_rows % 128 == 0
  int tile_rows;             // accumulator rows of one tile: 128 (one CTA) or 256 (CTA pair)
  int lag;                   // phases between the G tiles of a chunk and its X / W tiles (default 2)
  int nbuf;                  // scratch buffers of chunk_rows rows each (default 4); chunk k uses buffer k % nbuf
  int Cs, B, Bp, D, Dp;
  int g_tiles_n;             // Bp / 256
  int g_kb;                  // Dp / 64
  int x_tiles_m, x_tiles_n;  // ceil(B / tile_rows), ceil(D / 256)
  int x_kb_per_split;        // 64-class blocks per X tile
  int w_tiles_n, w_kb;       // ceil(D / 256), ceil(B / 64)
  unsigned int* g_done;      // [n_chunks], zeroed before the launch
  unsigned int* xw_done;     // [n_chunks], zeroed before the launch
  int kind_mask;             // debug: bit k set = tiles of kind k run (7 = all); anything else gives wrong results
  unsigned int* claim;       // dynamic feed: next unclaimed list index (zeroed before the launch)
  int rot;                   // slot rotation per dealing round (see ItemWalker), 0 = plain round-robin
  int debug_epi;             // debug (ARCB_DEBUG_EPI): 1 = G epilogue skips math + smem stores, 2 = also skips the TMEM loads
  const int4* items;         // [total_items][2] the global tile list, written by bwd_setup_kernel before the launch
  int total_items;
};

struct BwdArgs {
  const float* lse2;
  const float* gscale;
  const int* lbl_local;
  const float* zt2;
  const float* dphi;
  const float* cos_t;
  float* tproj;
  float s2;
  const float* grad_out;  // EXT only
  int64_t ldgo;
  float s_ext;
  float* dw;
  const __nv_bfloat16* what;
  const float* inv_nw;
  int dw_tma;
  unsigned long long* trace;  // TRACE builds only: [8 CTAs][kTraceItems][kTraceWords] clock64 stamps
};

constexpr int kTraceItems = 1024;
constexpr int kTraceWords = 12;
template <bool TRACE>
__device__ __forceinline__ unsigned long long* trace_slot(const BwdArgs& ar, int item_no) {
  if constexpr (!TRACE) return nullptr;
  if (!ar.trace || blockIdx.x >= 8 || item_no >= kTraceItems) return nullptr;
  return ar.trace + (static_cast<size_t>(blockIdx.x) * kTraceItems + item_no) * kTraceWords;
}

struct Item {
  int kind, chunk, m_blk, n_blk, kb0, kb1;
  unsigned int target;  // X / W tile: value of g_done[chunk] that makes its G^T complete;
                        // G tile: value of xw_done[chunk - nbuf] that frees the scratch buffer it overwrites
  int buf;              // scratch buffer of the chunk (chunk % nbuf)
};

__host__ __device__ inline int plan_rows(const BwdPlan& pl, int k) {
  const int rest = pl.Cs - k * pl.chunk_rows;
  return rest < pl.chunk_rows ? rest : pl.chunk_rows;
}
__host__ __device__ inline int plan_gm(const BwdPlan& pl, int k) {
  return (plan_rows(pl, k) + pl.tile_rows - 1) / pl.tile_rows;
}
__host__ __device__ inline int plan_nG(const BwdPlan& pl, int k) { return plan_gm(pl, k) * pl.g_tiles_n; }
__host__ __device__ inline int plan_xsplits(const BwdPlan& pl, int k) {
  const int kbt = (plan_rows(pl, k) + kBK - 1) / kBK;
  return (kbt + pl.x_kb_per_split - 1) / pl.x_kb_per_split;
}
__host__ __device__ inline int plan_nX(const BwdPlan& pl, int k) {
  return pl.x_tiles_m * pl.x_tiles_n * plan_xsplits(pl, k);
}
__host__ __device__ inline int plan_nW(const BwdPlan& pl, int k) { return plan_gm(pl, k) * pl.w_tiles_n; }

// The global list is a sequence of n_chunks + lag phases; phase p holds the G tiles of chunk p (if any) mixed
// PROPORTIONALLY with the X / W tiles of chunk p - lag (if any), and X with W likewise.  Mixing matters: if all
// CTAs reached the dW tiles together, the burst of fp32 dW stores (~3 TB/s) would saturate HBM writes and
// stall those epilogues; spread over the whole phase the same bytes need ~0.7 TB/s.  With a lag of two phases
// the G tiles an X / W tile depends on were all issued a full phase earlier (and, with 4 scratch buffers, so
// were the X / W tiles whose buffer a G tile overwrites), so a CTA rarely waits on a counter.
// (the lag is pl.lag, default 2)
__device__ __forceinline__ int phase_nG(const BwdPlan& pl, int p) { return p < pl.n_chunks ? plan_nG(pl, p) : 0; }
__device__ __forceinline__ int phase_nXW(const BwdPlan& pl, int p) {
  return (p >= pl.lag && p - pl.lag < pl.n_chunks) ? plan_nX(pl, p - pl.lag) + plan_nW(pl, p - pl.lag) : 0;
}
__device__ __forceinline__ int phase_items(const BwdPlan& pl, int p) { return phase_nG(pl, p) + phase_nXW(pl, p); }

// j-th of `total` slots, of which `n_b` belong to stream B spread evenly: is slot j a B slot, and which one?
__device__ __forceinline__ bool pick_b(int j, int n_b, int total, int& idx_b, int& idx_a) {
  const int before = static_cast<int>((static_cast<long long>(j) * n_b) / total);
  const int after = static_cast<int>((static_cast<long long>(j + 1) * n_b) / total);
  idx_b = before;
  idx_a = j - before;
  return after > before;
}

// Item j of phase `phase` (the arithmetic of the global list: integer divisions, done once per item by
// bwd_setup_kernel instead of by every role of every CTA in front of every tile).
__device__ inline void make_item(const BwdPlan& pl, int phase, int j, Item& it) {
  const int ctas = pl.tile_rows / kBM;
  const int phase_size = phase_items(pl, phase);
  const int n_xw = phase_nXW(pl, phase);
  int i_xw, i_g;
  if (!pick_b(j, n_xw, phase_size, i_xw, i_g)) {  // a G tile of chunk `phase`
    const int k = phase;
    const int gm = plan_gm(pl, k);
    it.kind = kItemG;
    it.chunk = k;
    it.m_blk = i_g % gm;
    it.n_blk = i_g / gm;
    it.kb0 = 0;
    it.kb1 = pl.g_kb;
    const int kp = k - pl.nbuf;
    it.target = kp >= 0 ? static_cast<unsigned>(ctas * (plan_nX(pl, kp) + plan_nW(pl, kp))) : 0u;
  } else {  // an X or W tile of chunk `phase - lag`
    const int k = phase - pl.lag;
    const int gm = plan_gm(pl, k);
    const int nx = plan_nX(pl, k);
    it.chunk = k;
    int i_x, i_w;
    if (pick_b(i_xw, nx, n_xw, i_x, i_w)) {
      it.kind = kItemX;
      it.m_blk = i_x % pl.x_tiles_m;
      const int r = i_x / pl.x_tiles_m;
      it.n_blk = r % pl.x_tiles_n;
      const int ks = r / pl.x_tiles_n;
      const int kbt = (plan_rows(pl, k) + kBK - 1) / kBK;
      it.kb0 = ks * pl.x_kb_per_split;
      it.kb1 = min(it.kb0 + pl.x_kb_per_split, kbt);
    } else {
      it.kind = kItemW;
      it.m_blk = i_w % gm;
      it.n_blk = i_w / gm;
      it.kb0 = 0;
      it.kb1 = pl.w_kb;
    }
    it.target = 2u * static_cast<unsigned>(ctas * plan_nG(pl, k));
  }
  it.buf = it.chunk % pl.nbuf;
}

// Items in phases [0, p): every chunk but the last has the same tile counts.
__host__ __device__ inline long long phase_begin_of(const BwdPlan& pl, int p) {
  const int nc = pl.n_chunks;
  auto prefix = [&](int k, bool g) -> long long {  // tiles of chunks [0, k)
    if (k <= 0) return 0;
    if (k > nc) k = nc;
    const long long full = g ? plan_nG(pl, 0) : plan_nX(pl, 0) + plan_nW(pl, 0);
    if (k < nc) return full * k;
    const long long tail = g ? plan_nG(pl, nc - 1) : plan_nX(pl, nc - 1) + plan_nW(pl, nc - 1);
    return full * (nc - 1) + tail;
  };
  return prefix(p, true) + prefix(p - pl.lag, false);
}

// Everything the backward needs set up before its GEMMs start, in ONE launch (it used to be two kernels and three
// memsets, i.e. five dependent launches in front of a kernel that takes 0.3 ms on an 8-way shard):
//   * the tile list: one block per phase writes that phase's records (two int4 each);
//   * per-sample inputs of the recompute epilogue: lse2 = lse * log2(e) (+inf on padding), zt2 likewise, and
//     gscale = upstream grad * s / B                                  (skipped when grad_logits is given);
//   * zeroes: the per-class projections t_j, the dxhat accumulator, the chunk counters + claim counter.
struct BwdSetup {
  const float* lse;        // nullptr: no per-sample preparation (grad_logits path)
  const float* zt;
  const float* grad_loss;  // device scalar or nullptr (= 1)
  float s;
  int B, Bp;
  float* lse2;
  float* zt2;
  float* gscale;
  float* zero_a;           // t_j
  long long n_a;
  float* zero_b;           // dxhat
  long long n_b;
  unsigned int* zero_c;    // counters
  int n_c;
  int n_phases;            // 0: no tile list (per-chunk schedule)
};

__global__ void __launch_bounds__(256) bwd_setup_kernel(const BwdPlan pl, int4* __restrict__ out, const BwdSetup su) {
  for (int p = blockIdx.x; p < su.n_phases; p += gridDim.x) {
    const long long begin = phase_begin_of(pl, p);
    const int n = phase_items(pl, p);
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
      Item it;
      make_item(pl, p, j, it);
      out[2 * (begin + j)] = make_int4(it.kind, it.chunk, it.m_blk, it.n_blk);
      out[2 * (begin + j) + 1] = make_int4(it.kb0, it.kb1, static_cast<int>(it.target), it.buf);
    }
  }
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthr = static_cast<long long>(gridDim.x) * blockDim.x;
  if (su.lse) {
    if (tid == 0) *su.gscale = (su.grad_loss ? *su.grad_loss : 1.f) * su.s / static_cast<float>(su.B);
    for (long long i = tid; i < su.Bp; i += nthr) {
      su.lse2[i] = (i < su.B) ? su.lse[i] * kLog2e : INFINITY;
      su.zt2[i] = (i < su.B) ? su.zt[i] * kLog2e : -INFINITY;
    }
  }
  for (long long i = tid; i < su.n_a; i += nthr) su.zero_a[i] = 0.f;
  for (long long i = tid; i < su.n_b; i += nthr) su.zero_b[i] = 0.f;
  for (long long i = tid; i < su.n_c; i += nthr) su.zero_c[i] = 0u;
}

// Every role of a CTA (or CTA pair) walks the same sub-sequence of the global list: tiles first, first + step, ...
// The record of the following tile is requested while the current one is being worked on.
struct ItemWalker {
  int round_base, slot, stride, rot_step;  // next record = round_base + slot; the slot rotates from round to round
  int4 a, b;
  __device__ __forceinline__ void fetch(const BwdPlan& pl) {
    const int n = round_base + slot;
    if (n < pl.total_items) {
      a = __ldg(pl.items + 2 * static_cast<size_t>(n));
      b = __ldg(pl.items + 2 * static_cast<size_t>(n) + 1);
    }
  }
  __device__ __forceinline__ void init(const BwdPlan& pl, int first, int step) {
    round_base = 0;
    slot = first;
    stride = step;
    // the list is dealt in rounds of `stride` consecutive tiles; which tile of a round a CTA (pair) takes moves
    // by a fixed odd-ish step every round, so that the periodic G / X / W pattern of the list cannot line up with
    // the dealing and hand one CTA the expensive kind round after round (pl.rot = 0: plain round-robin)
    rot_step = pl.rot % step;
    fetch(pl);
  }
  // false once this CTA's share of the list is exhausted (the last, partial round may not reach its slot)
  __device__ __forceinline__ bool next(const BwdPlan& pl, Item& it) {
    while (round_base < pl.total_items && round_base + slot >= pl.total_items) advance(pl);
    if (round_base >= pl.total_items) return false;
    it.kind = a.x; it.chunk = a.y; it.m_blk = a.z; it.n_blk = a.w;
    it.kb0 = b.x; it.kb1 = b.y; it.target = static_cast<unsigned>(b.z); it.buf = b.w;
    advance(pl);
    return true;
  }
  __device__ __forceinline__ void advance(const BwdPlan& pl) {
    round_base += stride;
    slot += rot_step;
    if (slot >= stride) slot -= stride;
    fetch(pl);
  }
};

// Where a role gets its tiles from.
//   static  (DYN = false): the CTA's fixed share of the list, see ItemWalker.
//   dynamic (DYN = true, CTA pairs): one thread of the pair (the follower's otherwise idle warp 1) claims the next
//     list index with an atomic whenever a ring slot is free, fetches the record and pushes it into the ring of
//     BOTH CTAs (st.async counted on each CTA's feed_full barrier).  Every role reads the records in ring order and
//     reports on the claimer's feed_empty barrier.  Tiles are still started in list order, so every dependency
//     (a counter a tile waits for) belongs to a tile with a smaller index that some pair has already claimed and
//     will reach without needing anything later in the list: no deadlock.  A pair that falls behind simply claims
//     less, which is what the fixed dealing could not do.
template <bool DYN>
struct TileFeed {
  ItemWalker w;            // static
  const int4* ring;        // dynamic: this CTA's ring [kRing][2]
  uint64_t* full;          // this CTA's feed_full[kRing]
  uint32_t empty_remote;   // shared::cluster address of feed_empty[0] in the claimer's CTA
  uint64_t* taken;         // claimer's CTA, producer only: feed_taken[kRing]
  uint32_t n;
  bool armer, collective;
  // armer: the one thread per CTA that announces the 32 bytes of every record on feed_full (the TMA producer);
  // collective: all 32 lanes of a warp call next() together (epilogue warps) and count as ONE reader.
  __device__ __forceinline__ void init(const BwdPlan& pl, int first, int step, uint8_t* ring_smem, uint64_t* full_bars,
                                       uint64_t* empty_bars, bool armer_, bool collective_,
                                       uint64_t* taken_bars = nullptr) {
    if constexpr (!DYN) {
      w.init(pl, first, step);
    } else {
      ring = reinterpret_cast<const int4*>(ring_smem);
      full = full_bars;
      empty_remote = mapa_u32(smem_u32(empty_bars), 1);
      n = 0;
      armer = armer_;
      collective = collective_;
      taken = taken_bars;
    }
  }
  __device__ __forceinline__ bool next(const BwdPlan& pl, Item& it) {
    if constexpr (!DYN) {
      return w.next(pl, it);
    } else {
      const uint32_t slot = n % PCfg<true>::kRing, use = n / PCfg<true>::kRing;
      ++n;
      if (armer) mbar_arrive_expect_tx(&full[slot], 32);
      mbar_wait(&full[slot], use & 1u, 0xb00u + slot);
      const int4 a = ring[2 * slot], b = ring[2 * slot + 1];
      if (collective) __syncwarp();
      if (!collective || lane_id() == 0) mbar_arrive_cluster(empty_remote + slot * 8);
      if (taken) mbar_arrive(&taken[slot]);  // the pair has started on this tile: the claimer may fetch the next one
      it.kind = a.x; it.chunk = a.y; it.m_blk = a.z; it.n_blk = a.w;
      it.kb0 = b.x; it.kb1 = b.y; it.target = static_cast<unsigned>(b.z); it.buf = b.w;
      return a.x >= 0;
    }
  }
};

// The claimer of a CTA pair (dynamic feed): runs until it has handed out the end-of-list record.
// A tile is claimed only once the pair's producer has started on the previous one: claiming further ahead would
// park tiles that other pairs' tiles depend on behind this pair's backlog (measured: counter waits x4).  The ring
// is deeper than that look-ahead only because the slowest role (the epilogue) reads a record a few tiles later.
__device__ __forceinline__ void feed_claim_loop(const BwdPlan& pl, uint8_t* ring_smem, uint64_t* full_bars,
                                                uint64_t* empty_bars, uint64_t* taken_bars) {
  constexpr int kRing = PCfg<true>::kRing;
  for (uint32_t n = 0;; ++n) {
    const uint32_t slot = n % kRing, use = n / kRing;
    mbar_wait(&empty_bars[slot], (use & 1u) ^ 1u, 0xc00u + slot);  // every role of both CTAs has read the old record
    if (n >= 1) {
      const uint32_t ps = (n - 1) % kRing, pu = (n - 1) / kRing;
      mbar_wait(&taken_bars[ps], pu & 1u, 0xd00u + ps);
    }
    const unsigned int idx = atomicAdd(pl.claim, 1u);
    int4 a = make_int4(-1, 0, 0, 0), b = make_int4(0, 0, 0, 0);
    if (idx < static_cast<unsigned int>(pl.total_items)) {
      a = __ldg(pl.items + 2 * static_cast<size_t>(idx));
      b = __ldg(pl.items + 2 * static_cast<size_t>(idx) + 1);
    }
    const uint32_t dst = smem_u32(ring_smem) + slot * 32, bar = smem_u32(&full_bars[slot]);
#pragma unroll
    for (uint32_t cta = 0; cta < 2; ++cta) {
      const uint32_t d = mapa_u32(dst, cta), m = mapa_u32(bar, cta);
      st_async_v4(d, a, m);
      st_async_v4(d + 16, b, m);
    }
    if (a.x < 0) break;
  }
}

// TMEM stage drained: in PAIR mode both CTAs' epilogue warps report to the leader, whose MMA issuer waits
template <bool PAIR>
__device__ __forceinline__ void arrive_tempty(uint64_t* bar) {
  if constexpr (PAIR) mbar_arrive_cluster(mapa_u32(smem_u32(bar), 0)); else mbar_arrive(bar);
}

// operand loads: `bar_addr` is the shared::cluster address of the full barrier (the pair leader's in PAIR mode)
template <bool PAIR>
__device__ __forceinline__ void ld2(void* dst, const CUtensorMap* tm, uint32_t bar_addr, int c0, int c1, uint64_t pol) {
  if constexpr (PAIR) {
    tma_load_2d_pair(dst, tm, bar_addr, c0, c1, pol);
  } else {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar_addr), "r"(c0), "r"(c1), "l"(pol)
        : "memory");
  }
}
template <bool PAIR>
__device__ __forceinline__ void ld3(void* dst, const CUtensorMap* tm, uint32_t bar_addr, int c0, int c1, int c2,
                                    uint64_t pol) {
  if constexpr (PAIR) {
    tma_load_3d_pair(dst, tm, bar_addr, c0, c1, c2, pol);
  } else {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar_addr), "r"(c0), "r"(c1), "r"(c2), "l"(pol)
        : "memory");
  }
}

template <bool EXT, bool PAIR, bool TRACE, bool DYN>
__global__ void __launch_bounds__(kGemmThreads, 1)
bwd_persistent_kernel(const __grid_constant__ CUtensorMap tmWhatK,   // w_hat [Cs, Dp]   box 64 x 128  (A of G)
                      const __grid_constant__ CUtensorMap tmXhatK,   // x_hat [B, Dp]    box 64 x 256  (B of G)
                      const __grid_constant__ CUtensorMap tmGtSt,    // G^T   [2R, Bp]   box 64 x 128  (store of G)
                      const __grid_constant__ CUtensorMap tmGtMN,    // G^T   [2R, Bp]   box 64 x 64   (A of X)
                      const __grid_constant__ CUtensorMap tmWhatMN,  // w_hat [Cs, Dp]   box 64 x 64   (B of X)
                      const __grid_constant__ CUtensorMap tmGtK,     // G^T   [2R, Bp]   box 64 x 128  (A of W)
                      const __grid_constant__ CUtensorMap tmXhatMN,  // x_hat [B, Dp]    box 64 x 64   (B of W)
                      const __grid_constant__ CUtensorMap tmDx,      // dxhat [B, D] f32 box 32 x 128
                      const __grid_constant__ CUtensorMap tmDw,      // dW    [Cs, D] f32 box 32 x 128
                      const BwdPlan pl, const BwdArgs ar) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0u) watchdog_abort(0xa11u, 0u);  // 128 B swizzle needs 1024 B aligned tiles; no slack is reserved
  using C = PCfg<PAIR>;
  constexpr int kPStages = C::kStages;
  constexpr int kPStageBytes = C::kStageBytes;
  uint8_t* epi_smem = smem + kPStages * kPStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kBarOff);
  uint64_t* full = bars;
  uint64_t* empty = bars + kPStages;
  uint64_t* tfull = bars + 2 * kPStages;
  uint64_t* tempty = bars + 2 * kPStages + 2;
  constexpr int kBufs = C::kBufs;
  uint64_t* slot_full = bars + 2 * kPStages + 4;              // [wg][buf] warpgroup -> its store issuer: box is in smem
  uint64_t* slot_free = bars + 2 * kPStages + 4 + 2 * kBufs;  // [wg][buf] store issuer -> warpgroup: box has been read
  uint64_t* feed_full = bars + 2 * kPStages + 4 + 4 * kBufs;  // dynamic feed: record n is in ring slot n % kRing
  uint64_t* feed_empty = feed_full + C::kRing;                // ... and every role has read it (claimer's CTA only)
  uint64_t* feed_taken = feed_empty + C::kRing;               // ... and the claimer's own producer has started on it
  uint8_t* ring_smem = smem + C::kRingOff;
  static_assert(!DYN || PAIR, "the dynamic feed uses the follower CTA's idle warp 1 as claimer");
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + C::kNumBars);
  const int rank = PAIR ? static_cast<int>(cluster_ctarank()) : 0;   // position inside the CTA pair
  const bool leader = rank == 0;
  const int walk_first = static_cast<int>(blockIdx.x) / C::kCtas;   // tile group of this CTA
  const int walk_step = static_cast<int>(gridDim.x) / C::kCtas;

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmWhatK);
    tma_prefetch_desc(&tmXhatK);
    tma_prefetch_desc(&tmGtSt);
    tma_prefetch_desc(&tmGtMN);
    tma_prefetch_desc(&tmWhatMN);
    tma_prefetch_desc(&tmGtK);
    tma_prefetch_desc(&tmXhatMN);
    tma_prefetch_desc(&tmDx);
    tma_prefetch_desc(&tmDw);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kPStages; ++i) {
      mbar_init(&full[i], C::kCtas);  // pair: leader producer (+tx bytes of both CTAs) and follower producer
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], C::kCtas * (kEpiThreads / 32));  // pair: both CTAs' epilogue warps, on the leader
    }
    for (int i = 0; i < 2 * kBufs; ++i) {
      mbar_init(&slot_full[i], kWgThreads / 32);
      mbar_init(&slot_free[i], 1);
    }
    for (int i = 0; i < C::kRing; ++i) {
      mbar_init(&feed_full[i], 1);  // the CTA's producer thread (+ the 32 record bytes)
      // readers of a record: producer, 2 store issuers and 8 epilogue warps per CTA + the leader's MMA issuer
      mbar_init(&feed_empty[i], 2 * (3 + kEpiThreads / 32) + 1);
      mbar_init(&feed_taken[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    if constexpr (PAIR) {
      tmem_alloc_pair(tmem_slot, 2 * kPBN);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(tmem_slot, 2 * kPBN);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------ TMA producer ------------------------------
      TileFeed<DYN> w;
      w.init(pl, walk_first, walk_step, ring_smem, feed_full, feed_empty, true, false, leader ? nullptr : feed_taken);
      Item it;
      int stage = 0;
      uint32_t phase = 0;
      int item_no = -1;
      while (w.next(pl, it)) {
        if (it.kb0 >= it.kb1 || !((pl.kind_mask >> it.kind) & 1)) continue;
        ++item_no;
        unsigned long long* tr = trace_slot<TRACE>(ar, item_no);
        long long w_empty = 0;
        const long long t_flag = TRACE ? clock64() : 0;
        if (it.kind != kItemG && (pl.kind_mask & 1)) {  // G^T of this chunk must be complete and visible
          wait_counter(pl.g_done + it.chunk, it.target, 0x500u);
          fence_proxy_async_all();
        }
        if (tr) {
          tr[8] = clock64() - t_flag;
          tr[10] = clock64();
        }
        const int mblk = it.m_blk * C::kCtas + rank;            // this CTA's 128-row block of the tile
        const int c0 = it.chunk * pl.chunk_rows;                // first class of the chunk in the shard
        const int buf_row = it.buf * pl.chunk_rows;             // first row of the chunk in the G^T scratch
        for (int kb = it.kb0; kb < it.kb1; ++kb) {
          if (tr) {
            const long long t0 = clock64();
            mbar_wait(&empty[stage], phase ^ 1u, 0x100u + stage);
            w_empty += clock64() - t0;
          } else {
            mbar_wait(&empty[stage], phase ^ 1u, 0x100u + stage);
          }
          // pair: both CTAs' bytes are counted on the leader's barrier (count 2: leader +tx, follower plain)
          const uint32_t full_addr = PAIR ? mapa_u32(smem_u32(&full[stage]), 0) : smem_u32(&full[stage]);
          if (leader) mbar_arrive_expect_tx(&full[stage], C::kCtas * kPStageBytes);
          uint8_t* sa = smem + stage * kPStageBytes;
          uint8_t* sb = sa + kPABytes;
          const int half = rank * (kPBN / 2);  // this CTA's half of the B tile (pair), else 0
          if (it.kind == kItemG) {
            // w_hat rows are read again by X(k) and by the dW epilogue, x_hat by every tile: keep both in L2
            ld2<PAIR>(sa, &tmWhatK, full_addr, kb * kBK, c0 + mblk * kBM, kEvictLast);
            ld2<PAIR>(sb, &tmXhatK, full_addr, kb * kBK, it.n_blk * kPBN + half, kEvictLast);
          } else if (it.kind == kItemX) {
            // A = G (class index slowest): 2 blocks of 64 samples; B = w_hat: blocks of 64 features; W(k) re-reads G^T
            ld3<PAIR>(sa, &tmGtMN, full_addr, 0, buf_row + kb * kBK, mblk * (kBM / 64), kEvictLast);
            ld3<PAIR>(sb, &tmWhatMN, full_addr, 0, c0 + kb * kBK, (it.n_blk * kPBN + half) / 64, kEvictNormal);
          } else {
            // last reader of this part of G^T: let it leave L2 first instead of being written back later
            ld2<PAIR>(sa, &tmGtK, full_addr, kb * kBK, buf_row + mblk * kBM, kEvictFirst);
            ld3<PAIR>(sb, &tmXhatMN, full_addr, 0, kb * kBK, (it.n_blk * kPBN + half) / 64, kEvictLast);
          }
          if (PAIR && !leader) mbar_arrive_cluster(full_addr);
          if (++stage == kPStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (tr) {
          tr[7] = w_empty;
          tr[11] = clock64();
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {
      // ------------------------------ MMA issuer (pair: leader only) --------------------------------
      TileFeed<DYN> w;
      w.init(pl, walk_first, walk_step, ring_smem, feed_full, feed_empty, false, false);
      Item it;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      int item_no = -1;
      while (w.next(pl, it)) {
        if (it.kb0 >= it.kb1 || !((pl.kind_mask >> it.kind) & 1)) continue;
        ++item_no;
        unsigned long long* tr = trace_slot<TRACE>(ar, item_no);
        long long w_full = 0;
        if (tr) tr[0] = clock64();
        const bool a_mn = it.kind == kItemX;
        const bool b_mn = it.kind != kItemG;
        const uint32_t idesc = make_idesc_bf16(C::kCtas * kBM, kPBN, a_mn, b_mn);
        // descriptor of k-slice 0 at smem offset 0; per slice only the (16-byte granular) start address moves
        const uint64_t da0 = a_mn ? make_smem_desc(0, kBK * 128, 1024) : make_smem_desc(0, 16, 1024);
        const uint64_t db0 = b_mn ? make_smem_desc(0, kBK * 128, 1024) : make_smem_desc(0, 16, 1024);
        const uint32_t a_step = a_mn ? (2048u >> 4) : (32u >> 4);
        const uint32_t b_step = b_mn ? (2048u >> 4) : (32u >> 4);
        mbar_wait(&tempty[acc], acc_phase ^ 1u, 0x200u + acc);
        tc_fence_after();
        if (tr) tr[1] = clock64();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * kPBN);
        for (int kb = it.kb0; kb < it.kb1; ++kb) {
          if (tr) {
            const long long t0 = clock64();
            mbar_wait(&full[stage], phase, 0x300u + stage);
            w_full += clock64() - t0;
          } else {
            mbar_wait(&full[stage], phase, 0x300u + stage);
          }
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem + stage * kPStageBytes);
          const uint64_t da = da0 + (a_base >> 4);
          const uint64_t db = db0 + ((a_base + kPABytes) >> 4);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint32_t accum = (kb > it.kb0 || k > 0) ? 1u : 0u;
            if constexpr (PAIR) tc_mma_bf16_pair(tmem_d, da + k * a_step, db + k * b_step, idesc, accum);
            else tc_mma_bf16(tmem_d, da + k * a_step, db + k * b_step, idesc, accum);
          }
          if constexpr (PAIR) tc_commit_pair(&empty[stage], 0x3); else tc_commit(&empty[stage]);
          if (++stage == kPStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if constexpr (PAIR) tc_commit_pair(&tfull[acc], 0x3); else tc_commit(&tfull[acc]);
        if (tr) {
          tr[2] = clock64();
          tr[3] = w_full;
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    } else if (DYN && lane == 0 && !leader) {
      // ------------------------------ tile claimer of the pair (dynamic feed) ------------------------
      feed_claim_loop(pl, ring_smem, feed_full, feed_empty, feed_taken);
    }
  } else if (warp == 2 || warp == 3) {
    if (lane == 0) {
      // ---------------- G^T store issuer of epilogue warpgroup (3 - warp) ----------------
      const int wg = 3 - warp;
      TileFeed<DYN> w;
      w.init(pl, walk_first, walk_step, ring_smem, feed_full, feed_empty, false, false);
      Item it;
      uint32_t n_box = 0;  // boxes this issuer has sent so far (two per G tile); box n lives in buffer n % kBufs
      while (w.next(pl, it)) {
        if (it.kind != kItemG || it.kb0 >= it.kb1 || !(pl.kind_mask & 1)) continue;
        const int buf_row = it.buf * pl.chunk_rows;
        const int mblk = it.m_blk * C::kCtas + rank;
        for (int bx = 0; bx < 2; ++bx, ++n_box) {
          const uint32_t buf = n_box % kBufs, use = n_box / kBufs;
          uint8_t* slot = epi_smem + (wg * kBufs + buf) * kPSlotBytes;
          mbar_wait(&slot_full[wg * kBufs + buf], use & 1u, 0x600u + wg);
          if (bx == 0 && it.chunk >= pl.nbuf && (pl.kind_mask & 6) == 6) {  // the X / W tiles that last read this scratch buffer must be done
            wait_counter(pl.xw_done + (it.chunk - pl.nbuf), it.target, 0x700u);
            fence_proxy_async_all();
          }
          tma_store_2d_hint(&tmGtSt, slot, it.n_blk * kPBN + wg * 128 + bx * 64, buf_row + mblk * kBM, kEvictLast);
          tma_store_commit();
          tma_store_wait_read0();
          mbar_arrive(&slot_free[wg * kBufs + buf]);  // the warpgroup may overwrite this buffer
        }
        tma_store_wait_all0();  // both boxes are written (and, through the barrier chain, the t_j atomics are done)
        __threadfence();
        atomicAdd(pl.g_done + it.chunk, 1u);  // each G tile counts twice: once per warpgroup
      }
    }
  } else if (warp >= 4) {
    // -------------------------------- epilogue ----------------------------------
    const int quad = warp & 3;
    const int row_in_tile = quad * 32 + lane;
    const int epi_tid = static_cast<int>(threadIdx.x) - (kGemmThreads - kEpiThreads);
    const int wg = epi_tid >> 7;
    const int wg_tid = epi_tid & 127;
    float* lse_s = reinterpret_cast<float*>(epi_smem + C::kRowsOff) + wg * 128;  // this warpgroup's 128 samples
    int* lbl_s = reinterpret_cast<int*>(epi_smem + C::kRowsOff + 256 * 4) + wg * 128;
    TileFeed<DYN> w;
    w.init(pl, walk_first, walk_step, ring_smem, feed_full, feed_empty, false, true);
    Item it;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t n_box = 0;  // G^T boxes this warpgroup has handed to its store issuer
    int prev_kind = -1;
    EpiOutState ost;
    ost.it = 0;
    const float gscale = EXT ? 0.f : __ldg(ar.gscale);
    int item_no = -1;
    while (w.next(pl, it)) {
      if (it.kb0 >= it.kb1 || !((pl.kind_mask >> it.kind) & 1)) continue;
      ++item_no;
      unsigned long long* tr = (epi_tid == 0) ? trace_slot<TRACE>(ar, item_no) : nullptr;
      if (tr) {
        tr[4] = clock64();
        tr[9] = it.kind;
      }
      // per-tile row data is fetched before waiting for the accumulator, so its latency is hidden
      float lse_v = 0.f;
      int lbl_v = -1;
      EpiDW::Params pw;
      EpiDW::RowData rdw;
      if (it.kind == kItemG) {
        if (!EXT) lse_v = __ldg(ar.lse2 + it.n_blk * kPBN + epi_tid);
        lbl_v = __ldg(ar.lbl_local + it.n_blk * kPBN + epi_tid);
      } else if (it.kind == kItemW) {
        // (t_j of this chunk is final: the producer of this tile only loads operands once g_done[chunk] is
        //  complete, but the epilogue may run ahead of it -- so wait for the same counter here)
        pw.dw = ar.dw;
        pw.what = ar.what;
        pw.t = ar.tproj;
        pw.inv_norm = ar.inv_nw;
        pw.row_off = it.chunk * pl.chunk_rows;
        pw.rows_valid = plan_rows(pl, it.chunk);
        pw.D = pl.D;
        pw.Dp = pl.Dp;
        pw.use_tma = ar.dw_tma;
        if (pl.kind_mask & 1)
          wait_counter(pl.g_done + it.chunk, it.target, 0x510u);
        TileCoord tcw;
        tcw.m_blk = it.m_blk * C::kCtas + rank;
        tcw.n_blk = it.n_blk;
        tcw.kb0 = it.kb0;
        tcw.kb1 = it.kb1;
        rdw = EpiDW::template prefetch<kPBN>(pw, tcw, row_in_tile, epi_tid);
      }
      mbar_wait(&tfull[acc], acc_phase, 0x400u + acc);
      tc_fence_after();
      if (tr) tr[5] = clock64();
      if (it.kind != kItemG && epi_tid == 0) {  // this tile has finished reading its G^T chunk
        fence_proxy_async_all();
        __threadfence();
        atomicAdd(pl.xw_done + it.chunk, 1u);
      }
      // the slot is used through two different mechanisms (mbarrier hand-off to the store issuer for
      // G tiles, bulk-group waits for X / W chunks): drain the previous one when the kind changes
      if (prev_kind != -1 && (it.kind == kItemG) != (prev_kind == kItemG)) {
        if (prev_kind == kItemG) {  // every buffer's last G^T box has been read by its bulk store
          const uint32_t uses = n_box / kBufs;  // n_box is even: the same number of boxes went through each buffer
#pragma unroll
          for (int b = 0; b < kBufs; ++b) mbar_wait(&slot_free[wg * kBufs + b], (uses & 1u) ^ 1u, 0x800u + wg);
        } else {
          if (wg_tid == 0) tma_store_wait_read<0>();
          named_bar_sync(2 + wg, kWgThreads);
        }
      }
      const uint32_t taddr = tmem_base + static_cast<uint32_t>(acc * kPBN) + (static_cast<uint32_t>(quad * 32) << 16);
      TileCoord tc;
      const int mblk = it.m_blk * C::kCtas + rank;  // this CTA's 128 accumulator rows of the tile
      tc.m_blk = mblk;
      tc.n_blk = it.n_blk;
      tc.kb0 = it.kb0;
      tc.kb1 = it.kb1;
      const int c0 = it.chunk * pl.chunk_rows;
      const int rows = plan_rows(pl, it.chunk);
      if (it.kind == kItemG) {
        typename EpiBwdG<EXT>::Params p;
        p.lse2 = ar.lse2;
        p.gscale = ar.gscale;
        p.lbl_local = ar.lbl_local;
        p.zt2 = ar.zt2;
        p.dphi = ar.dphi;
        p.cos_t = ar.cos_t;
        p.t_out = ar.tproj;
        p.row_off = c0;
        p.rows_valid = rows;
        p.B = pl.B;
        p.Bp = pl.Bp;
        p.s2 = ar.s2;
        p.grad_out = EXT ? ar.grad_out + c0 : nullptr;
        p.ldgo = ar.ldgo;
        p.s_ext = ar.s_ext;
        // stage this warpgroup's 128 log-sum-exps / labels (the previous tile's readers are all past
        // their last chunk: they arrived on slot_full, which this thread observes through slot_free below,
        // but a plain warpgroup barrier is cheaper and sufficient)
        named_bar_sync(2 + wg, kWgThreads);
        if (!EXT) lse_s[wg_tid] = lse_v;
        lbl_s[wg_tid] = lbl_v;
        named_bar_sync(2 + wg, kWgThreads);
        const int row_local = mblk * kBM + row_in_tile;
        const int cls = c0 + row_local;
        const int tile_cls0 = c0 + mblk * kBM;
        const bool row_ok = row_local < rows;
        float tsum = 0.f;
        const uint32_t sw = static_cast<uint32_t>(row_in_tile & 7);
#pragma unroll 1
        for (int bx = 0; bx < 2; ++bx, ++n_box) {
          // the bulk store of the box that last used this buffer has finished reading it
          const uint32_t buf = n_box % kBufs, use = n_box / kBufs;
          mbar_wait(&slot_free[wg * kBufs + buf], (use & 1u) ^ 1u, 0x900u + wg);
          const uint32_t srow = smem_u32(epi_smem + (wg * kBufs + buf) * kPSlotBytes) + row_in_tile * 128;
#pragma unroll 1
          for (int h = 0; h < 2; ++h) {
            const int cl = bx * 64 + h * 32;      // column inside this warpgroup's half
            const int c = wg * 128 + cl;          // column inside the tile
            const int col0 = it.n_blk * kPBN + c;
            float v[32];
            if (pl.debug_epi >= 2) continue;
            tmem_ld32(taddr + c, v);
            if (pl.debug_epi == 1) {
              tsum += v[0] + v[31];
              continue;
            }
            float g[32];
            EpiBwdG<EXT>::chunk_math(p, gscale, lse_s + cl, lbl_s[cl + lane], col0, cls, tile_cls0, row_ok, row_local,
                                     v, g);
#pragma unroll
            for (int j = 0; j < 32; ++j) tsum = fmaf(g[j], v[j], tsum);
            const uint32_t q0 = static_cast<uint32_t>(h * 4);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint32_t a = srow + (((q0 + q) ^ sw) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a),
                           "r"(pack_bf16x2(g[q * 8 + 0], g[q * 8 + 1])), "r"(pack_bf16x2(g[q * 8 + 2], g[q * 8 + 3])),
                           "r"(pack_bf16x2(g[q * 8 + 4], g[q * 8 + 5])), "r"(pack_bf16x2(g[q * 8 + 6], g[q * 8 + 7]))
                           : "memory");
            }
          }
          if (bx == 1) {  // all TMEM reads of this tile are done; t_j must precede the last hand-off
            tc_fence_before();
            if (row_ok) atomicAdd(ar.tproj + cls, tsum);
          }
          fence_proxy_async_smem();  // box -> visible to the TMA engine
          __syncwarp();
          if (lane == 0) {
            if (bx == 1) arrive_tempty<PAIR>(&tempty[acc]);
            mbar_arrive(&slot_full[wg * kBufs + buf]);
          }
        }
      } else if (it.kind == kItemX) {
        EpiRedAdd::Params p;
        p.M = pl.B;
        p.N = pl.D;
        p.row_scale = nullptr;
        EpiRedAdd::template consume<kPBN, kBufs>(p, ost, tc, tmDx, taddr, row_in_tile, epi_tid, epi_smem);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) arrive_tempty<PAIR>(&tempty[acc]);
      } else {
        EpiDW::template consume_with<kPBN, kBufs>(pw, ost, tc, tmDw, taddr, row_in_tile, epi_tid, epi_smem, rdw);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) arrive_tempty<PAIR>(&tempty[acc]);
      }
      if (tr) tr[6] = clock64();
      prev_kind = it.kind;
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
    }
    if (wg_tid == 0) tma_store_wait_all0();
  }

  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all(); else __syncthreads();  // pair: the peer may still signal our barriers
  if (warp == 2) {
    tc_fence_after();
    if constexpr (PAIR) tmem_dealloc_pair(tmem_base, 2 * kPBN); else tmem_dealloc(tmem_base, 2 * kPBN);
  }
}

}  // namespace arcb
