// Epilogue policies for gemm_sm100_kernel.  Each epilogue thread owns ONE accumulator row (a TMEM
// lane); the two epilogue warpgroups split the tile's BN columns in halves and walk their half in
// chunks of 32 fp32 values pulled with tcgen05.ld (epi_tid >> 7 = warpgroup, & 127 = row).
//
//   EpiStore      plain fp32 store (unit tests of the GEMM core)
//   EpiFwdStats   forward: margin substitution at the label column, online (max, sum-exp)
//                 -> the B x C logits never leave the SM            (arc_main.py:160-172 + :245/:327)
//   EpiLogits     forward, materialising: s * (phi at label, cos elsewhere) -> [B, C] fp32
//                 (the strict drop-in ArcMarginProduct.forward, arc_main.py:157-176)
//   EpiBwdG       backward recompute: G^T = dLoss/dcos for a class chunk, bf16, through a
//                 swizzled smem tile + TMA store; also the per-class projection term used by
//                 the weight-normalisation backward                 (autograd behind arc_main.py:329)
//   EpiRedAdd     split-K accumulate into fp32 global (dX_hat partial sums), TMA reduce-add
//   EpiDW         dW = inv_norm * (dW_hat - t * w_hat): F.normalize backward fused into the store
#pragma once

#include <cuda_bf16.h>
#include <math.h>

#include "gemm_sm100.cuh"

namespace arcb {

struct EpiNoState {};
struct EpiOutState {
  int it;  // chunks this warpgroup has sent so far: selects the smem buffer, survives across tiles
};

constexpr int kWgThreads = 128;

// One 32-column fp32 chunk of the tile (128 rows x 128 B) goes through a 128B-swizzled smem buffer
// and leaves as ONE bulk tensor store / reduce-add instead of 32-sector-per-instruction row stores.
// Each warpgroup owns two 16 KB buffers and alternates between them.
constexpr int kChunkBufBytes = kBM * 128;
constexpr int kChunkOutSmem = 2 * 2 * kChunkBufBytes;

template <bool REDUCE, int NBUF = 2>
__device__ __forceinline__ void chunk_out_tma(const CUtensorMap& tm, uint8_t* epi_smem, int wg, int wg_tid,
                                              int row_in_tile, int iter, const float (&r)[32], int c_inner,
                                              int c_outer) {
  static_assert(NBUF == 1 || NBUF == 2, "one or two chunk buffers per warpgroup");
  uint8_t* buf = epi_smem + (wg * NBUF + (NBUF == 2 ? (iter & 1) : 0)) * kChunkBufBytes;
  // the store that last used this buffer (NBUF chunks ago) must have finished reading it
  if (wg_tid == 0) tma_store_wait_read<NBUF - 1>();
  named_bar_sync(2 + wg, kWgThreads);
  const uint32_t base = smem_u32(buf) + row_in_tile * 128;
  const uint32_t sw = static_cast<uint32_t>(row_in_tile & 7);
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const uint32_t a = base + ((static_cast<uint32_t>(q) ^ sw) << 4);
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(r[4 * q]), "f"(r[4 * q + 1]),
                 "f"(r[4 * q + 2]), "f"(r[4 * q + 3])
                 : "memory");
  }
  fence_proxy_async_smem();
  named_bar_sync(2 + wg, kWgThreads);
  if (wg_tid == 0) {
    // plain stores are the final fp32 dW: streamed out once, never read again by these kernels
    if (REDUCE) tma_reduce_add_2d(&tm, buf, c_inner, c_outer); else tma_store_2d_hint(&tm, buf, c_inner, c_outer, kEvictFirst);
    tma_store_commit();
  }
}

// ------------------------------------------------------------------------------------
struct EpiStore {
  static constexpr int kSmemBytes = 0;
  static constexpr bool kUsesTmC = false;
  struct Params {
    float* out;
    int64_t ldo;
    int M, N;
    int accumulate;  // 1: atomicAdd (split-K), 0: store
  };
  using State = EpiNoState;
  __device__ static void init(const Params&, State&) {}
  __device__ static void drain(int) {}
  template <int BN>
  __device__ static void consume(const Params& p, State&, const TileCoord& tc, const CUtensorMap&, uint32_t taddr,
                                 int row_in_tile, int epi_tid, uint8_t*) {
    const int wg = epi_tid >> 7;
    const int row = tc.m_blk * kBM + row_in_tile;
    for (int c = wg * (BN / 2); c < (wg + 1) * (BN / 2); c += 32) {
      const int col0 = tc.n_blk * BN + c;
      if (col0 >= p.N) break;
      float v[32];
      tmem_ld32(taddr + c, v);
      if (row < p.M) {
        float* o = p.out + static_cast<int64_t>(row) * p.ldo + col0;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < p.N) {
            if (p.accumulate) atomicAdd(o + j, v[j]); else o[j] = v[j];
          }
      }
    }
  }
  template <int BN>
  __device__ static void finish(const Params&, State&, const TileCoord&, const CUtensorMap&, int, int, uint8_t*) {}
};

// ------------------------------------------------------------------------------------
// Forward statistics.  Rows = batch samples, columns = classes of this shard.
// Works in base-2 units: u = cos * s2 with s2 = s * log2(e); a partial is (max u, sum 2^(u-max)).
// Each (tile column block, warpgroup) writes one partial per row: part[(n_blk * 2 + wg)][row].
struct EpiFwdStats {
  static constexpr int kSmemBytes = 0;
  static constexpr bool kUsesTmC = false;
  struct Params {
    const int* lbl_local;  // [Bp] label - class_offset if this shard owns it, else -1
    const float* phi;      // [Bp] value that replaces the target cosine (phi or the fallback)
    float2* part;          // [2 * tiles_n][ldp]
    int ldp;
    int B, C;
    float s2;
  };
  using State = EpiNoState;
  __device__ static void init(const Params&, State&) {}
  __device__ static void drain(int) {}
  template <int BN>
  __device__ static void consume(const Params& p, State&, const TileCoord& tc, const CUtensorMap&, uint32_t taddr,
                                 int row_in_tile, int epi_tid, uint8_t*) {
    const int wg = epi_tid >> 7;
    const int row = tc.m_blk * kBM + row_in_tile;
    int lbl = -1;
    float phi = 0.f;
    if (row < p.B) {
      lbl = __ldg(p.lbl_local + row);
      phi = __ldg(p.phi + row);
    }
    float run_m = -INFINITY, run_l = 0.f;
    const int col_base = tc.n_blk * BN;
    for (int c = wg * (BN / 2); c < (wg + 1) * (BN / 2); c += 32) {
      const int col0 = col_base + c;
      if (col0 >= p.C) break;
      float v[32];
      tmem_ld32(taddr + c, v);
      if (col0 + 32 > p.C) {  // ragged last tile: padded classes must not enter the softmax
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j >= p.C) v[j] = -INFINITY;
      }
      const unsigned d = static_cast<unsigned>(lbl - col0);
      if (d < 32u) {  // rare: this chunk holds the row's label column
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j == static_cast<int>(d)) v[j] = phi;
      }
      float cm = v[0];
#pragma unroll
      for (int j = 1; j < 32; ++j) cm = fmaxf(cm, v[j]);
      const float nm = fmaxf(run_m, cm);
      run_l *= ex2_approx((run_m - nm) * p.s2);
      const float off = nm * p.s2;
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        a0 += ex2_approx(fmaf(v[j + 0], p.s2, -off));
        a1 += ex2_approx(fmaf(v[j + 1], p.s2, -off));
        a2 += ex2_approx(fmaf(v[j + 2], p.s2, -off));
        a3 += ex2_approx(fmaf(v[j + 3], p.s2, -off));
      }
      run_l += (a0 + a1) + (a2 + a3);
      run_m = nm;
    }
    if (row < p.B)
      p.part[static_cast<int64_t>(tc.n_blk * 2 + wg) * p.ldp + row] = make_float2(run_m * p.s2, run_l);
  }
  template <int BN>
  __device__ static void finish(const Params&, State&, const TileCoord&, const CUtensorMap&, int, int, uint8_t*) {}
};

// ------------------------------------------------------------------------------------
// Forward statistics that also KEEP the softmax numerators (stored-probabilities schedule):
// P'[j, i] = bf16(2^(s2 (cos_ij - 1))), the fixed reference s2 = s log2(e) (cos <= 1, so P' <= 1: no running maximum),
// written CLASS-major, P'^T [Cs, Bp] -- a class row holds the whole batch (2 KB at B = 1024), which is what both backward
// GEMMs want to stream: dW reads [128 classes x all samples] as one contiguous 256 KB region, dX reads the rows of its
// class slice as an MN-major operand.  The tile's accumulator is [128 samples x 256 classes] with one SAMPLE per thread,
// so the store is a transpose through shared memory: thread s writes its 2-byte values down column s of a
// [64 classes x 128 samples] staging tile (two 128B-swizzled boxes of 64 samples; the 32 lanes of a warp fill 64
// contiguous bytes of one class row: no bank conflicts), one bulk tensor store per box.  The partial of a (tile column
// block, warpgroup) is (s2, sum of P') -- the (reference, sum) format the merge kernels take; the sum is taken before
// the bf16 rounding (the rounding errors of a row average out far below the loss tolerance).  Block maxima for the
// exact top-3 go to their own array when asked for.
struct EpiFwdStatsP {
  static constexpr int kSmemBytes = 2 * 64 * 256;   // per warpgroup: [64 classes x 128 samples] bf16 = 16 KB
  static constexpr bool kUsesTmC = true;            // tmC: P'^T [Cs, Bp] bf16 (pitch Bp), box [64 classes x 64 samples]
  struct Params {
    const int* lbl_local;
    const float* phi;
    float2* part;
    float* bmax;  // [2 * tiles_n][ldp] or nullptr
    int ldp;
    int B, C;
    float s2;
  };
  using State = EpiNoState;
  __device__ static void init(const Params&, State&) {}
  __device__ static void drain(int epi_tid) {
    if ((epi_tid & 127) == 0) tma_store_wait_all0();
  }
  template <int BN>
  __device__ static void consume(const Params& p, State&, const TileCoord& tc, const CUtensorMap& tmC, uint32_t taddr,
                                 int row_in_tile, int epi_tid, uint8_t* smem) {
    const int wg = epi_tid >> 7;
    const int wg_tid = epi_tid & 127;
    const int row = tc.m_blk * kBM + row_in_tile;  // sample
    const bool row_ok = row < p.B;
    int lbl = -1;
    float phi = 0.f;
    if (row_ok) {
      lbl = __ldg(p.lbl_local + row);
      phi = __ldg(p.phi + row);
    }
    float run_l = 0.f, blk_m = -INFINITY;
    bool any = false;
    const int col_base = tc.n_blk * BN;
    uint8_t* buf = smem + wg * (64 * 256);
    // my column inside the staging tile: box (64 samples) and the 2 bytes inside a 128-byte class row; the 16-byte chunk
    // is XOR-swizzled with (class row & 7): eight addresses cover every row
    uint32_t soff[8];
    {
      const uint32_t sbox = smem_u32(buf) + static_cast<uint32_t>(row_in_tile >> 6) * (64 * 128);
      const uint32_t scol = static_cast<uint32_t>(row_in_tile & 63) * 2;  // byte inside the row
#pragma unroll
      for (uint32_t k = 0; k < 8; ++k) soff[k] = sbox + (((scol >> 4) ^ k) << 4) + (scol & 15);
    }
#pragma unroll 1
    for (int bx = 0; bx < BN / 128; ++bx) {  // this warpgroup's two groups of 64 classes
      const int cls0 = col_base + wg * (BN / 2) + bx * 64;
      if (cls0 >= p.C) break;  // warpgroup-uniform
      // the bulk stores that last used the staging tile have finished reading it
      if (wg_tid == 0) tma_store_wait_read0();
      named_bar_sync(2 + wg, kWgThreads);
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const int c = wg * (BN / 2) + bx * 64 + h * 32;
        const int col0 = col_base + c;
        float v[32];
        if (col0 < p.C) {
          tmem_ld32(taddr + c, v);
          any = true;
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = -INFINITY;
        }
        if (col0 + 32 > p.C) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j >= p.C) v[j] = -INFINITY;
        }
        const unsigned d = static_cast<unsigned>(lbl - col0);
        if (d < 32u) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j == static_cast<int>(d)) v[j] = phi;
        }
        if (p.bmax) {
          float cm = v[0];
#pragma unroll
          for (int j = 1; j < 32; ++j) cm = fmaxf(cm, v[j]);
          blk_m = fmaxf(blk_m, cm);
        }
        float a0 = 0.f, a1 = 0.f;
        const uint32_t hrow = static_cast<uint32_t>(h) * (32 * 128);  // class rows h * 32 + j of the staging tile
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float e0 = ex2_approx(fmaf(v[j], p.s2, -p.s2)), e1 = ex2_approx(fmaf(v[j + 1], p.s2, -p.s2));
          a0 += e0;
          a1 += e1;
          // padding samples (row >= B) must contribute nothing to dW: their column of P'^T is zero
          const uint32_t pk = row_ok ? pack_bf16x2(e0, e1) : 0u;
          // (the row offsets are compile-time constants after unrolling: they fold into the store's immediate)
          asm volatile("st.shared.u16 [%0], %1;" ::"r"(soff[j & 7] + hrow + static_cast<uint32_t>(j * 128)),
                       "h"(static_cast<unsigned short>(pk & 0xffffu))
                       : "memory");
          asm volatile("st.shared.u16 [%0], %1;" ::"r"(soff[(j + 1) & 7] + hrow + static_cast<uint32_t>((j + 1) * 128)),
                       "h"(static_cast<unsigned short>(pk >> 16))
                       : "memory");
        }
        run_l += a0 + a1;
      }
      fence_proxy_async_smem();
      named_bar_sync(2 + wg, kWgThreads);
      if (wg_tid == 0) {
        // rows (classes) >= Cs are clipped by the tensor map; sample columns always lie inside [0, Bp)
        tma_store_2d(&tmC, buf, tc.m_blk * kBM, cls0);
        tma_store_2d(&tmC, buf + 64 * 128, tc.m_blk * kBM + 64, cls0);
        tma_store_commit();
      }
    }
    if (row_ok) {
      p.part[static_cast<int64_t>(tc.n_blk * 2 + wg) * p.ldp + row] =
          any ? make_float2(p.s2, run_l) : make_float2(-INFINITY, 0.f);
      if (p.bmax) p.bmax[static_cast<int64_t>(tc.n_blk * 2 + wg) * p.ldp + row] = blk_m * p.s2;
    }
  }
  template <int BN>
  __device__ static void finish(const Params&, State&, const TileCoord&, const CUtensorMap&, int, int, uint8_t*) {}
};

// ------------------------------------------------------------------------------------
struct EpiLogits {
  static constexpr int kSmemBytes = 0;
  static constexpr bool kUsesTmC = false;
  struct Params {
    const int* lbl_local;
    const float* phi;
    float* out;  // [B, ldo]
    int64_t ldo;
    int B, C;
    float s;
  };
  using State = EpiNoState;
  __device__ static void init(const Params&, State&) {}
  __device__ static void drain(int) {}
  template <int BN>
  __device__ static void consume(const Params& p, State&, const TileCoord& tc, const CUtensorMap&, uint32_t taddr,
                                 int row_in_tile, int epi_tid, uint8_t*) {
    const int wg = epi_tid >> 7;
    const int row = tc.m_blk * kBM + row_in_tile;
    int lbl = -1;
    float phi = 0.f;
    if (row < p.B) {
      lbl = __ldg(p.lbl_local + row);
      phi = __ldg(p.phi + row);
    }
    const bool vec_ok = (p.ldo & 3) == 0;
    for (int c = wg * (BN / 2); c < (wg + 1) * (BN / 2); c += 32) {
      const int col0 = tc.n_blk * BN + c;
      if (col0 >= p.C) break;
      float v[32];
      tmem_ld32(taddr + c, v);
      const unsigned d = static_cast<unsigned>(lbl - col0);
      if (d < 32u) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j == static_cast<int>(d)) v[j] = phi;
      }
      if (row < p.B) {
        float* o = p.out + static_cast<int64_t>(row) * p.ldo + col0;
        if (vec_ok && col0 + 32 <= p.C) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<float4*>(o + j) = make_float4(v[j] * p.s, v[j + 1] * p.s, v[j + 2] * p.s, v[j + 3] * p.s);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j < p.C) o[j] = v[j] * p.s;
        }
      }
    }
  }
  template <int BN>
  __device__ static void finish(const Params&, State&, const TileCoord&, const CUtensorMap&, int, int, uint8_t*) {}
};

// ------------------------------------------------------------------------------------
// Backward recompute.  Rows = classes of the current chunk, columns = batch samples.
//   g_ij = gscale * (softmax_ij - onehot_ij) * d logit / d cos     with gscale = dLoss * s / B
// Non-target: g = gscale * 2^(cos*s2 - lse2_i).  Target: g = gscale * (2^(zt2_i - lse2_i) - 1) * dphi_i.
// The tile is written as bf16 G^T[class, sample] through a 128B-swizzled smem tile and TMA.
// t_j = sum_i g_ij * cos_ij is the projection w_hat_j . dW_hat_j needed by F.normalize's backward.
// EXT = true is the materialising path: dLoss/dlogits arrives from autograd as grad_out[B, ldgo]
// and g_ij = grad_out_ij * s * (dphi_i at the label column, 1 elsewhere).
// The labels and log-sum-exps of the tile's 256 samples are staged in smem once per tile; a warp
// ballot finds the (rare) chunks that contain a label of this 128-class tile.
template <bool EXT>
struct EpiBwdG {
  static constexpr int kStageBytes = kBM * 256 * 2;             // bf16 tile, 4 boxes of [128 x 64]
  static constexpr int kSmemBytes = kStageBytes + 256 * 4 * 2;  // + lse2 and labels of the tile's samples
  static constexpr bool kUsesTmC = true;
  struct Params {
    const float* lse2;       // [Bp] lse * log2(e); +inf beyond B
    const float* gscale;     // device scalar: upstream grad * s / B
    const int* lbl_local;    // [Bp], -1 = not in this shard / padding
    const float* zt2;        // [Bp] target logit * log2(e)
    const float* dphi;       // [Bp] d(target value)/d cos
    const float* cos_t;      // [Bp] exact fp32 target cosine
    float* t_out;            // [Cs] accumulates t_j (atomicAdd across batch tiles)
    int row_off;             // first class of this chunk inside the shard (multiple of 128)
    int rows_valid;          // classes in this chunk
    int B, Bp;
    float s2;
    const float* grad_out;   // EXT only: [B, ldgo], column 0 = first class of this chunk
    int64_t ldgo;
    float s_ext;
  };
  struct State {
    float gscale;
  };
  __device__ static void init(const Params& p, State& st) { st.gscale = EXT ? 0.f : __ldg(p.gscale); }
  __device__ static void drain(int epi_tid) {
    if (epi_tid == 0) tma_store_wait_all0();
  }
  // Stage the log-sum-exps and labels of the tile's 256 samples (one element per epilogue thread).
  __device__ static void stage_rows(uint8_t* smem, int epi_tid, float lse_v, int lbl_v) {
    float* lse_s = reinterpret_cast<float*>(smem + kStageBytes);
    int* lbl_s = reinterpret_cast<int*>(smem + kStageBytes + 256 * 4);
    if constexpr (!EXT) lse_s[epi_tid] = lse_v;
    lbl_s[epi_tid] = lbl_v;
  }
  template <int BN>
  __device__ static void consume(const Params& p, State& st, const TileCoord& tc, const CUtensorMap&, uint32_t taddr,
                                 int row_in_tile, int epi_tid, uint8_t* smem) {
    static_assert(BN == 256, "staging tile is sized for BN = 256");
    // the previous tile's TMA stores must have finished reading the staging tile
    if (epi_tid == 0) tma_store_wait_read0();
    const float lse_v = EXT ? 0.f : __ldg(p.lse2 + tc.n_blk * BN + epi_tid);
    const int lbl_v = __ldg(p.lbl_local + tc.n_blk * BN + epi_tid);
    stage_rows(smem, epi_tid, lse_v, lbl_v);
    named_bar_sync(1, kEpiThreads);
    body<BN>(p, st.gscale, tc, taddr, row_in_tile, epi_tid, smem);
  }
  // One 32-sample chunk of one class row: v = recomputed cosines (in/out: the label column gets the
  // exact fp32 cosine), g = dLoss/dcos.  lse_c -> the 32 staged log-sum-exps, my_lbl = the staged
  // label of sample col0 + lane.  All 32 lanes of the warp must call this together.
  __device__ static __forceinline__ void chunk_math(const Params& p, float gscale, const float* lse_c, int my_lbl,
                                                    int col0, int cls, int tile_cls0, bool row_ok, int row_local,
                                                    float (&v)[32], float (&g)[32]) {
    if constexpr (!EXT) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const float4 l4 = *reinterpret_cast<const float4*>(lse_c + j);
        g[j + 0] = ex2_approx(fmaf(v[j + 0], p.s2, -l4.x)) * gscale;
        g[j + 1] = ex2_approx(fmaf(v[j + 1], p.s2, -l4.y)) * gscale;
        g[j + 2] = ex2_approx(fmaf(v[j + 2], p.s2, -l4.z)) * gscale;
        g[j + 3] = ex2_approx(fmaf(v[j + 3], p.s2, -l4.w)) * gscale;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int i = col0 + j;
        float gj = 0.f;
        if (row_ok && i < p.B) gj = __ldg(p.grad_out + static_cast<int64_t>(i) * p.ldgo + row_local) * p.s_ext;
        g[j] = gj;
      }
    }
    // label columns of this chunk that fall into this tile's 128 classes (warp-uniform, rare)
    unsigned hit = __ballot_sync(0xffffffffu, static_cast<unsigned>(my_lbl - tile_cls0) < static_cast<unsigned>(kBM));
    while (hit) {
      const int j = __ffs(hit) - 1;
      hit &= hit - 1;
      const int tl = __shfl_sync(0xffffffffu, my_lbl, j);
      if (tl == cls) {  // the one thread that owns the label's class row
        const int i = col0 + j;
        const float ct = __ldg(p.cos_t + i);
        const float dp = __ldg(p.dphi + i);
        float gt;
        if constexpr (!EXT) {
          gt = gscale * (ex2_approx(__ldg(p.zt2 + i) - lse_c[j]) - 1.f) * dp;
        } else {
          gt = __ldg(p.grad_out + static_cast<int64_t>(i) * p.ldgo + row_local) * p.s_ext * dp;
        }
#pragma unroll
        for (int jj = 0; jj < 32; ++jj)
          if (jj == j) {
            g[jj] = gt;
            v[jj] = ct;
          }
      }
    }
  }
  // The tile proper: needs the staging tile free and stage_rows() visible to all epilogue threads.
  template <int BN>
  __device__ static void body(const Params& p, float gscale, const TileCoord& tc, uint32_t taddr, int row_in_tile,
                              int epi_tid, uint8_t* smem) {
    const int wg = epi_tid >> 7;
    const int lane = epi_tid & 31;
    const float* lse_s = reinterpret_cast<const float*>(smem + kStageBytes);
    const int* lbl_s = reinterpret_cast<const int*>(smem + kStageBytes + 256 * 4);
    const int row_local = tc.m_blk * kBM + row_in_tile;
    const int cls = p.row_off + row_local;
    const int tile_cls0 = p.row_off + tc.m_blk * kBM;
    const bool row_ok = row_local < p.rows_valid;
    float tsum = 0.f;
    const uint32_t srow = smem_u32(smem) + row_in_tile * 128;
    const uint32_t sw = static_cast<uint32_t>(row_in_tile & 7);
    for (int c = wg * (BN / 2); c < (wg + 1) * (BN / 2); c += 32) {
      const int col0 = tc.n_blk * BN + c;
      float v[32];
      tmem_ld32(taddr + c, v);
      float g[32];
      chunk_math(p, gscale, lse_s + c, lbl_s[c + lane], col0, cls, tile_cls0, row_ok, row_local, v, g);
#pragma unroll
      for (int j = 0; j < 32; ++j) tsum = fmaf(g[j], v[j], tsum);
      // 32 bf16 = 64 B = four 16 B chunks of this row inside box (c / 64)
      const uint32_t box = srow + static_cast<uint32_t>(c >> 6) * (kBM * 128);
      const uint32_t q0 = static_cast<uint32_t>((c & 63) >> 3);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t a = box + (((q0 + q) ^ sw) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pack_bf16x2(g[q * 8 + 0], g[q * 8 + 1])),
                     "r"(pack_bf16x2(g[q * 8 + 2], g[q * 8 + 3])), "r"(pack_bf16x2(g[q * 8 + 4], g[q * 8 + 5])),
                     "r"(pack_bf16x2(g[q * 8 + 6], g[q * 8 + 7]))
                     : "memory");
      }
    }
    if (row_ok) atomicAdd(p.t_out + cls, tsum);
  }
  template <int BN>
  __device__ static void finish(const Params& p, State&, const TileCoord& tc, const CUtensorMap& tmC, int, int epi_tid,
                                uint8_t* smem) {
    fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the TMA engine
    named_bar_sync(1, kEpiThreads);
    if (epi_tid == 0) {
#pragma unroll
      for (int bx = 0; bx < BN / 64; ++bx) {
        const int col0 = tc.n_blk * BN + bx * 64;
        if (col0 < p.Bp) tma_store_2d(&tmC, smem + bx * (kBM * 128), col0, tc.m_blk * kBM);
      }
      tma_store_commit();
    }
  }
};

// ------------------------------------------------------------------------------------
// Split-K partial sums: global[M, N] += tile, as bulk tensor reduce-adds (fp32) in L2.
struct EpiRedAdd {
  static constexpr int kSmemBytes = kChunkOutSmem;
  static constexpr bool kUsesTmC = true;  // tmC: fp32 [M, N] destination, box [128 x 32]
  struct Params {
    int M, N;
    const float* row_scale;  // optional [>= tiles * 128]: the tile's rows are multiplied by it before the add
  };
  using State = EpiOutState;
  __device__ static void init(const Params&, State& st) { st.it = 0; }
  __device__ static void drain(int epi_tid) {
    if ((epi_tid & 127) == 0) tma_store_wait_all0();
  }
  template <int BN, int NBUF = 2>
  __device__ static void consume(const Params& p, State& st, const TileCoord& tc, const CUtensorMap& tmC, uint32_t taddr,
                                 int row_in_tile, int epi_tid, uint8_t* smem) {
    const int wg = epi_tid >> 7;
    const float sc = p.row_scale ? __ldcg(p.row_scale + tc.m_blk * kBM + row_in_tile) : 1.f;
    for (int c = wg * (BN / 2); c < (wg + 1) * (BN / 2); c += 32) {
      const int col0 = tc.n_blk * BN + c;
      if (col0 >= p.N) break;
      float v[32];
      tmem_ld32(taddr + c, v);
      if (p.row_scale) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] *= sc;
      }
      chunk_out_tma<true, NBUF>(tmC, smem, wg, epi_tid & 127, row_in_tile, st.it++, v, col0, tc.m_blk * kBM);
    }
  }
  template <int BN>
  __device__ static void finish(const Params&, State&, const TileCoord&, const CUtensorMap&, int, int, uint8_t*) {}
};

// ------------------------------------------------------------------------------------
// Rows = classes of the chunk, columns = feature index.  Final dW for those rows:
//   dW_j = inv_norm_j * (dW_hat_j - t_j * w_hat_j)          (||w_j|| above the F.normalize eps)
//   dW_j = inv_norm_j * dW_hat_j                            (clamped rows: y = w / eps)
// use_tma: dW leaves as bulk tensor stores (needs D % 4 == 0); otherwise plain row stores.
struct EpiDW {
  static constexpr int kSmemBytes = kChunkOutSmem;
  static constexpr bool kUsesTmC = true;  // tmC: fp32 dW [Cs, D], box [128 x 32]
  struct Params {
    float* dw;                  // [Cs, D] fp32, pitch D
    const __nv_bfloat16* what;  // [Cs, Dp] bf16
    const float* t;             // [Cs]
    const float* inv_norm;      // [Cs] 1 / max(||w||, eps)
    int row_off, rows_valid;
    int D, Dp;
    int use_tma;
  };
  using State = EpiOutState;
  __device__ static void init(const Params&, State& st) { st.it = 0; }
  __device__ static void drain(int epi_tid) {
    if ((epi_tid & 127) == 0) tma_store_wait_all0();
  }
  // Per-row data of a tile that does not depend on the accumulator: fetched BEFORE the epilogue waits for the
  // MMAs of the tile, so the global-load latency (inv_norm, t_j, first chunk of w_hat) is off the critical path.
  struct RowData {
    float tj, inv;
    uint4 w0[4];
  };
  template <int BN>
  __device__ static RowData prefetch(const Params& p, const TileCoord& tc, int row_in_tile, int epi_tid) {
    RowData rd;
    const int wg = epi_tid >> 7;
    const int row_local = tc.m_blk * kBM + row_in_tile;
    const bool valid = row_local < p.rows_valid;
    const int cls = p.row_off + row_local;
    rd.tj = 0.f;
    rd.inv = 0.f;
    if (valid) {
      rd.inv = __ldg(p.inv_norm + cls);
      rd.tj = __ldcg(p.t + cls);  // written by G tiles of this launch: coherent load
    }
    const __nv_bfloat16* wrow = p.what + static_cast<int64_t>(valid ? cls : 0) * p.Dp;
    const int c_begin = wg * (BN / 2);
#pragma unroll
    for (int q = 0; q < 4; ++q) rd.w0[q] = make_uint4(0u, 0u, 0u, 0u);
    if (tc.n_blk * BN + c_begin < p.D) {
#pragma unroll
      for (int q = 0; q < 4; ++q) rd.w0[q] = __ldg(reinterpret_cast<const uint4*>(wrow + tc.n_blk * BN + c_begin) + q);
    }
    return rd;
  }
  template <int BN, int NBUF = 2>
  __device__ static void consume(const Params& p, State& st, const TileCoord& tc, const CUtensorMap& tmC, uint32_t taddr,
                                 int row_in_tile, int epi_tid, uint8_t* smem) {
    const RowData rd = prefetch<BN>(p, tc, row_in_tile, epi_tid);
    consume_with<BN, NBUF>(p, st, tc, tmC, taddr, row_in_tile, epi_tid, smem, rd);
  }
  template <int BN, int NBUF = 2>
  __device__ static void consume_with(const Params& p, State& st, const TileCoord& tc, const CUtensorMap& tmC,
                                      uint32_t taddr, int row_in_tile, int epi_tid, uint8_t* smem, const RowData& rd) {
    const int wg = epi_tid >> 7;
    const int row_local = tc.m_blk * kBM + row_in_tile;
    const bool valid = row_local < p.rows_valid;
    const int cls = p.row_off + row_local;
    const float inv = rd.inv;
    const float tj = (inv < 0.5e12f) ? rd.tj : 0.f;  // clamped rows take no projection
    const __nv_bfloat16* wrow = p.what + static_cast<int64_t>(valid ? cls : 0) * p.Dp;
    const int c_begin = wg * (BN / 2), c_end = (wg + 1) * (BN / 2);
    uint4 wnext[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) wnext[q] = rd.w0[q];
    for (int c = c_begin; c < c_end; c += 32) {
      const int col0 = tc.n_blk * BN + c;
      if (col0 >= p.D) break;
      uint4 wcur[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) wcur[q] = wnext[q];
      if (c + 32 < c_end && col0 + 32 < p.D) {  // prefetch the next chunk of w_hat before touching TMEM
#pragma unroll
        for (int q = 0; q < 4; ++q) wnext[q] = __ldg(reinterpret_cast<const uint4*>(wrow + col0 + 32) + q);
      }
      float v[32];
      tmem_ld32(taddr + c, v);
      float r[32];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t wr[4] = {wcur[q].x, wcur[q].y, wcur[q].z, wcur[q].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float w0 = __uint_as_float(wr[e] << 16);
          const float w1 = __uint_as_float(wr[e] & 0xffff0000u);
          r[q * 8 + 2 * e + 0] = inv * fmaf(-tj, w0, v[q * 8 + 2 * e + 0]);
          r[q * 8 + 2 * e + 1] = inv * fmaf(-tj, w1, v[q * 8 + 2 * e + 1]);
        }
      }
      if (p.use_tma) {
        chunk_out_tma<false, NBUF>(tmC, smem, wg, epi_tid & 127, row_in_tile, st.it++, r, col0, p.row_off + tc.m_blk * kBM);
      } else if (valid) {
        // straight from registers: each thread owns 128 contiguous bytes of its dW row (no shared-memory round trip)
        float* o = p.dw + static_cast<int64_t>(cls) * p.D + col0;
        if ((p.D & 3) == 0 && col0 + 32 <= p.D) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + j), "f"(r[j]), "f"(r[j + 1]),
                         "f"(r[j + 2]), "f"(r[j + 3])
                         : "memory");
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j < p.D) o[j] = r[j];
        }
      }
    }
  }
  template <int BN>
  __device__ static void finish(const Params&, State&, const TileCoord&, const CUtensorMap&, int, int, uint8_t*) {}
};

// Epilogue of the dW GEMM of the stored-probabilities backward: the accumulator is
//   dW_hat[j, n_blk * BN ..) = sum_i P'^T[j, i] (c_i x_hat_i)        (one BN-wide slice of the feature row)
// and t_j = <w_hat_j, dW_hat_j> (the F.normalize backward, arc_main.py:159) is NOT known before this GEMM, unlike in
// EpiDW.  The tiles run n-fastest (GemmSched::n_fastest) on a cooperative grid that is a multiple of tiles_n, so the
// tiles_n slices of one class tile are in flight on neighbouring CTAs at the same time.  Pass 1 reduces this slice's
// share of t_j and adds it to t[j] in global memory; the class tile's arrival counter tells when all slices have
// contributed; pass 2 is EpiDW's.  A CTA waits only for its neighbours' pass 1 of the SAME tile index, which never
// depends on this CTA's pass 2: no cycle.
// STAGED: dW leaves through swizzled staging buffers and bulk tensor stores (64 KB of shared memory); otherwise every
// thread stores its 512 contiguous bytes of the row itself and the shared memory goes to two more operand stages.
template <bool STAGED>
struct EpiDWT {
  // w_hat slice of a tile, one row per class: 128 rows of up to 512 B, pitch 528 B (bank-conflict-free 16 B row reads)
  static constexpr int kWPitch = 528;
  static constexpr int kWOff = STAGED ? kChunkOutSmem : 0;
  static constexpr int kWBarOff = kWOff + kBM * kWPitch;
  static constexpr int kSmemBytes = kWBarOff + 16;
  static constexpr bool kUsesTmC = true;
  struct Params {
    EpiDW::Params base;      // base.t == t_acc
    float* t_acc;            // [Cs] accumulators of t_j, zeroed before the launch
    unsigned int* arrivals;  // [tiles_m], zeroed before the launch
    int slices;              // tiles_n
    int next_tile_blocks;    // 128-row blocks between a CTA's consecutive tiles
    int m_blocks;            // 128-row blocks the launch walks over (pair launches: rounded up to even)
    int pair;                // 1: launched on the pair core (m_blk parity = CTA rank)
  };
  struct State {
    int it;
    int primed;
    uint32_t wphase;
  };
  __device__ static void init(const Params&, State& st) {
    st.it = 0;
    st.primed = 0;
    st.wphase = 0;
  }
  __device__ static void drain(int epi_tid) {
    if (STAGED) EpiDW::drain(epi_tid);
  }
  // Bulk copies of the w_hat rows of class block `m_blk` into shared memory: the first warpgroup's thread r fetches row r.
  // The barrier counts 128 arrivals; valid rows add their bytes to the transaction count.
  template <int BN>
  __device__ static void fetch_w(const Params& q, int m_blk, int n_blk, int epi_tid, uint8_t* smem) {
    if (epi_tid >= kBM) return;
    const EpiDW::Params& p = q.base;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + kWBarOff);
    const int row_local = m_blk * kBM + epi_tid;
    const int cols = min(BN, p.Dp - n_blk * BN);
    if (row_local < p.rows_valid && cols > 0) {
      mbar_arrive_expect_tx(bar, static_cast<uint32_t>(cols * 2));
      bulk_load_1d(smem + kWOff + epi_tid * kWPitch,
                   p.what + static_cast<int64_t>(p.row_off + row_local) * p.Dp + n_blk * BN, static_cast<uint32_t>(cols * 2),
                   bar);
    } else {
      mbar_arrive(bar);
    }
  }
  template <int BN>
  __device__ static void consume(const Params& q, State& st, const TileCoord& tc, const CUtensorMap& tmC, uint32_t taddr,
                                 int row_in_tile, int epi_tid, uint8_t* smem) {
    static_assert(BN == 256, "each epilogue thread keeps its 128 columns of w_hat in registers");
    const EpiDW::Params& p = q.base;
    const int wg = epi_tid >> 7;
    const int row_local = tc.m_blk * kBM + row_in_tile;
    const bool valid = row_local < p.rows_valid;
    const int cls = p.row_off + row_local;
    const int col_base = tc.n_blk * BN + wg * (BN / 2);  // first of this thread's 128 columns
    uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + kWBarOff);
    if (!st.primed) {  // first tile of this CTA: nobody has fetched its w_hat rows yet
      if (epi_tid == 0) {
        mbar_init(wbar, kBM);
        fence_mbar_init();
      }
      named_bar_sync(1, kEpiThreads);
      fetch_w<BN>(q, tc.m_blk, tc.n_blk, epi_tid, smem);
      st.primed = 1;
    }
    const float inv = valid ? __ldg(p.inv_norm + cls) : 0.f;
    mbar_wait(wbar, st.wphase, 0xd10u);
    st.wphase ^= 1u;
    uint4 w[16];
    {
      const uint4* ws = reinterpret_cast<const uint4*>(smem + kWOff + row_in_tile * kWPitch + wg * (BN / 2) * 2);
#pragma unroll
      for (int e = 0; e < 16; ++e) w[e] = (valid && col_base + e * 8 < p.Dp) ? ws[e] : make_uint4(0u, 0u, 0u, 0u);
    }
    named_bar_sync(1, kEpiThreads);  // the buffer has been read out: refill it with the rows of this CTA's NEXT tile
    {
      const int next_blk = tc.m_blk + q.next_tile_blocks;
      if (next_blk - (q.pair ? (tc.m_blk & 1) : 0) < q.m_blocks) fetch_w<BN>(q, next_blk, tc.n_blk, epi_tid, smem);
    }
    float tpart = 0.f;
#pragma unroll
    for (int c = 0; c < BN / 2; c += 32) {
      if (col_base + c >= p.D) break;
      float v[32];
      tmem_ld32(taddr + wg * (BN / 2) + c, v);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const uint4 we = w[(c >> 3) + e];
        const uint32_t wr[4] = {we.x, we.y, we.z, we.w};
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          tpart = fmaf(v[e * 8 + 2 * h + 0], __uint_as_float(wr[h] << 16), tpart);
          tpart = fmaf(v[e * 8 + 2 * h + 1], __uint_as_float(wr[h] & 0xffff0000u), tpart);
        }
      }
    }
    if (valid) atomicAdd(q.t_acc + cls, tpart);
    named_bar_sync(1, kEpiThreads);  // every epilogue thread of the CTA has added its share
    if (epi_tid == 0) {
      __threadfence();
      atomicAdd(q.arrivals + tc.m_blk, 1u);
      wait_counter(q.arrivals + tc.m_blk, static_cast<unsigned>(q.slices), 0xd00u);
    }
    named_bar_sync(1, kEpiThreads);
    float tj = valid ? __ldcg(q.t_acc + cls) : 0.f;
    if (!(inv < 0.5e12f)) tj = 0.f;  // clamped rows take no projection
#pragma unroll
    for (int c = 0; c < BN / 2; c += 32) {
      const int col0 = col_base + c;
      if (col0 >= p.D) break;
      float v[32];
      tmem_ld32(taddr + wg * (BN / 2) + c, v);
      float r[32];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const uint4 we = w[(c >> 3) + e];
        const uint32_t wr[4] = {we.x, we.y, we.z, we.w};
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          r[e * 8 + 2 * h + 0] = inv * fmaf(-tj, __uint_as_float(wr[h] << 16), v[e * 8 + 2 * h + 0]);
          r[e * 8 + 2 * h + 1] = inv * fmaf(-tj, __uint_as_float(wr[h] & 0xffff0000u), v[e * 8 + 2 * h + 1]);
        }
      }
      if (STAGED && p.use_tma) {
        chunk_out_tma<false, 2>(tmC, smem, wg, epi_tid & 127, row_in_tile, st.it++, r, col0, p.row_off + tc.m_blk * kBM);
      } else if (valid) {
        float* o = p.dw + static_cast<int64_t>(cls) * p.D + col0;
        if ((p.D & 3) == 0 && col0 + 32 <= p.D) {
#pragma unroll
          for (int j2 = 0; j2 < 32; j2 += 4)
            asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + j2), "f"(r[j2]), "f"(r[j2 + 1]),
                         "f"(r[j2 + 2]), "f"(r[j2 + 3])
                         : "memory");
        } else {
#pragma unroll
          for (int j2 = 0; j2 < 32; ++j2)
            if (col0 + j2 < p.D) o[j2] = r[j2];
        }
      }
    }
  }
  template <int BN>
  __device__ static void finish(const Params&, State&, const TileCoord&, const CUtensorMap&, int, int, uint8_t*) {}
};

}  // namespace arcb
