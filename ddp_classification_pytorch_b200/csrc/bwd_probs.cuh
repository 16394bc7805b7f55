// Backward of the ArcFace head from STORED softmax numerators (ARCB_KEEP_PROBS; autograd behind arc_main.py:329).
//
// The forward (EpiFwdStatsP) left P'[j, i] = bf16(2^(s2 (cos_ij - 1))) for the whole shard, CLASS-major: P'^T [Cs, Bp], a
// class row holds the whole batch.  With c_i = gscale * 2^(s2 - lse2_i) the gradient of the loss w.r.t. the cosines is g_ij = c_i P'_ij (the label
// entry is patched to g_target / c_i by probs_setup_kernel), so the backward needs NO recompute GEMM: it is two plain GEMMs
//     dX_hat[B, D]  = diag(c) * P' * w_hat            (gemm core with EpiRedAdd, split along the classes: every CTA keeps ONE
//                                                      output tile in TMEM for its whole slice and reduce-adds it once)
//     dW_hat[Cs, D] = P'^T * (c (.) x_hat)            (this file; A = P'^T is K-major: 128 class rows x 64 samples per k-block,
//                                                      the tile's whole K extent is one contiguous 256 KB region of HBM)
// 6 B C D executed flops for 6 B C D algorithmic ones, against 8 B C D with the recompute.
//
// What the recompute epilogue used to deliver on the side is the projection t_j = w_hat_j . dW_hat_j that F.normalize's
// backward subtracts (dW_j = (dW_hat_j - t_j w_hat_j) / ||w_j||).  It needs the WHOLE feature row of the accumulator, but
// a [128 x 512] fp32 accumulator is all of TMEM and would leave no second buffer to overlap the epilogue with the next
// tile's MMAs.  So a cluster of two CTAs works on the same 128 classes, each on one half of the features
// (its own cta_group::1 pipeline, a [128 x 256] accumulator, double buffered), and the epilogue has two passes:
//   pass 1  t_part = sum over my half of the columns of acc * w_hat   (TMEM loads + FMAs, nothing leaves the SM)
//   swap    every epilogue thread sends its partial to BOTH CTAs' shared memory (st.async, counted on an mbarrier);
//           four partials per class row (2 CTAs x 2 column halves per CTA) add up to t_j
//   pass 2  dW = inv_norm * (acc - t_j * w_hat) -> 128B-swizzled staging -> TMA store
// The swap is the only coupling of the two CTAs; they run the same tile sequence, so neither waits long.
//
// Roles (384 threads): warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator, warps 4..11 epilogue.
#pragma once

#include "epilogues.cuh"

namespace arcb {

constexpr int kPwBN = 256;      // feature columns per CTA (half of a 512-wide row)
constexpr int kPwStages = 3;    // [128 x 64] of P'^T + [64 x 256] of c.x_hat = 48 KB per stage

struct PwSmem {
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = kPwBN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;             // 48 KB
  static constexpr int kEpiOff = kPwStages * kStageBytes;           // 144 KB
  static constexpr int kEpiBytes = kChunkOutSmem;                   // two 16 KB chunk buffers per warpgroup
  static constexpr int kSwapOff = kEpiOff + kEpiBytes;              // t partials: [2 parities][4 senders][128 rows] fp32
  static constexpr int kSwapBytes = 2 * 4 * kBM * 4;                // 4 KB
  static constexpr int kBarOff = kSwapOff + kSwapBytes;
  static constexpr int kNumBars = 2 * kPwStages + 4 + 2;            // ring, tfull/tempty, swap[2]
  static constexpr int kTotal = kBarOff + kNumBars * 8 + 16;        // dynamic smem base is 1024 B aligned (checked)
};
static_assert(PwSmem::kTotal <= 232448, "dW-from-probabilities kernel exceeds the shared memory budget");

struct PwArgs {
  float* dw;                  // [Cs, D] fp32 (plain-store fallback)
  const __nv_bfloat16* what;  // [Cs, Dp]
  const float* inv_nw;        // [Cs]
  int Cs, D, Dp;
  int kb_total;               // ceil(B / 64)
  int tiles_m;                // ceil(Cs / 128)
  int debug;                  // ARCB_DEBUG_EPI (timing experiments, wrong results): 1 = no stores in pass 2, 2 = no pass 1 /
                              // swap either, 3 = no epilogue work at all
};

// per-sample inputs of this backward: c_i, the dW operand c (.) x_hat, the exact label entries, and the zeroed dX_hat
struct ProbsSetup {
  const float* lse;        // [B] global log-sum-exp (natural units)
  const float* zt;         // [B] global target logit (natural units)
  const float* grad_loss;  // device scalar or nullptr (= 1)
  const float* dphi;       // [Bp] d(target value) / d cos
  const int* lbl_local;    // [Bp] label - class_offset, -1 = other shard
  const __nv_bfloat16* xhat;
  __nv_bfloat16* xc;       // [B, Dp]
  float* cvec;             // [Bp]
  __nv_bfloat16* probs;    // P'^T [Cs, ldB]
  int64_t ldB;
  float s;
  int B, Bp, Dp;
  float* zero;             // dxhat
  long long n_zero;
};

__global__ void __launch_bounds__(256) probs_setup_kernel(const ProbsSetup su) {
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthr = static_cast<long long>(gridDim.x) * blockDim.x;
  const int lane = threadIdx.x & 31;
  const float gs = (su.grad_loss ? *su.grad_loss : 1.f) * su.s / static_cast<float>(su.B);
  const float s2 = su.s * kLog2e;
  for (long long i = tid >> 5; i < su.Bp; i += nthr >> 5) {  // one warp per sample
    float c = 0.f;
    if (i < su.B) {
      const float lse2 = su.lse[i] * kLog2e;
      c = gs * ex2_approx(s2 - lse2);
      if (lane == 0) {
        const int lbl = su.lbl_local[i];
        if (lbl >= 0) {
          // exact target gradient g_t = gscale (p_t - 1) dphi; stored as g_t / c_i = (2^(zt2 - s2) - 2^(lse2 - s2)) dphi
          const float zt2 = su.zt[i] * kLog2e;
          const float v = (ex2_approx(zt2 - s2) - ex2_approx(lse2 - s2)) * su.dphi[i];
          su.probs[static_cast<int64_t>(lbl) * su.ldB + i] = __float2bfloat16_rn(v);
        }
      }
      for (int d = lane; d < su.Dp; d += 32)
        su.xc[i * su.Dp + d] = __float2bfloat16_rn(c * __bfloat162float(su.xhat[i * su.Dp + d]));
    }
    if (lane == 0) su.cvec[i] = c;
  }
  for (long long i = tid; i < su.n_zero; i += nthr) su.zero[i] = 0.f;
}

__global__ void __launch_bounds__(kGemmThreads, 1)
dw_probs_kernel(const __grid_constant__ CUtensorMap tmPt,   // P'^T [Cs, Bp] as K-major A: box {64 samples, 128 classes}
                const __grid_constant__ CUtensorMap tmXc,   // c.x_hat [B, Dp] as MN-major B: {64 features, 64 samples, 4 blocks}
                const __grid_constant__ CUtensorMap tmDw,   // dW [Cs, D] f32, box 32 x 128
                const PwArgs ar, const int use_tma) {
  using L = PwSmem;
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0u) watchdog_abort(0xa13u, 0u);
  uint8_t* epi_smem = smem + L::kEpiOff;
  float* swap = reinterpret_cast<float*>(smem + L::kSwapOff);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* full = bars;
  uint64_t* empty = bars + kPwStages;
  uint64_t* tfull = bars + 2 * kPwStages;
  uint64_t* tempty = bars + 2 * kPwStages + 2;
  uint64_t* swap_bar = bars + 2 * kPwStages + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + L::kNumBars);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());
  const int rank = static_cast<int>(cluster_ctarank());   // = feature half this CTA computes
  const int cluster = static_cast<int>(blockIdx.x >> 1);
  const int n_clusters = static_cast<int>(gridDim.x >> 1);
  const int n_blk = (ar.debug == 8 || ar.debug == 9) ? 1 - rank : rank;  // debug 8 / 9: feature halves swapped between the CTAs
  const bool half_ok = n_blk * kPwBN < ar.D;              // D <= 256: the second CTA has no columns, it only swaps zeros

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmPt);
    tma_prefetch_desc(&tmXc);
    tma_prefetch_desc(&tmDw);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kPwStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], kEpiThreads / 32);
      mbar_init(&swap_bar[i], 1);  // armed by one local thread; the 2 x 256 partials arrive as transaction bytes
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 2 * kPwBN);
    tmem_relinquish();
  }
  tc_fence_before();
  cluster_sync_all();  // the peer's swap barriers are initialised before anybody sends to them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0 && half_ok) {
      // ------------------------------ TMA producer ------------------------------
      int stage = 0;
      uint32_t phase = 0;
      for (int t = cluster; t < ar.tiles_m; t += n_clusters) {
        for (int kb = 0; kb < ar.kb_total; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1u, 0x100u + stage);
          mbar_arrive_expect_tx(&full[stage], L::kStageBytes);
          uint8_t* sa = smem + stage * L::kStageBytes;
          uint8_t* sb = sa + L::kABytes;
          // P'^T: the same tile is fetched by both CTAs of the cluster (second one hits L2); last use of these bytes
          tma_load_2d_hint(sa, &tmPt, &full[stage], kb * kBK, t * kBM, kEvictFirst);
          // c.x_hat: read by every tile of every CTA -> keep in L2
          tma_load_3d_hint(sb, &tmXc, &full[stage], 0, kb * kBK, n_blk * (kPwBN / 64), kEvictLast);
          if (++stage == kPwStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && half_ok) {
      // ------------------------------ MMA issuer --------------------------------
      constexpr uint32_t idesc = make_idesc_bf16(kBM, kPwBN, false, true);
      const uint64_t da0 = make_smem_desc(0, 16, 1024);         // A K-major: +16 K elements = +32 B in the swizzle line
      const uint64_t db0 = make_smem_desc(0, kBK * 128, 1024);  // B MN-major: 64-element blocks kBK * 128 B apart
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = cluster; t < ar.tiles_m; t += n_clusters) {
        mbar_wait(&tempty[acc], acc_phase ^ 1u, 0x200u + acc);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * kPwBN);
        for (int kb = 0; kb < ar.kb_total; ++kb) {
          mbar_wait(&full[stage], phase, 0x300u + stage);
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem + stage * L::kStageBytes);
          const uint64_t da = da0 + (a_base >> 4);
          const uint64_t db = db0 + ((a_base + L::kABytes) >> 4);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k)  // A: +32 B, B: +16 K lines = +2048 B
            tc_mma_bf16(tmem_d, da + k * (32u >> 4), db + k * (2048u >> 4), idesc, (kb > 0 || k > 0) ? 1u : 0u);
          tc_commit(&empty[stage]);
          if (++stage == kPwStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        tc_commit(&tfull[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else if (warp >= 4) {
    // -------------------------------- epilogue ----------------------------------
    const int quad = warp & 3;
    const int row_in_tile = quad * 32 + lane;
    const int epi_tid = static_cast<int>(threadIdx.x) - (kGemmThreads - kEpiThreads);
    const int wg = epi_tid >> 7;
    const int wg_tid = epi_tid & 127;
    const int c_begin = wg * (kPwBN / 2), c_end = (wg + 1) * (kPwBN / 2);
    const uint32_t swap_local = smem_u32(swap);
    const uint32_t bar_local = smem_u32(swap_bar);
    int acc = 0;
    uint32_t acc_phase = 0;
    int it = 0;       // chunks sent through this warpgroup's staging buffer so far
    uint32_t n_tile = 0;
    for (int t = cluster; t < ar.tiles_m; t += n_clusters, ++n_tile) {
      const uint32_t par = n_tile & 1u, use = n_tile >> 1;
      const int cls = t * kBM + row_in_tile;
      const bool valid = cls < ar.Cs;
      const __nv_bfloat16* wrow = ar.what + static_cast<int64_t>(valid ? cls : 0) * ar.Dp + n_blk * kPwBN;
      const float inv = valid ? __ldg(ar.inv_nw + cls) : 0.f;
      if (epi_tid == 0 && ar.debug < 2) mbar_arrive_expect_tx(&swap_bar[par], 2u * kEpiThreads * 4u);  // 2 CTAs x 256 threads x 4 B
      // ---- pass 1: this thread's share of t_j ----
      float tpart = 0.f;
      if (ar.debug >= 2) {  // timing experiment: no projection, no swap
        if (half_ok) {
          mbar_wait(&tfull[acc], acc_phase, 0x400u + acc);
          tc_fence_after();
        }
      } else if (half_ok) {
        mbar_wait(&tfull[acc], acc_phase, 0x400u + acc);
        tc_fence_after();
        const uint32_t taddr = tmem_base + static_cast<uint32_t>(acc * kPwBN) + (static_cast<uint32_t>(quad * 32) << 16);
#pragma unroll 1
        for (int c = c_begin; c < c_end; c += 32) {
          if (n_blk * kPwBN + c >= ar.D) break;
          uint4 wv[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) wv[q] = __ldg(reinterpret_cast<const uint4*>(wrow + c) + q);
          float v[32];
          tmem_ld32(taddr + c, v);
          float a0 = 0.f, a1 = 0.f;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t wr[4] = {wv[q].x, wv[q].y, wv[q].z, wv[q].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              a0 = fmaf(v[q * 8 + 2 * e + 0], __uint_as_float(wr[e] << 16), a0);
              a1 = fmaf(v[q * 8 + 2 * e + 1], __uint_as_float(wr[e] & 0xffff0000u), a1);
            }
          }
          tpart += a0 + a1;
        }
      }
      // ---- swap: my partial goes to slot (rank * 2 + wg) of BOTH CTAs ----
      float tj = 0.f;
      if (ar.debug < 2) {
        const uint32_t off = ((par * 4u + static_cast<uint32_t>(rank * 2 + wg)) * kBM + static_cast<uint32_t>(row_in_tile)) * 4u;
#pragma unroll
        for (uint32_t dst = 0; dst < 2; ++dst) {
          const uint32_t a = mapa_u32(swap_local + off, dst), b = mapa_u32(bar_local + par * 8u, dst);
          asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(a),
                       "r"(__float_as_uint(tpart)), "r"(b)
                       : "memory");
        }
        mbar_wait(&swap_bar[par], use & 1u, 0xf00u + par);
        const float* sp = swap + (par * 4u) * kBM + row_in_tile;
        tj = (sp[0] + sp[kBM]) + (sp[2 * kBM] + sp[3 * kBM]);
        if (!(inv < 0.5e12f) || ar.debug == 4) tj = 0.f;  // clamped rows (||w|| <= eps) take no projection
      }
      // ---- pass 2: dW = inv * (acc - t_j * w_hat) ----
      if (half_ok) {
        const uint32_t taddr = tmem_base + static_cast<uint32_t>(acc * kPwBN) + (static_cast<uint32_t>(quad * 32) << 16);
#pragma unroll 1
        for (int c = c_begin; c < c_end; c += 32) {
          const int col0 = n_blk * kPwBN + c;
          if (col0 >= ar.D || ar.debug == 3) break;
          uint4 wv[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) wv[q] = __ldg(reinterpret_cast<const uint4*>(wrow + c) + q);
          float v[32];
          tmem_ld32(taddr + c, v);
          float r[32];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t wr[4] = {wv[q].x, wv[q].y, wv[q].z, wv[q].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              r[q * 8 + 2 * e + 0] = inv * fmaf(-tj, __uint_as_float(wr[e] << 16), v[q * 8 + 2 * e + 0]);
              r[q * 8 + 2 * e + 1] = inv * fmaf(-tj, __uint_as_float(wr[e] & 0xffff0000u), v[q * 8 + 2 * e + 1]);
            }
          }
          if (ar.debug >= 5) {  // debugging aids: 5 = constant, 6 = raw accumulator, 7 = the swapped projection t_j
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = ar.debug == 5 ? 1.f : ((ar.debug == 6 || ar.debug == 8) ? v[j] : tj);
          }
          if (ar.debug >= 1 && ar.debug <= 3) {
            if (r[0] == 123.456f && r[31] == 654.321f) tj += 1.f;  // keep the arithmetic alive
          } else if (use_tma) {
            chunk_out_tma<false, 2>(tmDw, epi_smem, wg, wg_tid, row_in_tile, it++, r, col0, t * kBM);
          } else if (valid) {
            float* o = ar.dw + static_cast<int64_t>(cls) * ar.D + col0;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < ar.D) o[j] = r[j];
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
    if (wg_tid == 0) tma_store_wait_all0();
  }

  tc_fence_before();
  cluster_sync_all();  // the peer may still be sending partials into this CTA's shared memory
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * kPwBN);
  }
}

}  // namespace arcb
