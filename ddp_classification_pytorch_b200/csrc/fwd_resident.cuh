// Forward of the ArcFace head with the batch block RESIDENT in shared memory (arc_main.py:159-172 + the softmax
// statistics of :245 / :327).
//
// gemm_sm100_kernel streams both operands of cos = x_hat * w_hat^T: every 128 x 256 tile re-fetches its 128 rows of
// x_hat (16 KB per k-block) next to the 256 class rows (32 KB per k-block), and leaves one (max, sum-exp) partial per
// row and 128-class block for a merge kernel to reduce (64 MB written and re-read at 1M classes x 1024 rows).
// Here a CTA owns ONE block of 128 samples for its whole life:
//   * its x_hat rows (128 x Dp bf16, up to 128 KB) are loaded once and stay in shared memory: only w_hat tiles stream
//     (a third less L2 -> SM operand traffic, which also leaves more of the fabric to the normaliser warps);
//   * the online (max, sum-exp) of a row lives in the registers of the epilogue thread that owns it across ALL class
//     tiles of the CTA: one partial per row and CTA leaves the kernel (2 * lanes per row instead of 2 * C / 256), so
//     the merge is a single small launch;
//   * the block maxima the exact top-3 re-score needs (arcb_shard_top3) are written only when asked for.
// CTA c works on sample block c % tiles_m and class tiles lane, lane + lanes, ... with lane = c / tiles_m and
// lanes = grid / tiles_m: at any time the CTAs are on neighbouring class tiles, so a w_hat tile is fetched from HBM
// once and found in L2 by the other sample blocks, and the in-kernel normaliser (NORM, see gemm_sm100.cuh) still
// produces rows in the order they are consumed.
//
// Roles as in gemm_sm100_kernel: warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator, warps 4..11 epilogue
// (two warpgroups, half of the 256 columns each), NORM: warps 12..19 normalise weight rows.
#pragma once

#include "gemm_sm100.cuh"

namespace arcb {

constexpr int kResMaxKb = 8;    // resident A: up to 8 k-blocks of [128 x 64] bf16 = 128 KB  (Dp <= 512)
constexpr int kResStages = 3;   // streamed B: 3 x [256 x 64] bf16 = 96 KB
constexpr int kResBN = 256;

struct ResSmem {
  static constexpr int kABlock = kBM * kBK * 2;            // 16 KB
  static constexpr int kABytes = kResMaxKb * kABlock;      // 128 KB
  static constexpr int kBBytes = kResBN * kBK * 2;         // 32 KB
  static constexpr int kBarOff = kABytes + kResStages * kBBytes;
  static constexpr int kNumBars = 2 * kResStages + 4 + 1;  // full/empty ring, tfull/tempty, a_full
  static constexpr int kTotal = kBarOff + kNumBars * 8 + 16 + 1024;  // +1024: manual alignment slack
};
static_assert(ResSmem::kTotal <= 232448, "resident forward exceeds the shared memory budget");

struct ResSched {
  int tiles_m, tiles_n, kb_total;
  int lanes;  // CTAs per sample block (class tiles are dealt round-robin to them)
};

struct ResParams {
  const int* lbl_local;  // [Bp] label - class_offset if this shard owns it, else -1
  const float* phi;      // [Bp] value that replaces the target cosine (phi or the fallback)
  float2* part;          // [2 * lanes][ldp]: one (max * s2, sum 2^(u - max)) per row, CTA and warpgroup
  float* bmax;           // [2 * tiles_n][ldp] block maxima (base-2 units) or nullptr
  int ldp;
  int B, C;
  float s2;
};

template <bool NORM>
__global__ void __launch_bounds__(kGemmThreads + (NORM ? 32 * kNormWarps : 0), 1)
fwd_resident_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const ResSched sc,
                    const ResParams ep, const NormArgs na) {
  using L = ResSmem;
  __shared__ int norm_progress;  // NORM: class tile the CTA's producer is working on (read by the normaliser warps)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_b = smem + L::kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* full = bars;
  uint64_t* empty = bars + kResStages;
  uint64_t* tfull = bars + 2 * kResStages;
  uint64_t* tempty = bars + 2 * kResStages + 2;
  uint64_t* a_full = bars + 2 * kResStages + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + L::kNumBars);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());
  const int cta = static_cast<int>(blockIdx.x);
  const int m_blk = cta % sc.tiles_m;
  const int my_lane = cta / sc.tiles_m;
  const bool has_work = my_lane < sc.lanes && my_lane < sc.tiles_n;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kResStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], kEpiThreads / 32);
    }
    mbar_init(a_full, 1);
    // a CTA without class tiles never advances: its normaliser warps must not wait for it
    if (NORM) norm_progress = has_work ? 0 : 0x3fffffff;
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 2 * kResBN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0 && has_work) {
      // ------------------------------ TMA producer ------------------------------
      // the sample block: once, all k-blocks on one barrier (x_hat is read by every CTA of this sample block: keep in L2)
      mbar_arrive_expect_tx(a_full, static_cast<uint32_t>(sc.kb_total) * L::kABlock);
      for (int kb = 0; kb < sc.kb_total; ++kb)
        tma_load_2d_hint(smem + kb * L::kABlock, &tmA, a_full, kb * kBK, m_blk * kBM, kEvictLast);
      int stage = 0;
      uint32_t phase = 0;
      for (int n = my_lane; n < sc.tiles_n; n += sc.lanes) {
        if constexpr (NORM) {  // all 256 rows of this class tile must be normalised and visible to the bulk loads
          *reinterpret_cast<volatile int*>(&norm_progress) = n;
          wait_counter(na.done + n, static_cast<unsigned>(kNormWarps * (kResBN / kNormUnitRows)), 0xe00u);
          fence_proxy_async_all();
        }
        for (int kb = 0; kb < sc.kb_total; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1u, 0x100u + stage);
          mbar_arrive_expect_tx(&full[stage], L::kBBytes);
          tma_load_2d(smem_b + stage * L::kBBytes, &tmB, &full[stage], kb * kBK, n * kResBN);
          if (++stage == kResStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
      if constexpr (NORM) *reinterpret_cast<volatile int*>(&norm_progress) = 0x3fffffff;  // done: never hold them back
    }
  } else if (warp == 1) {
    if (lane == 0 && has_work) {
      // ------------------------------ MMA issuer --------------------------------
      constexpr uint32_t idesc = make_idesc_bf16(kBM, kResBN, false, false);
      const uint64_t d0 = make_smem_desc(0, 16, 1024);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      mbar_wait(a_full, 0u, 0xa00u);
      tc_fence_after();
      const uint32_t a_res = smem_u32(smem);
      for (int n = my_lane; n < sc.tiles_n; n += sc.lanes) {
        mbar_wait(&tempty[acc], acc_phase ^ 1u, 0x200u + acc);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * kResBN);
        for (int kb = 0; kb < sc.kb_total; ++kb) {
          mbar_wait(&full[stage], phase, 0x300u + stage);
          tc_fence_after();
          const uint64_t da = d0 + ((a_res + kb * L::kABlock) >> 4);
          const uint64_t db = d0 + (smem_u32(smem_b + stage * L::kBBytes) >> 4);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k)  // +16 elements of K = +32 B inside the swizzle line = +2 in the descriptor
            tc_mma_bf16(tmem_d, da + k * 2, db + k * 2, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          tc_commit(&empty[stage]);  // frees the B slot when these MMAs retire
          if (++stage == kResStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        tc_commit(&tfull[acc]);  // accumulator complete -> epilogue
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else if (NORM && warp >= 4 + kEpiThreads / 32) {
    // ------------------ normaliser warps: F.normalize(weight) for the tiles still to come ------------------
    const int nw = warp - (4 + kEpiThreads / 32);
    const int64_t units = (static_cast<int64_t>(sc.tiles_n) * kResBN) / kNormUnitRows;
    for (int64_t u = blockIdx.x; u < units; u += gridDim.x) {
      const int blk = static_cast<int>(u / (kResBN / kNormUnitRows));
      while (blk > *reinterpret_cast<volatile int*>(&norm_progress) + na.lookahead) __nanosleep(200);
      const int64_t row0 = u * kNormUnitRows + nw * (kNormUnitRows / kNormWarps);
      int nrows = kNormUnitRows / kNormWarps;
      if (row0 + nrows > na.rows) nrows = static_cast<int>(na.rows > row0 ? na.rows - row0 : 0);
      if (nrows > 0) {
        if (na.Dp <= 128) l2norm_rows_warp<1>(na.w, na.ldw, row0, nrows, na.D, na.what, na.Dp, na.inv_nw, lane);
        else if (na.Dp <= 256) l2norm_rows_warp<2>(na.w, na.ldw, row0, nrows, na.D, na.what, na.Dp, na.inv_nw, lane);
        else l2norm_rows_warp<4>(na.w, na.ldw, row0, nrows, na.D, na.what, na.Dp, na.inv_nw, lane);
      }
      __syncwarp();
      if (lane == 0) {
        __threadfence();
        atomicAdd(na.done + blk, 1u);
      }
    }
  } else if (warp >= 4 && warp < 4 + kEpiThreads / 32) {
    // -------------------------------- epilogue ----------------------------------
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const int row_in_tile = quad * 32 + lane;
    const int epi_tid = static_cast<int>(threadIdx.x) - (kGemmThreads - kEpiThreads);  // 0..255
    const int wg = epi_tid >> 7;
    const int row = m_blk * kBM + row_in_tile;
    int lbl = -1;
    float phi = 0.f;
    if (row < ep.B) {
      lbl = __ldg(ep.lbl_local + row);
      phi = __ldg(ep.phi + row);
    }
    float run_m = -INFINITY, run_l = 0.f;  // this row's online softmax over ALL class tiles of the CTA (this half)
    if (has_work) {
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int n = my_lane; n < sc.tiles_n; n += sc.lanes) {
        mbar_wait(&tfull[acc], acc_phase, 0x400u + acc);
        tc_fence_after();
        const uint32_t taddr = tmem_base + static_cast<uint32_t>(acc * kResBN) + (static_cast<uint32_t>(quad * 32) << 16);
        const int col_base = n * kResBN;
        float blk_m = -INFINITY;
#pragma unroll 1
        for (int c = wg * (kResBN / 2); c < (wg + 1) * (kResBN / 2); c += 32) {
          const int col0 = col_base + c;
          if (col0 >= ep.C) break;
          float v[32];
          tmem_ld32(taddr + c, v);
          if (col0 + 32 > ep.C) {  // ragged last tile: padded classes must not enter the softmax
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j >= ep.C) v[j] = -INFINITY;
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
          blk_m = fmaxf(blk_m, cm);
          const float nm = fmaxf(run_m, cm);
          run_l *= ex2_approx((run_m - nm) * ep.s2);
          const float off = nm * ep.s2;
          float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            a0 += ex2_approx(fmaf(v[j + 0], ep.s2, -off));
            a1 += ex2_approx(fmaf(v[j + 1], ep.s2, -off));
            a2 += ex2_approx(fmaf(v[j + 2], ep.s2, -off));
            a3 += ex2_approx(fmaf(v[j + 3], ep.s2, -off));
          }
          run_l += (a0 + a1) + (a2 + a3);
          run_m = nm;
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[acc]);  // TMEM stage may be overwritten
        if (ep.bmax && row < ep.B)  // what arcb_shard_top3 prunes with: the maximum of this 128-class block
          ep.bmax[static_cast<int64_t>(n * 2 + wg) * ep.ldp + row] = blk_m * ep.s2;
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
    // one partial per row, CTA and warpgroup; CTAs without tiles (and rows never touched) contribute the empty partial
    if (row < ep.B && my_lane < sc.lanes)
      ep.part[static_cast<int64_t>(my_lane * 2 + wg) * ep.ldp + row] =
          (run_m == -INFINITY) ? make_float2(-INFINITY, 0.f) : make_float2(run_m * ep.s2, run_l);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * kResBN);
  }
}

// When the resident schedule applies: the sample block must fit (Dp <= 512) and every CTA must have a few class
// tiles to amortise loading it.
inline bool resident_forward_applies(int64_t B, int64_t Cs, int64_t Dp) {
  const int64_t tiles_m = (B + kBM - 1) / kBM, tiles_n = (Cs + kResBN - 1) / kResBN;
  const int sms = device_sm_count();
  if (Dp > kResMaxKb * kBK || tiles_m > sms) return false;
  const int64_t lanes = sms / tiles_m;
  return tiles_n >= 4 * lanes;
}

inline int resident_lanes(int64_t B) {
  const int64_t tiles_m = (B + kBM - 1) / kBM;
  return static_cast<int>(device_sm_count() / tiles_m);
}

// Launch.  NORM: the weight is normalised inside the kernel (cooperative launch: a tile may wait for rows normalised by
// any other CTA, so the whole grid has to be resident; a refused launch returns the error with nothing enqueued).
template <bool NORM>
inline cudaError_t launch_fwd_resident(const Operand& a, const Operand& b, int64_t M, int64_t N, int64_t K,
                                       const ResParams& ep, NormArgs na, cudaStream_t stream) {
  CUtensorMap tmA, tmB;
  if (!make_operand_tmap<false>(&tmA, a, kBM)) return cudaErrorInvalidValue;
  if (!make_operand_tmap<false>(&tmB, b, kResBN)) return cudaErrorInvalidValue;
  ResSched sc;
  sc.tiles_m = static_cast<int>((M + kBM - 1) / kBM);
  sc.tiles_n = static_cast<int>((N + kResBN - 1) / kResBN);
  sc.kb_total = static_cast<int>((K + kBK - 1) / kBK);
  if (sc.kb_total > kResMaxKb || sc.tiles_m <= 0 || sc.tiles_n <= 0) return cudaErrorInvalidValue;
  const int sms = device_sm_count();
  sc.lanes = sms / sc.tiles_m;
  if (sc.lanes < 1) return cudaErrorInvalidValue;
  // NORM: every SM runs a CTA (the ones without class tiles still normalise); else only the CTAs that have tiles
  const int grid = NORM ? sms : sc.lanes * sc.tiles_m;
  auto kern = fwd_resident_kernel<NORM>;
  static std::atomic<bool> attr_set[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!attr_set[dev].load(std::memory_order_acquire)) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, ResSmem::kTotal);
    if (e != cudaSuccess) return e;
    attr_set[dev].store(true, std::memory_order_release);
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads + (NORM ? 32 * kNormWarps : 0));
  cfg.dynamicSmemBytes = ResSmem::kTotal;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = NORM ? 1 : 0;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmA, tmB, sc, ep, na);
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e;
}

}  // namespace arcb
