// CTA-pair variant of the tcgen05 GEMM core:  D[M,N] = A[M,K] * B[N,K]^T  with cta_group::2.
//
// Why: a single CTA feeding M=128 x N=256 MMAs from shared memory needs 48 KB of operands written by
// TMA AND read by the tensor core per 512 MMA cycles = 187 B/clk, but an SM's shared memory moves
// 128 B/clk -- the 1-CTA core (gemm_sm100.cuh) saturates at ~2/3 of the tensor peak (profiles/r1_*).
// Two CTAs of a cluster (same TPC) issue ONE M=256 x N=256 MMA: each CTA stages its own 128 rows of A
// and only HALF of the B tile (128 of the 256 rows); the tensor cores of both SMs read both halves.
// Per CTA that is 32 KB per k-block in and 32 KB out of shared memory = 125 B/clk.
//
// Pair protocol (rank 0 = leader):
//   full[s]    lives in the leader, count 2: the leader's producer arrives with expect_tx(bytes of BOTH
//              CTAs), the follower's producer arrives remotely; both CTAs' TMA loads (cta_group::2)
//              count their bytes on the leader's barrier.
//   empty[s]   one per CTA, count 1: the leader's tcgen05.commit multicasts the arrive to both CTAs.
//   tfull[a]   one per CTA, count 1: multicast commit when the accumulator stage is complete.
//   tempty[a]  lives in the leader, count 2 x epilogue warps: both CTAs' epilogue warps arrive on it.
// Only the leader's warp 1 issues MMAs; every CTA runs its own producer and its own epilogue over its own
// 128 accumulator rows (TMEM lanes) x 256 columns -- the epilogue policies are those of the 1-CTA core.
#pragma once

#include "gemm_sm100.cuh"

namespace arcb {

// NSUB = 2 ("wide" tiles): a pair works on TWO neighbouring BN-wide column tiles at once -- one A stage feeds two MMAs,
// the two accumulators fill all 512 TMEM columns (no double buffering: meant for split-K launches where a pair has ONE
// output tile for its whole life), and A is fetched once instead of once per column tile.
template <class Epi, int BN, int STAGES, int NSUB = 1>
struct Gemm2Smem {
  static constexpr int kABytes = kBM * kBK * 2;        // this CTA's 128 rows of A
  static constexpr int kBSubBytes = (BN / 2) * kBK * 2;  // this CTA's half of one B tile
  static constexpr int kBBytes = NSUB * kBSubBytes;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiOff = STAGES * kStageBytes;
  static constexpr int kBarOff = kEpiOff + Epi::kSmemBytes;
  static constexpr int kNumBars = 2 * STAGES + 4;
  static constexpr int kTotal = kBarOff + kNumBars * 8 + 16;  // dynamic smem base is 1024 B aligned (checked)
};

// Pair-level tile decode: a pair owns two consecutive 128-row blocks (m_blk = 2 * m_pair + rank).
__device__ __forceinline__ TileCoord decode_tile_pair(const GemmSched& sc, int t, int rank) {
  const int pairs_m = (sc.tiles_m + 1) >> 1;
  TileCoord c;
  int r;
  if (sc.n_fastest) {
    c.n_blk = t % sc.tiles_n;
    r = t / sc.tiles_n;
    c.m_blk = (r % pairs_m) * 2 + rank;
    r /= pairs_m;
  } else {
    c.m_blk = (t % pairs_m) * 2 + rank;
    r = t / pairs_m;
    c.n_blk = r % sc.tiles_n;
    r /= sc.tiles_n;
  }
  const int ks = r;
  c.kb0 = ks * sc.kb_per_split;
  c.kb1 = min(c.kb0 + sc.kb_per_split, sc.kb_total);
  return c;
}

template <class Epi, int BN, int STAGES, bool A_MN, bool B_MN, int NSUB = 1>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm2_sm100_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                   const __grid_constant__ CUtensorMap tmC, const GemmSched sc, const typename Epi::Params ep) {
  using L = Gemm2Smem<Epi, BN, STAGES, NSUB>;
  static_assert(BN == 256, "pair MMA: N = 256 (128 rows of B per CTA)");
  static_assert(NSUB == 1 || NSUB == 2, "one or two column tiles per pair tile");
  constexpr int kAccStages = 2 / NSUB;  // NSUB = 2: both accumulators belong to the same tile
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* tfull = bars + 2 * STAGES;
  uint64_t* tempty = bars + 2 * STAGES + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + L::kNumBars);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if ((smem_u32(smem) & 1023u) != 0u) watchdog_abort(0xa12u, 0u);  // no alignment slack is reserved
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (Epi::kUsesTmC) tma_prefetch_desc(&tmC);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 2);   // used in the leader only: leader producer (+tx) and follower producer
      mbar_init(&empty[i], 1);  // multicast commit
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);                          // multicast commit
      mbar_init(&tempty[i], 2 * (kEpiThreads / 32));    // used in the leader only: both CTAs' epilogue warps
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_pair(tmem_slot, 2 * BN);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();  // the peer's barriers are initialised and its TMEM is allocated
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int pairs_m = (sc.tiles_m + 1) >> 1;
  const int total_tiles = pairs_m * sc.tiles_n * sc.k_splits;
  const int pair_id = static_cast<int>(blockIdx.x >> 1);
  const int num_pairs = static_cast<int>(gridDim.x >> 1);

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------ TMA producer (both CTAs) ------------------------------
      int stage = 0;
      uint32_t phase = 0;
      for (int t = pair_id; t < total_tiles; t += num_pairs) {
        const TileCoord tc = decode_tile_pair(sc, t, static_cast<int>(rank));
        for (int kb = tc.kb0; kb < tc.kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1u, 0x100u + stage);
          const uint32_t full_leader = mapa_u32(smem_u32(&full[stage]), 0);
          if (leader) mbar_arrive_expect_tx(&full[stage], 2 * L::kStageBytes);
          uint8_t* sa = smem + stage * L::kStageBytes;
          uint8_t* sb = sa + L::kABytes;
          if constexpr (!A_MN) {
            tma_load_2d_pair(sa, &tmA, full_leader, kb * kBK, tc.m_blk * kBM, kEvictNormal);
          } else {
            tma_load_3d_pair(sa, &tmA, full_leader, 0, kb * kBK, tc.m_blk * (kBM / 64), kEvictNormal);
          }
#pragma unroll
          for (int j = 0; j < NSUB; ++j) {  // column tile n_blk * NSUB + j (out-of-range boxes arrive as zeros)
            const int nb = tc.n_blk * NSUB + j;
            if constexpr (!B_MN) {
              tma_load_2d_pair(sb + j * L::kBSubBytes, &tmB, full_leader, kb * kBK,
                               nb * BN + static_cast<int>(rank) * (BN / 2), kEvictNormal);
            } else {
              tma_load_3d_pair(sb + j * L::kBSubBytes, &tmB, full_leader, 0, kb * kBK,
                               nb * (BN / 64) + static_cast<int>(rank) * (BN / 128), kEvictNormal);
            }
          }
          if (!leader) mbar_arrive_cluster(full_leader);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {
      // ------------------------------ MMA issuer (leader only) --------------------------------
      constexpr uint32_t idesc = make_idesc_bf16(2 * kBM, BN, A_MN, B_MN);
      const uint64_t da0 = A_MN ? make_smem_desc(0, kBK * 128, 1024) : make_smem_desc(0, 16, 1024);
      const uint64_t db0 = B_MN ? make_smem_desc(0, kBK * 128, 1024) : make_smem_desc(0, 16, 1024);
      constexpr uint32_t a_step = A_MN ? (2048u >> 4) : (32u >> 4);
      constexpr uint32_t b_step = B_MN ? (2048u >> 4) : (32u >> 4);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = pair_id; t < total_tiles; t += num_pairs) {
        const TileCoord tc = decode_tile_pair(sc, t, 0);
        if (tc.kb0 >= tc.kb1) continue;
        mbar_wait(&tempty[acc], acc_phase ^ 1u, 0x200u + acc);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = tc.kb0; kb < tc.kb1; ++kb) {
          mbar_wait(&full[stage], phase, 0x300u + stage);
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem + stage * L::kStageBytes);
          const uint64_t da = da0 + (a_base >> 4);
          const uint64_t db = db0 + ((a_base + L::kABytes) >> 4);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
#pragma unroll
            for (int j = 0; j < NSUB; ++j)
              tc_mma_bf16_pair(tmem_d + static_cast<uint32_t>(j * BN), da + k * a_step,
                               db + static_cast<uint64_t>(j * (L::kBSubBytes >> 4)) + k * b_step, idesc,
                               (kb > tc.kb0 || k > 0) ? 1u : 0u);
          }
          tc_commit_pair(&empty[stage], 0x3);  // frees the slot in both CTAs
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
        tc_commit_pair(&tfull[acc], 0x3);  // accumulator complete -> both epilogues
        if (kAccStages == 2) acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else if (warp >= 4) {
    // -------------------------------- epilogue (both CTAs) ----------------------------------
    const int quad = warp & 3;
    const int row_in_tile = quad * 32 + lane;
    const int epi_tid = static_cast<int>(threadIdx.x) - (kGemmThreads - kEpiThreads);
    uint8_t* epi_smem = smem + L::kEpiOff;
    int acc = 0;
    uint32_t acc_phase = 0;
    typename Epi::State st;
    Epi::init(ep, st);
    for (int t = pair_id; t < total_tiles; t += num_pairs) {
      const TileCoord tc = decode_tile_pair(sc, t, static_cast<int>(rank));
      if (tc.kb0 >= tc.kb1) continue;
      mbar_wait(&tfull[acc], acc_phase, 0x400u + acc);
      tc_fence_after();
      const uint32_t taddr = tmem_base + static_cast<uint32_t>(acc * BN) + (static_cast<uint32_t>(quad * 32) << 16);
      if constexpr (NSUB == 1) {
        Epi::template consume<BN>(ep, st, tc, tmC, taddr, row_in_tile, epi_tid, epi_smem);
      } else {
#pragma unroll 1
        for (int j = 0; j < NSUB; ++j) {
          TileCoord tj = tc;
          tj.n_blk = tc.n_blk * NSUB + j;
          Epi::template consume<BN>(ep, st, tj, tmC, taddr + static_cast<uint32_t>(j * BN), row_in_tile, epi_tid, epi_smem);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty[acc]), 0));  // leader's barrier
      Epi::template finish<BN>(ep, st, tc, tmC, row_in_tile, epi_tid, epi_smem);
      if (kAccStages == 2) acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
    }
    Epi::drain(epi_tid);
  }

  tc_fence_before();
  cluster_sync_all();  // nobody exits (or frees TMEM) while the peer may still signal its barriers / read its smem
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 2 * BN);
  }
}

// Launch one pair GEMM (grid = 2 * floor(SMs / 2) CTAs in clusters of 2).
template <class Epi, int BN, int STAGES, bool A_MN, bool B_MN, int NSUB = 1>
inline cudaError_t launch_gemm_pair(const Operand& a, const Operand& b, const CUtensorMap* tmC_opt, int64_t M,
                                    int64_t N, int64_t K, int split_k, const typename Epi::Params& ep,
                                    cudaStream_t stream) {
  using L = Gemm2Smem<Epi, BN, STAGES, NSUB>;
  static_assert(L::kTotal <= 232448, "shared memory budget exceeded");
  CUtensorMap tmA, tmB, tmC;
  bool ok = A_MN ? make_tmap_mn3(&tmA, a.ptr, a.mn_extent, a.k_extent, a.pitch_elems * 2, kBM / 64)
                 : make_operand_tmap<false>(&tmA, a, kBM);
  ok = ok && (B_MN ? make_tmap_mn3(&tmB, b.ptr, b.mn_extent, b.k_extent, b.pitch_elems * 2, BN / 128)
                   : make_operand_tmap<false>(&tmB, b, BN / 2));
  if (!ok) return cudaErrorInvalidValue;
  if (tmC_opt) tmC = *tmC_opt; else tmC = tmA;
  GemmSched sc;
  sc.tiles_m = static_cast<int>((M + kBM - 1) / kBM);
  sc.tiles_n = static_cast<int>((N + BN * NSUB - 1) / (BN * NSUB));  // tiles of NSUB column tiles each
  sc.kb_total = static_cast<int>((K + kBK - 1) / kBK);
  if (split_k < 1) split_k = 1;
  if (split_k > sc.kb_total) split_k = sc.kb_total;
  sc.kb_per_split = (sc.kb_total + split_k - 1) / split_k;
  sc.k_splits = (sc.kb_total + sc.kb_per_split - 1) / sc.kb_per_split;
  sc.n_fastest = 0;
  const int total = ((sc.tiles_m + 1) / 2) * sc.tiles_n * sc.k_splits;
  if (total <= 0) return cudaSuccess;
  int pairs = device_sm_count() / 2;
  if (pairs > total) pairs = total;
  if (pairs < 1) pairs = 1;
  auto kern = gemm2_sm100_kernel<Epi, BN, STAGES, A_MN, B_MN, NSUB>;
  static std::atomic<bool> attr_set[64];  // zero-initialised; forward and backward call in from different threads
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!attr_set[dev].load(std::memory_order_acquire)) {  // idempotent: a race only repeats the call
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
    if (e != cudaSuccess) return e;
    attr_set[dev].store(true, std::memory_order_release);
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = L::kTotal;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmC, sc, ep);
}

// Pair GEMM whose tiles_n column slices of one 256-row tile must be in flight TOGETHER on neighbouring pairs (EpiDWT
// exchanges a per-row reduction between them): n-fastest tile order, a number of pairs that is a multiple of tiles_n,
// no split-K, and a cooperative launch (on top of the cluster) so that every CTA a tile may wait for is resident.
// cudaErrorCooperativeLaunchTooLarge & co.: nothing was enqueued (the caller retries with fewer pairs).
template <class Epi, int BN, int STAGES, bool A_MN, bool B_MN>
inline cudaError_t launch_gemm_pair_row_slices(const Operand& a, const Operand& b, const CUtensorMap* tmC_opt, int64_t M,
                                               int64_t N, int64_t K, const typename Epi::Params& ep,
                                               cudaStream_t stream, bool cooperative, int pairs) {
  using L = Gemm2Smem<Epi, BN, STAGES>;
  static_assert(L::kTotal <= 232448, "shared memory budget exceeded");
  CUtensorMap tmA, tmB, tmC;
  bool ok = A_MN ? make_tmap_mn3(&tmA, a.ptr, a.mn_extent, a.k_extent, a.pitch_elems * 2, kBM / 64)
                 : make_operand_tmap<false>(&tmA, a, kBM);
  ok = ok && (B_MN ? make_tmap_mn3(&tmB, b.ptr, b.mn_extent, b.k_extent, b.pitch_elems * 2, BN / 128)
                   : make_operand_tmap<false>(&tmB, b, BN / 2));
  if (!ok) return cudaErrorInvalidValue;
  if (tmC_opt) tmC = *tmC_opt; else tmC = tmA;
  GemmSched sc;
  sc.tiles_m = static_cast<int>((M + kBM - 1) / kBM);
  sc.tiles_n = static_cast<int>((N + BN - 1) / BN);
  sc.kb_total = static_cast<int>((K + kBK - 1) / kBK);
  sc.kb_per_split = sc.kb_total;
  sc.k_splits = 1;
  sc.n_fastest = 1;
  const int total = ((sc.tiles_m + 1) / 2) * sc.tiles_n;
  if (total <= 0) return cudaSuccess;
  // `pairs` is the caller's (the epilogue parameters depend on it): a multiple of tiles_n, at most one pair per two SMs
  if (pairs > total || pairs < sc.tiles_n || pairs % sc.tiles_n) return cudaErrorInvalidValue;
  auto kern = gemm2_sm100_kernel<Epi, BN, STAGES, A_MN, B_MN>;
  static std::atomic<bool> attr_set[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!attr_set[dev].load(std::memory_order_acquire)) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
    if (e != cudaSuccess) return e;
    attr_set[dev].store(true, std::memory_order_release);
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = L::kTotal;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeCooperative;
  attr[1].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = cooperative ? 2 : 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmC, sc, ep);
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e;
}

}  // namespace arcb
