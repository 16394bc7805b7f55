// Persistent, warp-specialised tcgen05 GEMM for sm_100a:  D[M,N] = A[M,K] * B[N,K]^T
//   bf16 operands staged in shared memory by TMA (SWIZZLE_128B), fp32 accumulators in TMEM,
//   double-buffered so the epilogue of tile i overlaps the MMAs of tile i+1.
//
// Roles (384 threads, one CTA per SM):
//   warp 0  lane 0 : TMA producer          (full/empty mbarrier ring of STAGES slots)
//   warp 1  lane 0 : tcgen05.mma issuer    (one thread issues for the whole CTA)
//   warp 2         : TMEM allocator / deallocator
//   warps 4..11    : epilogue, two warpgroups -- warp w owns TMEM lanes 32*(w%4) .. +31, i.e. one
//                    accumulator row per thread; warpgroup g = (w-4)/4 walks columns
//                    [g*BN/2, (g+1)*BN/2) in chunks of 32, so every SM sub-partition has two
//                    epilogue warps to hide tcgen05.ld / MUFU latency.
//
// Either operand may be K-major (K contiguous in global memory) or MN-major (the M / N index
// contiguous); both are loaded as 128 B-wide swizzled boxes and described to the tensor core by
// the matching canonical shared-memory descriptor (ptx_sm100.cuh).
//
// The epilogue is a policy class (epilogues.cuh); the four hot-path GEMMs of the ArcFace head
// (F.linear at arc_main.py:159 and the two autograd GEMMs behind :329) are this
// kernel with different epilogues.
#pragma once

#include <stdlib.h>

#include <atomic>

#include "ptx_sm100.cuh"
#include "rowwise.cuh"

namespace arcb {

constexpr int kBM = 128;  // accumulator rows per tile = TMEM lanes
constexpr int kBK = 64;   // bf16 elements per 128 B swizzle line
constexpr int kEpiThreads = 256;            // two epilogue warpgroups
constexpr int kGemmThreads = 128 + kEpiThreads;

struct GemmSched {
  int tiles_m;
  int tiles_n;
  int k_splits;
  int kb_total;      // number of kBK-blocks over the whole K extent
  int kb_per_split;  // ceil(kb_total / k_splits)
  int n_fastest;     // pair core only: 1 = n fastest (EpiDWT: the column slices of a row tile run side by side)
};

// Work of the optional normaliser warps (NORM = true: warps 12..15, forward GEMM only): F.normalize(self.weight)
// (arc_main.py:159) runs INSIDE the GEMM kernel that consumes it, so the HBM-bound pass hides behind the tensor-bound
// one.  The rows are cut into units of 64; CTA c normalises units c, c + grid, ... (16 rows per warp) and bumps
// done[unit / 4]; the TMA producer of a tile waits until done[n_blk] has reached 16 (all 256 rows of its B tile).
// The normalisers never wait for anything but the pace of their own CTA's GEMM (they stay at most `lookahead`
// B tiles ahead of it, so that w_hat is still in L2 when the bulk loads come for it): no cyclic dependency.
struct NormArgs {
  const float* w;
  int64_t ldw, rows;
  int D, Dp;
  __nv_bfloat16* what;
  float* inv_nw;
  unsigned int* done;  // [tiles_n], zeroed before the launch
  int lookahead;       // B tiles
};
constexpr int kNormUnitRows = 64;
constexpr int kNormWarps = 8;  // normaliser warps per CTA (NORM kernels): 8 x 4 rows x 2 KB = 64 KB of loads in flight per SM

struct TileCoord {
  int m_blk, n_blk, kb0, kb1;
};

__device__ __forceinline__ TileCoord decode_tile(const GemmSched& sc, int t) {
  TileCoord c;
  c.m_blk = t % sc.tiles_m;  // m fastest: CTAs running side by side share the same B tile via L2
  int r = t / sc.tiles_m;
  c.n_blk = r % sc.tiles_n;
  int ks = r / sc.tiles_n;
  c.kb0 = ks * sc.kb_per_split;
  c.kb1 = min(c.kb0 + sc.kb_per_split, sc.kb_total);
  return c;
}

template <class Epi, int BN, int STAGES>
struct GemmSmem {
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiOff = STAGES * kStageBytes;
  static constexpr int kBarOff = kEpiOff + Epi::kSmemBytes;
  static constexpr int kNumBars = 2 * STAGES + 4;
  static constexpr int kTotal = kBarOff + kNumBars * 8 + 16 + 1024;  // +1024: manual alignment slack
};

template <class Epi, int BN, int STAGES, bool A_MN, bool B_MN, bool NORM = false>
__global__ void __launch_bounds__(kGemmThreads + (NORM ? 32 * kNormWarps : 0), 1)
gemm_sm100_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmC, const GemmSched sc, const typename Epi::Params ep,
                  const NormArgs na) {
  using L = GemmSmem<Epi, BN, STAGES>;
  static_assert(!NORM || (BN == 256 && !B_MN), "the normaliser produces K-major B tiles of 256 rows");
  __shared__ int norm_progress;  // NORM: B tile the CTA's producer is working on (read by the normaliser warps)
  static_assert(BN == 128 || BN == 256, "accumulator double-buffer must be a power of two <= 512 columns");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* tfull = bars + 2 * STAGES;
  uint64_t* tempty = bars + 2 * STAGES + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + L::kNumBars);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (Epi::kUsesTmC) tma_prefetch_desc(&tmC);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], kEpiThreads / 32);
    }
    if (NORM) norm_progress = 0;
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 2 * BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_tiles = sc.tiles_m * sc.tiles_n * sc.k_splits;

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------ TMA producer ------------------------------
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const TileCoord tc = decode_tile(sc, t);
        if constexpr (NORM) {  // all 256 rows of this B tile must be normalised and visible to the bulk loads
          *reinterpret_cast<volatile int*>(&norm_progress) = tc.n_blk;
          wait_counter(na.done + tc.n_blk, static_cast<unsigned>(kNormWarps * (BN / kNormUnitRows)), 0xe00u);
          fence_proxy_async_all();
        }
        for (int kb = tc.kb0; kb < tc.kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1u, 0x100u + stage);
          mbar_arrive_expect_tx(&full[stage], L::kStageBytes);
          uint8_t* sa = smem + stage * L::kStageBytes;
          uint8_t* sb = sa + L::kABytes;
          if constexpr (!A_MN) {
            tma_load_2d(sa, &tmA, &full[stage], kb * kBK, tc.m_blk * kBM);
          } else {
#pragma unroll
            for (int j = 0; j < kBM / 64; ++j)
              tma_load_2d(sa + j * (kBK * 128), &tmA, &full[stage], tc.m_blk * kBM + j * 64, kb * kBK);
          }
          if constexpr (!B_MN) {
            tma_load_2d(sb, &tmB, &full[stage], kb * kBK, tc.n_blk * BN);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sb + j * (kBK * 128), &tmB, &full[stage], tc.n_blk * BN + j * 64, kb * kBK);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ------------------------------ MMA issuer --------------------------------
      constexpr uint32_t idesc = make_idesc_bf16(kBM, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const TileCoord tc = decode_tile(sc, t);
        if (tc.kb0 >= tc.kb1) continue;
        mbar_wait(&tempty[acc], acc_phase ^ 1u, 0x200u + acc);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = tc.kb0; kb < tc.kb1; ++kb) {
          mbar_wait(&full[stage], phase, 0x300u + stage);
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem + stage * L::kStageBytes);
          const uint32_t b_base = a_base + L::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            // K-major: +16 elements = +32 B inside the swizzle line.
            // MN-major: +16 K lines = +2048 B; 64-element MN blocks are kBK*128 B apart (LBO).
            const uint64_t da = A_MN ? make_smem_desc(a_base + k * 2048, kBK * 128, 1024)
                                     : make_smem_desc(a_base + k * 32, 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc(b_base + k * 2048, kBK * 128, 1024)
                                     : make_smem_desc(b_base + k * 32, 16, 1024);
            tc_mma_bf16(tmem_d, da, db, idesc, (kb > tc.kb0 || k > 0) ? 1u : 0u);
          }
          tc_commit(&empty[stage]);  // frees the smem slot when these MMAs retire
          if (++stage == STAGES) {
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
    const int64_t units = (static_cast<int64_t>(sc.tiles_n) * BN) / kNormUnitRows;
    for (int64_t u = blockIdx.x; u < units; u += gridDim.x) {
      const int blk = static_cast<int>(u / (BN / kNormUnitRows));
      while (blk > *reinterpret_cast<volatile int*>(&norm_progress) + na.lookahead) __nanosleep(200);
      const int64_t row0 = u * kNormUnitRows + nw * (kNormUnitRows / kNormWarps);
      int nrows = kNormUnitRows / kNormWarps;
      if (row0 + nrows > na.rows) nrows = static_cast<int>(na.rows > row0 ? na.rows - row0 : 0);
      if (nrows > 0) {
        if (na.Dp <= 128) l2norm_rows_warp<1>(na.w, na.ldw, row0, nrows, na.D, na.what, na.Dp, na.inv_nw, lane);
        else if (na.Dp <= 256) l2norm_rows_warp<2>(na.w, na.ldw, row0, nrows, na.D, na.what, na.Dp, na.inv_nw, lane);
        else if (na.Dp <= 512) l2norm_rows_warp<4>(na.w, na.ldw, row0, nrows, na.D, na.what, na.Dp, na.inv_nw, lane);
        else if (na.Dp <= 1024) l2norm_rows_warp<8>(na.w, na.ldw, row0, nrows, na.D, na.what, na.Dp, na.inv_nw, lane);
        else l2norm_rows_warp<16>(na.w, na.ldw, row0, nrows, na.D, na.what, na.Dp, na.inv_nw, lane);
      }
      __syncwarp();
      if (lane == 0) {
        __threadfence();
        atomicAdd(na.done + blk, 1u);
      }
    }
  } else if (warp >= 4) {
    // -------------------------------- epilogue ----------------------------------
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const int row_in_tile = quad * 32 + lane;
    const int epi_tid = static_cast<int>(threadIdx.x) - (kGemmThreads - kEpiThreads);  // 0..255, wg = epi_tid >> 7
    uint8_t* epi_smem = smem + L::kEpiOff;
    int acc = 0;
    uint32_t acc_phase = 0;
    typename Epi::State st;
    Epi::init(ep, st);
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const TileCoord tc = decode_tile(sc, t);
      if (tc.kb0 >= tc.kb1) continue;
      mbar_wait(&tfull[acc], acc_phase, 0x400u + acc);
      tc_fence_after();
      const uint32_t taddr = tmem_base + static_cast<uint32_t>(acc * BN) + (static_cast<uint32_t>(quad * 32) << 16);
      Epi::template consume<BN>(ep, st, tc, tmC, taddr, row_in_tile, epi_tid, epi_smem);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);  // TMEM stage may be overwritten
      Epi::template finish<BN>(ep, st, tc, tmC, row_in_tile, epi_tid, epi_smem);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
    }
    Epi::drain(epi_tid);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

// ------------------------------------ host side ------------------------------------

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// Resolved once per process through the runtime, so the library does not link libcuda.
inline PFN_encodeTiled get_encode_fn() {
  static std::atomic<PFN_encodeTiled> cached{nullptr};
  PFN_encodeTiled fn = cached.load(std::memory_order_acquire);
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<PFN_encodeTiled>(p);
  cached.store(fn, std::memory_order_release);
  return fn;
}

// 2-D row-major tensor [outer, inner] with `pitch_bytes` between rows, box = [box_outer, box_inner],
// 128 B swizzle, out-of-bounds elements read as zero / are not written.
inline bool make_tmap_2d(CUtensorMap* tm, const void* base, CUtensorMapDataType dt, int elem_bytes, uint64_t inner,
                         uint64_t outer, uint64_t pitch_bytes, uint32_t box_inner, uint32_t box_outer) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return false;
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || (pitch_bytes & 15u) || box_inner * elem_bytes > 128) return false;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {pitch_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// MN-major bf16 operand stored row-major as [k_rows, mn_extent] (mn contiguous, mn_extent % 64 == 0) seen as the
// rank-3 tensor {64 elements, k_rows, mn_extent / 64 blocks}: ONE bulk copy of box {64, 64, nblk} lands in shared
// memory as nblk consecutive [64 k-rows x 128 B] swizzled blocks -- the canonical MN-major SWIZZLE_128B layout.
inline bool make_tmap_mn3(CUtensorMap* tm, const void* base, uint64_t mn_extent, uint64_t k_rows, uint64_t pitch_bytes,
                          uint32_t nblk) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return false;
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || (pitch_bytes & 15u) || (mn_extent % 64)) return false;
  cuuint64_t dims[3] = {64, k_rows, mn_extent / 64};
  cuuint64_t strides[2] = {pitch_bytes, 128};
  cuuint32_t box[3] = {64, kBK, nblk};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// Operand description for the launcher: a bf16 matrix in global memory.
//   K-major : stored [mn_extent rows, k_extent cols]   (k contiguous)
//   MN-major: stored [k_extent rows, mn_extent cols]   (m / n contiguous)
struct Operand {
  const void* ptr;
  int64_t mn_extent;
  int64_t k_extent;
  int64_t pitch_elems;
};

template <bool MN_MAJOR>
inline bool make_operand_tmap(CUtensorMap* tm, const Operand& op, int box_mn) {
  if (!MN_MAJOR)
    return make_tmap_2d(tm, op.ptr, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, op.k_extent, op.mn_extent, op.pitch_elems * 2,
                        kBK, box_mn);
  return make_tmap_2d(tm, op.ptr, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, op.mn_extent, op.k_extent, op.pitch_elems * 2,
                      64, kBK);
}

inline int device_sm_count() {
  static std::atomic<int> sms[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  int n = sms[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    sms[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

// Launch one GEMM.  M/N/K are the logical extents; split_k > 1 partitions K across CTAs (the
// epilogue must then be additive).  Returns cudaSuccess or the first error.
template <class Epi, int BN, int STAGES, bool A_MN, bool B_MN>
inline cudaError_t launch_gemm(const Operand& a, const Operand& b, const CUtensorMap* tmC_opt, int64_t M, int64_t N,
                               int64_t K, int split_k, const typename Epi::Params& ep, cudaStream_t stream,
                               int max_ctas = 0) {
  using L = GemmSmem<Epi, BN, STAGES>;
  static_assert(L::kTotal <= 232448, "shared memory budget exceeded");
  CUtensorMap tmA, tmB, tmC;
  if (!make_operand_tmap<A_MN>(&tmA, a, kBM)) return cudaErrorInvalidValue;
  if (!make_operand_tmap<B_MN>(&tmB, b, BN)) return cudaErrorInvalidValue;
  if (tmC_opt) tmC = *tmC_opt; else tmC = tmA;
  GemmSched sc;
  sc.tiles_m = static_cast<int>((M + kBM - 1) / kBM);
  sc.tiles_n = static_cast<int>((N + BN - 1) / BN);
  sc.kb_total = static_cast<int>((K + kBK - 1) / kBK);
  if (split_k < 1) split_k = 1;
  if (split_k > sc.kb_total) split_k = sc.kb_total;
  sc.kb_per_split = (sc.kb_total + split_k - 1) / split_k;
  sc.k_splits = (sc.kb_total + sc.kb_per_split - 1) / sc.kb_per_split;
  sc.n_fastest = 0;
  const int total = sc.tiles_m * sc.tiles_n * sc.k_splits;
  if (total <= 0) return cudaSuccess;
  int grid = device_sm_count();
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  if (grid > total) grid = total;
  auto kern = gemm_sm100_kernel<Epi, BN, STAGES, A_MN, B_MN>;
  static std::atomic<bool> attr_set[64];  // zero-initialised; forward and backward call in from different threads
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!attr_set[dev].load(std::memory_order_acquire)) {  // idempotent: a race only repeats the call
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
    if (e != cudaSuccess) return e;
    attr_set[dev].store(true, std::memory_order_release);
  }
  kern<<<grid, kGemmThreads, L::kTotal, stream>>>(tmA, tmB, tmC, sc, ep, NormArgs{});
  return cudaGetLastError();
}

// Forward GEMM with the weight normalisation fused in (NormArgs): K-major operands, one 256-column B tile per
// 256 rows of the weight.  Cooperative launch: a tile may wait for rows normalised by any other CTA, so the whole
// grid has to be resident.  cudaErrorCooperativeLaunchTooLarge & co. mean "nothing was enqueued, use the two-kernel
// path".  na.done must have been zeroed on `stream`.
template <class Epi, int STAGES>
inline cudaError_t launch_gemm_norm(const Operand& a, const Operand& b, int64_t M, int64_t N, int64_t K,
                                    const typename Epi::Params& ep, NormArgs na, cudaStream_t stream,
                                    const CUtensorMap* tmC_opt = nullptr) {
  constexpr int BN = 256;
  using L = GemmSmem<Epi, BN, STAGES>;
  static_assert(L::kTotal <= 232448, "shared memory budget exceeded");
  CUtensorMap tmA, tmB;
  if (!make_operand_tmap<false>(&tmA, a, kBM)) return cudaErrorInvalidValue;
  if (!make_operand_tmap<false>(&tmB, b, BN)) return cudaErrorInvalidValue;
  GemmSched sc;
  sc.tiles_m = static_cast<int>((M + kBM - 1) / kBM);
  sc.tiles_n = static_cast<int>((N + BN - 1) / BN);
  sc.kb_total = static_cast<int>((K + kBK - 1) / kBK);
  sc.kb_per_split = sc.kb_total;
  sc.k_splits = 1;
  sc.n_fastest = 0;
  const int total = sc.tiles_m * sc.tiles_n;
  if (total <= 0) return cudaSuccess;
  int grid = device_sm_count();
  if (grid > total) grid = total;
  auto kern = gemm_sm100_kernel<Epi, BN, STAGES, false, false, true>;
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
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads + 32 * kNormWarps);
  cfg.dynamicSmemBytes = L::kTotal;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmC_opt ? *tmC_opt : tmA, sc, ep, na);
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e;
}

}  // namespace arcb
