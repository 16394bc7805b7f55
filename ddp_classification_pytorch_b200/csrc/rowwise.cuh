// HBM-bound row kernels of the ArcFace head (one warp per row, 128-bit accesses, shuffle reduce).
//   l2norm_rows_kernel     F.normalize(x) / F.normalize(weight)      arc_main.py:159
//   target_prep_kernel     exact fp32 target cosine + margin branch   arc_main.py:160-165 at the label column
//   merge_stats_kernel     (max, sum-exp) partials -> one record per row
//   loss_kernel            nn.CrossEntropyLoss() mean                 arc_main.py:245,:327
//   l2norm_rows_bwd_kernel F.normalize backward for the embeddings    autograd behind arc_main.py:329
#pragma once

#include <cuda_bf16.h>
#include <math.h>
#include <stdint.h>

#include "ptx_sm100.cuh"

namespace arcb {

constexpr float kNormEps = 1e-12f;  // F.normalize default eps
// which forward kernel filled a workspace (written by target_prep_kernel, read by topk_rescore_kernel)
enum : int { kFwdStreaming = 0, kFwdResidentBmax = 1, kFwdResident = 2 };
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

// rows x D fp32 (pitch ldx) -> rows x Dp bf16 (zero padded to Dp) + inv_norm[rows].
// NV = float4 per lane kept in registers (D <= 128 * NV); NV == 0: generic two-pass path.
template <int NV>
__global__ void __launch_bounds__(256) l2norm_rows_kernel(const float* __restrict__ x, int64_t rows, int D, int64_t ldx,
                                                          __nv_bfloat16* __restrict__ out, int Dp,
                                                          float* __restrict__ inv_norm) {
  const int lane = threadIdx.x & 31;
  const int64_t warps_total = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows;
       row += warps_total) {
    const float* xr = x + row * ldx;
    __nv_bfloat16* orow = out + row * Dp;
    if constexpr (NV > 0) {
      float4 v[NV];
      float ss = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int e = (i * 32 + lane) * 4;
        v[i] = (e < D) ? ld_stream_f4(xr + e) : make_float4(0.f, 0.f, 0.f, 0.f);
        ss += v[i].x * v[i].x + v[i].y * v[i].y + v[i].z * v[i].z + v[i].w * v[i].w;
      }
      ss = warp_sum(ss);
      const float inv = 1.f / fmaxf(sqrtf(ss), kNormEps);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int e = (i * 32 + lane) * 4;
        if (e < Dp) {
          uint2 pk;
          pk.x = pack_bf16x2(v[i].x * inv, v[i].y * inv);
          pk.y = pack_bf16x2(v[i].z * inv, v[i].w * inv);
          *reinterpret_cast<uint2*>(orow + e) = pk;
        }
      }
      if (lane == 0) inv_norm[row] = inv;
    } else {
      float ss = 0.f;
      for (int e = lane; e < D; e += 32) {
        const float a = xr[e];
        ss += a * a;
      }
      ss = warp_sum(ss);
      const float inv = 1.f / fmaxf(sqrtf(ss), kNormEps);
      for (int e = lane; e < Dp; e += 32) orow[e] = __float2bfloat16_rn(e < D ? xr[e] * inv : 0.f);
      if (lane == 0) inv_norm[row] = inv;
    }
  }
}

// The same normalisation for `nrows` consecutive rows handled by ONE warp, RIF rows in flight so that a handful of
// warps per SM keeps enough bytes outstanding (used by the normaliser warps inside the forward GEMM kernel).
// Same arithmetic and summation order as l2norm_rows_kernel<NV>, hence bit-identical outputs.
template <int NV>
__device__ __forceinline__ void l2norm_rows_warp(const float* __restrict__ x, int64_t ldx, int64_t row0, int nrows, int D,
                                                 __nv_bfloat16* __restrict__ out, int Dp, float* __restrict__ inv_norm,
                                                 int lane) {
  constexpr int RIF = NV <= 4 ? 4 : (NV <= 8 ? 2 : 1);
  for (int r = 0; r < nrows; r += RIF) {
    float4 v[RIF][NV];
#pragma unroll
    for (int j = 0; j < RIF; ++j) {
      const float* xr = x + (row0 + r + j) * ldx;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int e = (i * 32 + lane) * 4;
        v[j][i] = (r + j < nrows && e < D) ? ld_stream_f4(xr + e) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
#pragma unroll
    for (int j = 0; j < RIF; ++j) {
      if (r + j >= nrows) break;
      float ss = 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i)
        ss += v[j][i].x * v[j][i].x + v[j][i].y * v[j][i].y + v[j][i].z * v[j][i].z + v[j][i].w * v[j][i].w;
      ss = warp_sum(ss);
      const float inv = 1.f / fmaxf(sqrtf(ss), kNormEps);
      __nv_bfloat16* orow = out + (row0 + r + j) * Dp;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int e = (i * 32 + lane) * 4;
        if (e < Dp) {
          uint2 pk;
          pk.x = pack_bf16x2(v[j][i].x * inv, v[j][i].y * inv);
          pk.y = pack_bf16x2(v[j][i].z * inv, v[j][i].w * inv);
          *reinterpret_cast<uint2*>(orow + e) = pk;
        }
      }
      if (lane == 0) inv_norm[row0 + r + j] = inv;
    }
  }
}

// One warp per sample: exact fp32 target cosine and everything derived from it.

//   lbl_local : label - class_offset when this shard owns the class, else -1
//   sub       : value substituted for the target cosine (phi, or the fallback branch)
//   zt        : s * sub (natural units) or -inf when the label is not in this shard
//   dphi      : d sub / d cos   (finite limit cos_m where the reference produces NaN)
//   cos_t     : the fp32 cosine itself
// The same warp first normalises its sample (F.normalize(input), arc_main.py:159: bf16 unit row, zero padded to Dp,
// and 1 / max(||x||, eps)) -- one launch for everything that is per sample.  `vec`: rows are 16-byte aligned and
// D % 4 == 0 (128-bit loads, same summation order as l2norm_rows_kernel).
__global__ void __launch_bounds__(256)
target_prep_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ w, int64_t ldw,
                   const int64_t* __restrict__ labels, __nv_bfloat16* __restrict__ xhat, int Dp,
                   float* __restrict__ inv_nx, int vec, int B, int Bp, int D, int64_t Cs, int64_t class_offset, float s,
                   float cos_m, float sin_m, float th, float mm, int easy, int* __restrict__ lbl_local,
                   float* __restrict__ sub, float* __restrict__ zt, float* __restrict__ dphi,
                   float* __restrict__ cos_t, unsigned int* __restrict__ zero, int n_zero, int* __restrict__ mode_out,
                   int mode) {
  // (also clears the per-tile counters of the forward kernel that normalises the weight itself, and records which
  //  forward schedule is about to fill this workspace)
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n_zero; k += gridDim.x * blockDim.x) zero[k] = 0u;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    mode_out[0] = mode;
    mode_out[1] = 0;  // block counter of merge_publish_kernel
  }
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= Bp) return;
  float inx = 0.f;
  if (i < B) {
    const float* xr = x + static_cast<int64_t>(i) * ldx;
    __nv_bfloat16* orow = xhat + static_cast<int64_t>(i) * Dp;
    float ss = 0.f;
    if (vec) {
      for (int e = lane * 4; e < D; e += 128) {
        const float4 v = *reinterpret_cast<const float4*>(xr + e);
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
      }
      ss = warp_sum(ss);
      inx = 1.f / fmaxf(sqrtf(ss), kNormEps);
      for (int e = lane * 4; e < Dp; e += 128) {
        const float4 v = (e < D) ? *reinterpret_cast<const float4*>(xr + e) : make_float4(0.f, 0.f, 0.f, 0.f);
        uint2 pk;
        pk.x = pack_bf16x2(v.x * inx, v.y * inx);
        pk.y = pack_bf16x2(v.z * inx, v.w * inx);
        *reinterpret_cast<uint2*>(orow + e) = pk;
      }
    } else {
      for (int e = lane; e < D; e += 32) ss += xr[e] * xr[e];
      ss = warp_sum(ss);
      inx = 1.f / fmaxf(sqrtf(ss), kNormEps);
      for (int e = lane; e < Dp; e += 32) orow[e] = __float2bfloat16_rn(e < D ? xr[e] * inx : 0.f);
    }
    if (lane == 0) inv_nx[i] = inx;
  }
  const int64_t loc = (i < B) ? labels[i] - class_offset : -1;
  if (loc < 0 || loc >= Cs) {  // padding rows, labels owned by another shard (or out of range)
    if (lane == 0) {
      lbl_local[i] = -1;
      sub[i] = 0.f;
      zt[i] = -INFINITY;
      dphi[i] = 0.f;
      cos_t[i] = 0.f;
    }
    return;
  }
  const float* xr = x + static_cast<int64_t>(i) * ldx;
  const float* wr = w + loc * ldw;
  float acc = 0.f, ss = 0.f;
  for (int e = lane; e < D; e += 32) {
    const float wv = wr[e];
    acc = fmaf(xr[e], wv, acc);
    ss = fmaf(wv, wv, ss);
  }
  acc = warp_sum(acc);
  ss = warp_sum(ss);
  if (lane == 0) {
    // the class row's norm is taken from the row itself: the weight may still be being normalised (inside the
    // forward GEMM kernel), and the target logit must not depend on which path produced w_hat
    const float inw = 1.f / fmaxf(sqrtf(ss), kNormEps);
    const float c = acc * inx * inw;
    const float om = 1.f - c * c;
    const float sine = sqrtf(fminf(fmaxf(om, 0.f), 1.f));
    const float phi = c * cos_m - sine * sin_m;
    const bool take = easy ? (c > 0.f) : (c > th);
    const float v = take ? phi : (easy ? c : c - mm);
    float d = 1.f;
    if (take) d = (om > 0.f && om < 1.f && sine > 0.f) ? (cos_m + sin_m * c / sine) : cos_m;
    lbl_local[i] = static_cast<int>(loc);
    sub[i] = v;
    zt[i] = s * v;
    dphi[i] = d;
    cos_t[i] = c;
  }
}

// Online merge of the (base-2 max, sum) partials part[k0 .. k1)[row]: one pass; eight partials are fetched before any
// of them is used so that the (strided, latency-bound) loads overlap.
__device__ __forceinline__ float2 merge_partials(const float2* __restrict__ part, int k0, int k1, int64_t ldp, int row) {
  float m = -INFINITY, l = 0.f;
  for (int k = k0; k < k1; k += 8) {
    float2 p[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      p[j] = (k + j < k1) ? part[static_cast<int64_t>(k + j) * ldp + row] : make_float2(-INFINITY, 0.f);
    float gm = p[0].x;
#pragma unroll
    for (int j = 1; j < 8; ++j) gm = fmaxf(gm, p[j].x);
    if (gm == -INFINITY) continue;  // only empty partials: nothing to add (and -inf - -inf would be NaN)
    const float nm = fmaxf(m, gm);
    float acc = l * exp2f(m - nm);  // m = -inf: l is 0 and exp2f(-inf) = 0
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += p[j].y * exp2f(p[j].x - nm);  // empty partial: 0 * 2^-inf = 0
    l = acc;
    m = nm;
  }
  return make_float2(m, l);
}

// part[n][ldp] -> out[slice][ldo] = merged (max, sum) over the slice's share of n.
// grid = (ceil(B / 256), slices): thread = one row, block y = one slice of the partial list, so the
// 8 B loads of a warp are coalesced and n / slices of them are in flight per thread.
__global__ void __launch_bounds__(256)
merge_stats_kernel(const float2* __restrict__ part, int n, int64_t ldp, int B, float2* __restrict__ out,
                   int64_t ldo) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= B) return;
  const int per = (n + gridDim.y - 1) / gridDim.y;
  const int k0 = blockIdx.y * per;
  out[blockIdx.y * ldo + row] = merge_partials(part, k0, min(n, k0 + per), ldp, row);
}

// Last merge level + the shard record: part[n][ldp] (all of it) and zt -> stats[B][4] = (m2, l, zt, 0).
__global__ void __launch_bounds__(128)
merge_pack_kernel(const float2* __restrict__ part, int n, int64_t ldp, const float* __restrict__ zt, int B,
                  float* __restrict__ stats) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= B) return;
  const float2 r = merge_partials(part, 0, n, ldp, row);
  *reinterpret_cast<float4*>(stats + static_cast<int64_t>(row) * 4) = make_float4(r.x, r.y, zt[row], 0.f);
}

// stats[P][B][4] = (m2, l, zt, -) per shard -> lse[B] (natural log units), loss = mean(lse - zt).
// Single block: deterministic reduction order.
__global__ void __launch_bounds__(1024)
loss_kernel(const float* __restrict__ stats, int P, int B, float* __restrict__ lse, float* __restrict__ zt_out,
            float* __restrict__ loss) {
  __shared__ double red[32];
  double acc = 0.0;
  for (int row = threadIdx.x; row < B; row += blockDim.x) {
    float m = -INFINITY, zt = -INFINITY;
    for (int p = 0; p < P; ++p) {
      const float4 r = *reinterpret_cast<const float4*>(stats + (static_cast<int64_t>(p) * B + row) * 4);
      m = fmaxf(m, r.x);
      zt = fmaxf(zt, r.z);
    }
    float l = 0.f;
    for (int p = 0; p < P; ++p) {
      const float4 r = *reinterpret_cast<const float4*>(stats + (static_cast<int64_t>(p) * B + row) * 4);
      l += r.y * exp2f(r.x - m);
    }
    const float v = (m + log2f(l)) * kLn2;
    lse[row] = v;
    if (zt_out) zt_out[row] = zt;
    acc += static_cast<double>(v) - static_cast<double>(zt);
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    double v = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.0;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) *loss = static_cast<float>(v / B);
  }
}

// ---------------------------------------------------------------------------------------------
// Class-sharded forward exchange fused with its producer (stage 2).  Instead of "merge kernel -> NCCL
// all_gather -> loss kernel", the merge kernel itself stores this rank's [B, 4] record into slot `rank` of EVERY rank's
// gather buffer (plain stores through NVLink peer mappings: symmetric memory) and then raises flag[rank] on every rank to
// the step's epoch (release, system scope); the loss kernel of each rank waits until all P flags have reached the epoch
// (acquire, system scope) and reads its local buffer.  No collective launch sits on the forward's critical path.
// Flags only grow (epoch = step number), the gather buffer is double buffered by the caller (slot = epoch & 1): a rank
// can run at most one step ahead of its peers, because its loss kernel needs every peer's flag of the same step.
// ---------------------------------------------------------------------------------------------
struct PeerTable {
  float* buf[16];          // [p] base of rank p's gather buffer slot for this step: [P, B, 4] fp32
  unsigned int* flag[16];  // [p] base of rank p's flag array: [P] uint32
};

__device__ __forceinline__ void st_release_sys_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Last merge level + publication: one thread per row (like merge_pack_kernel); the record goes to the local `stats`
// (kept for aux / tests) and to all P ranks; the block that finishes last raises the flags.  `counter`: zero before the
// launch (target_prep_kernel clears it).
__global__ void __launch_bounds__(128)
merge_publish_kernel(const float2* __restrict__ part, int n, int64_t ldp, const float* __restrict__ zt, int B,
                     float* __restrict__ stats, const PeerTable peers, int rank, int P, unsigned int epoch,
                     unsigned int* __restrict__ counter) {
  __shared__ unsigned int is_last;
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row < B) {
    const float2 r = merge_partials(part, 0, n, ldp, row);
    const float4 rec = make_float4(r.x, r.y, zt[row], 0.f);
    *reinterpret_cast<float4*>(stats + static_cast<int64_t>(row) * 4) = rec;
    for (int p = 0; p < P; ++p)
      *reinterpret_cast<float4*>(peers.buf[p] + (static_cast<int64_t>(rank) * B + row) * 4) = rec;
  }
  __threadfence_system();  // every thread's peer stores are performed system-wide ...
  __syncthreads();         // ... before this block is counted as done
  if (threadIdx.x == 0) is_last = (atomicAdd(counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (is_last && threadIdx.x < P) {
    __threadfence_system();
    st_release_sys_u32(peers.flag[threadIdx.x] + rank, epoch);
  }
}

// loss_kernel that first waits for the P records of this step (see above).  `flags` is this rank's own flag array.
__global__ void __launch_bounds__(1024)
loss_wait_kernel(const float* __restrict__ stats, const unsigned int* __restrict__ flags, unsigned int epoch, int P, int B,
                 float* __restrict__ lse, float* __restrict__ zt_out, float* __restrict__ loss) {
  __shared__ double red[32];
  if (threadIdx.x < P) {
    const long long t0 = clock64();
    unsigned int spins = 0;
    // flags only grow; a peer that is a step ahead has a larger value (its data for THIS step sits in the other slot)
    while (static_cast<int>(ld_acquire_sys_u32(flags + threadIdx.x) - epoch) < 0) {
      __nanosleep(200);
#if ARCB_WATCHDOG_CYCLES > 0
      if ((++spins & 0xffu) == 0 && clock64() - t0 > ARCB_WATCHDOG_CYCLES) watchdog_abort(0x9e0u + threadIdx.x, epoch);
#endif
    }
  }
  __syncthreads();
  double acc = 0.0;
  for (int row = threadIdx.x; row < B; row += blockDim.x) {
    float m = -INFINITY, zt = -INFINITY;
    for (int p = 0; p < P; ++p) {
      const float4 r = __ldcv(reinterpret_cast<const float4*>(stats + (static_cast<int64_t>(p) * B + row) * 4));
      m = fmaxf(m, r.x);
      zt = fmaxf(zt, r.z);
    }
    float l = 0.f;
    for (int p = 0; p < P; ++p) {
      const float4 r = __ldcv(reinterpret_cast<const float4*>(stats + (static_cast<int64_t>(p) * B + row) * 4));
      l += r.y * exp2f(r.x - m);
    }
    const float v = (m + log2f(l)) * kLn2;
    lse[row] = v;
    if (zt_out) zt_out[row] = zt;
    acc += static_cast<double>(v) - static_cast<double>(zt);
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    double v = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.0;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) *loss = static_cast<float>(v / B);
  }
}

// ---------------------------------------------------------------------------------------------
// Class-sharded backward exchange fused with its consumer (stage 2): the sum of the shards' dX_hat partials
// and F.normalize's backward for the embeddings in ONE kernel over peer memory, instead of "NCCL all_reduce ->
// l2norm_rows_bwd_kernel".  Every rank's partial [B, Dp] lives in symmetric memory.  Rank r owns the rows
// [r * rows_per, (r + 1) * rows_per) of the batch:
//   announce   raise ready[rank] on every rank (my partial is complete: the backward kernel ran before this one)
//   reduce     wait for all ready flags; for my rows: sum the P partials (peer loads over NVLink), apply
//              dx = (dxh - xh (xh . dxh)) / ||x|| and store the finished rows in my slice buffer (symmetric, local)
//   publish    when all my blocks are done: raise done[rank] on every rank
//   collect    wait for all done flags; copy every rank's slice into the local output dx [B, D]  (all-gather by pull)
// own_only (data-parallel trainer, reduce_scatter semantics): the reduce step writes my rows straight to `dx` and the
// kernel ends there.  Flags only grow (epoch = step number); partials and slices are double buffered by the caller.
// ---------------------------------------------------------------------------------------------
struct PeerReduceArgs {
  const float* partial[16];   // [p] rank p's dX_hat partial of this step, [B, Dp] fp32
  const float* slice[16];     // [p] rank p's finished rows of this step, [rows_per, D] fp32
  unsigned int* ready[16];    // [p] rank p's ready flags [P]
  unsigned int* done[16];     // [p] rank p's done flags [P]
  float* my_slice;            // = slice[rank], writable
  unsigned int* counter;      // local, zero between launches: blocks that have finished their rows
  int rank, P;
  unsigned int epoch;
  int B, D, Dp, rows_per;     // rows_per = ceil(B / P)
  int own_only;
};

__device__ __forceinline__ void wait_flags_sys(const unsigned int* flags, int P, unsigned int epoch, unsigned int tag) {
  if (threadIdx.x < static_cast<unsigned>(P)) {
    const long long t0 = clock64();
    unsigned int spins = 0;
    while (static_cast<int>(ld_acquire_sys_u32(flags + threadIdx.x) - epoch) < 0) {
      __nanosleep(100);
#if ARCB_WATCHDOG_CYCLES > 0
      if ((++spins & 0xffu) == 0 && clock64() - t0 > ARCB_WATCHDOG_CYCLES) watchdog_abort(tag + threadIdx.x, epoch);
#endif
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256)
dx_peer_reduce_kernel(const float* __restrict__ x, int64_t ldx, float* __restrict__ dx, int64_t ldo, const PeerReduceArgs a) {
  const int lane = threadIdx.x & 31;
  // announce: the backward kernel that wrote my partial completed before this launch; make it visible system-wide
  if (blockIdx.x == 0 && threadIdx.x < a.P) {
    __threadfence_system();
    st_release_sys_u32(a.ready[threadIdx.x] + a.rank, a.epoch);
  }
  wait_flags_sys(a.ready[a.rank], a.P, a.epoch, 0x9f00u);
  // reduce + normalise-backward for my rows: one row per block at a time, one float4 (x P peer loads, all in flight at
  // once: a single NVLink round trip per row) per thread; D <= 2048, D % 4 == 0 (checked by the launcher)
  __shared__ float red_ss[8], red_dot[8];
  const int r0 = a.rank * a.rows_per;
  const int r1 = min(a.B, r0 + a.rows_per);
  const int n4 = a.D >> 2;
  for (int row = r0 + blockIdx.x; row < r1; row += gridDim.x) {
    const float4* xr = reinterpret_cast<const float4*>(x + static_cast<int64_t>(row) * ldx);
    float4* out = reinterpret_cast<float4*>(a.own_only ? dx + static_cast<int64_t>(row - r0) * ldo
                                                        : a.my_slice + static_cast<int64_t>(row - r0) * a.D);
    float4 sum[2], xv[2];
    float ss = 0.f, dot = 0.f;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int idx = threadIdx.x + k * 256;
      sum[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      xv[k] = sum[k];
      if (idx < n4) {
        // all P loads are issued before the first one is used: one NVLink round trip, not P of them
        float4 v[16];
#pragma unroll
        for (int p = 0; p < 16; ++p)
          if (p < a.P) v[p] = __ldcv(reinterpret_cast<const float4*>(a.partial[p] + static_cast<int64_t>(row) * a.Dp) + idx);
#pragma unroll
        for (int p = 0; p < 16; ++p)
          if (p < a.P) {
            sum[k].x += v[p].x; sum[k].y += v[p].y; sum[k].z += v[p].z; sum[k].w += v[p].w;
          }
        xv[k] = xr[idx];
        ss += xv[k].x * xv[k].x + xv[k].y * xv[k].y + xv[k].z * xv[k].z + xv[k].w * xv[k].w;
        dot += xv[k].x * sum[k].x + xv[k].y * sum[k].y + xv[k].z * sum[k].z + xv[k].w * sum[k].w;
      }
    }
    ss = warp_sum(ss);
    dot = warp_sum(dot);
    __syncthreads();  // the previous row's readers of red_* are done
    if (lane == 0) {
      red_ss[threadIdx.x >> 5] = ss;
      red_dot[threadIdx.x >> 5] = dot;
    }
    __syncthreads();
    ss = 0.f;
    dot = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      ss += red_ss[i];
      dot += red_dot[i];
    }
    const float nrm = sqrtf(ss);
    const bool ok = nrm > kNormEps;  // same arithmetic as l2norm_rows_bwd_kernel
    const float inv = ok ? 1.f / nrm : 1.f / kNormEps;
    const float proj = ok ? dot * inv * inv : 0.f;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int idx = threadIdx.x + k * 256;
      if (idx < n4)
        out[idx] = make_float4(inv * (sum[k].x - xv[k].x * proj), inv * (sum[k].y - xv[k].y * proj),
                               inv * (sum[k].z - xv[k].z * proj), inv * (sum[k].w - xv[k].w * proj));
    }
  }
  if (a.own_only) return;
  // publish: all of this rank's blocks have finished their rows -> done[rank] on every rank
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(a.counter, 1u);
    if (prev == gridDim.x - 1) {
      *a.counter = 0u;  // for the next launch (stream order: it starts after this kernel has ended)
      __threadfence_system();
      for (int p = 0; p < a.P; ++p) st_release_sys_u32(a.done[p] + a.rank, a.epoch);
    }
  }
  // collect: every rank's finished rows -> local dx
  wait_flags_sys(a.done[a.rank], a.P, a.epoch, 0x9f40u);
  for (int p = 0; p < a.P; ++p) {
    const int p0 = p * a.rows_per;
    const int rows = min(a.B, p0 + a.rows_per) - p0;
    if (rows <= 0) continue;
    const float4* src = reinterpret_cast<const float4*>(a.slice[p]);
    const int total = rows * n4;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
      const int r = i / n4, c = i - r * n4;
      const float4 v = __ldcv(src + i);
      *reinterpret_cast<float4*>(dx + static_cast<int64_t>(p0 + r) * ldo + c * 4) = v;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Exact top-3 / argmax of the scaled logits (consumers at arc_main.py:338-345, :377-388) without
// the logits: the forward GEMM left, per row, the bf16-accurate maximum of every 128-class block
// (`part`).  |cos_bf16 - cos_fp32| <= 2^-8 (two operand roundings, |x_hat||w_hat| = 1), so every
// class that can be among the true top 3 lies in a block whose maximum is >= (third largest block
// maximum) - 2 * 2^-8 * s2.  Those few blocks are re-scored in fp32 from the ORIGINAL operands,
// with the margin value at the label column, and the top 3 of the re-scored logits are returned.
// One block per row; val = logits in natural units, idx = global class ids.
// ---------------------------------------------------------------------------------------------
struct Top3 {
  float v[3];
  int i[3];
};
__device__ __forceinline__ void top3_insert(Top3& t, float v, int i) {
  // order: larger value first; equal values: smaller index first (what a stable descending sort gives)
  auto better = [](float va, int ia, float vb, int ib) { return va > vb || (va == vb && ia < ib); };
  if (!better(v, i, t.v[2], t.i[2])) return;
  if (better(v, i, t.v[0], t.i[0])) {
    t.v[2] = t.v[1]; t.i[2] = t.i[1]; t.v[1] = t.v[0]; t.i[1] = t.i[0]; t.v[0] = v; t.i[0] = i;
  } else if (better(v, i, t.v[1], t.i[1])) {
    t.v[2] = t.v[1]; t.i[2] = t.i[1]; t.v[1] = v; t.i[1] = i;
  } else {
    t.v[2] = v; t.i[2] = i;
  }
}

constexpr int kTopkThreads = 256;

__global__ void __launch_bounds__(kTopkThreads)
topk_rescore_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ w, int64_t ldw,
                    const float* __restrict__ inv_nx, const float* __restrict__ inv_nw,
                    const int* __restrict__ lbl_local, const float* __restrict__ sub, const float2* __restrict__ part,
                    const float* __restrict__ bmax, const int* __restrict__ fwd_mode, int nparts, int64_t ldp, int B,
                    int D, int64_t Cs, int64_t class_offset, float s, float* __restrict__ top_val,
                    int64_t* __restrict__ top_idx) {
  // where the forward left the block maxima: the per-tile partials themselves (streaming schedule) or the separate
  // array of the resident schedule; a resident forward that was not asked to keep them cannot serve this kernel
  const int mode = *fwd_mode & 0xff;  // (bit 8: the forward also kept its probabilities)
  if (mode == kFwdResident) __trap();  // arcb_shard_top3 after a forward without ARCB_FWD_BLOCK_MAXIMA: fail loudly
  const float* bm = (mode == kFwdStreaming) ? reinterpret_cast<const float*>(part) : bmax;
  const int bm_step = (mode == kFwdStreaming) ? 2 : 1;
  extern __shared__ float xs[];  // D floats: the normalised embedding of this row
  __shared__ Top3 red[kTopkThreads / 32];
  __shared__ float thr_s;
  const int row = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float inx = inv_nx[row];
  for (int e = tid; e < D; e += blockDim.x) xs[e] = x[static_cast<int64_t>(row) * ldx + e] * inx;

  // (1) third largest block maximum (base-2 units u = s * log2e * cos)
  Top3 mine;
  for (int k = 0; k < 3; ++k) { mine.v[k] = -INFINITY; mine.i[k] = 0x7fffffff; }
  for (int p = tid; p < nparts; p += blockDim.x) top3_insert(mine, bm[(p * ldp + row) * bm_step], p);
  for (int o = 16; o > 0; o >>= 1) {
    Top3 other;
    for (int k = 0; k < 3; ++k) {
      other.v[k] = __shfl_xor_sync(0xffffffffu, mine.v[k], o);
      other.i[k] = __shfl_xor_sync(0xffffffffu, mine.i[k], o);
    }
    for (int k = 0; k < 3; ++k) top3_insert(mine, other.v[k], other.i[k]);
  }
  if (lane == 0) red[warp] = mine;
  __syncthreads();
  if (tid == 0) {
    Top3 all = red[0];
    for (int wv = 1; wv < kTopkThreads / 32; ++wv)
      for (int k = 0; k < 3; ++k) top3_insert(all, red[wv].v[k], red[wv].i[k]);
    // a row with fewer than 3 blocks: every block is a candidate
    const float third = (nparts >= 3) ? all.v[2] : -INFINITY;
    thr_s = third - 2.f * 0.00390625f * s * kLog2e;
  }
  __syncthreads();
  const float thr = thr_s;

  // (2) re-score every candidate block in fp32: one warp per class, lanes split the feature dimension
  const int lbl = lbl_local[row];
  const float sub_v = sub[row];
  Top3 best;
  for (int k = 0; k < 3; ++k) { best.v[k] = -INFINITY; best.i[k] = 0x7fffffff; }
  for (int p = 0; p < nparts; ++p) {
    if (!(bm[(p * ldp + row) * bm_step] >= thr)) continue;  // block-uniform
    const int c_begin = p * 128;
    const int c_end = min(static_cast<int64_t>(c_begin + 128), Cs);
    for (int c = c_begin + warp; c < c_end; c += kTopkThreads / 32) {
      const float* wr = w + static_cast<int64_t>(c) * ldw;
      float acc = 0.f;
      for (int e = lane; e < D; e += 32) acc = fmaf(xs[e], wr[e], acc);
      acc = warp_sum(acc);
      if (lane == 0) {
        const float cosv = (c == lbl) ? sub_v : acc * inv_nw[c];
        top3_insert(best, s * cosv, c);
      }
    }
  }
  // (3) merge the warps' candidates
  __syncthreads();
  if (lane == 0) red[warp] = best;
  __syncthreads();
  if (tid == 0) {
    Top3 all = red[0];
    for (int wv = 1; wv < kTopkThreads / 32; ++wv)
      for (int k = 0; k < 3; ++k) top3_insert(all, red[wv].v[k], red[wv].i[k]);
    for (int k = 0; k < 3; ++k) {
      top_val[row * 3 + k] = all.v[k];
      top_idx[row * 3 + k] = (all.v[k] == -INFINITY) ? -1 : (class_offset + all.i[k]);
    }
  }
}

// dx = (dxh - xh * (xh . dxh)) / ||x||   with xh = x / ||x||;  clamped rows (||x|| <= eps): dx = dxh / eps.
__global__ void __launch_bounds__(256)
l2norm_rows_bwd_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ dxh, int64_t ldg, int rows,
                       int D, float* __restrict__ dx, int64_t ldo) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float* xr = x + static_cast<int64_t>(row) * ldx;
  const float* gr = dxh + static_cast<int64_t>(row) * ldg;
  float ss = 0.f, dot = 0.f;
  for (int e = lane; e < D; e += 32) {
    const float a = xr[e];
    ss = fmaf(a, a, ss);
    dot = fmaf(a, gr[e], dot);
  }
  ss = warp_sum(ss);
  dot = warp_sum(dot);
  const float nrm = sqrtf(ss);
  float* o = dx + static_cast<int64_t>(row) * ldo;
  if (nrm > kNormEps) {
    const float inv = 1.f / nrm;
    const float proj = dot * inv * inv;  // (xh . dxh) / ||x|| expressed on raw x
    for (int e = lane; e < D; e += 32) o[e] = inv * (gr[e] - xr[e] * proj);
  } else {
    for (int e = lane; e < D; e += 32) o[e] = gr[e] * (1.f / kNormEps);
  }
}

}  // namespace arcb
