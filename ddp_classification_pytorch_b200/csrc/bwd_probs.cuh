// Backward of the ArcFace head from STORED softmax numerators (the default for large shards, include/arcface_b200.h: ARCB_BWD_RECOMPUTE; autograd behind arc_main.py:329).
//
// The forward (EpiFwdStatsP) left P'[j, i] = bf16(2^(s2 (cos_ij - 1))) for the whole shard, CLASS-major: P'^T [Cs, Bp], a
// class row holds the whole batch.  With c_i = gscale * 2^(s2 - lse2_i) the gradient of the loss w.r.t. the cosines is
// g_ij = c_i P'_ij (the label entry is patched to g_target / c_i by probs_setup_kernel), so the backward needs NO
// recompute GEMM: it is two plain GEMMs on the generic cores
//     dX_hat[B, D]  = diag(c) * P' * w_hat        (EpiRedAdd, split along the classes: every CTA keeps ONE output tile
//                                                  in TMEM for its whole slice and reduce-adds it once)
//     dW_hat[Cs, D] = P'^T * (c (.) x_hat)        (EpiDWT; A = P'^T is K-major: a tile's whole K extent is one contiguous
//                                                  256 KB region of HBM)
// 6 B C D executed flops for 6 B C D algorithmic ones, against 8 B C D with the recompute.
//
// What the recompute epilogue used to deliver on the side is the projection t_j = w_hat_j . dW_hat_j that F.normalize's
// backward subtracts (dW_j = (dW_hat_j - t_j w_hat_j) / ||w_j||).  It needs the WHOLE feature row of the accumulator, but
// a [128 x 512] fp32 accumulator is all of TMEM and would leave no second buffer to overlap the epilogue with the next
// tile's MMAs.  EpiDWT (epilogues.cuh) therefore works on [128 x 256] slices and exchanges the per-row partial sums of
// the two slices through global memory.  This file holds the set-up kernel.
#pragma once

#include "epilogues.cuh"

namespace arcb {

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
  float* zero_t;           // [n_t] t_j accumulators of EpiDWT
  long long n_t;
  unsigned int* zero_c;    // [n_c] class-tile arrival counters of EpiDWT
  long long n_c;
  const int* fwd_mode;     // the workspace's forward-schedule word: bit 8 = this forward stored P'^T
};

__global__ void __launch_bounds__(256) probs_setup_kernel(const ProbsSetup su) {
  // a forward run with other flags left no probabilities in this workspace: abort instead of using stale memory
  if (!(*su.fwd_mode & 0x100)) watchdog_abort(0xbad0u, static_cast<uint32_t>(*su.fwd_mode));
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
  for (long long i = tid; i < su.n_t; i += nthr) su.zero_t[i] = 0.f;
  for (long long i = tid; i < su.n_c; i += nthr) su.zero_c[i] = 0u;
}

}  // namespace arcb
