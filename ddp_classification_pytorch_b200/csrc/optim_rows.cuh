// Optimizer step for a weight shard, fused with the row normalisation the NEXT forward needs.
//   torch.optim.SGD(lr, momentum=0.9, weight_decay=5e-4)   arc_main.py:252-253
//   torch.optim.Adam(lr)                                    arc_main.py:249-250
//   F.normalize(self.weight) of the following step          arc_main.py:159
// One warp owns one class row: it reads the row of w / dw / optimizer state once, applies the update
// rule in registers, writes the new w and state, reduces the new row norm with shuffles and emits
// the bf16 unit row + 1/norm straight into the forward workspace.  HBM-bound:
//   SGD-momentum 22 B per weight element (r w,dw,buf 12 + w w,buf 8 + w w_hat 2), Adam 30 B.
#pragma once

#include <cuda_bf16.h>
#include <math.h>
#include <stdint.h>

#include "rowwise.cuh"

namespace arcb {

// Plain (coherent) streaming load: w and the state rows are rewritten by the same warp later on.
__device__ __forceinline__ float4 ld_na_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

// torch.optim.SGD, single-tensor formulation (torch/optim/sgd.py _single_tensor_sgd):
//   g = dw + wd * w;  buf = g (first step) | momentum * buf + (1 - dampening) * g;
//   g = g + momentum * buf (nesterov) | buf;  w = w - lr * g.
struct SgdRule {
  float lr, momentum, one_minus_damp, wd;
  int nesterov, first, has_buf;
  static constexpr int kStates = 1;
  __device__ __forceinline__ void apply(float& w, float g, float& buf, float& /*unused*/) const {
    g = fmaf(wd, w, g);
    if (has_buf) {
      buf = first ? g : fmaf(one_minus_damp, g, momentum * buf);
      g = nesterov ? fmaf(momentum, buf, g) : buf;
    }
    w = fmaf(-lr, g, w);
  }
};

// torch.optim.Adam, single-tensor formulation (torch/optim/adam.py _single_tensor_adam, amsgrad off):
//   g = dw + wd * w;  m = lerp(m, g, 1 - beta1);  v = beta2 * v + (1 - beta2) * g * g;
//   w = w - (lr / (1 - beta1^t)) * m / (sqrt(v) / sqrt(1 - beta2^t) + eps).
struct AdamRule {
  float one_minus_b1, b2, one_minus_b2, eps, wd, step_size, bc2_sqrt;
  static constexpr int kStates = 2;
  __device__ __forceinline__ void apply(float& w, float g, float& m, float& v) const {
    g = fmaf(wd, w, g);
    m = fmaf(one_minus_b1, g - m, m);
    v = fmaf(one_minus_b2 * g, g, b2 * v);
    const float denom = sqrtf(v) / bc2_sqrt + eps;
    w = fmaf(-step_size, m / denom, w);
  }
};

// NV = float4 per lane held in registers (D <= 128 * NV, D % 4 == 0, 16-byte aligned rows);
// NV == 0: generic scalar path that re-reads the freshly written row for the normalised copy.
template <int NV, class Rule>
__global__ void __launch_bounds__(256)
optim_rows_kernel(float* __restrict__ w, int64_t ldw, const float* __restrict__ g, int64_t ldg,
                  float* __restrict__ s0, float* __restrict__ s1, int64_t lds, int64_t rows, int D, Rule rule,
                  __nv_bfloat16* __restrict__ what, int Dp, float* __restrict__ inv_norm) {
  const int lane = threadIdx.x & 31;
  const int64_t warps_total = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows;
       row += warps_total) {
    float* wr = w + row * ldw;
    const float* gr = g + row * ldg;
    float* ar = s0 ? s0 + row * lds : nullptr;
    float* br = (Rule::kStates > 1 && s1) ? s1 + row * lds : nullptr;
    float ss = 0.f;
    if constexpr (NV > 0) {
      float4 wv[NV];
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int e = (i * 32 + lane) * 4;
        wv[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (e < D) {
          wv[i] = ld_na_f4(wr + e);
          const float4 gv = ld_stream_f4(gr + e);
          float4 a = ar ? ld_na_f4(ar + e) : make_float4(0.f, 0.f, 0.f, 0.f);
          float4 b = br ? ld_na_f4(br + e) : make_float4(0.f, 0.f, 0.f, 0.f);
          rule.apply(wv[i].x, gv.x, a.x, b.x);
          rule.apply(wv[i].y, gv.y, a.y, b.y);
          rule.apply(wv[i].z, gv.z, a.z, b.z);
          rule.apply(wv[i].w, gv.w, a.w, b.w);
          *reinterpret_cast<float4*>(wr + e) = wv[i];
          if (ar) *reinterpret_cast<float4*>(ar + e) = a;
          if (br) *reinterpret_cast<float4*>(br + e) = b;
          ss += wv[i].x * wv[i].x + wv[i].y * wv[i].y + wv[i].z * wv[i].z + wv[i].w * wv[i].w;
        }
      }
      if (what) {  // same arithmetic as l2norm_rows_kernel, so a later stand-alone normalise agrees bit for bit
        ss = warp_sum(ss);
        const float inv = 1.f / fmaxf(sqrtf(ss), kNormEps);
        __nv_bfloat16* orow = what + row * Dp;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int e = (i * 32 + lane) * 4;
          if (e < Dp) {
            uint2 pk;
            pk.x = pack_bf16x2(wv[i].x * inv, wv[i].y * inv);
            pk.y = pack_bf16x2(wv[i].z * inv, wv[i].w * inv);
            *reinterpret_cast<uint2*>(orow + e) = pk;
          }
        }
        if (lane == 0) inv_norm[row] = inv;
      }
    } else {
      for (int e = lane; e < D; e += 32) {
        float wv = wr[e];
        float a = ar ? ar[e] : 0.f, b = br ? br[e] : 0.f;
        rule.apply(wv, gr[e], a, b);
        wr[e] = wv;
        if (ar) ar[e] = a;
        if (br) br[e] = b;
        ss += wv * wv;
      }
      if (what) {
        ss = warp_sum(ss);
        const float inv = 1.f / fmaxf(sqrtf(ss), kNormEps);
        __nv_bfloat16* orow = what + row * Dp;
        // each lane re-reads only the elements it wrote itself
        for (int e = lane; e < Dp; e += 32) orow[e] = __float2bfloat16_rn(e < D ? wr[e] * inv : 0.f);
        if (lane == 0) inv_norm[row] = inv;
      }
    }
  }
}

}  // namespace arcb
