// Kernels of the conditioning front half that are not GEMMs:
//   reflect_pad_rows   ReflectionPad1d over the frame axis (+ LeakyReLU) into a padded channels-last operand tensor
//                      (Conv1dResnet, reference serenade.py:289-303, :361-376)
//   im2col_k3s2        patches of Conv2d(k3, s2, p1) on an NHWC tensor -> [M][K] GEMM operand
//                      (ReferenceEncoder conv stack, modules/gst/style_encoder.py:136-153)
//   gru_stl_kernel     GRU over the conv features, last hidden state -> multi-head attention over the style
//                      tokens -> style embedding (style_encoder.py:184-190, :236-252; gst/attention.py:111-147)
// The convolutions themselves run on the tcgen05 GEMM core (gemm.cuh) with BatchNorm folded into weight and bias.
#pragma once
#include <cuda_bf16.h>
#include <cstdint>

#include "rowops.cuh"

namespace srn {

__device__ __forceinline__ void load8_any(const float* p, float (&v)[8]) { load8(p, v); }
__device__ __forceinline__ void load8_any(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}

// dst[b][i][c] = act(src[b][reflect(i - pad)][c]),  i in [0, L + 2 pad), reflect(j) = |j| for j < 0, 2(L-1) - j for j >= L.
// slope 1 = identity, 0.2 = LeakyReLU.  C % 8 == 0.  grid = (ceil((L + 2 pad) * C / 8 / 256), B).
template <typename IT, typename OT>
__global__ void __launch_bounds__(256) reflect_pad_rows(const IT* __restrict__ src, long long src_bstride, int L, int C,
                                                        int pad, float slope, OT* __restrict__ dst,
                                                        long long dst_bstride) {
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.y;
  const int c8 = C / 8;
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)(L + 2 * pad) * c8) return;
  const int i = (int)(idx / c8), c = (int)(idx % c8) * 8;
  int j = i - pad;
  j = j < 0 ? -j : (j >= L ? 2 * (L - 1) - j : j);
  float v[8];
  load8_any(src + b * src_bstride + (long long)j * C + c, v);
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = v[k] > 0.f ? v[k] : v[k] * slope;
  store8_op(dst + b * dst_bstride + (long long)i * C + c, v);
}

// out[m][(kh * 3 + kw) * Cin + ci] = in[b][2 ho - 1 + kh][2 wo - 1 + kw][ci] (0 outside), m = (b * Ho + ho) * Wo + wo,
// columns [9 Cin, Kp) zero.  One block per output position; Kp % 8 == 0.
template <typename IT, typename OT>
__global__ void __launch_bounds__(128) im2col_k3s2(const IT* __restrict__ in, int H, int W, int Cin, int Ho, int Wo,
                                                   int Kp, OT* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  const long long m = blockIdx.x;
  const int wo = (int)(m % Wo);
  const int ho = (int)((m / Wo) % Ho);
  const int b = (int)(m / ((long long)Wo * Ho));
  OT* o = out + m * Kp;
  if (Cin % 8 == 0) {
    const int c8 = Cin / 8;
    for (int g = threadIdx.x; g < Kp / 8; g += blockDim.x) {
      float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      const int tap = g / c8;
      if (tap < 9) {
        const int hi = 2 * ho - 1 + tap / 3, wi = 2 * wo - 1 + tap % 3;
        if (hi >= 0 && hi < H && wi >= 0 && wi < W)
          load8_any(in + (((long long)b * H + hi) * W + wi) * Cin + (g % c8) * 8, v);
      }
      store8_op(o + g * 8, v);
    }
  } else {
    for (int k = threadIdx.x; k < Kp; k += blockDim.x) {
      float v = 0.f;
      const int tap = k / Cin;
      if (tap < 9) {
        const int hi = 2 * ho - 1 + tap / 3, wi = 2 * wo - 1 + tap % 3;
        if (hi >= 0 && hi < H && wi >= 0 && wi < W) v = to_f(in[(((long long)b * H + hi) * W + wi) * Cin + k % Cin]);
      }
      from_f_op(o + k, v);
    }
  }
}

// One block of 384 threads (= 3 * gru_units) per batch element.
//   h_t = GRU(x_t, h_{t-1}), gate order r, z, n (torch.nn.GRU);  x_t = hs[b][t][0..G)
//   q = Wq h_T + bq;  per head (4 x 64): p = softmax(q_h . K_h^T / 8) over ntok tokens;  ctx = p V_h;  out = Wo ctx + bo
// kmat / vmat = linear_k / linear_v applied to tanh(gst_embs) (input independent: folded on the host).
constexpr int kGruH = 128;
constexpr int kGruChunk = 8;
template <typename OT>
__global__ void __launch_bounds__(384) gru_stl_kernel(const OT* __restrict__ hs, int Lp, int G,
                                                      const float* __restrict__ w_ihT /*[G][384]*/,
                                                      const float* __restrict__ b_ih,
                                                      const float* __restrict__ w_hhT /*[128][384]*/,
                                                      const float* __restrict__ b_hh, const float* __restrict__ wq,
                                                      const float* __restrict__ bq, const float* __restrict__ kmat,
                                                      const float* __restrict__ vmat, int ntok,
                                                      const float* __restrict__ wo, const float* __restrict__ bo,
                                                      float* __restrict__ out /*[B][256]*/) {
  extern __shared__ float sm[];
  float* xs = sm;                          // [kGruChunk][G]
  float* hsm = xs + kGruChunk * G;         // [128]
  float* gis = hsm + kGruH;                // [384]
  float* ghs = gis + 3 * kGruH;            // [384]
  float* qs = ghs + 3 * kGruH;             // [256]
  float* ps = qs + 256;                    // [4][64] (ntok <= 64)
  float* cs = ps + 256;                    // [256]
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.x, j = threadIdx.x;
  if (j < kGruH) hsm[j] = 0.f;
  const float bi = b_ih[j], bh = b_hh[j];
  for (int t0 = 0; t0 < Lp; t0 += kGruChunk) {
    const int nt = min(kGruChunk, Lp - t0);
    __syncthreads();
    for (int i = j; i < nt * G; i += 384) xs[i] = to_f(hs[((long long)b * Lp + t0) * G + i]);
    __syncthreads();
    float gi[kGruChunk];
#pragma unroll
    for (int t = 0; t < kGruChunk; ++t) gi[t] = bi;
    for (int k = 0; k < G; ++k) {
      const float w = w_ihT[(long long)k * 384 + j];
#pragma unroll
      for (int t = 0; t < kGruChunk; ++t) gi[t] = fmaf(w, xs[t * G + k], gi[t]);
    }
#pragma unroll
    for (int t = 0; t < kGruChunk; ++t) {
      if (t < nt) {
        float gh = bh;
        for (int k = 0; k < kGruH; ++k) gh = fmaf(w_hhT[k * 384 + j], hsm[k], gh);
        gis[j] = gi[t];
        ghs[j] = gh;
        __syncthreads();
        if (j < kGruH) {
          const float r = 1.f / (1.f + expf(-(gis[j] + ghs[j])));
          const float z = 1.f / (1.f + expf(-(gis[kGruH + j] + ghs[kGruH + j])));
          const float n = tanhf(gis[2 * kGruH + j] + r * ghs[2 * kGruH + j]);
          hsm[j] = (1.f - z) * n + z * hsm[j];
        }
        __syncthreads();
      }
    }
  }
  // style token layer
  if (j < 256) {
    float a = bq[j];
    for (int k = 0; k < kGruH; ++k) a = fmaf(wq[j * kGruH + k], hsm[k], a);
    qs[j] = a;
  }
  __syncthreads();
  if (j < 256) {
    const int hh = j >> 6, n = j & 63;
    float s = -INFINITY;
    if (n < ntok) {
      s = 0.f;
      for (int d = 0; d < 64; ++d) s = fmaf(qs[hh * 64 + d], kmat[n * 256 + hh * 64 + d], s);
      s *= 0.125f;       // 1 / sqrt(64)
    }
    ps[j] = s;
  }
  __syncthreads();
  if (j < 256) {
    const int hh = j >> 6;
    float mx = -INFINITY;
    for (int n = 0; n < ntok; ++n) mx = fmaxf(mx, ps[hh * 64 + n]);
    float den = 0.f;
    for (int n = 0; n < ntok; ++n) den += expf(ps[hh * 64 + n] - mx);
    float c = 0.f;
    for (int n = 0; n < ntok; ++n) c = fmaf(expf(ps[hh * 64 + n] - mx) / den, vmat[n * 256 + j], c);
    cs[j] = c;
  }
  __syncthreads();
  if (j < 256) {
    float a = bo[j];
    for (int k = 0; k < 256; ++k) a = fmaf(wo[j * 256 + k], cs[k], a);
    out[(long long)b * 256 + j] = a;
  }
}

}  // namespace srn
