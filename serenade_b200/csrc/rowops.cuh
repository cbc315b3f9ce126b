// Memory-bound row kernels of the estimator (channels-last [B][T][C] activations):
//   pack_cf_to_cl      (B,C,T) f32 channels-first -> masked channels-last operand columns
//   gn_apply_kernel    GroupNorm(8)+Mish(+time bias)+mask            (decoder.py:66-77, :97)
//   gn_res_spk_ln      GroupNorm+Mish+mask + residual + SpeakerAdapter LN*scale+bias, then the
//                      transformer's LayerNorm1 on the same row       (decoder.py:14-45, :99-100; transformer.py:286)
//   layernorm_kernel   LayerNorm(512) affine                          (transformer.py:322)
//   softmax_kernel     key-masked softmax rows                        (diffusers Attention / SDPA bool mask)
//   gemv_kernel        small fp32 linears hoisted out of the step loop (time MLP, time bias, speaker scale/bias)
// All single pass over HBM, 16-byte vector accesses, warp-shuffle reductions, fp32 math.
#pragma once
#include <cuda_bf16.h>
#include <cstdint>

#include "ptx.cuh"

namespace srn {

constexpr int kC = 512;         // hidden width of the ssc1 decoder
constexpr int kGroups = 8;      // GroupNorm groups (decoder.py:67)

#if SRN_TRACE
// -DSRN_TRACE=1 builds: per-block stamps of gn_res_spk_ln (8 slots per block: 0-4 clock64 at the phase boundaries,
// 5 SM id, 6/7 globaltimer at entry/exit); see Engine::trace_op
static __device__ unsigned long long* g_row_trace;
__device__ __forceinline__ void row_stamp(int k) {
  if (g_row_trace == nullptr || threadIdx.x != 0) return;
  unsigned long long* t = g_row_trace + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 16;
  t[k] = (unsigned long long)clock64();
  unsigned long long gt;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt));
  if (k == 0) {
    unsigned smid;
    asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
    t[5] = smid;
    t[6] = gt;
  }
  if (k == 4) t[7] = gt;
}
#define ROW_STAMP(k) row_stamp(k)
#else
#define ROW_STAMP(k) do { } while (0)
#endif

__device__ __forceinline__ float mish_f(float x) {
  // x * tanh(softplus(x)) with w = e^x: tanh(ln(1 + w)) = (w^2 + 2w) / (w^2 + 2w + 2)  -- one exp and one
  // division instead of expf + log1pf + tanhf (the row kernels were instruction-bound on those, not HBM-bound).
  // x > 20: tanh(softplus(x)) == 1 in fp32 (torch's softplus threshold); x << 0: n -> 2w keeps full relative precision.
  const float w = __expf(fminf(x, 20.f));
  const float n = w * (w + 2.0f);
  return x > 20.f ? x : x * __fdividef(n, n + 2.0f);
}
// The same without the denormal handling that `__expf` / `__fdividef` carry when the file is not compiled with
// -ftz=true (ncu, round 2: 31 warp instructions per element in gn_apply, most of them predicated range fix-ups around
// the two MUFU ops): flush-to-zero ex2 / rcp, 9 instructions, and no select for x > 20 (n + 2 == n there, so the ratio is 1).
__device__ __forceinline__ float mish_fast(float x) {
  float w, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(w) : "f"(fminf(x, 20.f) * 1.4426950408889634f));
  const float a = w + 2.0f;
  const float n = w * a;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaf(w, a, 2.0f)));
  return x * (n * r);
}
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }

__device__ __forceinline__ void load8(const float* p, float (&v)[8]) {
  const float4 a = reinterpret_cast<const float4*>(p)[0];
  const float4 b = reinterpret_cast<const float4*>(p)[1];
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {      // bf16 -> f32 is a 16-bit shift
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ void store8(float* p, const float (&v)[8]) {
  reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&v)[8]) {
  uint4 u;
  __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]);
  __nv_bfloat162 b = __floats2bfloat162_rn(v[2], v[3]);
  __nv_bfloat162 c = __floats2bfloat162_rn(v[4], v[5]);
  __nv_bfloat162 d = __floats2bfloat162_rn(v[6], v[7]);
  u.x = *reinterpret_cast<uint32_t*>(&a); u.y = *reinterpret_cast<uint32_t*>(&b);
  u.z = *reinterpret_cast<uint32_t*>(&c); u.w = *reinterpret_cast<uint32_t*>(&d);
  *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ void from_f(float* p, float v) { *p = v; }
__device__ __forceinline__ void from_f(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// Operand stores: tensors that feed a GEMM as TF32 are rounded to nearest here (the tensor core truncates fp32
// inputs to TF32, which is biased and twice the error); bf16 operands round in the conversion.
__device__ __forceinline__ float tf32_rn(float v) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
  return __uint_as_float(u);
}
__device__ __forceinline__ void store8_op(float* p, const float (&v)[8]) {
  reinterpret_cast<float4*>(p)[0] = make_float4(tf32_rn(v[0]), tf32_rn(v[1]), tf32_rn(v[2]), tf32_rn(v[3]));
  reinterpret_cast<float4*>(p)[1] = make_float4(tf32_rn(v[4]), tf32_rn(v[5]), tf32_rn(v[6]), tf32_rn(v[7]));
}
__device__ __forceinline__ void store8_op(__nv_bfloat16* p, const float (&v)[8]) { store8(p, v); }
__device__ __forceinline__ void from_f_op(float* p, float v) { *p = tf32_rn(v); }
__device__ __forceinline__ void from_f_op(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ------------------------------------------------------------------------------------------------
// (B, C, T) f32 channels-first  ->  dst[b][t][col0 + c] = src[b][c][t] * mask[b][t]    (OT, channels-last)
template <typename OT>
__global__ void pack_cf_to_cl(const float* __restrict__ src, int C, int T, const uint8_t* __restrict__ mask,
                              int mask_bstride, OT* __restrict__ dst, long long dst_bstride, int ldd, int col0) {
  __shared__ float tile[32][33];
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.z;
  const int t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, t = t0 + threadIdx.x;
    tile[i][threadIdx.x] = (c < C && t < T) ? src[((long long)b * C + c) * T + t] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int t = t0 + i, c = c0 + threadIdx.x;
    if (t < T && c < C) {
      const float m = mask[(long long)b * mask_bstride + t] ? 1.f : 0.f;
      from_f_op(dst + b * dst_bstride + (long long)t * ldd + col0 + c, tile[threadIdx.x][i] * m);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Valid-frame GroupNorm mode: number of unmasked frames of every batch entry at full rate (out[b]) and at half
// rate (out[B + b], frames 0, 2, 4, ...).  One block per batch entry.
__global__ void __launch_bounds__(256) count_valid_kernel(const uint8_t* __restrict__ mask, int T, int mask_bstride,
                                                          int B, int* __restrict__ out) {
  __shared__ int red[2][8];
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.x;
  int full = 0, half = 0;
  for (int t = threadIdx.x; t < T; t += 256) {
    const int m = mask[(long long)b * mask_bstride + t] != 0;
    full += m;
    half += m & ((t & 1) == 0);
  }
  for (int o = 16; o > 0; o >>= 1) {
    full += __shfl_xor_sync(0xffffffffu, full, o);
    half += __shfl_xor_sync(0xffffffffu, half, o);
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = full; red[1][threadIdx.x >> 5] = half; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int f = 0, h = 0;
    for (int i = 0; i < 8; ++i) { f += red[0][i]; h += red[1][i]; }
    out[b] = f;
    out[B + b] = h;
  }
}

// ------------------------------------------------------------------------------------------------
// Padded-frame skipping: out[b] = last unmasked frame of batch entry b + 1 at full rate, out[B + b] the same at half
// rate (frames 0, 2, 4, ...; 0 = no valid frame).  Works for any mask, not only length prefixes.  One block per entry.
__global__ void __launch_bounds__(256) row_extent_kernel(const uint8_t* __restrict__ mask, int T, int mask_bstride,
                                                         int B, int* __restrict__ out) {
  __shared__ int red[2][8];
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.x;
  int full = 0, half = 0;
  for (int t = threadIdx.x; t < T; t += 256) {
    if (mask[(long long)b * mask_bstride + t] != 0) {
      full = max(full, t + 1);
      if ((t & 1) == 0) half = max(half, (t >> 1) + 1);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    full = max(full, __shfl_xor_sync(0xffffffffu, full, o));
    half = max(half, __shfl_xor_sync(0xffffffffu, half, o));
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = full; red[1][threadIdx.x >> 5] = half; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int f = 0, h = 0;
    for (int i = 0; i < 8; ++i) { f = max(f, red[0][i]); h = max(h, red[1][i]); }
    out[b] = f;
    out[B + b] = h;
  }
}

// ------------------------------------------------------------------------------------------------
// Reduce the conv epilogue's per-tile partial sums to (mean, rstd) for the 8 groups of batch b.
// part: [B][mtiles][8][2].  Result in smem stats[16] = {mean_g, rstd_g}.  Call with >= 256 threads.
__device__ __forceinline__ void gn_reduce_stats(const float* __restrict__ part, int b, int mtiles, int T,
                                                const int* __restrict__ nvalid /* per batch entry, or null: T */,
                                                float* stats /*smem [16]*/, double* red /*smem [256]*/) {
  const int t = threadIdx.x;
  if (t < 256) {
    const int vi = t & 15, tl = t >> 4;
    double acc = 0.0;
    for (int m = tl; m < mtiles; m += 16) acc += (double)part[((long long)b * mtiles + m) * 16 + vi];
    red[t] = acc;
  }
  __syncthreads();
  if (t < 16) {
    double acc = 0.0;
    for (int i = 0; i < 16; ++i) acc += red[i * 16 + t];
    red[t] = acc;
  }
  __syncthreads();
  if (t < 8) {
    // (double divisions and square roots are long software sequences on this chip and sit on every block's critical
    //  path: a float seed plus one Newton step in double gives 1/n and 1/sqrt(var + eps) to ~1e-14)
    const double n = 64.0 * (double)(nvalid != nullptr ? max(nvalid[b], 1) : T);
    double inv_n = (double)(1.0f / (float)n);
    inv_n = inv_n * (2.0 - n * inv_n);
    const double mean = red[2 * t] * inv_n;
    double var = red[2 * t + 1] * inv_n - mean * mean;
    if (var < 0.0) var = 0.0;
    const double x = var + 1e-5;
    double r = (double)rsqrtf((float)x);
    r = r * (1.5 - 0.5 * x * r * r);
    stats[2 * t] = (float)mean;
    stats[2 * t + 1] = (float)r;
  }
  __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// The two GroupNorm row kernels.  Neither is bound by HBM bandwidth: ncu (profiles/r02_ncu_full_rowkernels.txt) shows
// them issue- and latency-bound (31 and 748 warp instructions per element / row in the round-1 forms, most of them the
// denormal fix-ups around Mish's two MUFU ops and per-row shared-memory constant loads).  Round 2:
//   gn_apply_kernel   out = (Mish(GN(y)) + tbias) * mask; 256 threads = 4 rows x 64 threads x 8 channels; every thread
//                     keeps U rows' loads in flight (issued before the statistics prologue / before the previous batch
//                     is consumed); flush-to-zero Mish (mish_fast)
//   gn_res_spk_ln     u = Mish(GN(y2)) * mask + r (ResnetBlock1D, decoder.py:98-99); s = LN_c(u) * scale[b] + bias[b]
//                     (SpeakerAdapter, decoder.py:34-45) -> hres (f32 residual stream); ln = LayerNorm(s) * g1 + b1
//                     (transformer.py:286) -> OT.  A warp takes two rows per iteration, so the per-channel constants are
//                     fetched from shared memory once for both and the rows' reduction chains interleave; each LayerNorm
//                     is one fused (sum, sum of squares) butterfly instead of two dependent ones; the next iteration's
//                     rows are loaded into registers while the current ones are reduced and stored.
//                     (A variant that streamed the rows through a shared-memory ring filled by cp.async.bulk was 5 %
//                     faster in isolation but made the SOLVE irreproducible - about 2 solves in 1000 differed in one clip,
//                     first differing tensor = this kernel's output, only in the full pipeline: the kernel alone was
//                     bit-stable over 200 000 launches, tools/stress/rowkernel_stress.cu.  Not understood, so not shipped:
//                     profiles/r02_determinism.txt.)
// Same-box A/B against the round-1 forms: profiles/r02_ab_rowkernels_widepv.txt (117 -> 89 us and 76 -> 59 us per C2 call).
template <typename OT>
struct Raw8;
template <>
struct Raw8<__nv_bfloat16> {
  uint4 u;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { u = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void unpack(float (&v)[8]) const {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
};
template <>
struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) {
    a = reinterpret_cast<const float4*>(p)[0];
    b = reinterpret_cast<const float4*>(p)[1];
  }
  __device__ __forceinline__ void unpack(float (&v)[8]) const {
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};

template <typename OT>
__global__ void __launch_bounds__(256, 4) gn_apply_kernel(const OT* __restrict__ y, const float* __restrict__ part, int mtiles,
                                                      int T, long long bstride, const float* __restrict__ gamma,
                                                      const float* __restrict__ beta, const float* __restrict__ tbias,
                                                      const uint8_t* __restrict__ mask, int mask_bstride, int mask_step,
                                                      OT* __restrict__ out, int rows_per_block,
                                                      const int* __restrict__ nvalid, int tbias_bstride) {
  constexpr int U = sizeof(OT) == 2 ? 4 : 2;      // rows in flight per thread
  __shared__ float stats[16];
  __shared__ double red[256];
  pdl_launch_dependents();
  const int c = (threadIdx.x & 63) * 8;
  const int g = c >> 6;
  float ga[8], be[8];
  load8(gamma + c, ga);       // weights: not produced by an earlier kernel, fetched under the previous kernel's tail
  load8(beta + c, be);
  pdl_wait();
  const int b = blockIdx.y;
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(r0 + rows_per_block, T);
  const long long boff = b * bstride + c;
  const uint8_t* mrow = mask + (long long)b * mask_bstride;
  Raw8<OT> raw[U];
  uint8_t mk[U];
  auto issue = [&](int rb) {
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int r = rb + 4 * j;
      if (r < r1) {
        raw[j].load(y + boff + (long long)r * kC);
        mk[j] = mrow[(long long)r * mask_step];
      }
    }
  };
  int rb = r0 + (threadIdx.x >> 6);
  issue(rb);                  // in flight during the statistics prologue
  gn_reduce_stats(part, b, mtiles, T, nvalid, stats, red);
  const float mean = stats[2 * g], rstd = stats[2 * g + 1];
  float tb[8];
  if (tbias != nullptr) load8(tbias + (long long)b * tbias_bstride + c, tb);      // (stride 0: one t for the batch)
  else {
#pragma unroll
    for (int i = 0; i < 8; ++i) tb[i] = 0.f;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) { ga[i] *= rstd; be[i] -= mean * ga[i]; }
  for (; rb < r1; rb += 4 * U) {
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int r = rb + 4 * j;
      if (r < r1) {
        float v[8];
        raw[j].unpack(v);
        const bool on = mk[j] != 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = on ? mish_fast(fmaf(v[i], ga[i], be[i])) + tb[i] : 0.f;   // (select: y may be stale at skipped padded tiles)
        store8_op(out + boff + (long long)r * kC, v);
      }
    }
    if (rb + 4 * U < r1) issue(rb + 4 * U);
  }
}

// rows per warp and iteration of gn_res_spk_ln (two for 2-byte rows: the per-channel constants are fetched from shared
// memory once for both, and the two rows' reduction chains interleave)
template <typename OT>
__host__ __device__ constexpr int gnres_rpw() { return sizeof(OT) == 2 ? 2 : 1; }
// (y / res rows are loaded from global memory straight into registers, one iteration ahead of their use)
template <typename OT>
__global__ void __launch_bounds__(256, 2) gn_res_spk_ln(const OT* __restrict__ y, const float* __restrict__ part, int mtiles,
                                                               int T, long long bstride, const float* __restrict__ gamma,
                                                               const float* __restrict__ beta, const OT* __restrict__ res,
                                                               const uint8_t* __restrict__ mask, int mask_bstride, int mask_step,
                                                               const float* __restrict__ spk_scale, const float* __restrict__ spk_bias,
                                                               const float* __restrict__ ln_g, const float* __restrict__ ln_b,
                                                               float* __restrict__ hres, OT* __restrict__ ln_out,
                                                               int rows_per_block, const int* __restrict__ nvalid) {
  constexpr int RPW = gnres_rpw<OT>();
  constexpr int SROWS = 8 * RPW;                         // rows per block iteration
  __shared__ float stats[16];
  __shared__ double red[256];
  __shared__ __align__(16) float cst[6][kC];     // folded GN scale / shift, speaker scale / bias, LN gain / bias
  pdl_launch_dependents();
  float wg[2], wb[2], wlg[2], wlb[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int c = threadIdx.x + k * 256;
    wg[k] = gamma[c]; wb[k] = beta[c]; wlg[k] = ln_g[c]; wlb[k] = ln_b[c];
  }
  pdl_wait();
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(r0 + rows_per_block, T);
  const long long boff = b * bstride;
  Raw8<OT> ry[RPW][2], rr[RPW][2];
  uint8_t mk[RPW];
  auto fetch = [&](int rbase) {       // this warp's rows of the given iteration -> registers
#pragma unroll
    for (int j = 0; j < RPW; ++j) {
      if (rbase + j < r1) {
        const long long o = boff + (long long)(rbase + j) * kC + lane * 8;
        ry[j][0].load(y + o); ry[j][1].load(y + o + 256);
        rr[j][0].load(res + o); rr[j][1].load(res + o + 256);
        mk[j] = mask[(long long)b * mask_bstride + (long long)(rbase + j) * mask_step];
      }
    }
  };
  int rbase = r0 + warp * RPW;
  fetch(rbase);               // in flight during the statistics prologue
  float wsc[2], wsb[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int c = threadIdx.x + k * 256;
    wsc[k] = spk_scale[(long long)b * kC + c];
    wsb[k] = spk_bias[(long long)b * kC + c];
  }
  gn_reduce_stats(part, b, mtiles, T, nvalid, stats, red);
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int c = threadIdx.x + k * 256;
    const float a = stats[2 * (c >> 6) + 1] * wg[k];
    cst[0][c] = a;
    cst[1][c] = wb[k] - stats[2 * (c >> 6)] * a;
    cst[2][c] = wsc[k];
    cst[3][c] = wsb[k];
    cst[4][c] = wlg[k];
    cst[5][c] = wlb[k];
  }
  __syncthreads();
  for (; rbase < r1; rbase += SROWS) {
    float u[RPW][16], m[RPW], s1[RPW], q1[RPW];
#pragma unroll
    for (int j = 0; j < RPW; ++j) {
      m[j] = (rbase + j < r1 && mk[j]) ? 1.f : 0.f;
      s1[j] = 0.f; q1[j] = 0.f;
    }
#pragma unroll
    for (int hlf = 0; hlf < 2; ++hlf) {
      const int c = hlf * 256 + lane * 8;
      float ga[8], be[8];
      load8(&cst[0][c], ga);
      load8(&cst[1][c], be);
#pragma unroll
      for (int j = 0; j < RPW; ++j) {
        float v[8], rv[8];
        if (rbase + j < r1) { ry[j][hlf].unpack(v); rr[j][hlf].unpack(rv); }
        else {
#pragma unroll
          for (int i = 0; i < 8; ++i) { v[i] = 0.f; rv[i] = 0.f; }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float x = fmaf(mish_fast(fmaf(v[i], ga[i], be[i])), m[j], rv[i]);
          u[j][hlf * 8 + i] = x;
          s1[j] += x;
          q1[j] = fmaf(x, x, q1[j]);
        }
      }
    }
    const int cur = rbase;
    if (rbase + SROWS < r1) fetch(rbase + SROWS);      // the next rows: in flight during the reductions and stores below
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int j = 0; j < RPW; ++j) {
        s1[j] += __shfl_xor_sync(0xffffffffu, s1[j], o);
        q1[j] += __shfl_xor_sync(0xffffffffu, q1[j], o);
      }
    }
    float mu[RPW], inv[RPW], s2[RPW], q2[RPW];
#pragma unroll
    for (int j = 0; j < RPW; ++j) {
      mu[j] = s1[j] * (1.0f / kC);
      inv[j] = 1.0f / sqrtf(fmaxf(q1[j] * (1.0f / kC) - mu[j] * mu[j], 0.f) + 1e-5f);
      s2[j] = 0.f; q2[j] = 0.f;
    }
#pragma unroll
    for (int hlf = 0; hlf < 2; ++hlf) {
      const int c = hlf * 256 + lane * 8;
      float sc[8], sb[8];
      load8(&cst[2][c], sc);
      load8(&cst[3][c], sb);
#pragma unroll
      for (int j = 0; j < RPW; ++j) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          o[i] = fmaf((u[j][hlf * 8 + i] - mu[j]) * inv[j], sc[i], sb[i]);
          u[j][hlf * 8 + i] = o[i];
          s2[j] += o[i];
          q2[j] = fmaf(o[i], o[i], q2[j]);
        }
        if (cur + j < r1) store8(hres + boff + (long long)(cur + j) * kC + c, o);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int j = 0; j < RPW; ++j) {
        s2[j] += __shfl_xor_sync(0xffffffffu, s2[j], o);
        q2[j] += __shfl_xor_sync(0xffffffffu, q2[j], o);
      }
    }
#pragma unroll
    for (int j = 0; j < RPW; ++j) {
      mu[j] = s2[j] * (1.0f / kC);
      inv[j] = rsqrtf(fmaxf(q2[j] * (1.0f / kC) - mu[j] * mu[j], 0.f) + 1e-5f);
    }
#pragma unroll
    for (int hlf = 0; hlf < 2; ++hlf) {
      const int c = hlf * 256 + lane * 8;
      float g1[8], b1[8];
      load8(&cst[4][c], g1);
      load8(&cst[5][c], b1);
#pragma unroll
      for (int j = 0; j < RPW; ++j) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = fmaf((u[j][hlf * 8 + i] - mu[j]) * inv[j], g1[i], b1[i]);
        if (cur + j < r1) store8_op(ln_out + boff + (long long)(cur + j) * kC + c, o);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// LayerNorm over 512 channels, affine, eps 1e-5; one warp per row.
template <typename OT>
__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ x, long long nrows,
                                                        const float* __restrict__ g, const float* __restrict__ bta,
                                                        OT* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= nrows) return;
  const long long rowoff = r * kC;
  float u[16];
#pragma unroll
  for (int hlf = 0; hlf < 2; ++hlf) {
    float v[8];
    load8(x + rowoff + hlf * 256 + lane * 8, v);
#pragma unroll
    for (int i = 0; i < 8; ++i) u[hlf * 8 + i] = v[i];
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += u[i];
  const float mu = warp_sum(s) * (1.0f / kC);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) { u[i] -= mu; q += u[i] * u[i]; }
  const float inv = rsqrtf(warp_sum(q) * (1.0f / kC) + 1e-5f);
#pragma unroll
  for (int hlf = 0; hlf < 2; ++hlf) {
    const int c = hlf * 256 + lane * 8;
    float gg[8], bb[8], o[8];
    load8(g + c, gg);
    load8(bta + c, bb);
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = u[hlf * 8 + i] * inv * gg[i] + bb[i];
    store8_op(out + rowoff + c, o);
  }
}

// ------------------------------------------------------------------------------------------------
// P[z][q][k] = softmax_k(scores[z][q][k]) with keys where mask[b][k*step] == 0 excluded.  One block per row.
// Single pass over HBM: each 128-thread half-block owns one row and keeps it in registers (NV float4 per thread,
// S <= 512 * NV); scores are read once (4 B/elem) and P written once.  ld must be a multiple of 4.
__device__ __forceinline__ void store4(float* p, float a, float b, float c, float d) {   // P is a GEMM operand
  *reinterpret_cast<float4*>(p) = make_float4(tf32_rn(a), tf32_rn(b), tf32_rn(c), tf32_rn(d));
}
__device__ __forceinline__ void store4(__nv_bfloat16* p, float a, float b, float c, float d) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&lo);
  u.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(p) = u;
}

template <typename OT, int NV>
__global__ void __launch_bounds__(256) softmax_kernel(const float* __restrict__ sc, int S, int ld, int HB,
                                                      long long nrows, const uint8_t* __restrict__ mask,
                                                      int mask_bstride, int mask_step, OT* __restrict__ P) {
  __shared__ float redf[2][4];
  pdl_launch_dependents();
  pdl_wait();
  const int half = threadIdx.x >> 7, t = threadIdx.x & 127;
  const int lane = threadIdx.x & 31, warp = (threadIdx.x >> 5) & 3;
  const long long rowi = (long long)blockIdx.x * 2 + half;        // z * S + q
  const bool live = rowi < nrows;
  const long long rr = live ? rowi : 0;
  const int b = (int)((rr / S) / HB);
  const float* row = sc + rr * ld;
  const uint8_t* mk = mask + (long long)b * mask_bstride;
  float v[NV][4];
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int k = (i * 128 + t) * 4;
    float4 x4 = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    if (live && k < S) x4 = *reinterpret_cast<const float4*>(row + k);   // k + 3 < ld always (ld % 4 == 0)
    v[i][0] = x4.x; v[i][1] = x4.y; v[i][2] = x4.z; v[i][3] = x4.w;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool ok = live && (k + j < S) && mk[(long long)(k + j) * mask_step] != 0;
      v[i][j] = ok ? v[i][j] : -INFINITY;
      mx = fmaxf(mx, v[i][j]);
    }
  }
  mx = warp_max(mx);
  if (lane == 0) redf[half][warp] = mx;
  __syncthreads();
  mx = fmaxf(fmaxf(redf[half][0], redf[half][1]), fmaxf(redf[half][2], redf[half][3]));
  __syncthreads();
  if (mx == -INFINITY) mx = 0.f;          // fully masked row: all probabilities 0 instead of NaN
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[i][j] = expf(v[i][j] - mx);       // exp(-inf) = 0 for masked / out-of-range keys
      sum += v[i][j];
    }
  sum = warp_sum(sum);
  if (lane == 0) redf[half][warp] = sum;
  __syncthreads();
  const float tot = redf[half][0] + redf[half][1] + redf[half][2] + redf[half][3];
  const float inv = tot > 0.f ? 1.0f / tot : 0.f;
  if (!live) return;
  OT* prow = P + rowi * ld;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int k = (i * 128 + t) * 4;
    if (k < S) store4(prow + k, v[i][0] * inv, v[i][1] * inv, v[i][2] * inv, v[i][3] * inv);   // pad cols < ld
  }
}

// ------------------------------------------------------------------------------------------------
// out[s][o] = act_out( W[o] . act_in(in[s]) + bias[o] ),  fp32, one warp per (s, o).  I % 4 == 0 not required.
enum GemvAct : int { ACT_NONE = 0, ACT_SILU = 1, ACT_MISH = 2 };
__device__ __forceinline__ float apply_act(float v, int a) {
  return a == ACT_SILU ? silu_f(v) : (a == ACT_MISH ? mish_f(v) : v);
}
__global__ void __launch_bounds__(256) gemv_kernel(const float* __restrict__ W, const float* __restrict__ bias,
                                                   const float* __restrict__ in, float* __restrict__ out, int I, int O,
                                                   int act_in, int act_out) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int o = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int s = blockIdx.y;
  if (o >= O) return;
  const float* w = W + (long long)o * I;
  const float* x = in + (long long)s * I;
  float acc = 0.f;
  for (int i = lane; i < I; i += 32) acc += w[i] * apply_act(x[i], act_in);
  acc = warp_sum(acc);
  if (lane == 0) out[(long long)s * O + o] = apply_act(acc + (bias ? bias[o] : 0.f), act_out);
}

// *a = va, *b = vb: scalars handed to the device by value (no host staging buffer, hence no host synchronisation)
__global__ void set_scalars_kernel(float* __restrict__ a, float va, float* __restrict__ b, float vb) {
  pdl_launch_dependents();
  pdl_wait();
  if (threadIdx.x == 0) { *a = va; *b = vb; }
}

// The (t, dt) tables of the Euler loop from a t_span that lives on the device, in the reference's own fp32 arithmetic
// (flow_matching.py:79-91: t += dt; dt = t_span[k + 1] - t - the running t makes dt wobble in the last ulp, and the host
// path of solve_staged does exactly the same additions): one thread, n sequential steps.
__global__ void tspan_tables_kernel(const float* __restrict__ t_span, int n, float* __restrict__ ts, float* __restrict__ dts) {
  pdl_launch_dependents();
  pdl_wait();
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float t = t_span[0];
  float dt = __fsub_rn(t_span[1], t_span[0]);
  for (int s = 1; s <= n; ++s) {
    ts[s - 1] = t;
    dts[s - 1] = dt;
    t = __fadd_rn(t, dt);
    if (s < n) dt = __fsub_rn(t_span[s + 1], t);
  }
}

// emb[s][j] = sin(1000 t_s f_j), emb[s][half + j] = cos(...),  f_j = exp(-j ln(1e4)/(half-1))  (decoder.py:54-63)
__global__ void sinusoid_kernel(const float* __restrict__ ts, int dim, float e /* ln(1e4)/(half-1) */,
                                float* __restrict__ emb) {
  pdl_launch_dependents();
  pdl_wait();
  const int s = blockIdx.x;
  const int half = dim / 2;
  for (int j = threadIdx.x; j < half; j += blockDim.x) {
    const float f = expf((float)j * -e);
    const float a = 1000.f * ts[s] * f;
    emb[(long long)s * dim + j] = sinf(a);
    emb[(long long)s * dim + half + j] = cosf(a);
  }
}

// ------------------------------------------------------------------------------------------------
// CFM.compute_loss, forward value (flow_matching.py:112-133), channels-first f32 [B][C][T].
// y = (1 - (1 - sigma) t_b) z + t_b x1 with torch's eager rounding (separate multiplies and adds, no FMA).
__global__ void __launch_bounds__(256) cfm_mix_kernel(const float* __restrict__ x1, const float* __restrict__ z,
                                                      const float* __restrict__ t, long long per_clip, long long n,
                                                      float one_minus_sigma, float* __restrict__ y) {
  pdl_wait();
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const float tb = t[i / per_clip];
  const float a = __fsub_rn(1.0f, __fmul_rn(one_minus_sigma, tb));
  y[i] = __fadd_rn(__fmul_rn(a, z[i]), __fmul_rn(tb, x1[i]));
}

// partial[2 * block] = sum over the block's elements of ((v - u) * m)^2 with u = x1 - (1 - sigma) z and m = sel (null:
// every frame - without mask_l the reference sums over the padded frames too, where v = 0 but u is not);
// partial[2 * block + 1] = number of frames with cnt != 0 it saw in channel 0 (the divisor).  Doubles, fixed order.
__global__ void __launch_bounds__(256) cfm_loss_partial_kernel(const float* __restrict__ v, const float* __restrict__ x1,
                                                               const float* __restrict__ z,
                                                               const uint8_t* __restrict__ sel,
                                                               const uint8_t* __restrict__ cntm, int lmask_bstride, int C,
                                                               int T, long long n, float one_minus_sigma,
                                                               double* __restrict__ partial) {
  __shared__ double red[2][256];
  pdl_wait();      // v comes from the estimator's last GEMM, which releases its dependents early
  double acc = 0.0, cnt = 0.0;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const int tt = (int)(i % T);
    const long long bc = i / T;
    const int c = (int)(bc % C), b = (int)(bc / C);
    const long long mi = (long long)b * lmask_bstride + tt;
    if (sel == nullptr || sel[mi] != 0) {
      const float u = __fsub_rn(x1[i], __fmul_rn(one_minus_sigma, z[i]));
      const double d = (double)v[i] - (double)u;
      acc += d * d;
    }
    if (c == 0 && cntm[mi] != 0) cnt += 1.0;
  }
  red[0][threadIdx.x] = acc;
  red[1][threadIdx.x] = cnt;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      red[0][threadIdx.x] += red[0][threadIdx.x + s];
      red[1][threadIdx.x] += red[1][threadIdx.x + s];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    partial[2 * blockIdx.x] = red[0][0];
    partial[2 * blockIdx.x + 1] = red[1][0];
  }
}

// loss = sum / (frames * C)   (flow_matching.py:128-132)
__global__ void cfm_loss_final_kernel(const double* __restrict__ partial, int nblocks, int C, float* __restrict__ loss) {
  pdl_wait();
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double s = 0.0, c = 0.0;
  for (int i = 0; i < nblocks; ++i) { s += partial[2 * i]; c += partial[2 * i + 1]; }
  *loss = (float)(s / (c * (double)C));
}

}  // namespace srn
