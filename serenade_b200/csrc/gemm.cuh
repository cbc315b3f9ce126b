// Warp-specialised tcgen05 GEMM core for sm_100a.
//
//   D[z][m][n] = sum_k A[z][m + shift(k)][col(k)] * B[z][n][k]        (both operands K-major)
//
// A is described by up to 24 K "segments" (tap shift, source tensor, column base), which is how the
// masked Conv1d k3 / stride-2 / ConvTranspose1d / 1x1 / channel-concat forms of the estimator all map
// onto one kernel: TMA fetches the shifted activation rows directly from the channels-last tensor and
// zero-fills rows outside [0, T) (the conv padding), so no im2col buffer ever exists.
//
// Roles (64 + 32 EW threads per CTA, CTA pairs): warp 0 = TMA producer, warp 1 = tcgen05.mma issuer (leader CTA),
// warps 2.. = EW epilogue warps (8, or 16 for the epilogue-bound K = 512 GEMMs): TMEM -> registers -> fused
// epilogue -> swizzled smem box -> TMA store.  Accumulators live in TMEM (128 lanes x 2 stages x BN columns).
#pragma once
#include <cuda_bf16.h>

#include "ptx.cuh"

// -DSRN_TRACE=1: gemm_kernel stamps clock64 at its phase boundaries into GemmParams::trace (8 slots per CTA).
#ifndef SRN_TRACE
#define SRN_TRACE 0
#endif

namespace srn {

enum EpiMode : int {
  EPI_STORE_F32 = 0,   // out f32 = acc*alpha + bias
  EPI_STORE_OT = 1,    // out OT  = (acc + bias) [* rowmask]
  EPI_RESID_F32 = 2,   // out f32 = acc + bias + resid           (in place allowed)
  EPI_RESID_OT = 3,    // out OT  = (acc + bias + resid) [* rowmask]
  EPI_GEGLU = 4,       // out OT[n/2] = (acc[2i]+b) * gelu_erf(acc[2i+1]+b)   (interleaved weights)
  EPI_QKV = 5,         // n <  split: OT row-major store; n >= split: transposed store into V^T
  EPI_EULER = 6,       // channels-first f32: v = (acc+bias)*mask ; x_out = x_in + dt * v
  // Attention without a materialised score matrix (two passes over Q K^T, then P~ V):
  EPI_ROWMAX = 7,      // no tensor output: stat_max[z][key slice of 64][row] = max over the slice's valid keys of acc*alpha
  EPI_EXP = 8,         // out OT = exp(acc*alpha - rowmax) (0 at masked keys), stat_sum[...] = partial row sums (f32)
  EPI_PV = 9,          // out OT = acc / rowsum
  EPI_RESID_ACT = 10,  // x f32 (mapD) = [resid (mapR) +] acc + bias; second output (mapD2) OT = leaky_relu(x, act_slope)
  EPI_STORE_GN = 11,   // out OT (bf16, or unrounded f32 in the TF32-class mode) = acc + bias, GroupNorm partial sums
                       // (sum, sum of squares per 64-channel group and 128-row tile) taken from the f32 accumulators
};

constexpr int kWinRows = 192;      // rows of an A window (window mode): 128 + the largest shift span

struct KSeg {
  int src;    // 0 / 1: which A tensor map
  int col0;   // first column (elements) inside that tensor
  int shift;  // row shift (conv tap offset)
  int nblk;   // number of 128-byte K blocks in this segment
};

struct GemmParams {
  CUtensorMap mapA[2];
  CUtensorMap mapB;
  CUtensorMap mapD;    // output tile store: box = 32 rows x 128 bytes, SWIZZLE_128B (all epilogues but EULER)
  CUtensorMap mapD2;   // EPI_QKV: V^T store, box = 32 (d rows) x 32 (s, contiguous), no swizzle; tile group 1: its output
  CUtensorMap mapB2;   // tile group 1: its B operand
  CUtensorMap mapR;      // EPI_RESID_OT (2-byte operands): the f32 residual as 32 x 32 boxes, like an f32 mapD
  KSeg seg[24];
  int nseg;
  int nkb;          // total K blocks
  int ntn, ntm, nz; // tile grid: N tiles, M tiles, batch entries (persistent CTAs loop over ntn*ntm*nz tiles)
  int M;            // valid rows per batch entry
  int N;            // valid output columns
  int HB;           // inner batch (heads): blockIdx.z = b * HB + h
  int a_hcol;       // A column offset per h (elements)
  int a_zbatch;     // A batch coordinate: 1 -> z, 0 -> b
  int b_col0;       // B column base (elements)
  int b_hcol;       // B column offset per h
  int b_hrow;       // B row offset per h
  int b_batch_mode; // 0 -> 0 (weights), 1 -> b, 2 -> z
  int epi;
  float alpha;
  float act_slope;  // EPI_STORE_OT: out = (v > 0 ? v : act_slope * v) after the bias (1 = identity, 0.2 = LeakyReLU, 0 = ReLU)
  const float* bias;
  const uint8_t* rowmask;  // may be null
  int mask_bstride;        // bytes per batch entry
  int mask_step;           // mask index = row * mask_step + mask_off  (step 2 at half rate: mask[:, ::2])
  int mask_off;
  void* out;
  long long out_bstride, out_hstride;  // elements
  int ldo;
  int out_cols;     // inner extent of the output tensor (columns >= out_cols are never written)
  const float* resid;
  long long resid_bstride;
  int ldr;
  float* gn_part;   // [B][m_tiles][8][2] partial (sum, sum of squares) per 64-channel group, or null
  const uint8_t* gn_mask;   // non-null: only rows with gn_mask[b * mask_bstride + row * mask_step] != 0 enter gn_part
  // EPI_QKV
  int split_n;
  void* out2;       // V^T [B][split rows][ld2]
  long long out2_bstride;
  int ld2;
  // EPI_EULER
  const float* x_in;
  float* x_out;
  float* v_out;     // may be null
  float dt;
  const float* dt_ptr;         // if non-null, dt is read from device memory (graph replay)
  const uint8_t* blend_mask;   // optional infill blend (B, T_cf): x_out = blend_src where non-zero
  const float* blend_src;
  int T_cf;         // row length of the channels-first tensors
  int d_hcol;       // output column offset per h (elements), e.g. 512 for the per-head attention output
  int d_zbatch;     // output batch coordinate: 1 -> z, 0 -> b
  unsigned long long* trace;   // -DSRN_TRACE=1 builds: per-CTA clock64 stamps of the kernel's phases (srn_profile_ops)
  // Window mode (convolutions: all K segments read the same source columns at different row shifts).  Instead of one
  // 128-row A box per (segment, K block), each CTA loads ONE kWinRows-row window per K block - rows
  // m0 + win_min .. - and the segments' MMAs address shifted 128-row views of it through the smem descriptor.
  // Cuts the L2 -> SM operand traffic of a k-tap conv by up to k.  mapA[0] must have kWinRows-row boxes.
  int win;          // 0 = off; 1 = on; 3 = on, descriptors carry the matrix base offset ((addr >> 7) & 7)
  int win_min;      // smallest row shift of the segments (max - min <= kWinRows - 128)
  // Tile groups: ONE launch for two GEMMs that read the same A tensors (a resnet's k3 conv + its 1x1 res_conv; the two
  // output parities of the transposed conv).  N tiles [0, grp_ntn0) are group 0 (segments [0, nseg), mapB, mapD, bias,
  // mask_off), N tiles [grp_ntn0, ntn) group 1 (segments [grp_seg0, grp_seg0 + grp_nseg), mapB2, mapD2, bias2,
  // mask_off2); both have N columns.  grp_ntn0 = 0: no groups.  GroupNorm partial sums: group 0 only.
  int grp_ntn0, grp_seg0, grp_nseg, grp_nkb;
  const float* bias2;
  int mask_off2;
  // EPI_ROWMAX / EPI_EXP / EPI_PV: per-row softmax statistics, one entry per 64-key slice:
  // stat[(z * stat_n + j) * stat_ld + row], j < stat_n = ceil(keys / 64)
  float* stat_max;
  float* stat_sum;
  int stat_n;
  int stat_ld;
  // Shift-free softmax (test in the EPI_ROWMAX prologue, gemm_mainloop.cuh).  EPI_QKV: qk_norm2[(b * 4 + h) * 16 + (K ? 8 : 0) + part] collects
  // (atomic max of float bits) the largest squared norm of a 64..128-column part of any Q / K row of head h.
  // EPI_ROWMAX evaluates the test per (batch, head) entry (gemm_mainloop.cuh), writes noshift[z] and skips the tiles of
  // entries whose |score| is provably <= 64; EPI_EXP runs those entries with shift 0 (and clears the maxima for the
  // next attention block).
  // Padded-frame skipping (transformer-block GEMMs of ragged batches).  row_extent[b] = last unmasked frame of batch
  // entry b + 1 (at this level's rate), or null.  Row tiles that start at or beyond it are skipped by all three warp
  // roles: their outputs are only ever consumed at padded frames, which every consumer masks.  ext_keys = 1 (Q K^T
  // passes): the N tiles are keys of the same entry and are skipped likewise; ext_keys = 2 (P~ V): the K loop and the
  // row-sum slices stop at the extent, so no skipped (stale) P~ / V^T / statistics entry is ever read.
  const int* row_extent;
  int ext_keys;
  int ext_halo;     // convolutions: a row tile is skipped only if it starts >= extent + ext_halo (1 for 3 taps: the row next
                    // to the last valid frame still sees it); a skipped tile's output is the bias, and its GroupNorm
                    // partial sums are written from the bias (the reference's statistics span the padded frames)
  float* qk_norm2;
  int* noshift;
  unsigned long long* ns_count;   // EPI_ROWMAX: [0] += shift-free entries, [1] += entries (statistics only), or null
  int nz_norm;      // number of (batch, head) entries of qk_norm2
  int rev;          // 1: process the tiles in descending order (gemm_mainloop.cuh: logical)
  float noshift_max; // the bound below which an entry runs shift-free (64; SRN_NOSHIFT_MAX overrides it for tests)
  const uint8_t* keymask;   // (B, keys): key k of batch b takes part iff keymask[b*keymask_bstride + k*keymask_step] != 0
  int keymask_bstride;
  int keymask_step;
};

template <typename OT>
struct OpTraits;
template <>
struct OpTraits<__nv_bfloat16> {
  static constexpr bool kTF32 = false;
  static constexpr int kKind = 0;      // tcgen05.mma kind::f16
  static constexpr uint32_t kFmt = 1;
  static constexpr int kBK = 64;   // elements per 128-byte K block
};
template <>
struct OpTraits<float> {
  static constexpr bool kTF32 = true;
  static constexpr int kKind = 1;      // kind::tf32
  static constexpr uint32_t kFmt = 2;
  static constexpr int kBK = 32;
};
__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Exact-erf GELU (diffusers GEGLU uses F.gelu, erf form).  erf by Abramowitz-Stegun 7.1.26 (|abs err| <= 1.5e-7):
// the GEGLU epilogue runs 16 of these per 32-column chunk and was instruction-bound on erff.
__device__ __forceinline__ float gelu_erf(float g) {
  const float x = fabsf(g) * 0.70710678118654752440f;
  const float t = __fdividef(1.0f, fmaf(0.3275911f, x, 1.0f));
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  const float e = 1.0f - poly * t * __expf(-x * x);      // erf(|g| / sqrt 2)
  return 0.5f * g * (1.0f + copysignf(e, g));
}

// GELU for 2-byte outputs (the GEGLU epilogue was issue-bound on the 18 instructions of gelu_erf per element - ncu:
// selected + not_selected = 44 % of the samples).  With z = |g| and q(z) = Phi(-z) = erfc(z / sqrt 2) / 2,
// gelu(g) = max(g, 0) - z q(z); log2 q is smooth, and a degree-6 polynomial on [0, 5.5] (Chebyshev fit, max error
// 3.8e-5 in the exponent) gives |gelu error| <= 3.9e-6 absolute and <= 2.8e-5 relative on either side - two orders
// below the bf16 rounding of the result (2e-3).  z is clamped at 5.5 for the exponent only (z q < 1.1e-7 there).
// 11 instructions, one MUFU.  The TF32-class mode (f32 outputs, 1e-3 per-step gate) keeps gelu_erf.
__device__ __forceinline__ float gelu_tail(float g) {
  const float z = fabsf(g);
  const float zc = fminf(z, 5.5f);
  float e = fmaf(2.6304731363779865e-05f, zc, -0.0006637311307713389f);
  e = fmaf(e, zc, 0.007507435977458954f);
  e = fmaf(e, zc, -0.05203327164053917f);
  e = fmaf(e, zc, -0.46023255586624146f);
  e = fmaf(e, zc, -1.1506421566009521f);
  e = fmaf(e, zc, -1.0000295639038086f);
  float q;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(q) : "f"(e));
  return fmaf(-z, q, fmaxf(g, 0.f));
}
#ifndef SRN_GELU_TAIL
#define SRN_GELU_TAIL 1
#endif

template <typename OT>
__device__ __forceinline__ void store_row32(OT* dst, const float (&v)[32]);
template <>
__device__ __forceinline__ void store_row32<float>(float* dst, const float (&v)[32]) {
#pragma unroll
  for (int i = 0; i < 8; ++i)
    reinterpret_cast<float4*>(dst)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
}
template <>
__device__ __forceinline__ void store_row32<__nv_bfloat16>(__nv_bfloat16* dst, const float (&v)[32]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint4 u;
    __nv_bfloat162 a = __floats2bfloat162_rn(v[8 * i + 0], v[8 * i + 1]);
    __nv_bfloat162 b = __floats2bfloat162_rn(v[8 * i + 2], v[8 * i + 3]);
    __nv_bfloat162 c = __floats2bfloat162_rn(v[8 * i + 4], v[8 * i + 5]);
    __nv_bfloat162 d = __floats2bfloat162_rn(v[8 * i + 6], v[8 * i + 7]);
    u.x = *reinterpret_cast<uint32_t*>(&a);
    u.y = *reinterpret_cast<uint32_t*>(&b);
    u.z = *reinterpret_cast<uint32_t*>(&c);
    u.w = *reinterpret_cast<uint32_t*>(&d);
    reinterpret_cast<uint4*>(dst)[i] = u;
  }
}
__device__ __forceinline__ void store_one(float* p, float v) { *p = v; }
__device__ __forceinline__ void store_one(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

}  // namespace srn

#include "gemm_kernel.cuh"
