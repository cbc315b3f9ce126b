// Warp-specialised tcgen05 GEMM core for sm_100a.
//
//   D[z][m][n] = sum_k A[z][m + shift(k)][col(k)] * B[z][n][k]        (both operands K-major)
//
// A is described by up to six K "segments" (tap shift, source tensor, column base), which is how the
// masked Conv1d k3 / stride-2 / ConvTranspose1d / 1x1 / channel-concat forms of the estimator all map
// onto one kernel: TMA fetches the shifted activation rows directly from the channels-last tensor and
// zero-fills rows outside [0, T) (the conv padding), so no im2col buffer ever exists.
//
// Roles (192 threads): warp 0 = TMA producer, warp 1 = tcgen05.mma issuer, warps 2..5 = epilogue
// (TMEM -> registers -> fused epilogue -> global).  Accumulators live in TMEM (128 lanes x BN columns).
#pragma once
#include <cuda_bf16.h>

#include "ptx.cuh"

namespace srn {

enum EpiMode : int {
  EPI_STORE_F32 = 0,   // out f32 = acc*alpha + bias [, GroupNorm partial sums]
  EPI_STORE_OT = 1,    // out OT  = (acc + bias) [* rowmask]
  EPI_RESID_F32 = 2,   // out f32 = acc + bias + resid           (in place allowed)
  EPI_RESID_OT = 3,    // out OT  = (acc + bias + resid) [* rowmask]
  EPI_GEGLU = 4,       // out OT[n/2] = (acc[2i]+b) * gelu_erf(acc[2i+1]+b)   (interleaved weights)
  EPI_QKV = 5,         // n <  split: OT row-major store; n >= split: transposed store into V^T
  EPI_EULER = 6,       // channels-first f32: v = (acc+bias)*mask ; x_out = x_in + dt * v
};

struct KSeg {
  int src;    // 0 / 1: which A tensor map
  int col0;   // first column (elements) inside that tensor
  int shift;  // row shift (conv tap offset)
  int nblk;   // number of 128-byte K blocks in this segment
};

struct GemmParams {
  CUtensorMap mapA[2];
  CUtensorMap mapB;
  KSeg seg[6];
  int nseg;
  int nkb;          // total K blocks
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
  const float* bias;
  const uint8_t* rowmask;  // may be null
  int mask_bstride;        // bytes per batch entry
  int mask_step;           // mask index = row * mask_step + mask_off  (step 2 at half rate: mask[:, ::2])
  int mask_off;
  void* out;
  long long out_bstride, out_hstride;  // elements
  int ldo;
  const float* resid;
  long long resid_bstride;
  int ldr;
  float* gn_part;   // [B][m_tiles][8][2] partial (sum, sum of squares) per 64-channel group, or null
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
};

template <typename OT>
struct OpTraits;
template <>
struct OpTraits<__nv_bfloat16> {
  static constexpr bool kTF32 = false;
  static constexpr uint32_t kFmt = 1;
  static constexpr int kBK = 64;   // elements per 128-byte K block
};
template <>
struct OpTraits<float> {
  static constexpr bool kTF32 = true;
  static constexpr uint32_t kFmt = 2;
  static constexpr int kBK = 32;
};

__device__ __forceinline__ float gelu_erf(float g) { return 0.5f * g * (1.0f + erff(g * 0.70710678118654752440f)); }

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

template <int BN, int STAGES>
constexpr int gemm_smem_bytes() {
  return STAGES * (128 * 128 + BN * 128) + 1024 /*align slack*/ + 256 /*barriers*/ + 4 * 8 * 2 * 4 /*gn red*/;
}

template <typename OT, int BN, int STAGES>
__global__ void __launch_bounds__(192, 1) gemm_kernel(const __grid_constant__ GemmParams p) {
  using TR = OpTraits<OT>;
  constexpr int A_BYTES = 128 * 128;
  constexpr int B_BYTES = BN * 128;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr int BK = TR::kBK;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  float* gn_red = reinterpret_cast<float*>(smem + STAGES * STAGE_BYTES + 256);  // [4 warps][BN/64][2]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN;
  const int m0 = blockIdx.y * 128;
  const int z = blockIdx.z;
  const int b = z / p.HB;
  const int h = z - b * p.HB;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&p.mapA[0]);
    tma_prefetch_desc(&p.mapB);
  }
  if (warp == 1 && elect_one()) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(accum_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<BN>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      // ------------------------------------------------------------ TMA producer
      const int a_batch = p.a_zbatch ? z : b;
      const int b_batch = p.b_batch_mode == 0 ? 0 : (p.b_batch_mode == 1 ? b : z);
      const int b_row = n0 + h * p.b_hrow;
      int seg = 0, j = 0;
      uint32_t stage = 0, phase = 0;
      for (int kb = 0; kb < p.nkb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * STAGE_BYTES;
        uint8_t* sb = sa + A_BYTES;
        mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
        const KSeg sg = p.seg[seg];
        tma_load_3d(sa, &p.mapA[sg.src], &full_bar[stage], sg.col0 + j * BK + h * p.a_hcol, m0 + sg.shift, a_batch);
        tma_load_3d(sb, &p.mapB, &full_bar[stage], p.b_col0 + kb * BK + h * p.b_hcol, b_row, b_batch);
        if (++j == sg.nblk) { j = 0; ++seg; }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      // ------------------------------------------------------------ MMA issuer
      constexpr uint32_t idesc = make_idesc(TR::kFmt, 128, BN);
      uint32_t stage = 0, phase = 0;
      for (int kb = 0; kb < p.nkb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
        const uint64_t da = make_kmajor_sw128_desc(sa);
        const uint64_t db = make_kmajor_sw128_desc(sa + A_BYTES);
#pragma unroll
        for (int k = 0; k < 4; ++k)   // 4 x 32 bytes of K per 128-byte swizzled row
          umma<TR::kTF32>(tmem_base, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
        tc_commit(&empty_bar[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      tc_commit(accum_bar);
    }
  } else {
    // -------------------------------------------------------------- epilogue (4 warps, one row per thread)
    const int q = warp & 3;                 // TMEM lane quarter this warp may access
    const int row = m0 + q * 32 + lane;     // row inside the batch entry
    const bool row_ok = row < p.M;
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    float mk = 1.0f;
    if (p.rowmask != nullptr && row_ok)
      mk = p.rowmask[(long long)b * p.mask_bstride + (long long)row * p.mask_step + p.mask_off] ? 1.0f : 0.0f;
    float gs = 0.f, gss = 0.f;
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      const int nc = n0 + c * 32;
      if (nc >= p.N) break;                  // warp-uniform
      uint32_t raw[32];
      tmem_ld_32x32b_x32(tmem_base + (uint32_t(q * 32) << 16) + uint32_t(c * 32), raw);
      tmem_ld_wait();
      float v[32];
      const bool full_chunk = nc + 32 <= p.N;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        float bi = 0.f;
        if (p.bias != nullptr && (full_chunk || nc + i < p.N)) bi = __ldg(p.bias + nc + i);
        v[i] = __uint_as_float(raw[i]) * p.alpha + bi;
      }
      switch (p.epi) {
        case EPI_STORE_F32: {
          if (p.gn_part != nullptr) {
            if (row_ok) {
#pragma unroll
              for (int i = 0; i < 32; ++i) { gs += v[i]; gss += v[i] * v[i]; }
            }
            if (c & 1) {   // a 64-channel group is complete
              float s = gs, ss = gss;
#pragma unroll
              for (int o = 16; o > 0; o >>= 1) {
                s += __shfl_xor_sync(0xffffffffu, s, o);
                ss += __shfl_xor_sync(0xffffffffu, ss, o);
              }
              if (lane == 0) {
                gn_red[(q * (BN / 64) + (c >> 1)) * 2 + 0] = s;
                gn_red[(q * (BN / 64) + (c >> 1)) * 2 + 1] = ss;
              }
              gs = 0.f; gss = 0.f;
            }
          }
          if (row_ok) {
            float* dst = reinterpret_cast<float*>(p.out) + b * p.out_bstride + h * p.out_hstride +
                         (long long)row * p.ldo + nc;
            if (full_chunk) {
              store_row32<float>(dst, v);
            } else {
              for (int i = 0; i < 32; ++i)
                if (nc + i < p.N) dst[i] = v[i];
            }
          }
        } break;
        case EPI_STORE_OT: {
          if (row_ok) {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] *= mk;
            OT* dst = reinterpret_cast<OT*>(p.out) + b * p.out_bstride + h * p.out_hstride +
                      (long long)row * p.ldo + nc;
            if (full_chunk) {
              store_row32<OT>(dst, v);
            } else {
              for (int i = 0; i < 32; ++i)
                if (nc + i < p.N) store_one(dst + i, v[i]);
            }
          }
        } break;
        case EPI_RESID_F32:
        case EPI_RESID_OT: {
          if (row_ok) {
            const float* rs = p.resid + b * p.resid_bstride + (long long)row * p.ldr + nc;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 r4 = reinterpret_cast<const float4*>(rs)[i];
              v[4 * i + 0] += r4.x; v[4 * i + 1] += r4.y; v[4 * i + 2] += r4.z; v[4 * i + 3] += r4.w;
            }
            if (p.epi == EPI_RESID_F32) {
              store_row32<float>(reinterpret_cast<float*>(p.out) + b * p.out_bstride + (long long)row * p.ldo + nc, v);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] *= mk;
              store_row32<OT>(reinterpret_cast<OT*>(p.out) + b * p.out_bstride + (long long)row * p.ldo + nc, v);
            }
          }
        } break;
        case EPI_GEGLU: {
          if (row_ok) {
            float o[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) o[i] = v[2 * i] * gelu_erf(v[2 * i + 1]);
            OT* dst = reinterpret_cast<OT*>(p.out) + b * p.out_bstride + (long long)row * p.ldo + (nc >> 1);
#pragma unroll
            for (int i = 0; i < 16; ++i) store_one(dst + i, o[i]);
          }
        } break;
        case EPI_QKV: {
          if (nc < p.split_n) {
            if (row_ok)
              store_row32<OT>(reinterpret_cast<OT*>(p.out) + b * p.out_bstride + (long long)row * p.ldo + nc, v);
          } else if (row_ok) {
            OT* dst = reinterpret_cast<OT*>(p.out2) + b * p.out2_bstride + (long long)(nc - p.split_n) * p.ld2 + row;
#pragma unroll
            for (int i = 0; i < 32; ++i) store_one(dst + (long long)i * p.ld2, v[i]);   // lanes -> consecutive rows
          }
        } break;
        case EPI_EULER: {
          if (row_ok) {
            const long long base = ((long long)b * p.N + nc) * p.T_cf + row;
            const float dt = p.dt_ptr != nullptr ? __ldg(p.dt_ptr) : p.dt;
            const bool blend = p.blend_mask != nullptr && p.blend_mask[(long long)b * p.T_cf + row] != 0;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              if (nc + i < p.N) {
                const float vv = v[i] * mk;
                const long long idx = base + (long long)i * p.T_cf;
                if (p.v_out != nullptr) p.v_out[idx] = vv;
                if (p.x_out != nullptr) p.x_out[idx] = blend ? p.blend_src[idx] : p.x_in[idx] + dt * vv;
              }
            }
          }
        } break;
        default: break;
      }
    }
    if (p.epi == EPI_STORE_F32 && p.gn_part != nullptr) {
      asm volatile("bar.sync 1, 128;" ::: "memory");   // the four epilogue warps only
      const int t = threadIdx.x - 64;
      if (t >= 0 && t < (BN / 64) * 2) {
        const int g = t >> 1, w = t & 1;
        const float s = gn_red[(0 * (BN / 64) + g) * 2 + w] + gn_red[(1 * (BN / 64) + g) * 2 + w] +
                        gn_red[(2 * (BN / 64) + g) * 2 + w] + gn_red[(3 * (BN / 64) + g) * 2 + w];
        const int group = (n0 >> 6) + g;
        p.gn_part[(((long long)b * gridDim.y + blockIdx.y) * 8 + group) * 2 + w] = s;
      }
    }
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<BN>(tmem_base);
}

}  // namespace srn
