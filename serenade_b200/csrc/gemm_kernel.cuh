// Persistent warp-specialised tcgen05 GEMM kernel (see gemm.cuh for the operand model).
//
// One CTA per SM loops over output tiles (128 x BN); 320 threads:
//   warp 0      TMA producer      (smem ring of STAGES x {A 128x128B, B BNx128B}, SWIZZLE_128B)
//   warp 1      tcgen05.mma issuer (one elected lane), accumulators in TMEM, two accumulator stages
//   warps 2..9  epilogue: TMEM -> registers -> fused epilogue -> global, overlapped with the next
//               tile's main loop through the tmem_full / tmem_empty barriers
#pragma once

namespace srn {

template <int BN>
struct GemmCfg {
  static constexpr int kStages = BN == 256 ? 4 : 6;
  static constexpr int kABytes = 128 * 128;
  static constexpr int kBBytes = BN * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarOff = kStages * kStageBytes;       // barriers + tmem slot (256 B)
  static constexpr int kRedOff = kBarOff + 256;               // GroupNorm partial reduction scratch (256 B)
  static constexpr int kSmemBytes = kRedOff + 256 + 1024;     // + alignment slack
  static constexpr int kThreads = 320;
  static constexpr int kEpiThreads = 256;
  static constexpr int kTmemCols = 2 * BN;
};

template <typename OT>
__device__ __forceinline__ void store_row16(OT* dst, const float (&v)[16]);
template <>
__device__ __forceinline__ void store_row16<float>(float* dst, const float (&v)[16]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
    reinterpret_cast<float4*>(dst)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
}
template <>
__device__ __forceinline__ void store_row16<__nv_bfloat16>(__nv_bfloat16* dst, const float (&v)[16]) {
#pragma unroll
  for (int i = 0; i < 2; ++i) {
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

// One 32-column chunk of one accumulator row: bias/alpha were already applied to v.
template <typename OT>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, float (&v)[32], int b, int h, int row, int nc,
                                               bool row_ok, float mk, float& gs, float& gss) {
  const bool full_chunk = nc + 32 <= p.N;
  switch (p.epi) {
    case EPI_STORE_F32: {
      if (p.gn_part != nullptr && row_ok) {
#pragma unroll
        for (int i = 0; i < 32; ++i) { gs += v[i]; gss += v[i] * v[i]; }
      }
      if (row_ok) {
        float* dst = reinterpret_cast<float*>(p.out) + b * p.out_bstride + h * p.out_hstride + (long long)row * p.ldo + nc;
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
        OT* dst = reinterpret_cast<OT*>(p.out) + b * p.out_bstride + h * p.out_hstride + (long long)row * p.ldo + nc;
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
        store_row16<OT>(reinterpret_cast<OT*>(p.out) + b * p.out_bstride + (long long)row * p.ldo + (nc >> 1), o);
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

template <typename OT, int BN>
__global__ void __launch_bounds__(320, 1) gemm_kernel(const __grid_constant__ GemmParams p) {
  using TR = OpTraits<OT>;
  using CFG = GemmCfg<BN>;
  constexpr int STAGES = CFG::kStages;
  constexpr int BK = TR::kBK;
  constexpr int NCH = BN / 32;          // 32-column chunks per tile
  constexpr int NG = BN / 64;           // GroupNorm groups per tile

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + CFG::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;     // [2] accumulator stage ready for the epilogue
  uint64_t* tempty_bar = tfull_bar + 2;         // [2] accumulator stage drained by the epilogue
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  float* gn_red = reinterpret_cast<float*>(smem + CFG::kRedOff);   // [2 buffers][4 quarters][NG][2]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_mn = p.ntn * p.ntm;
  const int total_tiles = tiles_mn * p.nz;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&p.mapA[0]);
    tma_prefetch_desc(&p.mapA[1]);
    tma_prefetch_desc(&p.mapB);
  }
  if (warp == 1 && elect_one()) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 8);     // one arrival per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<CFG::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      // ------------------------------------------------------------ TMA producer
      uint32_t stage = 0, phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int z = tile / tiles_mn;
        const int rem = tile - z * tiles_mn;
        const int m0 = (rem / p.ntn) * 128;
        const int n0 = (rem % p.ntn) * BN;
        const int b = z / p.HB;
        const int h = z - b * p.HB;
        const int a_batch = p.a_zbatch ? z : b;
        const int b_batch = p.b_batch_mode == 0 ? 0 : (p.b_batch_mode == 1 ? b : z);
        const int b_row = n0 + h * p.b_hrow;
        int seg = 0, j = 0;
        for (int kb = 0; kb < p.nkb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * CFG::kStageBytes;
          uint8_t* sb = sa + CFG::kABytes;
          mbar_expect_tx(&full_bar[stage], CFG::kStageBytes);
          const KSeg sg = p.seg[seg];
          tma_load_3d(sa, &p.mapA[sg.src], &full_bar[stage], sg.col0 + j * BK + h * p.a_hcol, m0 + sg.shift, a_batch);
          tma_load_3d(sb, &p.mapB, &full_bar[stage], p.b_col0 + kb * BK + h * p.b_hcol, b_row, b_batch);
          if (++j == sg.nblk) { j = 0; ++seg; }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      // ------------------------------------------------------------ MMA issuer
      constexpr uint32_t idesc = make_idesc(TR::kFmt, 128, BN);
      uint32_t stage = 0, phase = 0;
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const uint32_t as = it & 1;
        mbar_wait(&tempty_bar[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tacc = tmem_base + as * BN;
        for (int kb = 0; kb < p.nkb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * CFG::kStageBytes);
          const uint64_t da = make_kmajor_sw128_desc(sa);
          const uint64_t db = make_kmajor_sw128_desc(sa + CFG::kABytes);
#pragma unroll
          for (int k = 0; k < 4; ++k)   // 4 x 32 bytes of K per 128-byte swizzled row
            umma<TR::kTF32>(tacc, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          tc_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        tc_commit(&tfull_bar[as]);
      }
    }
  } else {
    // -------------------------------------------------------------- epilogue (8 warps)
    const int q = warp & 3;                 // TMEM lane quarter this warp may access (warp id % 4)
    const int sub = (warp - 2) >> 2;        // which half of the tile's columns
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int z = tile / tiles_mn;
      const int rem = tile - z * tiles_mn;
      const int mt = rem / p.ntn;
      const int m0 = mt * 128;
      const int n0 = (rem % p.ntn) * BN;
      const int b = z / p.HB;
      const int h = z - b * p.HB;
      const uint32_t as = it & 1;
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < p.M;
      float mk = 1.0f;
      if (p.rowmask != nullptr && row_ok)
        mk = p.rowmask[(long long)b * p.mask_bstride + (long long)row * p.mask_step + p.mask_off] ? 1.0f : 0.0f;
      float* red = gn_red + as * (4 * NG * 2);
      mbar_wait(&tfull_bar[as], (it >> 1) & 1);
      tc_fence_after();
      float gs = 0.f, gss = 0.f;
#pragma unroll 1
      for (int cc = 0; cc < NCH / 2; ++cc) {
        const int c = sub * (NCH / 2) + cc;
        const int nc = n0 + c * 32;
        if (nc >= p.N) break;                  // warp-uniform
        uint32_t raw[32];
        tmem_ld_32x32b_x32(tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * BN + c * 32), raw);
        tmem_ld_wait();
        float v[32];
        const bool full_chunk = nc + 32 <= p.N;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float bi = 0.f;
          if (p.bias != nullptr && (full_chunk || nc + i < p.N)) bi = __ldg(p.bias + nc + i);
          v[i] = __uint_as_float(raw[i]) * p.alpha + bi;
        }
        epilogue_chunk<OT>(p, v, b, h, row, nc, row_ok, mk, gs, gss);
        if (p.gn_part != nullptr && (c & 1)) {   // a 64-channel group is complete
          const float s = warp_sum_f(gs), ss = warp_sum_f(gss);
          if (lane == 0) {
            red[(q * NG + (c >> 1)) * 2 + 0] = s;
            red[(q * NG + (c >> 1)) * 2 + 1] = ss;
          }
          gs = 0.f; gss = 0.f;
        }
      }
      // release the accumulator stage to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
      if (p.gn_part != nullptr) {
        asm volatile("bar.sync 1, 256;" ::: "memory");   // the eight epilogue warps only
        const int t = threadIdx.x - 64;
        if (t < NG * 2) {
          const int g = t >> 1, w = t & 1;
          const float s = red[(0 * NG + g) * 2 + w] + red[(1 * NG + g) * 2 + w] + red[(2 * NG + g) * 2 + w] +
                          red[(3 * NG + g) * 2 + w];
          const int group = (n0 >> 6) + g;
          p.gn_part[(((long long)b * p.ntm + mt) * 8 + group) * 2 + w] = s;
        }
      }
    }
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<CFG::kTmemCols>(tmem_base);
}

}  // namespace srn
