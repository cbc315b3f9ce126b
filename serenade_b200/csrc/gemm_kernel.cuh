// Persistent warp-specialised tcgen05 GEMM kernel (see gemm.cuh for the operand model).
//
// CTA pairs (cluster of 2, tcgen05 cta_group::2) loop over 256 x BN output tiles: each CTA of the pair loads its
// own 128 rows of A and its own half of the B rows by TMA, the leader CTA issues UMMA M=256 for both, and each
// CTA's TMEM receives the accumulator rows of its own 128 A rows (halves B traffic per CTA; the single-CTA
// version was bound by L2->SM bandwidth).  Per CTA:
//   warp 0      TMA producer      (smem ring of STAGES x {A 128x128B, B (BN/2)x128B}, SWIZZLE_128B)
//   warp 1      tcgen05.mma issuer (leader CTA, one elected lane), two TMEM accumulator stages
//   warps 2..   epilogue (8 or 16 warps): TMEM -> registers -> fused epilogue -> swizzled smem box -> TMA store,
//               overlapped with the next tile's main loop through the tmem_full / tmem_empty barriers
// The epilogue mode is a template parameter: ncu showed the epilogue warps latency-bound on ~520 instructions
// per 32-column chunk when the mode was a runtime switch with per-element predicates.
#pragma once

#include "gemm_epilogue.cuh"

namespace srn {

// Eight epilogue warps, two per lane quarter (each drains half of the tile's columns), two staging boxes per warp.
// Sixteen warps (four per scheduler, 64 accumulator columns each, one staging box per warp, 96 registers) were built
// and measured in round 2 for the K = 512 GEMMs whose tile loop is longer than their MMA time: slower everywhere
// (QKV 368 vs 334 us, exp 261 vs 246 us, P~ V 255 vs 244 us per C2 estimator call; profiles/r02_ab_epilogue_warps.txt).
// "Wide" 256 x 512 tiles for P~ V (one accumulator filling all 512 TMEM columns, A fetched once for both column
// halves: a quarter less L2 -> SM traffic per FLOP, but no epilogue / main loop overlap) were built and measured in
// round 2: 69.7 vs 67.0 us at C2's full rate, 26.3 vs 27.1 us at half rate, 132.4 vs 134.5 ms per 10-step solve at C5's
// shape (profiles/r02_ab_rowkernels_widepv.txt) - within noise, so the kernel keeps two accumulator stages throughout.
template <int BN>
struct GemmCfg {
  static constexpr int EW = 8;
  static constexpr int kStages = BN == 256 ? 4 : 5;
  static constexpr int kABytes = 128 * 128;            // this CTA's 128 rows of the pair's 256-row A tile
  static constexpr int kBBytes = (BN / 2) * 128;       // this CTA's half of the B tile's BN rows
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kOpBytes = kStages * kStageBytes;
  static constexpr int kSlices = EW / 4;                      // column slices of a tile (one epilogue warp per lane quarter each)
  static constexpr int kChunksPerWarp = (BN / 32) / kSlices;  // 32-column chunks per epilogue warp and tile
  static constexpr int kStgWarp = 8192;                       // staging bytes per epilogue warp (two 4 KB boxes)
  static constexpr int kStgOff = kOpBytes;                    // epilogue staging: EW warps x kStgWarp = 64 KB
  static constexpr int kBarOff = kStgOff + EW * kStgWarp;     // barriers + tmem slot (256 B)
  static constexpr int kRedOff = kBarOff + 256;               // GroupNorm partial reduction scratch (256 B)
  static constexpr int kBiasOff = kRedOff + 256;              // per epilogue warp: its bias slice (4 x BN floats in all)
  static constexpr int kWinBarOff = kBiasOff + 8 * 512;       // window mode: its own barriers (256 B)
  static constexpr int kSmemBytes = kWinBarOff + 256 + 1024;  // + alignment slack
  // window mode re-partitions the operand stages: two A windows, the rest is a ring of B stages
  static constexpr int kWinABytes = kWinRows * 128;
  static constexpr int kWinBStages = (kStages * kStageBytes - 2 * kWinABytes) / kBBytes;
  static_assert(kWinBStages >= 3 && 4 + 2 * kWinBStages <= 32, "window-mode barriers fit their 256 bytes");
  static constexpr int kThreads = 64 + 32 * EW;
  static constexpr int kTmemCols = 2 * BN;
  static_assert(kStgOff % 1024 == 0, "staging boxes must be 1024-byte aligned for SWIZZLE_128B");
  static_assert(kSmemBytes <= 227 * 1024, "shared memory per CTA");
};

#ifndef SRN_KEY_FASTPATH
#define SRN_KEY_FASTPATH 1
#endif

// One 32-column chunk of one accumulator row (bias / alpha already applied to v).
// Softmax modes: `racc` is this thread's running row max (ROWMAX) or row sum (EXP), `rowc` the row constant
// (EXP: rowmax * log2 e, PV: 1 / rowsum), `keybits` bit i = key column nc + i takes part.
template <typename OT, int EPI>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, const CUtensorMap* dmap, float (&v)[32], Stager& st,
                                               int lane, int b, int h, int z, int row, int row0, int nc, bool row_ok,
                                               float mk, float& gs, float& gss, float& racc, float rowc,
                                               uint32_t keybits, bool gn_ok) {
  const int batch = p.d_zbatch ? z : b;
  constexpr int kOtCols = 128 / (int)sizeof(OT);     // output columns per 128-byte staging row
  // fp32-stored operand tensors are consumed as TF32: round to nearest here (the tensor core truncates)
  auto round_op = [](float (&w)[32]) {
    if constexpr (sizeof(OT) == 4) {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        uint32_t u;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(w[i]));
        w[i] = __uint_as_float(u);
      }
    }
  };
  if constexpr (EPI == EPI_STORE_F32) {
    stager_acquire(st, lane);
    stager_put<float, 32>(st, lane, v);
    stager_flush<float>(st, lane, p, nc + h * p.d_hcol, row0, batch);
  } else if constexpr (EPI == EPI_STORE_GN) {
    // conv output that feeds a GroupNorm: statistics from the f32 accumulators (gn_ok: this row and tile group enter
    // them), the tensor itself in operand width - half the bytes here and in the row kernel that applies the norm; it
    // is not a GEMM operand, so the TF32-class mode keeps it unrounded f32
    if (gn_ok) {
#pragma unroll
      for (int i = 0; i < 32; ++i) { gs += v[i]; gss = fmaf(v[i], v[i], gss); }
    }
    stager_acquire(st, lane);
    stager_put<OT, 32>(st, lane, v);
    stager_flush<OT>(st, lane, p, nc + 32 - kOtCols, row0, batch, false, dmap);
  } else if constexpr (EPI == EPI_STORE_OT) {
    if (p.act_slope != 1.0f) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * p.act_slope;
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] *= mk;
    round_op(v);
    stager_acquire(st, lane);
    stager_put<OT, 32>(st, lane, v);
    stager_flush<OT>(st, lane, p, nc + 32 - kOtCols + h * p.d_hcol, row0, batch, false, dmap);
  } else if constexpr (EPI == EPI_RESID_F32 || EPI == EPI_RESID_OT) {     // RESID_OT: residual already added by the caller
    if constexpr (EPI == EPI_RESID_OT && sizeof(OT) == 2) {
      // one output box only (the warp's other box receives the residual, gemm_mainloop.cuh: kResidTma)
      if (st.fill == 0) {
        if (lane == 0) tma_store_wait_read<0>();
        __syncwarp();
      }
    } else {
      stager_acquire(st, lane);
    }
    if constexpr (EPI == EPI_RESID_F32) {      // out and residual are the same f32 tensor: TMA reduce-add in place
      stager_put<float, 32>(st, lane, v);
      stager_flush<float>(st, lane, p, nc, row0, batch, true);
    } else {
      // (a select, not a product: with padded-frame skipping the rows of padded frames may hold stale data upstream,
      //  and 0 * NaN would carry a NaN of an earlier call into the next convolution's operand)
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = mk != 0.f ? v[i] : 0.f;
      round_op(v);
      stager_put<OT, 32>(st, lane, v);
      stager_flush<OT>(st, lane, p, nc + 32 - kOtCols, row0, batch);
      if constexpr (sizeof(OT) == 2) st.buf = 0;
    }
  } else if constexpr (EPI == EPI_RESID_ACT) {
    // v is the updated f32 stream (the caller added the residual and stored it from the residual box): emit its
    // activated OT copy, the next convolution's operand, through the warp's other box
    if (st.fill == 0) {
      if (lane == 0) tma_store_wait_read<0>();
      __syncwarp();
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * p.act_slope;
    round_op(v);
    stager_put<OT, 32>(st, lane, v);
    stager_flush<OT>(st, lane, p, nc + 32 - kOtCols, row0, batch, false, &p.mapD2);
    st.buf = 0;
  } else if constexpr (EPI == EPI_GEGLU) {
    float o[16];
#pragma unroll
    for (int i = 0; i < 16; ++i)
      o[i] = v[2 * i] * ((SRN_GELU_TAIL && sizeof(OT) == 2) ? gelu_tail(v[2 * i + 1]) : gelu_erf(v[2 * i + 1]));
    if constexpr (sizeof(OT) == 4) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        uint32_t u;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(o[i]));
        o[i] = __uint_as_float(u);
      }
    }
    stager_acquire(st, lane);
    stager_put<OT, 16>(st, lane, o);
    stager_flush<OT>(st, lane, p, (nc >> 1) + 16 - kOtCols, row0, batch);
  } else if constexpr (EPI == EPI_QKV) {
    round_op(v);
    if (p.qk_norm2 != nullptr && nc < p.split_n) {      // (gs: this row's squared norm over the warp's Q | K columns)
#pragma unroll
      for (int i = 0; i < 32; ++i) gs = fmaf(v[i], v[i], gs);
    }
    if (nc < p.split_n) {
      stager_acquire(st, lane);
      stager_put<OT, 32>(st, lane, v);
      stager_flush<OT>(st, lane, p, nc + 32 - kOtCols, row0, batch);
    } else {
      stager_store_transposed<OT>(st, lane, v, &p.mapD2, row0, nc - p.split_n, batch);
    }
  } else if constexpr (EPI == EPI_ROWMAX) {
    if (SRN_KEY_FASTPATH && keybits == 0xffffffffu) {      // all 32 keys of the chunk take part (warp-uniform): no selects
      float m0 = racc, m1 = -INFINITY;
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        m0 = fmaxf(m0, v[i] * p.alpha);
        m1 = fmaxf(m1, v[i + 1] * p.alpha);
      }
      racc = fmaxf(m0, m1);
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if ((keybits >> i) & 1u) racc = fmaxf(racc, v[i] * p.alpha);
    }
  } else if constexpr (EPI == EPI_EXP) {
    // exponent clamp at 64 log2 e: never active on a valid query row (shifted rows sit near 0, shift-free entries are
    // proven within +-64), it only keeps the garbage rows of PADDED query frames finite in the shift-free case (their
    // norms do not enter the test, so that a clip's decision does not depend on the padding its batch gives it)
    constexpr float kExpArgMax = 92.332482f;
    // (32768 exponentials per CTA and tile = 2048 clocks of MUFU.EX2; moving every other one to the FMA pipes was
    //  measured and is no faster: the MUFU is not what bounds this epilogue)
    const float sc = p.alpha * 1.4426950408889634f;
    if (SRN_KEY_FASTPATH && keybits == 0xffffffffu) {
      // all 32 keys take part (every chunk but a row's last ones and padded batches): no per-element select, and the
      // row sum as four interleaved partial sums instead of one 32-long dependent chain.  The order of the additions
      // is fixed per 64-key slice whatever the tile shape, so "alone == inside a batch" stays bit-exact.
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float x = fminf(fmaf(v[i + k], sc, -rowc), kExpArgMax);
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(v[i + k]) : "f"(x));
        }
        s0 += v[i]; s1 += v[i + 1]; s2 += v[i + 2]; s3 += v[i + 3];
      }
      racc += (s0 + s1) + (s2 + s3);
    } else {
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float e;
          const float x = fminf(fmaf(v[i + k], sc, -rowc), kExpArgMax);
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x));
          v[i + k] = ((keybits >> (i + k)) & 1u) ? e : 0.f;
        }
        s0 += v[i]; s1 += v[i + 1]; s2 += v[i + 2]; s3 += v[i + 3];
      }
      racc += (s0 + s1) + (s2 + s3);
    }
    round_op(v);
    stager_acquire(st, lane);
    stager_put<OT, 32>(st, lane, v);
    stager_flush<OT>(st, lane, p, nc + 32 - kOtCols, row0, batch);
  } else if constexpr (EPI == EPI_PV) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] *= rowc;
    round_op(v);
    stager_acquire(st, lane);
    stager_put<OT, 32>(st, lane, v);
    stager_flush<OT>(st, lane, p, nc + 32 - kOtCols + h * p.d_hcol, row0, batch);
  } else if constexpr (EPI == EPI_EULER) {
    if (row_ok) {
      const long long base = ((long long)b * p.N + nc) * p.T_cf + row;
      const float dt = p.dt_ptr != nullptr ? __ldg(p.dt_ptr) : p.dt;
      const bool blend = p.blend_mask != nullptr && p.blend_mask[(long long)b * p.T_cf + row] != 0;
      // x_in and x_out are the same buffer: issue all loads of the chunk before the first store, otherwise every
      // load waits behind the previous (possibly aliasing) store and the chunk costs 32 serial memory latencies
      float xin[32];
      if (p.x_out != nullptr) {
        const float* src = blend ? p.blend_src : p.x_in;
#pragma unroll
        for (int i = 0; i < 32; ++i) xin[i] = nc + i < p.N ? src[base + (long long)i * p.T_cf] : 0.f;
      }
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        if (nc + i < p.N) {
          const float vv = v[i] * mk;
          const long long idx = base + (long long)i * p.T_cf;
          if (p.v_out != nullptr) p.v_out[idx] = vv;
          if (p.x_out != nullptr) p.x_out[idx] = blend ? xin[i] : xin[i] + dt * vv;
        }
      }
    }
  }
}

}  // namespace srn

#include "gemm_mainloop.cuh"
