// The __global__ kernel: role dispatch, TMA/MMA pipelines, epilogue tile loop.  Included by gemm_kernel.cuh.
//
// Variants that were built, measured on the B200 and NOT kept (round 1; the numbers are in DESIGN.md 3.4 and under
// profiles/r01_ab_*.txt, the code is in the git history before this file's round-2 rewrite): A-operand multicast over
// four-CTA clusters (only 132 of 148 SMs can hold 4-CTA clusters: no gain), an A-stationary tile order for the K = 512
// GEMMs (half the operand bytes, yet slower), attention passes in (batch, head) slices, exp2 on the FMA pipes, direct
// st.global epilogue stores; round 2: sixteen epilogue warps and 256 x 512 tiles (gemm_kernel.cuh), the e4m3 row-max
// pass of round 1 (its Q | K copy cost more than it saved once most entries run shift-free).
#pragma once

namespace srn {

#ifndef SRN_EPI_PREFETCH
#define SRN_EPI_PREFETCH 1
#endif
#if SRN_TRACE
#define SRN_STAMP(k) do { if (p.trace != nullptr) p.trace[blockIdx.x * 8 + (k)] = (unsigned long long)clock64(); } while (0)
#else
#define SRN_STAMP(k) do { } while (0)
#endif

template <typename OT, int BN, int EPI>
__global__ void __launch_bounds__(GemmCfg<BN>::kThreads, 1) gemm_kernel(const __grid_constant__ GemmParams p) {
  using TR = OpTraits<OT>;
  using CFG = GemmCfg<BN>;
  constexpr int EW = CFG::EW;                   // epilogue warps
  constexpr int STAGES = CFG::kStages;
  constexpr int BK = TR::kBK;
  constexpr int NCH = BN / 32;                  // 32-column chunks per tile
  constexpr int CPW = CFG::kChunksPerWarp;      // ... per epilogue warp
  constexpr int NG = BN / 64;                   // GroupNorm groups per tile

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + CFG::kBarOff);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;     // [2] accumulator stage ready for the epilogue
  uint64_t* tempty_bar = tfull_bar + 2;         // [2] accumulator stage drained by the epilogue
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  uint64_t* resid_bar = reinterpret_cast<uint64_t*>(smem + CFG::kBarOff + 128);   // [8] one per epilogue warp (kResidTma)
  uint64_t* wa_full = reinterpret_cast<uint64_t*>(smem + CFG::kWinBarOff);      // window mode: [2] A windows
  uint64_t* wa_empty = wa_full + 2;
  uint64_t* wb_full = wa_empty + 2;                                             // [kWinBStages] B ring
  uint64_t* wb_empty = wb_full + CFG::kWinBStages;
  float* gn_red = reinterpret_cast<float*>(smem + CFG::kRedOff);   // [2 buffers][4 quarters][NG][2]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_mn = p.ntn * p.ntm;
  const int total_tiles = tiles_mn * p.nz;       // units of work per CTA pair
  const uint32_t rank = cluster_ctarank() & 1u;  // rank inside the CTA pair, 0 = leader (issues the MMAs)
  constexpr uint32_t leader = 0;
  const int pair0 = blockIdx.x / 2;
  const int npairs = gridDim.x / 2;
  // tiles of this pair: pair0, pair0 + npairs, ...
  const int t_begin = pair0, t_end = total_tiles, t_step = npairs;
  // p.rev: the same tiles in the opposite order (last (z, mi, ni) first).  A launch whose input is larger than the L2
  // and was written by the launch before it in ascending tile order finds the END of that tensor still cached and its
  // beginning evicted; reading in the writer's order is the worst case for an LRU-like cache (every line misses).
  auto logical = [&](int tile) -> int { return p.rev ? total_tiles - 1 - tile : tile; };

  pdl_launch_dependents();
  if (threadIdx.x == 0) SRN_STAMP(0);
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&p.mapA[0]);
    tma_prefetch_desc(&p.mapA[1]);
    tma_prefetch_desc(&p.mapB);
    tma_prefetch_desc(&p.mapD);
    if (EPI == EPI_RESID_OT && sizeof(OT) == 2) tma_prefetch_desc(&p.mapR);
    if (p.grp_ntn0 > 0) { tma_prefetch_desc(&p.mapB2); tma_prefetch_desc(&p.mapD2); }
  }
  if (warp == 1 && elect_one()) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 2 * EW);    // one arrival per epilogue warp of both CTAs (leader's copy is used)
    }
    for (int s = 0; s < 8; ++s) mbar_init(&resid_bar[s], 1);
    if (p.win) {
      for (int s = 0; s < 2; ++s) { mbar_init(&wa_full[s], 1); mbar_init(&wa_empty[s], 1); }
      for (int s = 0; s < CFG::kWinBStages; ++s) { mbar_init(&wb_full[s], 1); mbar_init(&wb_empty[s], 1); }
    }
    fence_barrier_init();
  }
  // Padded-frame skipping: the extents of this launch's batch entries -> shared memory, BEFORE the dependency wait.
  // That is safe although an earlier kernel of the same stream wrote them: the launch right after row_extent_kernel is
  // a fully serialised one (engine_run.inl: prepass), so no later kernel can even start before the extents are final.
  // (made visible to all warps by the cluster barrier below; host side guarantees <= 256 entries)
  __shared__ int s_ext[256];
  const bool use_ext = p.row_extent != nullptr;
  if (use_ext) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_ext[i] = i * p.HB < p.nz ? __ldcg(p.row_extent + i) : 0;
  }
  if (warp == 2) tmem_alloc_2sm<CFG::kTmemCols>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();     // everything above overlapped with the previous kernel's tail
  if (threadIdx.x == 0) SRN_STAMP(1);
  // Shift-free softmax, decided per (batch, head) entry z (so a clip's result never depends on what else is in the
  // batch): s_ns[z] != 0 means every score of that entry is provably within +-64 - its row-max tiles are skipped
  // and its exp tiles run with shift 0.
  __shared__ uint8_t s_ns[(EPI == EPI_ROWMAX || EPI == EPI_EXP) ? 1024 : 4];
  if constexpr (EPI == EPI_ROWMAX) {
    // The test itself (every CTA, redundantly: 16 L2 loads per entry).  qk_norm2[z * 16 + {0..7 Q, 8..15 K parts}] holds
    // the largest squared norm of each column part of any Q / K row of entry z (atomic maxima from the QKV epilogue); by
    // Cauchy-Schwarz |q . k| alpha <= sqrt(sum maxq * sum maxk) alpha.  Bound <= 64: exp(score) can neither overflow
    // (e^64 = 6.2e27: a row of 10^5 keys times |v| <= 10^3 stays below 1e36 in the f32 accumulators of P~ V) nor
    // underflow (e^-64 = 1.6e-28), and softmax is invariant to the shift.
    for (int z = threadIdx.x; z < p.nz && z < 1024; z += blockDim.x) {
      uint8_t ok = 0;
      if (p.qk_norm2 != nullptr) {
        float q = 0.f, k = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          q += p.qk_norm2[z * 16 + i];
          k += p.qk_norm2[z * 16 + 8 + i];
        }
        const float bound = sqrtf(q * k) * p.alpha * 1.01f;    // (1 %: the operands are rounded after the norms were taken)
        ok = bound <= p.noshift_max ? 1 : 0;                    // (NaN / Inf fail the test)
        if (blockIdx.x == 0) p.noshift[z] = ok;                 // for the exp pass
      }
      if (blockIdx.x == 0 && p.ns_count != nullptr) {           // (statistics for srn_softmax_stats)
        if (ok) atomicAdd(p.ns_count, 1ull);
        atomicAdd(p.ns_count + 1, 1ull);
      }
      s_ns[z] = ok;
    }
    __syncthreads();
  }
  if constexpr (EPI == EPI_EXP) {
    for (int z = threadIdx.x; z < p.nz && z < 1024; z += blockDim.x)
      s_ns[z] = p.noshift != nullptr ? (uint8_t)(*reinterpret_cast<const volatile int*>(p.noshift + z)) : (uint8_t)0;
    if (p.qk_norm2 != nullptr && blockIdx.x == 0)     // the row-max pass has read the maxima: clear them
      for (int i = threadIdx.x; i < p.nz_norm * 16; i += blockDim.x) p.qk_norm2[i] = 0.f;
    __syncthreads();
  }
  // (row-max pass: tiles of shift-free entries are skipped by all three roles)
  auto skip_tile = [&](int z) -> bool {
    if constexpr (EPI == EPI_ROWMAX) return p.qk_norm2 != nullptr && s_ns[z] != 0;     // (null beyond 1024 entries)
    else return false;
  };
  // Padded-frame skipping: extents of the batch entries of this launch (host side guarantees <= 256 of them)
  // (b = batch entry, mi / ni = row / column tile inside the entry: the callers have them anyway - integer divisions
  //  by run-time values are ~150-clock dependent chains, and the epilogue warps' per-tile chain is what bounds the
  //  K = 512 GEMMs)
  auto tile_ext = [&](int b) -> int { return use_ext ? s_ext[b] : 0x7fffffff; };
  auto skip_pad = [&](int ext, int mi, int ni) -> bool {
    if (!use_ext) return false;
    if (mi * 256 >= ext + p.ext_halo) return true;        // the pair's 256 rows are all padding
    return p.ext_keys == 1 && ni * BN >= ext;              // ... or all its keys are
  };
  // Tile coordinates kept incrementally (tile -> tile + t_step) for the MMA-issuing lane: three run-time integer
  // divisions per tile in front of its MMAs cost the mainloop-bound launches ~450 clocks per tile.
  struct TileCoord {
    int z, mi, ni, a, c, d, ntn, ntm;
    bool back;
    __device__ __forceinline__ void init(int tile, int step, int ntn_, int ntm_, bool back_) {
      ntn = ntn_; ntm = ntm_; back = back_;
      const int tmn = ntn * ntm;
      z = tile / tmn;
      const int rem = tile - z * tmn;
      mi = rem / ntn; ni = rem - mi * ntn;
      a = step / tmn;
      const int r1 = step - a * tmn;
      c = r1 / ntn; d = r1 - c * ntn;
    }
    __device__ __forceinline__ void advance() {
      if (!back) {
        ni += d;
        if (ni >= ntn) { ni -= ntn; ++mi; }
        mi += c;
        if (mi >= ntm) { mi -= ntm; ++z; }
        z += a;
      } else {
        ni -= d;
        if (ni < 0) { ni += ntn; --mi; }
        mi -= c;
        if (mi < 0) { mi += ntm; --z; }
        z -= a;
      }
    }
  };
  auto tile_nkb = [&](int ext, bool g1) -> int {
    if (g1) return p.grp_nkb;
    if (use_ext && p.ext_keys == 2) { const int n = (ext + BK - 1) / BK; return n < p.nkb ? n : p.nkb; }   // P~ V: K blocks up to the extent
    return p.nkb;
  };
  auto in_group1 = [&](int ni) -> bool { return p.grp_ntn0 > 0 && ni >= p.grp_ntn0; };

  if (warp == 0) {
    if (elect_one()) {
      // ------------------------------------------------------------ TMA producer
      uint32_t stage = 0, phase = 0;
      if (p.win) {
        // window mode: per K block one A window (all the segments' rows) + one B box per segment
        uint32_t wa = 0, wap = 0, wb = 0, wbp = 0;
        const int kcb = p.seg[0].nblk;
        for (int tile = t_begin; tile < t_end; tile += t_step) {
          const int lt = logical(tile);
          const int z = lt / tiles_mn;
          const int rem = lt % tiles_mn;
          const int mi = rem / p.ntn, ni = rem - mi * p.ntn;
          const int b = z / p.HB;
          if (skip_pad(tile_ext(b), mi, ni)) continue;
          const int m0 = mi * 256 + (int)rank * 128;
          const int n0 = ni * BN;
          const int h = z - b * p.HB;
          const int a_batch = p.a_zbatch ? z : b;
          const int b_batch = p.b_batch_mode == 0 ? 0 : (p.b_batch_mode == 1 ? b : z);
          const int b_row = n0 + (int)rank * (BN / 2) + h * p.b_hrow;
          for (int cb = 0; cb < kcb; ++cb) {
            mbar_wait(&wa_empty[wa], wap ^ 1);
            if (rank == 0) mbar_expect_tx(&wa_full[wa], 2 * CFG::kWinABytes);
            tma_load_3d_2sm(smem + wa * CFG::kWinABytes, &p.mapA[0], map_to_cta(&wa_full[wa], leader),
                            p.seg[0].col0 + cb * BK + h * p.a_hcol, m0 + p.win_min, a_batch);
            if (++wa == 2) { wa = 0; wap ^= 1; }
            for (int s = 0; s < p.nseg; ++s) {
              mbar_wait(&wb_empty[wb], wbp ^ 1);
              if (rank == 0) mbar_expect_tx(&wb_full[wb], 2 * CFG::kBBytes);
              tma_load_3d_2sm(smem + 2 * CFG::kWinABytes + wb * CFG::kBBytes, &p.mapB, map_to_cta(&wb_full[wb], leader),
                              p.b_col0 + (s * kcb + cb) * BK + h * p.b_hcol, b_row, b_batch);
              if (++wb == (uint32_t)CFG::kWinBStages) { wb = 0; wbp ^= 1; }
            }
          }
        }
      } else {      // (the per-segment boxes of the plain path)
        for (int tile = t_begin; tile < t_end; tile += t_step) {
          const int lt = logical(tile);
          const int z = lt / tiles_mn;
          if (skip_tile(z)) continue;
          const int rem = lt % tiles_mn;
          const int mi = rem / p.ntn, ni = rem - mi * p.ntn;
          const int b = z / p.HB;
          const int ext = tile_ext(b);
          if (skip_pad(ext, mi, ni)) continue;
          const bool g1 = in_group1(ni);
          const int nkb_t = tile_nkb(ext, g1);
          const int m0 = mi * 256 + (int)rank * 128;
          const int n0 = (g1 ? ni - p.grp_ntn0 : ni) * BN;     // row of this tile group's B operand
          const CUtensorMap* mB = g1 ? &p.mapB2 : &p.mapB;
          const int h = z - b * p.HB;
          const int a_batch = p.a_zbatch ? z : b;
          const int b_batch = p.b_batch_mode == 0 ? 0 : (p.b_batch_mode == 1 ? b : z);
          const int b_row = n0 + (int)rank * (BN / 2) + h * p.b_hrow;
          int seg = g1 ? p.grp_seg0 : 0, j = 0;
          for (int kb = 0; kb < nkb_t; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* sa = smem + stage * CFG::kStageBytes;
            uint8_t* sb = sa + CFG::kABytes;
            // both CTAs' bytes complete on the pair LEADER's full barrier; only the leader arms it (for both)
            const uint32_t fb = map_to_cta(&full_bar[stage], leader);
            const KSeg sg = p.seg[seg];
            if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * CFG::kStageBytes);
            tma_load_3d_2sm(sa, &p.mapA[sg.src], fb, sg.col0 + j * BK + h * p.a_hcol, m0 + sg.shift, a_batch);
            tma_load_3d_2sm(sb, mB, fb, p.b_col0 + kb * BK + h * p.b_hcol, b_row, b_batch);
            if (++j == sg.nblk) { j = 0; ++seg; }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0 && elect_one()) {
      // ------------------------------------------------------------ MMA issuer (leader CTA only, M = 256)
      constexpr uint32_t idesc = make_idesc(TR::kFmt, 256, BN);
      uint32_t stage = 0, phase = 0;
      uint32_t it = 0;
      TileCoord tc;
      tc.init(logical(t_begin), t_step, p.ntn, p.ntm, p.rev != 0);
      int cur_z = -1, cur_ext = 0x7fffffff;
      if (p.win) {
        uint32_t wa = 0, wap = 0, wb = 0, wbp = 0;
        const int kcb = p.seg[0].nblk;
        const bool boff = (p.win & 2) != 0;
        for (int tile = t_begin; tile < t_end; tile += t_step, ++it, tc.advance()) {
          if (use_ext) {
            if (tc.z != cur_z) { cur_z = tc.z; cur_ext = tile_ext(tc.z / p.HB); }
            if (skip_pad(cur_ext, tc.mi, tc.ni)) { --it; continue; }
          }
          const uint32_t as = it & 1;
          mbar_wait(&tempty_bar[as], ((it >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t tacc = tmem_base + as * BN;
          for (int cb = 0; cb < kcb; ++cb) {
            mbar_wait(&wa_full[wa], wap);
            tc_fence_after();
            const uint32_t wbase = smem_u32(smem + wa * CFG::kWinABytes);
            for (int s = 0; s < p.nseg; ++s) {
              mbar_wait(&wb_full[wb], wbp);
              tc_fence_after();
              // the segment's 128 rows start (shift - win_min) rows into the window
              const uint64_t da = make_kmajor_sw128_desc_rows(wbase + (uint32_t)(p.seg[s].shift - p.win_min) * 128u, boff);
              const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem + 2 * CFG::kWinABytes + wb * CFG::kBBytes));
#pragma unroll
              for (int k = 0; k < 4; ++k) umma_pair<TR::kKind>(tacc, da + 2 * k, db + 2 * k, idesc, (cb | s | k) != 0);
              tc_commit_mask(&wb_empty[wb], 0x3);
              if (++wb == (uint32_t)CFG::kWinBStages) { wb = 0; wbp ^= 1; }
            }
            tc_commit_mask(&wa_empty[wa], 0x3);
            if (++wa == 2) { wa = 0; wap ^= 1; }
          }
          tc_commit_mask(&tfull_bar[as], 0x3);
        }
      } else {    // (plain path)
        for (int tile = t_begin; tile < t_end; tile += t_step, ++it, tc.advance()) {
          if (skip_tile(tc.z)) { --it; continue; }
          if (use_ext) {
            if (tc.z != cur_z) { cur_z = tc.z; cur_ext = tile_ext(tc.z / p.HB); }
            if (skip_pad(cur_ext, tc.mi, tc.ni)) { --it; continue; }
          }
          const int nkb_t = tile_nkb(cur_ext, in_group1(tc.ni));
          const uint32_t as = it & 1;
          mbar_wait(&tempty_bar[as], ((it >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t tacc = tmem_base + as * BN;
          for (int kb = 0; kb < nkb_t; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            if (it == 0 && kb == 0) SRN_STAMP(2);
            const uint32_t sa = smem_u32(smem + stage * CFG::kStageBytes);
            const uint64_t da = make_kmajor_sw128_desc(sa);
            const uint64_t db = make_kmajor_sw128_desc(sa + CFG::kABytes);
#pragma unroll
            for (int k = 0; k < 4; ++k)   // 4 x 32 bytes of K per 128-byte swizzled row
              umma_pair<TR::kKind>(tacc, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            tc_commit_mask(&empty_bar[stage], 0x3);     // frees the stage in both CTAs
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          tc_commit_mask(&tfull_bar[as], 0x3);          // both CTAs' epilogues may drain their half
          SRN_STAMP(3);
        }
      }
    }
  } else {
    // -------------------------------------------------------------- epilogue (EW warps)
    const int q = warp & 3;                 // TMEM lane quarter this warp may access (warp id % 4)
    const int sub = (warp - 2) >> 2;        // which slice of the tile's columns (EW / 4 slices)
    Stager st;
    st.base = smem_u32(smem + CFG::kStgOff + (warp - 2) * CFG::kStgWarp);
    st.buf = 0;
    st.fill = 0;
    st.one = 0u;
    uint32_t rphase = 0;      // parity of this warp's residual barrier (kResidTma)
    const bool has_alpha = p.alpha != 1.0f;
    float* bias_s = reinterpret_cast<float*>(smem + CFG::kBiasOff) + (warp - 2) * (CPW * 32);   // this warp's bias slice
    const uint32_t tepi_addr = map_to_cta(&tempty_bar[0], leader);
    uint32_t it = 0;
    // Per-row / per-key side inputs of the softmax epilogues, requested one tile ahead (SRN_EPI_PREFETCH): the clock64
    // trace showed expscores at 9.9 k clocks per tile with a 4.6 k chunk loop - the rest was the serial chain of
    // dependent L2 round trips (16 slice maxima four at a time, then the key mask) in front of every tile.
    float pre_s[16];          // slice statistics 0..15 of this lane's row
    uint8_t pre_k[CPW];       // key-mask bytes of this lane's key in each of the warp's chunks (0 beyond N)
    auto stat_issue = [&](int zz, int rr, bool ok, int j0, int cnt, float (&s)[16]) {
      const float* sp = (EPI == EPI_EXP ? p.stat_max : p.stat_sum) + (long long)zz * p.stat_n * p.stat_ld + rr;
#pragma unroll
      for (int u = 0; u < 16; ++u)
        s[u] = (ok && j0 + u < cnt) ? sp[(long long)(j0 + u) * p.stat_ld] : (EPI == EPI_EXP ? -INFINITY : 0.f);
    };
    auto key_issue = [&](int bb, int key0, uint8_t (&k)[CPW]) {
#pragma unroll
      for (int cc = 0; cc < CPW; ++cc) {
        const int key = key0 + cc * 32 + lane;
        k[cc] = key < p.N ? p.keymask[(long long)bb * p.keymask_bstride + (long long)key * p.keymask_step] : (uint8_t)0;
      }
    };
    int pre_tile = -1;        // the tile pre_s / pre_k were requested for
    for (int tile = t_begin; tile < t_end; tile += t_step, ++it) {
      const int lt = logical(tile);
      const int z = lt / tiles_mn;
      if (skip_tile(z)) { --it; continue; }
      const bool noshift = EPI == EPI_EXP && p.noshift != nullptr && s_ns[z] != 0;
      const int rem = lt % tiles_mn;
      const int mi = rem / p.ntn, ni = rem - mi * p.ntn;
      const int b = z / p.HB;
      const int h = z - b * p.HB;
      const int ext = tile_ext(b);
      // tile group (gemm.cuh): group 1 has its own output tensor, bias and mask offset, and no GroupNorm statistics
      const bool g1 = in_group1(ni);
      const int n0 = (g1 ? ni - p.grp_ntn0 : ni) * BN;
      const float* bias = g1 ? p.bias2 : p.bias;
      const CUtensorMap* dmap = g1 ? &p.mapD2 : &p.mapD;
      const bool has_bias = bias != nullptr;
      const bool gn_on = EPI == EPI_STORE_GN && p.gn_part != nullptr && !g1;
      if (skip_pad(ext, mi, ni)) {
        if constexpr (EPI == EPI_STORE_GN) {
          // a skipped convolution tile is all bias: its GroupNorm partial sums (the reference's statistics include the
          // padded frames) are rows x sum(bias), rows x sum(bias^2) per 64-channel group; none in valid-frame mode
          const int t = threadIdx.x - 64;
          if (gn_on && t < NG * 2 && n0 < p.N) {
            const int g = t >> 1, w = t & 1;
            const int mts = mi * 2 + (int)rank;
            int rows = p.M - mts * 128;
            rows = rows < 0 ? 0 : (rows > 128 ? 128 : rows);
            if (p.gn_mask != nullptr) rows = 0;
            float sb = 0.f;
            for (int c = 0; c < 64; ++c) {
              const float bv = has_bias ? bias[n0 + g * 64 + c] : 0.f;
              sb += w ? bv * bv : bv;
            }
            p.gn_part[(((long long)b * (2 * p.ntm) + mts) * 8 + (n0 >> 6) + g) * 2 + w] = sb * (float)rows;
          }
        }
        --it;
        continue;
      }
      // statistics slices of this entry that were written / may be read: up to the extent
      const int stat_n_t = (use_ext && p.ext_keys != 0) ? min(p.stat_n, (ext + 63) >> 6) : p.stat_n;
      const int mt = mi * 2 + (int)rank;     // 128-row tile index inside the batch entry
      const int m0 = mt * 128;
      const uint32_t as = it & 1;
      const int row0 = m0 + q * 32;
      const int row = row0 + lane;
      const bool row_ok = row < p.M;
      float mk = 1.0f;
      if (p.rowmask != nullptr && row_ok)
        mk = p.rowmask[(long long)b * p.mask_bstride + (long long)row * p.mask_step + (g1 ? p.mask_off2 : p.mask_off)] ? 1.0f : 0.0f;
      bool gn_ok = row_ok && gn_on;      // does this row enter the GroupNorm statistics (valid-frame mode: unmasked rows only)
      if constexpr (EPI == EPI_STORE_GN) {
        if (p.gn_mask != nullptr && gn_ok)
          gn_ok = p.gn_mask[(long long)b * p.mask_bstride + (long long)row * p.mask_step] != 0;
      }
      float* red = gn_red + as * (4 * NG * 2);
      const int c_first = sub * CPW;
      int nvalid = (p.N - (n0 + c_first * 32) + 31) / 32;     // chunks of this warp that start inside N
      nvalid = nvalid < 0 ? 0 : (nvalid > CPW ? CPW : nvalid);
      // 2-byte outputs fill one 128-byte staging row per TWO chunks: never stop on a half-filled box (N = keys
      // need not be a multiple of 64 for EPI_EXP; columns >= N hold exact zeros and are clipped by the store)
      if constexpr (sizeof(OT) == 2 && EPI == EPI_EXP) nvalid = (nvalid + 1) & ~1;
      const uint32_t trow = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * BN + c_first * 32);
      // ---- everything that does not depend on the accumulator is fetched BEFORE waiting for it: with few epilogue
      // warps per scheduler any global-load latency left inside the chunk loop is fully exposed (ncu: long_sb)
      float gs = 0.f, gss = 0.f;
      float racc = EPI == EPI_ROWMAX ? -INFINITY : 0.f, rowc = 0.f;
      if constexpr (EPI == EPI_EXP || EPI == EPI_PV) {     // per-row constants from the statistics of the earlier pass
        // slices 0..15 of this row were requested while the previous tile's chunks were processed (pre_s); only the
        // first tile of a CTA and rows with more than 16 slices wait for loads here, 16 in flight at a time
        float a = EPI == EPI_EXP ? -INFINITY : 0.f;
        for (int j0 = 0; j0 < ((EPI == EPI_EXP && noshift) ? 0 : stat_n_t); j0 += 16) {
          if (j0 > 0 || !SRN_EPI_PREFETCH || pre_tile != tile) stat_issue(z, row, row_ok, j0, stat_n_t, pre_s);
#pragma unroll
          for (int u = 0; u < 16; ++u) a = EPI == EPI_EXP ? fmaxf(a, pre_s[u]) : a + pre_s[u];
        }
        if (row_ok) {
          if constexpr (EPI == EPI_EXP) rowc = (a == -INFINITY || noshift ? 0.f : a) * 1.4426950408889634f;   // fully masked row: P = 0
          else rowc = a > 0.f ? 1.0f / a : 0.f;
        }
      }
      uint32_t keybits[CPW];
      if constexpr (EPI == EPI_ROWMAX || EPI == EPI_EXP) {
        if (!SRN_EPI_PREFETCH || pre_tile != tile) key_issue(b, n0 + c_first * 32, pre_k);
#pragma unroll
        for (int cc = 0; cc < CPW; ++cc) keybits[cc] = __ballot_sync(0xffffffffu, pre_k[cc] != 0);
      }
      if (has_bias) {       // this warp's bias slice -> smem (read back as broadcasts, one vector per 4 columns)
        const int col = n0 + c_first * 32 + lane * 4;
        float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (lane < CPW * 8 && col < p.N) bv = __ldg(reinterpret_cast<const float4*>(bias + col));
        __syncwarp();       // the previous tile's reads of this slice are done
        if (lane < CPW * 8) *reinterpret_cast<float4*>(bias_s + lane * 4) = bv;
        __syncwarp();
      }
      // (EPI_RESID_F32 adds its residual in the L2 with a TMA reduce-add store: nothing to load)
      // EPI_RESID_OT with 2-byte outputs: lane = row makes per-lane loads of the f32 residual 32 cache lines per
      // instruction (clock64 trace: 2750 clk per chunk, all L1 wavefronts).  Its two output boxes per tile need only
      // ONE staging box in flight, so the warp's second box receives the residual chunk by chunk through TMA, one
      // chunk ahead (kResidTma); the other precisions keep the register path.
      // EPI_RESID_ACT: the same residual box, but the sum goes back INTO the box and is stored from there as the f32
      // stream (mapD) while its activated 2-byte copy leaves through the warp's other box (mapD2).
      constexpr bool kAct = EPI == EPI_RESID_ACT;
      constexpr bool kResidTma = (EPI == EPI_RESID_OT && sizeof(OT) == 2) || kAct;
      constexpr bool kResid = EPI == EPI_RESID_OT && !kResidTma;
      float4 r4[kResid ? 8 : 1];
      const float4* rs = nullptr;
      if constexpr (kResid) {
        rs = reinterpret_cast<const float4*>(p.resid + b * p.resid_bstride + (long long)row * p.ldr + n0 + c_first * 32);
        if (row_ok && nvalid > 0) {
#pragma unroll
          for (int i = 0; i < 8; ++i) r4[i] = rs[i];
        }
      }
      const bool has_res = !kAct || p.resid != nullptr;      // EPI_RESID_ACT without a residual: store + activated copy
      uint8_t* const rbox_ptr = smem + CFG::kStgOff + (warp - 2) * CFG::kStgWarp + 4096;      // (kResidTma: EW = 8, two boxes)
      if constexpr (kResidTma) {
        __syncwarp();      // the previous tile's reads of the residual box are done
        if (lane == 0 && nvalid > 0 && has_res) {
          if constexpr (kAct) tma_store_wait_read<0>();      // ... and so is the store of its last chunk
          mbar_expect_tx(&resid_bar[warp - 2], 4096);
          tma_load_3d(reinterpret_cast<void*>(rbox_ptr), &p.mapR, &resid_bar[warp - 2], n0 + c_first * 32, row0, b);
        }
      }
      mbar_wait(&tfull_bar[as], (it >> 1) & 1);
      tc_fence_after();
      if (threadIdx.x == 64) { if (it == 0) SRN_STAMP(4); SRN_STAMP(5); }
      uint32_t raw[32];
      if (nvalid > 0) tmem_ld_32x32b_x32(trow, raw);
      else {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(tepi_addr + as * 8);
      }
      if constexpr (SRN_EPI_PREFETCH && (EPI == EPI_EXP || EPI == EPI_PV || EPI == EPI_ROWMAX)) {
        // the next tile's side inputs: in flight during this tile's chunks
        const int tn = tile + t_step;
        if (tn < t_end) {
          const int zn = logical(tn) / tiles_mn;
          const int remn = logical(tn) % tiles_mn;
          pre_tile = tn;        // (a skipped tile's values are simply never used)
          if constexpr (EPI != EPI_ROWMAX) {
            const int rown = ((remn / p.ntn) * 2 + (int)rank) * 128 + q * 32 + lane;
            const int extn = tile_ext(zn / p.HB);
            const int cntn = (use_ext && p.ext_keys != 0) ? min(p.stat_n, (extn + 63) >> 6) : p.stat_n;
            if (!(EPI == EPI_EXP && p.noshift != nullptr && s_ns[zn] != 0)) stat_issue(zn, rown, rown < p.M, 0, cntn, pre_s);
          }
          if constexpr (EPI != EPI_PV) key_issue(zn / p.HB, (remn % p.ntn) * BN + c_first * 32, pre_k);
        }
      }
#pragma unroll 1
      for (int cc = 0; cc < nvalid; ++cc) {
        const int c = c_first + cc;
        const int nc = n0 + c * 32;
        tmem_ld_wait();
        float v[32];
        if (has_bias) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 bb = *reinterpret_cast<const float4*>(bias_s + cc * 32 + i * 4);
            v[4 * i + 0] = __uint_as_float(raw[4 * i + 0]) + bb.x;
            v[4 * i + 1] = __uint_as_float(raw[4 * i + 1]) + bb.y;
            v[4 * i + 2] = __uint_as_float(raw[4 * i + 2]) + bb.z;
            v[4 * i + 3] = __uint_as_float(raw[4 * i + 3]) + bb.w;
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);
        }
        if (cc + 1 < nvalid) {
          tmem_ld_32x32b_x32(trow + (cc + 1) * 32, raw);   // overlaps with this chunk's epilogue
        } else {
          // the last accumulator chunk is in registers: hand the TMEM stage back to the MMA warp now
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(tepi_addr + as * 8);
        }
        if constexpr (kResidTma) {
          const uint32_t rbox = st.base + 4096 + lane * 128;
          if (has_res) {
            mbar_wait(&resid_bar[warp - 2], rphase);
            rphase ^= 1;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              float4 rv;
              asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                           : "=f"(rv.x), "=f"(rv.y), "=f"(rv.z), "=f"(rv.w)
                           : "r"(rbox + ((i ^ (lane & 7)) << 4)));
              v[4 * i + 0] += rv.x; v[4 * i + 1] += rv.y; v[4 * i + 2] += rv.z; v[4 * i + 3] += rv.w;
            }
          } else {
            if (lane == 0) tma_store_wait_read<0>();      // the box's previous store has been read
            __syncwarp();
          }
          if constexpr (kAct) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              st_shared_v4(rbox + ((i ^ (lane & 7)) << 4), __float_as_uint(v[4 * i + 0]), __float_as_uint(v[4 * i + 1]),
                           __float_as_uint(v[4 * i + 2]), __float_as_uint(v[4 * i + 3]));
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                               reinterpret_cast<uint64_t>(&p.mapD)),
                           "r"(st.base + 4096), "r"(nc), "r"(row0), "r"(b)
                           : "memory");
              tma_store_commit();
            }
          } else {
            __syncwarp();      // every lane has its row: the box may receive the next chunk
            if (lane == 0 && cc + 1 < nvalid) {
              mbar_expect_tx(&resid_bar[warp - 2], 4096);
              tma_load_3d(reinterpret_cast<void*>(rbox_ptr), &p.mapR, &resid_bar[warp - 2], nc + 32, row0, b);
            }
          }
        }
        if constexpr (kResid) {
          if (row_ok) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              v[4 * i + 0] += r4[i].x; v[4 * i + 1] += r4[i].y; v[4 * i + 2] += r4[i].z; v[4 * i + 3] += r4[i].w;
            }
            if (cc + 1 < nvalid) {            // next chunk's residual: in flight during this chunk's stores
#pragma unroll
              for (int i = 0; i < 8; ++i) r4[i] = rs[(cc + 1) * 8 + i];
            }
          }
        }
        if constexpr (EPI == EPI_STORE_F32) {
          if (has_alpha) {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] *= p.alpha;
          }
        }
        uint32_t kbits = 0;
        if constexpr (EPI == EPI_ROWMAX || EPI == EPI_EXP) {
#pragma unroll
          for (int j = 0; j < CPW; ++j) kbits = j == cc ? keybits[j] : kbits;    // keeps keybits[] in registers
        }
        epilogue_chunk<OT, EPI>(p, dmap, v, st, lane, b, h, z, row, row0, nc, row_ok, mk, gs, gss, racc, rowc, kbits, gn_ok);
        if constexpr (kAct) {
          if (lane == 0 && cc + 1 < nvalid && has_res) {      // the box's store has been read: fetch the next residual chunk
            tma_store_wait_read<0>();
            mbar_expect_tx(&resid_bar[warp - 2], 4096);
            tma_load_3d(reinterpret_cast<void*>(rbox_ptr), &p.mapR, &resid_bar[warp - 2], nc + 32, row0, b);
          }
        }
        if constexpr (EPI == EPI_ROWMAX || EPI == EPI_EXP) {
          // one partial per 64-key slice, whatever the tile shape: the order in which the row sum is formed (here
          // and in the P~ V epilogue) is then independent of BN, i.e. of the batch size the plan was built for
          if ((cc & 1) || cc + 1 == nvalid) {
            if (row_ok)
              (EPI == EPI_ROWMAX ? p.stat_max : p.stat_sum)[((long long)z * p.stat_n + (nc >> 6)) * p.stat_ld + row] = racc;
            racc = EPI == EPI_ROWMAX ? -INFINITY : 0.f;
          }
        }
        if constexpr (EPI == EPI_QKV) {
          // a 64-column part of a head is complete (always 64, whatever the tile shape: the test must not depend on the
          // batch the plan was built for): largest squared norm over this warp's VALID rows -> atomic max (non-negative
          // floats order like their bit patterns); padded frames are no keys, and see EPI_EXP for padded queries
          if (p.qk_norm2 != nullptr && (c & 1) && nc < p.split_n) {
            float m = (row_ok && mk != 0.f) ? gs : 0.f;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
            if (lane == 0) {
              const int isk = nc >= (p.split_n >> 1) ? 1 : 0;
              const int cc0 = nc - isk * (p.split_n >> 1);
              atomicMax(reinterpret_cast<int*>(p.qk_norm2) + (b * 4 + (cc0 >> 9)) * 16 + isk * 8 + ((cc0 & 511) >> 6),
                        __float_as_int(m));
            }
            gs = 0.f;
          }
        }
        if constexpr (EPI == EPI_STORE_GN) {
          if (gn_on && (c & 1)) {   // a 64-channel group is complete
            const float s = warp_sum_f(gs), ss = warp_sum_f(gss);
            if (lane == 0) {
              red[(q * NG + (c >> 1)) * 2 + 0] = s;
              red[(q * NG + (c >> 1)) * 2 + 1] = ss;
            }
            gs = 0.f; gss = 0.f;
          }
        }
      }
      if constexpr (EPI == EPI_STORE_GN) {
        if (p.gn_part != nullptr) {     // (uniform over the CTA's epilogue warps: every tile, also group 1's, meets here)
          asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");   // the epilogue warps only
          const int t = threadIdx.x - 64;
          if (gn_on && t < NG * 2) {
            const int g = t >> 1, w = t & 1;
            const float s = red[(0 * NG + g) * 2 + w] + red[(1 * NG + g) * 2 + w] + red[(2 * NG + g) * 2 + w] +
                            red[(3 * NG + g) * 2 + w];
            const int group = (n0 >> 6) + g;
            if (n0 < p.N) p.gn_part[(((long long)b * (2 * p.ntm) + mt) * 8 + group) * 2 + w] = s;
          }
        }
      }
    }
    if (threadIdx.x == 64) SRN_STAMP(6);
    if (lane == 0) tma_store_wait_all();     // smem boxes must outlive their bulk stores
    if (threadIdx.x == 64) SRN_STAMP(7);
  }
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();        // the peer may still be reading this CTA's smem / signalling its barriers
  if (warp == 2) tmem_dealloc_2sm<CFG::kTmemCols>(tmem_base);
}

}  // namespace srn
