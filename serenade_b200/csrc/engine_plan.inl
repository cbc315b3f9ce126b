// Engine<OT>: per-(B,T) plan = workspace buffers + the op list with TMA maps.  Included by engine.cu.
namespace srn {

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
static inline int round_up(int a, int b) { return cdiv(a, b) * b; }

template <typename OT>
Plan<OT>& Engine<OT>::get_plan(int B, int T) {
  if (!finalized_) fail("weights not finalized: call srn_finalize_weights first");
  // bounded cache, least recently used first (a hit moves the plan to the back)
  for (auto it = plans_.begin(); it != plans_.end(); ++it)
    if (it->B == B && it->T == T) {
      if (std::next(it) != plans_.end()) plans_.splice(plans_.end(), plans_, it);
      return plans_.back();
    }
  if (plans_.size() >= kMaxPlans) {
    if (last_plan_ == &plans_.front()) last_plan_ = nullptr;
    plans_.pop_front();
  }
  plans_.emplace_back();
  Plan<OT>& pl = plans_.back();
  pl.B = B;
  pl.T = T;
  try {
    build_plan(pl);
  } catch (...) {
    plans_.pop_back();
    throw;
  }
  return pl;
}

template <typename OT>
GemmParams Engine<OT>::base_gemm(const Plan<OT>& pl, int level) const {
  GemmParams gp;
  memset(&gp, 0, sizeof gp);
  gp.HB = 1;
  gp.alpha = 1.0f;
  gp.act_slope = 1.0f;
  gp.M = pl.lv[level].T;
  gp.mask_step = 1;
  return gp;
}

namespace {
template <typename OT>
void set_A(GemmParams& gp, int idx, const OT* ptr, int cols, int rows, long long alloc_rows, int ld, int nbatch) {
  gp.mapA[idx] = make_tmap_3d(ptr, sizeof(OT), cols, rows, nbatch, ld, (uint64_t)alloc_rows * ld, OpTraits<OT>::kBK, 128);
  if (idx == 0) gp.mapA[1] = gp.mapA[0];
}
// Call after every add_seg of the op.  TF32 mode: the packed weight rows are [hi | lo] (engine_weights.inl: up_w),
// so the K loop runs twice over the same A segments while the B column keeps advancing: A.hi + A.lo.
template <typename OT>
void set_B_weight(GemmParams& gp, const OT* W, int N, int K, int bn) {
  if (gp.nkb * OpTraits<OT>::kBK != K) fail("set_B_weight: K = %d but the A segments cover %d", K, gp.nkb * OpTraits<OT>::kBK);
  const int Kw = OpTraits<OT>::kTF32 ? 2 * K : K;
  gp.mapB = make_tmap_3d(W, sizeof(OT), Kw, N, 1, Kw, (uint64_t)N * Kw, OpTraits<OT>::kBK, bn / 2);   // each CTA of a pair loads half the B rows
  gp.b_batch_mode = 0;
  if (OpTraits<OT>::kTF32) {
    const int n = gp.nseg;
    if (2 * n > 24) fail("too many K segments");
    for (int i = 0; i < n; ++i) gp.seg[n + i] = gp.seg[i];
    gp.nseg = 2 * n;
    gp.nkb *= 2;
  }
}
// Output store map: 32-row x 128-byte boxes, SWIZZLE_128B (see gemm_epilogue.cuh)
template <typename T>
void set_D(GemmParams& gp, const T* ptr, int cols, int rows, long long bstride, int ld, int nbatch) {
  gp.mapD = make_tmap_3d(ptr, sizeof(T), cols, rows, nbatch, ld, (uint64_t)bstride, 128 / sizeof(T), 32);
  gp.out_cols = cols;
}
inline void add_seg(GemmParams& gp, int src, int col0, int shift, int nblk) {
  if (gp.nseg >= 24) fail("too many K segments");
  gp.seg[gp.nseg++] = KSeg{src, col0, shift, nblk};
  gp.nkb += nblk;
}
// Tile group 1 (gemm.cuh: GemmParams::grp_*): a second GEMM over the same A tensors in the same launch.  Call begin_group1
// after set_B_weight of group 0 (the TF32-class mode has doubled its segments by then), then add_seg_g1 per segment, then
// set_B2_weight.  ntn0 = number of N tiles of group 0.
inline void begin_group1(GemmParams& gp, int ntn0) {
  gp.grp_ntn0 = ntn0;
  gp.grp_seg0 = gp.nseg;
  gp.grp_nseg = 0;
  gp.grp_nkb = 0;
}
inline void add_seg_g1(GemmParams& gp, int src, int col0, int shift, int nblk) {
  if (gp.grp_seg0 + gp.grp_nseg >= 24) fail("too many K segments");
  gp.seg[gp.grp_seg0 + gp.grp_nseg++] = KSeg{src, col0, shift, nblk};
  gp.grp_nkb += nblk;
}
template <typename OT>
void set_B2_weight(GemmParams& gp, const OT* W, int N, int K, int bn) {
  if (gp.grp_nkb * OpTraits<OT>::kBK != K) fail("set_B2_weight: K = %d but the A segments cover %d", K, gp.grp_nkb * OpTraits<OT>::kBK);
  const int Kw = OpTraits<OT>::kTF32 ? 2 * K : K;
  gp.mapB2 = make_tmap_3d(W, sizeof(OT), Kw, N, 1, Kw, (uint64_t)N * Kw, OpTraits<OT>::kBK, bn / 2);
  if (OpTraits<OT>::kTF32) {
    const int n = gp.grp_nseg;
    if (gp.grp_seg0 + 2 * n > 24) fail("too many K segments");
    for (int i = 0; i < n; ++i) gp.seg[gp.grp_seg0 + n + i] = gp.seg[gp.grp_seg0 + i];
    gp.grp_nseg = 2 * n;
    gp.grp_nkb *= 2;
  }
}
// SRN_GROUPED=0: one launch per GEMM as in round 1 (A/B switch)
inline bool grouped_enabled() {
  static const bool on = [] {
    const char* e = getenv("SRN_GROUPED");
    return !(e != nullptr && e[0] == '0');
  }();
  return on;
}
// Window mode (gemm.cuh: GemmParams::win) for a convolution whose K segments are row shifts of ONE operand with the
// weights packed [tap][channel]: one kWinRows-row A window per K block then serves every tap.  Call after the
// add_seg calls; (ptr, cols, rows, alloc_rows, ld, nbatch) as given to set_A.  2-byte operands only (the TF32-class
// mode doubles the segments for its hi / lo weights).  SRN_WIN=0 switches it off for A/B runs.
inline int win_mode() {
  static const int m = [] {
    const char* e = getenv("SRN_WIN");
    return e != nullptr ? atoi(e) : 1;
  }();
  return m;
}
template <typename OT>
void maybe_window(GemmParams& gp, const OT* ptr, int cols, int rows, long long alloc_rows, int ld, int nbatch) {
  static const int min_seg = [] {
    const char* e = getenv("SRN_WIN_MINSEG");
    return e != nullptr ? atoi(e) : 4;
  }();
  if (win_mode() == 0 || sizeof(OT) != 2 || gp.nseg < min_seg) return;
  int mn = gp.seg[0].shift, mx = mn;
  for (int i = 0; i < gp.nseg; ++i) {
    if (gp.seg[i].src != 0 || gp.seg[i].col0 != gp.seg[0].col0 || gp.seg[i].nblk != gp.seg[0].nblk) return;
    mn = std::min(mn, gp.seg[i].shift);
    mx = std::max(mx, gp.seg[i].shift);
  }
  if (mx - mn > kWinRows - 128) return;
  gp.mapA[0] = make_tmap_3d(ptr, sizeof(OT), cols, rows, nbatch, ld, (uint64_t)alloc_rows * ld, OpTraits<OT>::kBK, kWinRows);
  gp.mapA[1] = gp.mapA[0];
  gp.win = win_mode();
  gp.win_min = mn;
}
// Row tiles / key tiles / K blocks that lie entirely in the padding of a ragged batch are skipped in the transformer
// blocks (SRN_SKIP_PADDED=0: compute them like the reference does).
inline bool skip_padded() {
  static const bool on = [] {
    const char* e = getenv("SRN_SKIP_PADDED");
    return !(e != nullptr && e[0] == '0');
  }();
  return on;
}
// Shift-free softmax when every score is provably within +-64 (gemm_mainloop.cuh, EPI_ROWMAX prologue); SRN_NOSHIFT=0:
// always run the row-max pass.
inline bool noshift_enabled() {
  static const bool on = [] {
    const char* e = getenv("SRN_NOSHIFT");
    return !(e != nullptr && e[0] == '0');
  }();
  return on;
}
inline bool zigzag_enabled() {
  static const bool on = [] {
    const char* e = getenv("SRN_ZIGZAG");
    return !(e != nullptr && e[0] == '0');
  }();
  return on;
}
// row_tiles = number of 256-row pair tiles; one CTA pair per SM pair (74 on a B200)
inline int pick_bn(int N, long long row_tiles) {
  const long long nt256 = (N + 255) / 256, nt128 = (N + 127) / 128;
  const bool little_waste = nt256 * 256 * 10 <= (long long)N * 11;
  if (!little_waste) return 128;
  // waves x relative cost of one tile (a 256 x 128 tile costs a bit more than half a 256 x 256 one)
  const double c256 = (double)((nt256 * row_tiles + 73) / 74) * 1.0;
  static const double k128 = [] {      // SRN_BN128_COST: A/B switch for the cost ratio (profiles/r02_ab_bn128.txt)
    const char* e = getenv("SRN_BN128_COST");
    return e != nullptr && atof(e) > 0.0 ? atof(e) : 0.58;
  }();
  const double c128 = (double)((nt128 * row_tiles + 73) / 74) * k128;
  return c256 <= c128 ? 256 : 128;
}
}  // namespace

template <typename OT>
void Engine<OT>::add_resnet(Plan<OT>& pl, int ridx, int level, const OT* src0, int c0, int ld0, const OT* src1,
                            int c1, const std::string& tag, const TfmW<OT>& next) {
  const Level& L = pl.lv[level];
  const ResnetW<OT>& w = resnets_[ridx];
  const int B = pl.B;
  const long long hs = (long long)L.R * 512;
  const long long tiles = (long long)L.mtiles * B;
  // conv (taps = 3 or 1) of the resnet input; `res` != null: the 1x1 res_conv rides along as tile group 1 of the same
  // launch (same A tensors, centre-tap segments only, its own weights / bias / output)
  auto conv_op = [&](const char* nm, const OT* a0, int ac0, int ald0, const OT* a1, int ac1, const OT* W, int taps,
                     const float* bias, OT* out, bool gn, bool with_res) {
    Op op;
    op.kind = OP_GEMM;
    op.name = tag + "." + nm;
    op.gp = base_gemm(pl, level);
    GemmParams& gp = op.gp;
    set_A<OT>(gp, 0, a0, ac0, L.T, L.R, ald0, B);
    if (a1 != nullptr) set_A<OT>(gp, 1, a1, ac1, L.T, L.R, 512, B);
    for (int k = 0; k < taps; ++k) {
      const int shift = taps == 3 ? k - 1 : 0;
      add_seg(gp, 0, 0, shift, ac0 / BK);
      if (a1 != nullptr) add_seg(gp, 1, 0, shift, ac1 / BK);
    }
    if (a1 == nullptr && !with_res) maybe_window<OT>(gp, a0, ac0, L.T, L.R, ald0, B);
    op.bn = pick_bn(with_res ? 1024 : 512, tiles);
    set_B_weight<OT>(gp, W, 512, taps * (ac0 + ac1), op.bn);
    gp.N = 512;
    gp.epi = EPI_STORE_GN;
    gp.bias = bias;
    gp.out = out;
    gp.out_bstride = hs;
    gp.ldo = 512;
    gp.gn_part = gn ? pl.gnpart.template as<float>() : nullptr;
    if (skip_padded() && B <= 256) {      // padded-frame skipping: fully padded row tiles are all bias (gemm.cuh: ext_halo)
      gp.row_extent = pl.rowext.template as<int>() + (size_t)level * B;
      gp.ext_halo = taps == 3 ? 1 : 0;
    }
    set_D<OT>(gp, out, 512, L.T, hs, 512, B);
    int ntn = 512 / op.bn;
    if (with_res) {
      begin_group1(gp, ntn);
      add_seg_g1(gp, 0, 0, 0, ac0 / BK);
      if (a1 != nullptr) add_seg_g1(gp, 1, 0, 0, ac1 / BK);
      set_B2_weight<OT>(gp, w.wr, 512, ac0 + ac1, op.bn);
      gp.bias2 = w.br;
      gp.mapD2 = make_tmap_3d(pl.r.template as<OT>(), sizeof(OT), 512, L.T, B, 512, (uint64_t)hs, 128 / sizeof(OT), 32);
      ntn *= 2;
    }
    op.grid = dim3(ntn, L.mtiles, B);
    if (a0 == reinterpret_cast<const OT*>(pl.xin.p)) op.alg_scale = (double)cfg_.in_channels / (double)ac0;   // 242 of the 256 packed channels
    op.tap = Tap{nullptr, out, 1, L.T, 512, 512, hs, B, 0};
    op.level = level;
    pl.ops.push_back(op);
    return &pl.ops.back();
  };
  OT* ybuf = pl.y.template as<OT>();
  if (grouped_enabled()) {
    Op* o = conv_op("y1", src0, c0, ld0, src1, c1, w.w1, 3, w.b1, ybuf, true, true);
    o->tap.name = "y1";
    o->tap2 = Tap{"r", pl.r.p, 1, L.T, 512, 512, hs, B, 0};
  } else {
    conv_op("y1", src0, c0, ld0, src1, c1, w.w1, 3, w.b1, ybuf, true, false)->tap.name = "y1";
    conv_op("r", src0, c0, ld0, src1, c1, w.wr, 1, w.br, pl.r.template as<OT>(), false, false)->tap.name = "r";
  }
  {
    Op op;
    op.kind = OP_GN_APPLY;
    op.name = tag + ".a1";
    op.y = pl.y.p;
    op.part = pl.gnpart.template as<float>();
    op.gamma = w.g1;
    op.beta = w.be1;
    op.out = pl.a1.p;
    op.resnet_idx = ridx;
    op.use_tbias = 1;
    op.level = level;
    op.tap = Tap{"a1", pl.a1.p, 1, L.T, 512, 512, hs, B, 0};
    pl.ops.push_back(op);
  }
  conv_op("y2", pl.a1.template as<OT>(), 512, 512, nullptr, 0, w.w2, 3, w.b2, ybuf, true, false)->tap.name = "y2";
  {
    Op op;
    op.kind = OP_GN_RES_SPK_LN;
    op.name = tag + ".resnet";
    op.y = pl.y.p;
    op.part = pl.gnpart.template as<float>();
    op.gamma = w.g2;
    op.beta = w.be2;
    op.res = pl.r.p;
    op.ln_g = next.ln1g;
    op.ln_b = next.ln1b;
    op.hres = pl.hres.template as<float>();
    op.out = pl.ln.p;
    op.resnet_idx = ridx;
    op.level = level;
    op.tap = Tap{"resnet", pl.hres.p, 0, L.T, 512, 512, hs, B, 0};
    pl.ops.push_back(op);
  }
}

template <typename OT>
void Engine<OT>::add_tfm(Plan<OT>& pl, int tidx, int level, OT* out_hc, const std::string& tag) {
  const Level& L = pl.lv[level];
  const TfmW<OT>& w = tfms_[tidx];
  const int B = pl.B, S = L.T, Sp = L.Sp, H = 4;
  const long long tiles = (long long)L.mtiles * B;
  OT* ln = pl.ln.template as<OT>();
  OT* qk = pl.qk.template as<OT>();
  OT* vt = pl.vt.template as<OT>();
  OT* P = pl.P.template as<OT>();
  OT* ao = pl.ao.template as<OT>();
  OT* gg = pl.gg.template as<OT>();
  float* hres = pl.hres.template as<float>();
  auto push = [&](Op& op, const char* nm) {
    op.kind = OP_GEMM;
    op.name = tag + "." + nm;
    op.level = level;
    // Padded-frame skipping (gemm.cuh: row_extent).  Every GEMM of the transformer block but ff_out: its output is the
    // next convolution's (already masked) operand, so it writes its zeros itself.
    const std::string n(nm);
    if (skip_padded() && B <= 256 && n != "ff_out") {
      op.gp.row_extent = pl.rowext.template as<int>() + (size_t)level * B;
      op.gp.ext_keys = (n == "rowmax" || n == "expscores") ? 1 : (n == "pv" ? 2 : 0);
    }
    pl.ops.push_back(op);
  };
  {  // fused Q|K|V projection; V is written transposed (V^T [B][2048][Sp]) so that P.V sees a K-major operand
    Op op;
    op.gp = base_gemm(pl, level);
    GemmParams& gp = op.gp;
    set_A<OT>(gp, 0, ln, 512, S, L.R, 512, B);
    add_seg(gp, 0, 0, 0, 512 / BK);
    op.bn = pick_bn(6144, tiles);
    set_B_weight<OT>(gp, w.wqkv, 6144, 512, op.bn);
    gp.N = 6144;
    gp.epi = EPI_QKV;
    gp.out = qk;
    gp.out_bstride = (long long)L.R * 4096;
    gp.ldo = 4096;
    gp.split_n = 4096;
    gp.out2 = vt;
    gp.out2_bstride = (long long)2048 * Sp;
    gp.ld2 = Sp;
    set_D<OT>(gp, qk, 4096, S, gp.out_bstride, 4096, B);
    gp.mapD2 = make_tmap_3d(vt, sizeof(OT), S, 2048, B, Sp, (uint64_t)2048 * Sp, 32, 32, false);
    op.grid = dim3(6144 / op.bn, L.mtiles, B);
    op.tap = Tap{"qk", qk, 1, S, 4096, 4096, gp.out_bstride, B, 0};
    if (noshift_enabled() && B * H <= 1024) {
      gp.qk_norm2 = pl.qk_norm2.template as<float>();
      gp.rowmask = pl.mask.template as<uint8_t>();      // (EPI_QKV only uses it to keep padded frames out of the norms)
      gp.mask_bstride = pl.Tp;
      gp.mask_step = L.mask_step;
    }
    push(op, "qkv");
  }

  // Attention without a score matrix in HBM.  Q K^T runs twice on the tensor cores: pass 1 keeps only per-row
  // maxima (key-masked; skipped entirely for entries that are provably shift-free), pass 2 recomputes the tile, writes P~ = exp(s - rowmax) in operand precision and partial
  // row sums; P~ V then divides by the row sum in its epilogue.  Replaces scores f32 write + read and the softmax
  // kernel (12 B per score) by 2 + 2 B per score; 1/sqrt(d), the key mask and the normalisation follow
  // F.scaled_dot_product_attention with a boolean key mask.
  const int sc_bn = pick_bn(S, tiles * H);
  const int stat_n = cdiv(S, 64);          // one (max, sum) partial per row per 64-key slice
  if ((size_t)B * H * stat_n * L.Sp * 4 > pl.smax.bytes) fail("softmax statistics buffer too small");
  Op exp_op;
  for (int pass = 0; pass < 2; ++pass) {
    Op op;
    op.gp = base_gemm(pl, level);
    GemmParams& gp = op.gp;
    op.bn = sc_bn;
    gp.HB = H;
    set_A<OT>(gp, 0, qk, 4096, S, L.R, 4096, B);
    gp.a_hcol = 512;
    add_seg(gp, 0, 0, 0, 512 / BK);
    gp.mapB = make_tmap_3d(qk, sizeof(OT), 4096, S, B, 4096, (uint64_t)L.R * 4096, BK, op.bn / 2);
    gp.b_col0 = 2048;
    gp.b_hcol = 512;
    gp.b_batch_mode = 1;
    gp.N = S;
    gp.epi = pass == 0 ? EPI_ROWMAX : EPI_EXP;
    gp.alpha = 1.0f / sqrtf(512.0f);
    gp.stat_max = pl.smax.template as<float>();
    gp.stat_sum = pl.ssum.template as<float>();
    gp.stat_n = stat_n;
    gp.stat_ld = L.Sp;
    if (noshift_enabled() && B * H <= 1024) {      // shift-free softmax test: evaluated by the row-max pass, consumed by the exp pass
      gp.noshift = pl.noshift.template as<int>();
      gp.qk_norm2 = pl.qk_norm2.template as<float>();
      gp.nz_norm = B * H;
      gp.ns_count = pl.ns_count.template as<unsigned long long>();
      static const float nmax = getenv("SRN_NOSHIFT_MAX") ? (float)atof(getenv("SRN_NOSHIFT_MAX")) : 64.0f;
      gp.noshift_max = nmax < 64.0f ? nmax : 64.0f;      // (tests lower it to get shifted and shift-free entries in one launch)
    }
    gp.keymask = pl.mask.template as<uint8_t>();
    gp.keymask_bstride = pl.Tp;
    gp.keymask_step = L.mask_step;
    gp.out = P;
    gp.out_bstride = (long long)H * S * Sp;
    gp.out_hstride = (long long)S * Sp;
    gp.ldo = Sp;
    if (pass == 0) gp.mapD = gp.mapA[0];      // no tensor output; keeps the descriptor prefetch valid
    else set_D<OT>(gp, P, S, S, (long long)S * Sp, Sp, B * H);
    gp.d_zbatch = 1;
    op.grid = dim3(cdiv(S, op.bn), L.mtiles, B * H);
    if (pass == 1) op.tap = Tap{"P", P, 1, S, S, Sp, (long long)S * Sp, B * H, 0};
    if (pass == 0) op.alg_scale = 0.0;      // (F(L) counts Q K^T once: the exp pass)
    if (pass == 0) push(op, "rowmax");
    else exp_op = op;
  }
  {  // O[b, :, h*512:(h+1)*512] = (P~[b,h] V[b,h]) / rowsum
    Op op;
    op.gp = base_gemm(pl, level);
    GemmParams& gp = op.gp;
    gp.HB = H;
    gp.mapA[0] = make_tmap_3d(P, sizeof(OT), S, S, (uint64_t)B * H, Sp, (uint64_t)S * Sp, BK, 128);
    gp.mapA[1] = gp.mapA[0];
    gp.a_zbatch = 1;
    add_seg(gp, 0, 0, 0, cdiv(S, BK));
    op.bn = pick_bn(512, tiles * H);
    gp.mapB = make_tmap_3d(vt, sizeof(OT), S, 2048, B, Sp, (uint64_t)2048 * Sp, BK, op.bn / 2);
    gp.b_hrow = 512;
    gp.b_batch_mode = 1;
    gp.N = 512;
    gp.epi = EPI_PV;
    gp.stat_sum = pl.ssum.template as<float>();
    gp.stat_n = stat_n;
    gp.stat_ld = L.Sp;
    gp.out = ao;
    gp.out_bstride = (long long)L.R * 2048;
    gp.out_hstride = 512;
    gp.ldo = 2048;
    set_D<OT>(gp, ao, 2048, S, gp.out_bstride, 2048, B);
    gp.d_hcol = 512;
    op.grid = dim3(512 / op.bn, L.mtiles, B * H);
    op.tap = Tap{"ao", ao, 1, S, 2048, 2048, gp.out_bstride, B, 0};
    // (P~ of the whole (batch, head) range is B H S Sp elements - 131 MB at C2's full rate, 2.3 GB at C5 - more than the
    //  126 MB L2; running both passes slice by slice so that a slice's P~ stays in the L2 was measured in round 1 and
    //  lost to its extra launches: profiles/r01_ab_attn_slices.txt)
    push(exp_op, "expscores");
    push(op, "pv");
  }
  {  // h = attn_out W_o^T + b_o + h    (transformer.py:304)
    Op op;
    op.gp = base_gemm(pl, level);
    GemmParams& gp = op.gp;
    set_A<OT>(gp, 0, ao, 2048, S, L.R, 2048, B);
    add_seg(gp, 0, 0, 0, 2048 / BK);
    op.bn = pick_bn(512, tiles);
    set_B_weight<OT>(gp, w.wo, 512, 2048, op.bn);
    gp.N = 512;
    gp.epi = EPI_RESID_F32;
    gp.bias = w.bo;
    gp.resid = hres;
    gp.resid_bstride = (long long)L.R * 512;
    gp.ldr = 512;
    gp.out = hres;
    gp.out_bstride = gp.resid_bstride;
    gp.ldo = 512;
    set_D<float>(gp, hres, 512, S, gp.out_bstride, 512, B);
    op.grid = dim3(512 / op.bn, L.mtiles, B);
    op.tap = Tap{"h_attn", hres, 0, S, 512, 512, gp.out_bstride, B, 0};
    push(op, "attn_out");
  }
  {
    Op op;
    op.kind = OP_LAYERNORM;
    op.name = tag + ".ln3";
    op.ln_g = w.ln3g;
    op.ln_b = w.ln3b;
    op.level = level;
    op.tap = Tap{"ln3", ln, 1, S, 512, 512, (long long)L.R * 512, B, 0};
    pl.ops.push_back(op);
  }
  {  // GEGLU: (ln W_v^T + b_v) * gelu(ln W_g^T + b_g), value/gate interleaved in the packed weight
    Op op;
    op.gp = base_gemm(pl, level);
    GemmParams& gp = op.gp;
    set_A<OT>(gp, 0, ln, 512, S, L.R, 512, B);
    add_seg(gp, 0, 0, 0, 512 / BK);
    op.bn = 256;   // the GEGLU epilogue packs 4 chunks (128 accumulator columns) per 128-byte output row
    set_B_weight<OT>(gp, w.wgl, 4096, 512, op.bn);
    gp.N = 4096;
    gp.epi = EPI_GEGLU;
    gp.bias = w.bgl;
    gp.out = gg;
    gp.out_bstride = (long long)L.R * 2048;
    gp.ldo = 2048;
    set_D<OT>(gp, gg, 2048, S, gp.out_bstride, 2048, B);
    op.grid = dim3(4096 / op.bn, L.mtiles, B);
    op.tap = Tap{"gg", gg, 1, S, 2048, 2048, gp.out_bstride, B, 0};
    push(op, "geglu");
  }
  {  // h = gg W_2^T + b_2 + h, stored masked in operand precision for the convs that consume it
    Op op;
    op.gp = base_gemm(pl, level);
    GemmParams& gp = op.gp;
    set_A<OT>(gp, 0, gg, 2048, S, L.R, 2048, B);
    add_seg(gp, 0, 0, 0, 2048 / BK);
    op.bn = pick_bn(512, tiles);
    set_B_weight<OT>(gp, w.wf, 512, 2048, op.bn);
    gp.N = 512;
    gp.epi = EPI_RESID_OT;
    gp.bias = w.bf;
    gp.resid = hres;
    gp.resid_bstride = (long long)L.R * 512;
    gp.ldr = 512;
    gp.out = out_hc;
    gp.out_bstride = (long long)L.R * 512;
    gp.ldo = 512;
    gp.rowmask = pl.mask.template as<uint8_t>();
    gp.mask_bstride = pl.Tp;
    gp.mask_step = L.mask_step;
    set_D<OT>(gp, out_hc, 512, S, gp.out_bstride, 512, B);
    gp.mapR = make_tmap_3d(hres, sizeof(float), 512, S, B, 512, (uint64_t)gp.resid_bstride, 32, 32);   // residual boxes (kResidTma)
    op.grid = dim3(512 / op.bn, L.mtiles, B);
    op.tap = Tap{"tfm", out_hc, 1, S, 512, 512, gp.out_bstride, B, 0};
    push(op, "ff_out");
  }
}

template <typename OT>
void Engine<OT>::build_plan(Plan<OT>& pl) {
  const int B = pl.B, T = pl.T, H = 4, C = 512;
  const int T2 = (T + 1) / 2;
  pl.Tp = 2 * T2;
  // Sp: the row pitch of P~, V^T and the softmax statistics.  A multiple of 64 elements keeps every 128-byte box row of
  // the P~ V operands inside one 128-byte line (SRN_SP_ALIGN=8 restores the round-1 pitch for the A/B)
  static const int sp_align = [] {
    const char* e = getenv("SRN_SP_ALIGN");
    return e != nullptr && atoi(e) >= 8 ? atoi(e) : 64;
  }();
  pl.lv[0] = Level{T, pl.Tp, round_up(T, sp_align), cdiv(T, 256), 1};
  pl.lv[1] = Level{T2, T2, round_up(T2, sp_align), cdiv(T2, 256), 2};
  const size_t rows = (size_t)B * pl.Tp;
  const size_t es = sizeof(OT);
  pl.xin.alloc(rows * 256 * es);
  pl.y.alloc(rows * C * es);      // conv outputs that feed a GroupNorm, operand width (EPI_STORE_GN)
  pl.r.alloc(rows * C * es);
  pl.gnpart.alloc((size_t)B * 2 * pl.lv[0].mtiles * 16 * 4);   // one (sum, sumsq) x 8 groups per 128-row tile
  pl.a1.alloc(rows * C * es);
  pl.hres.alloc(rows * C * 4);
  pl.ln.alloc(rows * C * es);
  pl.qk.alloc(rows * 4096 * es);
  pl.qk_norm2.alloc((size_t)B * H * 16 * 4);
  pl.noshift.alloc((size_t)B * H * 4);
  pl.ns_count.alloc(16);
  pl.vt.alloc((size_t)B * 2048 * pl.lv[0].Sp * es);
  {  // softmax statistics: one (max, sum) partial per row per 128-column (or 64-column) slice of the score GEMM
    const size_t n128 = 2 * (size_t)cdiv(T, 128);
    pl.smax.alloc((size_t)B * H * n128 * pl.lv[0].Sp * 4);
    pl.ssum.alloc((size_t)B * H * n128 * pl.lv[0].Sp * 4);
  }
  pl.P.alloc((size_t)B * H * T * pl.lv[0].Sp * es);
  pl.ao.alloc(rows * 2048 * es);
  pl.gg.alloc(rows * 2048 * es);
  pl.hid0.alloc(rows * C * es);
  pl.hid1.alloc((size_t)B * T2 * C * es);
  pl.hc.alloc(rows * C * es);
  pl.cv.alloc(rows * C * es);
  pl.x_state.alloc((size_t)B * cfg_.out_channels * T * 4);
  pl.mask.alloc(rows);
  pl.spk_in.alloc((size_t)B * cfg_.spk_embed_dim * 4);
  pl.mu_in.alloc((size_t)B * cmu_ * T * 4);
  const int nres = (int)resnets_.size();
  pl.spk_scale.alloc((size_t)nres * B * C * 4);
  pl.spk_bias.alloc((size_t)nres * B * C * 4);
  pl.blend_mask.alloc((size_t)B * T);
  pl.nvalid.alloc((size_t)2 * B * sizeof(int));
  pl.rowext.alloc((size_t)2 * B * sizeof(int));
  pl.blend_src.alloc((size_t)B * cfg_.out_channels * T * 4);

  OT* xin = pl.xin.template as<OT>();
  OT* hid0 = pl.hid0.template as<OT>();
  OT* hid1 = pl.hid1.template as<OT>();
  OT* hc = pl.hc.template as<OT>();
  OT* cv = pl.cv.template as<OT>();
  const uint8_t* mask = pl.mask.template as<uint8_t>();
  const long long tiles0 = (long long)pl.lv[0].mtiles * B, tiles1 = (long long)pl.lv[1].mtiles * B;

  {  // x (B,80,T) -> xin[:, :, 0:80] (masked)
    Op op;
    op.kind = OP_PACKX;
    op.name = "pack_x";
    op.tap = Tap{"xin", xin, 1, T, 256, 256, (long long)pl.Tp * 256, B, 0};
    pl.ops.push_back(op);
  }
  auto plain_conv3 = [&](const char* nm, int level, const OT* src, const ConvW<OT>& w, OT* dst, int out_level) {
    const Level& L = pl.lv[level];
    Op op;
    op.kind = OP_GEMM;
    op.name = nm;
    op.level = level;
    op.gp = base_gemm(pl, level);
    GemmParams& gp = op.gp;
    set_A<OT>(gp, 0, src, 512, L.T, L.R, 512, B);
    for (int k = 0; k < 3; ++k) add_seg(gp, 0, 0, k - 1, 512 / BK);
    maybe_window<OT>(gp, src, 512, L.T, L.R, 512, B);
    op.bn = pick_bn(512, (long long)L.mtiles * B);
    set_B_weight<OT>(gp, w.w, 512, 1536, op.bn);
    gp.N = 512;
    gp.epi = EPI_STORE_OT;
    gp.bias = w.b;
    gp.out = dst;
    gp.out_bstride = (long long)L.R * 512;
    gp.ldo = 512;
    gp.rowmask = mask;
    gp.mask_bstride = pl.Tp;
    gp.mask_step = pl.lv[out_level].mask_step;
    set_D<OT>(gp, dst, 512, L.T, gp.out_bstride, 512, B);
    op.grid = dim3(512 / op.bn, L.mtiles, B);
    op.tap = Tap{"out", dst, 1, L.T, 512, 512, gp.out_bstride, B, 0};
    pl.ops.push_back(op);
  };

  int ri = 0;
  // ---- down 0 (full rate): resnet(242 -> 512) + transformer, skip saved in hid0, Downsample1D
  add_resnet(pl, ri, 0, xin, 256, 256, nullptr, 0, "down0", tfms_[ri]);
  add_tfm(pl, ri, 0, hid0, "down0");
  ++ri;
  {  // Conv1d k3 s2 p1 on the pair view [B][T2][1024] of hid0: out[m] = W0 x[2m-1] + W1 x[2m] + W2 x[2m+1]
    const Level& L = pl.lv[1];
    Op op;
    op.kind = OP_GEMM;
    op.name = "down0.out";
    op.level = 1;
    op.gp = base_gemm(pl, 1);
    GemmParams& gp = op.gp;
    set_A<OT>(gp, 0, hid0, 1024, T2, T2, 1024, B);
    add_seg(gp, 0, 512, -1, 512 / BK);
    add_seg(gp, 0, 0, 0, 512 / BK);
    add_seg(gp, 0, 512, 0, 512 / BK);
    op.bn = pick_bn(512, tiles1);
    set_B_weight<OT>(gp, down_conv_[0].w, 512, 1536, op.bn);
    gp.N = 512;
    gp.epi = EPI_STORE_OT;
    gp.bias = down_conv_[0].b;
    gp.out = cv;
    gp.out_bstride = (long long)T2 * 512;
    gp.ldo = 512;
    gp.rowmask = mask;
    gp.mask_bstride = pl.Tp;
    gp.mask_step = 2;
    set_D<OT>(gp, cv, 512, T2, gp.out_bstride, 512, B);
    op.grid = dim3(512 / op.bn, L.mtiles, B);
    op.tap = Tap{"out", cv, 1, T2, 512, 512, gp.out_bstride, B, 0};
    pl.ops.push_back(op);
  }
  // ---- down 1 (half rate)
  add_resnet(pl, ri, 1, cv, 512, 512, nullptr, 0, "down1", tfms_[ri]);
  add_tfm(pl, ri, 1, hid1, "down1");
  ++ri;
  plain_conv3("down1.out", 1, hid1, down_conv_[1], cv, 1);
  // ---- mid blocks (half rate)
  const OT* cur = cv;
  for (int i = 0; i < cfg_.num_mid_blocks; ++i) {
    const std::string tag = "mid" + std::to_string(i);
    add_resnet(pl, ri, 1, cur, 512, 512, nullptr, 0, tag, tfms_[ri]);
    add_tfm(pl, ri, 1, hc, tag);
    cur = hc;
    ++ri;
  }
  // ---- up 0 (half rate): cat([h, hid1]) -> resnet(1024 -> 512) + transformer, ConvTranspose1d(4, 2, 1)
  add_resnet(pl, ri, 1, cur, 512, 512, hid1, 512, "up0", tfms_[ri]);
  add_tfm(pl, ri, 1, hc, "up0");
  ++ri;
  // ConvTranspose1d(4, 2, 1): out[2m] = W1 x[m] + W3 x[m-1], out[2m+1] = W2 x[m] + W0 x[m+1].  The two output parities
  // are two GEMMs over the same operand (K = 2 x 512 each) that write the interleaved full-rate tensor viewed as
  // [B][T2][1024]; they run as the two tile groups of ONE launch (SRN_GROUPED=0: two launches).
  {
    const Level& L = pl.lv[1];
    const bool grouped = grouped_enabled();
    for (int par = 0; par < (grouped ? 1 : 2); ++par) {
      Op op;
      op.kind = OP_GEMM;
      op.name = (grouped || par == 1) ? "up0.out" : "up0.out_even";
      op.level = 1;
      op.gp = base_gemm(pl, 1);
      GemmParams& gp = op.gp;
      set_A<OT>(gp, 0, hc, 512, T2, T2, 512, B);
      add_seg(gp, 0, 0, par == 0 ? -1 : 0, 512 / BK);
      add_seg(gp, 0, 0, par == 0 ? 0 : 1, 512 / BK);
      op.bn = pick_bn(grouped ? 1024 : 512, tiles1);
      set_B_weight<OT>(gp, par == 0 ? up_conv_[0].w : up_conv_[0].w_odd, 512, 1024, op.bn);
      gp.N = 512;
      gp.epi = EPI_STORE_OT;
      gp.bias = up_conv_[0].b;
      gp.out = cv + par * 512;              // frame 2m + par of the full-rate tensor viewed as [B][T2][1024]
      gp.out_bstride = (long long)T2 * 1024;
      gp.ldo = 1024;
      gp.rowmask = mask;
      gp.mask_bstride = pl.Tp;
      gp.mask_step = 2;
      gp.mask_off = par;
      set_D<OT>(gp, cv + par * 512, 512, T2, gp.out_bstride, 1024, B);
      int ntn = 512 / op.bn;
      if (grouped) {      // group 1 = the odd output frames
        begin_group1(gp, ntn);
        add_seg_g1(gp, 0, 0, 0, 512 / BK);
        add_seg_g1(gp, 0, 0, 1, 512 / BK);
        set_B2_weight<OT>(gp, up_conv_[0].w_odd, 512, 1024, op.bn);
        gp.bias2 = up_conv_[0].b;
        gp.mask_off2 = 1;
        gp.mapD2 = make_tmap_3d(cv + 512, sizeof(OT), 512, T2, B, 1024, (uint64_t)gp.out_bstride, 128 / sizeof(OT), 32);
        ntn *= 2;
      }
      op.grid = dim3(ntn, L.mtiles, B);
      if (grouped || par == 1) op.tap = Tap{"out", cv, 1, T, 512, 512, (long long)pl.Tp * 512, B, 0};
      pl.ops.push_back(op);
    }
  }
  // ---- up 1 (full rate): cat([h[:, :, :T], hid0]) -> resnet + transformer, Conv1d k3
  add_resnet(pl, ri, 0, cv, 512, 512, hid0, 512, "up1", tfms_[ri]);
  add_tfm(pl, ri, 0, hc, "up1");
  ++ri;
  plain_conv3("up1.out", 0, hc, up_conv_[1], cv, 0);
  // ---- final Block1D + 1x1 projection fused with the Euler update
  {
    const Level& L = pl.lv[0];
    Op op;
    op.kind = OP_GEMM;
    op.name = "final.y";
    op.gp = base_gemm(pl, 0);
    GemmParams& gp = op.gp;
    set_A<OT>(gp, 0, cv, 512, T, pl.Tp, 512, B);
    for (int k = 0; k < 3; ++k) add_seg(gp, 0, 0, k - 1, 512 / BK);
    maybe_window<OT>(gp, cv, 512, T, pl.Tp, 512, B);
    op.bn = pick_bn(512, tiles0);
    set_B_weight<OT>(gp, final_conv_.w, 512, 1536, op.bn);
    gp.N = 512;
    gp.epi = EPI_STORE_GN;
    gp.bias = final_conv_.b;
    gp.out = pl.y.p;
    gp.out_bstride = (long long)pl.Tp * 512;
    gp.ldo = 512;
    gp.gn_part = pl.gnpart.template as<float>();
    set_D<OT>(gp, pl.y.template as<OT>(), 512, T, gp.out_bstride, 512, B);
    op.grid = dim3(512 / op.bn, L.mtiles, B);
    pl.ops.push_back(op);
  }
  {
    Op op;
    op.kind = OP_GN_APPLY;
    op.name = "final_block";
    op.y = pl.y.p;
    op.part = pl.gnpart.template as<float>();
    op.gamma = final_g_;
    op.beta = final_be_;
    op.out = pl.a1.p;
    op.level = 0;
    op.tap = Tap{"final_block", pl.a1.p, 1, T, 512, 512, (long long)pl.Tp * 512, B, 0};
    pl.ops.push_back(op);
  }
  {
    Op op;
    op.kind = OP_GEMM;
    op.name = "final_proj_euler";
    op.is_euler = true;
    op.gp = base_gemm(pl, 0);
    GemmParams& gp = op.gp;
    set_A<OT>(gp, 0, pl.a1.template as<OT>(), 512, T, pl.Tp, 512, B);
    add_seg(gp, 0, 0, 0, 512 / BK);
    op.bn = 128;
    set_B_weight<OT>(gp, proj_w_, cfg_.out_channels, 512, op.bn);
    gp.N = cfg_.out_channels;
    gp.epi = EPI_EULER;
    gp.bias = proj_b_;
    gp.rowmask = mask;
    gp.mask_bstride = pl.Tp;
    gp.mask_step = 1;
    gp.x_in = pl.x_state.template as<float>();
    gp.T_cf = T;
    gp.mapD = gp.mapA[0];     // unused by EPI_EULER (direct channels-first stores); keeps the descriptor prefetch valid
    op.grid = dim3(1, pl.lv[0].mtiles, B);
    pl.euler_op = (int)pl.ops.size();
    pl.ops.push_back(op);
  }
  // Zig-zag tile order (GemmParams::rev): every GEMM walks its tiles in the direction opposite to the one in which its
  // input was written, so that what the producer wrote last - the part of a larger-than-L2 tensor that is still cached -
  // is read first.  Row kernels write in ascending order; the row-max pass reads what the exp pass reads (the QKV
  // output) and is mostly skipped, so the exp pass takes the same direction.  SRN_ZIGZAG=0: all ascending.
  if (zigzag_enabled()) {
    int dir = 0;
    for (Op& op : pl.ops) {
      if (op.kind != OP_GEMM) { dir = 0; continue; }
      op.gp.rev = 1 - dir;
      if (op.gp.epi != EPI_ROWMAX) dir = op.gp.rev;
    }
  }
}

}  // namespace srn
