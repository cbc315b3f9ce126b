// Engine<OT>: HiFi-GAN vocoder, the stage that follows the CFM decoder in serenade-decode (reference
// serenade/vocoder/models/hifigan.py:21-186, layers/residual_block.py:143-258, vocoder.py:51-75).  Included by engine.cu.
//
// Activations are channels-last [B][T_i][C_i]; every Conv1d / ConvTranspose1d is a launch of the tcgen05 GEMM core:
//   Conv1d(k, dilation d)   k K-segments of the same operand tensor at row shifts (j - (k-1)/2) d; TMA zero-fills rows
//                           outside [0, T) = the conv's zero padding
//   ConvTranspose1d(2s, s)  ONE GEMM with N = s * C_out over the three input rows q-1, q, q+1: output row s q + phi is
//                           column block phi of row q of the [T][s * C_out] view of the output; each phase uses two of
//                           the three taps (the third weight block is zero)
//   LeakyReLU before a conv is applied where its operand is produced: in the epilogue of the previous conv
//   (EPI_STORE_OT, act_slope) or by the cast kernel that turns the f32 residual stream into an operand
//   x = xt + x              EPI_RESID_F32: TMA reduce-add into the f32 residual stream, in place
// Channel counts below one 128-byte K block (C = 32 in bf16) are zero-padded in the operand tensors and weights.
namespace srn {

namespace {
enum VocOut { VOC_F32 = 0, VOC_OT = 1, VOC_RESID = 2, VOC_RESID_ACT = 3 };
inline std::vector<float> pad_to(const std::vector<float>& v, size_t n) {
  std::vector<float> o(n, 0.f);
  std::copy(v.begin(), v.end(), o.begin());
  return o;
}
}  // namespace

template <typename OT>
void Engine<OT>::pack_vocoder() {
  voc_ = VocW{};
  if (host_w_.find("vocoder.input_conv.weight") == host_w_.end()) return;
  auto any = [&](const std::string& k) -> const HostTensor& {
    auto it = host_w_.find("vocoder." + k);
    if (it == host_w_.end()) fail("missing weight 'vocoder.%s' (weight norm must be folded before upload)", k.c_str());
    return it->second;
  };
  auto has = [&](const std::string& k) { return host_w_.find("vocoder." + k) != host_w_.end(); };
  // Conv1d weight (O, I, k) -> [Op][k][Ip] (zero padded), K-major rows
  auto conv_pack = [&](const HostTensor& w, int Op, int Ip) {
    const int O = (int)w.dims[0], I = (int)w.dims[1], k = (int)w.dims[2];
    std::vector<float> pk((size_t)Op * k * Ip, 0.f);
    for (int o = 0; o < O; ++o)
      for (int i = 0; i < I; ++i)
        for (int t = 0; t < k; ++t) pk[((size_t)o * k + t) * Ip + i] = w.data[((size_t)o * I + i) * k + t];
    return pk;
  };
  // The same conv (dilation d, 'same' zero padding) on the grouped view: GEMM row q = time steps g q .. g q + g - 1.
  // Output sub-step p, tap j reads time g q + p + (j - c) d = row q + shift, sub-step h; all (p, j) with one shift
  // share a K segment.  Returns [Np][nseg * g I] and the segment shifts.  g = 1 with I padded is the plain form.
  auto conv_pack_grouped = [&](const HostTensor& w, int g, int d, int Np, int kc, std::vector<int>& shifts) {
    const int O = (int)w.dims[0], I = (int)w.dims[1], k = (int)w.dims[2], c = (k - 1) / 2;
    auto fdiv = [](int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); };
    shifts.clear();
    for (int p = 0; p < g; ++p)
      for (int j = 0; j < k; ++j) shifts.push_back(fdiv(p + (j - c) * d, g));
    std::sort(shifts.begin(), shifts.end());
    shifts.erase(std::unique(shifts.begin(), shifts.end()), shifts.end());
    const int nseg = (int)shifts.size();
    std::vector<float> pk((size_t)Np * nseg * kc, 0.f);
    for (int p = 0; p < g; ++p)
      for (int j = 0; j < k; ++j) {
        const int tpos = p + (j - c) * d, sh = fdiv(tpos, g), h = tpos - sh * g;
        const int seg = (int)(std::lower_bound(shifts.begin(), shifts.end(), sh) - shifts.begin());
        for (int o = 0; o < O; ++o)
          for (int i = 0; i < I; ++i)
            pk[((size_t)(p * O + o) * nseg + seg) * kc + h * I + i] = w.data[((size_t)o * I + i) * k + j];
      }
    return pk;
  };
  auto tile_bias = [&](const std::vector<float>& b, int g) {
    std::vector<float> o((size_t)round_up((int)b.size() * g, 128), 0.f);
    for (int p = 0; p < g; ++p) std::copy(b.begin(), b.end(), o.begin() + (size_t)p * b.size());
    return o;
  };
  const HostTensor& win = any("input_conv.weight");
  if (win.dims.size() != 3 || win.dims[2] % 2 != 1) fail("vocoder.input_conv.weight must be (C, in, odd k)");
  VocW& v = voc_;
  v.c0 = (int)win.dims[0];
  v.in_ch = (int)win.dims[1];
  v.k_in = (int)win.dims[2];
  v.in_pad = round_up(v.in_ch, BK);
  if (v.in_ch % 8 != 0 || v.c0 % BK != 0) fail("vocoder: in_channels must be a multiple of 8 and channels of %d", BK);
  v.w_in = up_w(conv_pack(win, v.c0, v.in_pad), v.c0, v.k_in * v.in_pad);
  v.b_in = up_f32(pad_to(any("input_conv.bias").data, round_up(v.c0, 128)));
  int n_up = 0;
  while (has("upsamples." + std::to_string(n_up) + ".1.weight")) ++n_up;
  if (n_up < 1) fail("vocoder: no upsampling stages");
  int n_blocks = 0;
  while (has("blocks." + std::to_string(n_blocks) + ".convs1.0.1.weight")) ++n_blocks;
  if (n_blocks % n_up != 0 || n_blocks / n_up < 1 || n_blocks / n_up > 4) fail("vocoder: 1..4 residual blocks per stage expected");
  v.n_blk = n_blocks / n_up;
  int n_dil = 0;
  while (has("blocks.0.convs1." + std::to_string(n_dil) + ".1.weight")) ++n_dil;
  v.n_dil = n_dil;
  if ((int)voc_dilations.size() != n_dil) fail("vocoder: %d dilations configured but the blocks have %d convs", (int)voc_dilations.size(), n_dil);
  v.hop = 1;
  int cin = v.c0;
  for (int i = 0; i < n_up; ++i) {
    VocStage sg;
    const HostTensor& wu = any("upsamples." + std::to_string(i) + ".1.weight");     // ConvTranspose1d: (in, out, k)
    if ((int)wu.dims[0] != cin || wu.dims[2] % 2 != 0) fail("vocoder.upsamples.%d.1.weight: unexpected shape", i);
    const int co = (int)wu.dims[1], ks = (int)wu.dims[2], s = ks / 2, pad = s / 2 + s % 2;   // hifigan.py:95, :103-106
    sg.s = s;
    sg.c = co;
    sg.cp = round_up(co, BK);
    if (co % 8 != 0) fail("vocoder: channel counts must be multiples of 8");
    const int cip = round_up(cin, BK);
    // out[s q + phi] = sum_{j in -1,0,1} W[:, :, phi + pad + s j]^T x[q - j]; K segments: x[q-1] | x[q] | x[q+1]
    std::vector<float> pk((size_t)s * co * 3 * cip, 0.f);
    for (int phi = 0; phi < s; ++phi)
      for (int j = -1; j <= 1; ++j) {
        const int k = phi + pad + s * j;
        if (k < 0 || k >= ks) continue;
        const int segi = j == 1 ? 0 : (j == 0 ? 1 : 2);
        for (int o = 0; o < co; ++o)
          for (int c = 0; c < cin; ++c)
            pk[((size_t)(phi * co + o) * 3 + segi) * cip + c] = wu.data[((size_t)c * co + o) * ks + k];
      }
    sg.w_up = up_w(pk, s * co, 3 * cip);
    std::vector<float> bu((size_t)round_up(s * co, 128), 0.f);
    const std::vector<float>& b = any("upsamples." + std::to_string(i) + ".1.bias").data;
    for (int phi = 0; phi < s; ++phi)
      for (int o = 0; o < co; ++o) bu[(size_t)phi * co + o] = b[o];
    sg.b_up = up_f32(bu);
    // Row grouping.  The activations of a stage are flat [T][C] arrays, so a convolution may read and write them as
    // [T / g][g C] for any g that divides the stage length: g C must be whole K blocks (no padding), at most one
    // 256-column tile, and the conv's distinct row shifts must fit the GEMM's K segments.  Per conv the g with the
    // fewest K blocks per time step wins (ties: the larger g - fuller tiles, all epilogue warps busy).
    int hop_here = v.hop * s;
    const int seg_limit = OpTraits<OT>::kTF32 ? 12 : 24;
    auto n_shifts = [&](int g, int k, int d) {
      auto fdiv = [](int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); };
      std::vector<int> sh;
      for (int p = 0; p < g; ++p)
        for (int jj = 0; jj < k; ++jj) sh.push_back(fdiv(p + (jj - (k - 1) / 2) * d, g));
      std::sort(sh.begin(), sh.end());
      sh.erase(std::unique(sh.begin(), sh.end()), sh.end());
      return (int)sh.size();
    };
    auto pick_g = [&](int k, int d) {        // 0 = no unpadded grouping possible
      int best = 0;
      double best_cost = 1e30;
      for (int g : {1, 2, 4, 8}) {
        if ((g * co) % BK != 0 || g * co > 256 || hop_here % g != 0) continue;
        const int ns = n_shifts(g, k, d);
        if (ns > seg_limit) continue;
        const double cost = (double)ns * (g * co / BK) / g;
        if (cost <= best_cost) { best_cost = cost; best = g; }
      }
      return best;
    };
    sg.padded = false;
    for (int jb = 0; jb < v.n_blk; ++jb) {
      const int k = (int)any("blocks." + std::to_string(i * v.n_blk + jb) + ".convs1.0.1.weight").dims[2];
      for (int d : voc_dilations)
        if (pick_g(k, d) == 0 || pick_g(k, 1) == 0) sg.padded = true;
    }
    if (i < n_up - 1 && co % BK != 0) sg.padded = true;      // the next transposed conv reads rows padded to a K block
    for (int jb = 0; jb < v.n_blk; ++jb) {
      VocBlock blk;
      const std::string p = "blocks." + std::to_string(i * v.n_blk + jb);
      blk.sh1.resize(n_dil);
      blk.sh2.resize(n_dil);
      for (int d = 0; d < n_dil; ++d) {
        const HostTensor& w1 = any(p + ".convs1." + std::to_string(d) + ".1.weight");
        const HostTensor& w2 = any(p + ".convs2." + std::to_string(d) + ".1.weight");
        if ((int)w1.dims[0] != co || (int)w1.dims[1] != co || w1.dims[2] % 2 != 1 || w2.dims != w1.dims)
          fail("vocoder.%s: unexpected conv shapes", p.c_str());
        blk.k = (int)w1.dims[2];
        const int ga = sg.padded ? 1 : pick_g(blk.k, voc_dilations[d]), gb = sg.padded ? 1 : pick_g(blk.k, 1);
        const int kc1 = sg.padded ? sg.cp : ga * co, kc2 = sg.padded ? sg.cp : gb * co;
        const int n1 = sg.padded ? sg.cp : ga * co, n2 = gb * co;
        blk.g1.push_back(ga);
        blk.g2.push_back(gb);
        const std::vector<float> p1 = conv_pack_grouped(w1, ga, voc_dilations[d], n1, kc1, blk.sh1[d]);
        const std::vector<float> p2 = conv_pack_grouped(w2, gb, 1, n2, kc2, blk.sh2[d]);
        if ((int)std::max(blk.sh1[d].size(), blk.sh2[d].size()) > seg_limit)
          fail("vocoder: conv kernel size %d needs more K segments than the GEMM has", blk.k);
        blk.w1.push_back(up_w(p1, n1, (int)blk.sh1[d].size() * kc1));
        blk.b1.push_back(up_f32(tile_bias(any(p + ".convs1." + std::to_string(d) + ".1.bias").data, ga)));
        blk.w2.push_back(up_w(p2, n2, (int)blk.sh2[d].size() * kc2));
        blk.b2.push_back(up_f32(tile_bias(any(p + ".convs2." + std::to_string(d) + ".1.bias").data, gb)));
      }
      sg.blocks.push_back(std::move(blk));
    }
    v.stages.push_back(std::move(sg));
    v.hop *= s;
    cin = co;
  }
  const HostTensor& wo = any("output_conv.1.weight");
  if ((int)wo.dims[0] != 1 || (int)wo.dims[1] != cin || wo.dims[2] % 2 != 1) fail("vocoder.output_conv.1.weight must be (1, C_last, odd k)");
  v.k_out = (int)wo.dims[2];
  {   // output conv on the last stage's (possibly grouped) rows; N = g outputs per row, padded to 4 (16-byte f32 rows)
    const VocStage& last = v.stages.back();
    v.g_out = 1;
    v.kc_out = last.cp;
    if (!last.padded)
      for (int g : {1, 2, 4})
        if ((g * last.c) % BK == 0 && g * last.c <= 128 && v.hop % g == 0) {
          v.g_out = g;
          v.kc_out = g * last.c;
        }
    const std::vector<float> po = conv_pack_grouped(wo, v.g_out, 1, 4, v.kc_out, v.sh_out);
    v.w_out = up_w(po, 4, (int)v.sh_out.size() * v.kc_out);
    v.b_out = up_f32(tile_bias(any("output_conv.1.bias").data, v.g_out));
  }
  v.present = true;
}

// One GEMM of the vocoder.  The A segments of `gp` are already set.
template <typename OT>
void Engine<OT>::voc_gemm(GemmParams& gp, const OT* W, int N, int K, const float* bias, int M, int nbatch, int mode,
                          void* out, long long out_bstride, int ldo, float slope, cudaStream_t st, const float* resid,
                          OT* act_out) {
  const int mtiles = cdiv(M, 256);
  const int bn = pick_bn(N, (long long)mtiles * nbatch);
  set_B_weight<OT>(gp, W, N, K, bn);
  gp.M = M;
  gp.N = N;
  gp.bias = bias;
  gp.out = out;
  gp.out_bstride = out_bstride;
  gp.ldo = ldo;
  gp.act_slope = slope;
  if (mode == VOC_OT) {
    gp.epi = EPI_STORE_OT;
    set_D<OT>(gp, static_cast<OT*>(out), N, M, out_bstride, ldo, nbatch);
  } else if (mode == VOC_RESID_ACT) {
    // out f32 = resid + conv (resid may be `out` itself); act_out = LeakyReLU(out) as the next conv's operand
    if (N % (128 / (int)sizeof(OT)) != 0) fail("vocoder: fused residual + activation needs whole 128-byte operand rows");
    gp.epi = EPI_RESID_ACT;
    set_D<float>(gp, static_cast<float*>(out), N, M, out_bstride, ldo, nbatch);
    gp.resid = resid;      // null: no residual, only the two stores
    if (resid != nullptr) gp.mapR = make_tmap_3d(resid, sizeof(float), N, M, nbatch, ldo, (uint64_t)out_bstride, 32, 32);
    gp.mapD2 = make_tmap_3d(act_out, sizeof(OT), N, M, nbatch, ldo, (uint64_t)out_bstride, 128 / sizeof(OT), 32);
  } else {
    gp.epi = mode == VOC_RESID ? EPI_RESID_F32 : EPI_STORE_F32;
    if (mode == VOC_RESID) {
      gp.resid = static_cast<float*>(out);
      gp.resid_bstride = out_bstride;
      gp.ldr = ldo;
    }
    set_D<float>(gp, static_cast<float*>(out), N, M, out_bstride, ldo, nbatch);
  }
  launch_gemm(gp, bn, dim3(cdiv(N, bn), mtiles, nbatch), st);
}

// Public entry: one CUDA graph per (B, T).  Inputs are staged into engine-owned buffers (a graph bakes its pointers),
// the ~120 launches of the generator replay as one graph launch, the waveform is copied out.  For a single 5 s clip
// the per-launch host work (a tensor-map encode per operand) is otherwise longer than the kernels.
template <typename OT>
void Engine<OT>::vocoder_decode(const float* mel, const float* scale, const float* shift, int B, int T, float* wave,
                                cudaStream_t st) {
  const VocW& v = voc_;
  if (!v.present) fail("vocoder weights (vocoder.*) were not loaded");
  if (B < 1 || T < 1) fail("vocoder_decode: empty input");
  if ((scale == nullptr) != (shift == nullptr)) fail("vocoder_decode: scale and shift come together");
  // sizes of the largest stage
  long long rows_max = (long long)B * T, ua_max = (long long)B * T * std::max(v.in_pad, v.c0);
  long long f32_max = 0, op_max = 0;
  {
    long long t = T;
    for (const VocStage& sg : v.stages) {
      t *= sg.s;
      f32_max = std::max(f32_max, (long long)B * t * sg.c);
      op_max = std::max(op_max, (long long)B * t * sg.cp);
      rows_max = std::max(rows_max, (long long)B * t);
    }
  }
  ua_max = std::max(ua_max, op_max);
  bool moved = false;      // a scratch buffer was reallocated: graphs that baked its address are stale
  auto grow_t = [&](DevBuf& b, size_t bytes) {
    if (b.bytes < bytes) {
      b.alloc(bytes);
      moved = true;
    }
  };
  grow_t(vb_ua_[0], (size_t)ua_max * sizeof(OT));
  grow_t(vb_ua_[1], (size_t)ua_max * sizeof(OT));
  grow_t(vb_c_, (size_t)f32_max * sizeof(float));
  for (int j = 0; j < v.n_blk; ++j) grow_t(vb_x_[j], (size_t)f32_max * sizeof(float));
  grow_t(vb_xa_, (size_t)op_max * sizeof(OT));
  grow_t(vb_ta_, (size_t)op_max * sizeof(OT));
  grow_t(vb_y4_, (size_t)rows_max * 4 * sizeof(float));
  grow_t(vb_mel_, (size_t)B * T * v.in_ch * sizeof(float));
  grow_t(vb_wave_, (size_t)rows_max * sizeof(float));
  grow_t(vb_aff_, (size_t)2 * v.in_ch * sizeof(float));
  if (moved) {
    for (auto& g : voc_graphs_)
      if (g.exec) cudaGraphExecDestroy(g.exec);
    voc_graphs_.clear();
  }
  const bool affine = scale != nullptr;
  const size_t mel_bytes = (size_t)B * T * v.in_ch * sizeof(float);
  const size_t wave_bytes = (size_t)B * T * v.hop * sizeof(float);
  float* aff = vb_aff_.template as<float>();
  if (!use_graph || debug) {
    voc_body(mel, scale, shift, B, T, wave, st);
    return;
  }
  SRN_CUDA(cudaMemcpyAsync(vb_mel_.p, mel, mel_bytes, cudaMemcpyDeviceToDevice, st));
  if (affine) {
    SRN_CUDA(cudaMemcpyAsync(aff, scale, (size_t)v.in_ch * sizeof(float), cudaMemcpyDeviceToDevice, st));
    SRN_CUDA(cudaMemcpyAsync(aff + v.in_ch, shift, (size_t)v.in_ch * sizeof(float), cudaMemcpyDeviceToDevice, st));
  }
  VocGraph* vg = nullptr;
  for (auto& g : voc_graphs_)
    if (g.B == B && g.T == T && g.affine == affine) vg = &g;
  if (vg == nullptr) {
    if (voc_graphs_.size() >= 8) {
      if (voc_graphs_.front().exec) cudaGraphExecDestroy(voc_graphs_.front().exec);
      voc_graphs_.pop_front();
    }
    if (capture_stream_ == nullptr) SRN_CUDA(cudaStreamCreateWithFlags(&capture_stream_, cudaStreamNonBlocking));
    const int64_t before = launches;
    cudaGraph_t g = nullptr;
    SRN_CUDA(cudaStreamBeginCapture(capture_stream_, cudaStreamCaptureModeThreadLocal));
    try {
      voc_body(vb_mel_.template as<float>(), affine ? aff : nullptr, affine ? aff + v.in_ch : nullptr, B, T,
               vb_wave_.template as<float>(), capture_stream_);
    } catch (...) {
      cudaStreamEndCapture(capture_stream_, &g);
      if (g) cudaGraphDestroy(g);
      launches = before;
      throw;
    }
    SRN_CUDA(cudaStreamEndCapture(capture_stream_, &g));
    VocGraph ng;
    ng.B = B;
    ng.T = T;
    ng.affine = affine;
    ng.nodes = launches - before;
    launches = before;
    const cudaError_t e = cudaGraphInstantiate(&ng.exec, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) fail("cudaGraphInstantiate (vocoder) failed: %s", cudaGetErrorString(e));
    voc_graphs_.push_back(ng);
    vg = &voc_graphs_.back();
  }
  SRN_CUDA(cudaGraphLaunch(vg->exec, st));
  launches += vg->nodes;
  SRN_CUDA(cudaMemcpyAsync(wave, vb_wave_.p, wave_bytes, cudaMemcpyDeviceToDevice, st));
}

// The generator itself (launches only; every scratch buffer is already large enough).
template <typename OT>
void Engine<OT>::voc_body(const float* mel, const float* scale, const float* shift, int B, int T, float* wave,
                          cudaStream_t st) {
  const VocW& v = voc_;
  auto base = [&]() {
    GemmParams gp;
    memset(&gp, 0, sizeof gp);
    gp.HB = 1;
    gp.alpha = 1.0f;
    gp.act_slope = 1.0f;
    gp.mask_step = 1;
    return gp;
  };
  auto blocks_for = [](long long n) { return dim3((unsigned)((n + 255) / 256)); };
  // Window mode (engine_plan.inl: maybe_window): the segments of a conv are row shifts of one operand
  auto maybe_win = [&](GemmParams& gp, const OT* a, int cols, long long rows_per_clip) {
    maybe_window<OT>(gp, a, cols, (int)rows_per_clip, rows_per_clip, cols, B);
  };
  auto conv_segs = [&](GemmParams& gp, const OT* a, int cp, int t, int k, int d) {
    set_A<OT>(gp, 0, a, cp, t, t, cp, B);
    for (int j = 0; j < k; ++j) add_seg(gp, 0, 0, (j - (k - 1) / 2) * d, cp / BK);
    maybe_win(gp, a, cp, t);
  };
  auto shift_segs = [&](GemmParams& gp, const OT* a, int kc, long long rows_per_clip, const std::vector<int>& shifts) {
    set_A<OT>(gp, 0, a, kc, (int)rows_per_clip, rows_per_clip, kc, B);
    for (int sh : shifts) add_seg(gp, 0, 0, sh, kc / BK);
    maybe_win(gp, a, kc, rows_per_clip);
  };
  OT* ua = vb_ua_[0].template as<OT>();
  OT* ub = vb_ua_[1].template as<OT>();
  // normalised mel -> operand; input conv (hifigan.py:177), LeakyReLU of the first upsampling stage fused
  launch_k(voc_prepare_mel<OT>, blocks_for((long long)B * T * (v.in_pad / 8)), dim3(256), st, mel, (long long)B * T,
           v.in_ch, v.in_pad, scale, shift, ua);
  ++launches;
  {
    GemmParams gp = base();
    conv_segs(gp, ua, v.in_pad, T, v.k_in, 1);
    voc_gemm(gp, v.w_in, v.c0, v.k_in * v.in_pad, v.b_in, T, B, VOC_OT, ub, (long long)T * v.c0, v.c0, 0.1f, st);
  }
  std::swap(ua, ub);                    // ua = LeakyReLU(c_0) [B][T][C0]
  long long t = T;
  int cin_p = v.c0;
  float* c = vb_c_.template as<float>();
  OT* xa = vb_xa_.template as<OT>();
  OT* ta = vb_ta_.template as<OT>();
  for (size_t i = 0; i < v.stages.size(); ++i) {
    const VocStage& sg = v.stages[i];
    const long long tn = t * sg.s, rows = (long long)B * tn;
    if (tn > 0x7fffffffLL / 2) fail("vocoder: sequence too long");
    // Unpadded rows: every second conv adds its residual from a TMA-loaded box, stores the f32
    // stream AND its LeakyReLU'd operand copy (EPI_RESID_ACT) - no cast pass between dilations, and the blocks'
    // streams start from `c` directly (no copies).  Otherwise: cast kernels + in-place reduce-add.
    const bool fuse = !sg.padded;
    const bool up_act = fuse && (sg.s * sg.c) % (128 / (int)sizeof(OT)) == 0;
    {   // ConvTranspose1d as one GEMM over rows q-1, q, q+1; the output [T][s C] is [s T][C]
      GemmParams gp = base();
      set_A<OT>(gp, 0, ua, cin_p, (int)t, t, cin_p, B);
      add_seg(gp, 0, 0, -1, cin_p / BK);
      add_seg(gp, 0, 0, 0, cin_p / BK);
      add_seg(gp, 0, 0, 1, cin_p / BK);
      maybe_win(gp, ua, cin_p, t);
      if (up_act)      // c and xa = LeakyReLU(c), the operand of every block's first conv, from one epilogue
        voc_gemm(gp, sg.w_up, sg.s * sg.c, 3 * cin_p, sg.b_up, (int)t, B, VOC_RESID_ACT, c, t * sg.s * sg.c, sg.s * sg.c, 0.1f,
                 st, nullptr, xa);
      else
        voc_gemm(gp, sg.w_up, sg.s * sg.c, 3 * cin_p, sg.b_up, (int)t, B, VOC_F32, c, t * sg.s * sg.c, sg.s * sg.c, 1.0f, st);
    }
    float* x[4] = {nullptr, nullptr, nullptr, nullptr};
    for (int j = 0; j < v.n_blk; ++j) x[j] = vb_x_[j].template as<float>();
    // operand tensors of this stage: [T][cw] with cw = C (grouped rows, no padding) or C padded to a K block
    const int cw = sg.padded ? sg.cp : sg.c;
    // operand of every block's first conv [+ the blocks' own copies of the residual stream]
    if (!up_act) {
      launch_k(voc_lrelu_cast<OT>, blocks_for(rows * (cw / 8)), dim3(256), st, (const float*)c, rows, sg.c, cw, 0.1f, xa,
               fuse ? (float*)nullptr : x[0], !fuse && v.n_blk > 1 ? x[1] : (float*)nullptr,
               !fuse && v.n_blk > 2 ? x[2] : (float*)nullptr);
      ++launches;
    }
    if (!fuse && v.n_blk > 3)
      SRN_CUDA(cudaMemcpyAsync(x[3], c, (size_t)rows * sg.c * sizeof(float), cudaMemcpyDeviceToDevice, st));
    for (int j = 0; j < v.n_blk; ++j) {
      const VocBlock& blk = sg.blocks[j];
      for (int d = 0; d < v.n_dil; ++d) {
        // every block's first operand is LeakyReLU(c) = xa, shared; afterwards LeakyReLU of its own stream
        OT* a_in = xa;
        if (d > 0) a_in = ub;      // (free here: ua holds this stage's input only until the transposed conv above has run)
        if (d > 0 && !fuse) {
          launch_k(voc_lrelu_cast<OT>, blocks_for(rows * (cw / 8)), dim3(256), st, (const float*)x[j], rows, sg.c, cw,
                   0.1f, a_in, (float*)nullptr, (float*)nullptr, (float*)nullptr);
          ++launches;
        }
        {   // xt = convs1[d](x): LeakyReLU + dilated conv (residual_block.py:253); the LeakyReLU of convs2 is fused
          GemmParams gp = base();
          const int g = blk.g1[d], kc = sg.padded ? sg.cp : g * sg.c, n = kc;
          const long long tg = tn / g;
          shift_segs(gp, a_in, kc, tg, blk.sh1[d]);
          voc_gemm(gp, blk.w1[d], n, (int)blk.sh1[d].size() * kc, blk.b1[d], (int)tg, B, VOC_OT, ta, tg * n, n, 0.1f, st);
        }
        {   // x = convs2[d](xt) + x (residual_block.py:255-256)
          GemmParams gp = base();
          const int g = blk.g2[d], kc = sg.padded ? sg.cp : g * sg.c, n = g * sg.c;
          const long long tg = tn / g;
          shift_segs(gp, ta, kc, tg, blk.sh2[d]);
          if (fuse && (d == 0 || d + 1 < v.n_dil))      // (d == 0: the residual is c, not x[j]; the last operand copy is unused)
            voc_gemm(gp, blk.w2[d], n, (int)blk.sh2[d].size() * kc, blk.b2[d], (int)tg, B, VOC_RESID_ACT, x[j], tg * n, n, 0.1f,
                     st, d == 0 ? c : x[j], ub);
          else
            voc_gemm(gp, blk.w2[d], n, (int)blk.sh2[d].size() * kc, blk.b2[d], (int)tg, B, VOC_RESID, x[j], tg * n, n, 1.0f, st);
        }
      }
    }
    // c = mean of the blocks (hifigan.py:180-183); next consumer's LeakyReLU fused: 0.1 before an upsampling stage,
    // torch's default 0.01 before the output conv (hifigan.py:137-140)
    const bool last = i + 1 == v.stages.size();
    launch_k(voc_mean_lrelu<OT>, blocks_for(rows * (cw / 8)), dim3(256), st, (const float*)x[0], (const float*)x[1],
             (const float*)x[2], (const float*)x[3], v.n_blk, rows, sg.c, cw, last ? 0.01f : 0.1f, ua);
    ++launches;
    t = tn;
    cin_p = cw;
  }
  {   // output conv k7 -> 1 channel (g per grouped row, N padded to 4), then tanh
    GemmParams gp = base();
    const long long tg = t / v.g_out;
    shift_segs(gp, ua, v.kc_out, tg, v.sh_out);
    voc_gemm(gp, v.w_out, 4, (int)v.sh_out.size() * v.kc_out, v.b_out, (int)tg, B, VOC_F32, vb_y4_.template as<float>(), tg * 4,
             4, 1.0f, st);
  }
  launch_k(voc_tanh_out, blocks_for((long long)B * t), dim3(256), st, (const float*)vb_y4_.template as<float>(),
           (long long)B * t, v.g_out, 4, wave);
  ++launches;
  SRN_CUDA(cudaGetLastError());
}

}  // namespace srn
