// Engine<OT>: conditioning front half of Serenade.inference.  Included by engine.cu.
//   content_encode  Conv1dResnet (serenade.py:224-376): every convolution is a launch of the GEMM core over
//                   reflection-padded channels-last operands; the ResnetBlock's 1x1 shortcut is a second K segment
//                   of its last 1x1 conv (K = [LeakyReLU(c) | h], weights [W_1x1 | W_shortcut], biases summed).
//   style_encode    StyleEncoder (modules/gst/style_encoder.py): six Conv2d(k3, s2)+BatchNorm+ReLU as im2col GEMMs
//                   (BatchNorm folded into weight and bias, ReLU in the epilogue), then GRU + style-token attention
//                   in one small kernel.
namespace srn {

namespace {
inline std::vector<float> fold_weight_norm(const HostTensor& v, const HostTensor& g) {
  const int O = (int)v.dims[0];
  const size_t per = v.data.size() / O;
  std::vector<float> w(v.data.size());
  for (int o = 0; o < O; ++o) {
    double n2 = 0.0;
    for (size_t i = 0; i < per; ++i) n2 += (double)v.data[o * per + i] * v.data[o * per + i];
    const float s = (float)((double)g.data[o] / std::sqrt(n2));
    for (size_t i = 0; i < per; ++i) w[o * per + i] = v.data[o * per + i] * s;
  }
  return w;
}
}  // namespace

template <typename OT>
void Engine<OT>::pack_frontend() {
  front_.present = false;
  if (host_w_.find("encoder.model.1.weight_v") == host_w_.end()) return;     // decoder-only state dict
  const int C = 512;
  auto shape_of = [&](const std::string& k) -> const std::vector<int64_t>& {
    auto it = host_w_.find(k);
    if (it == host_w_.end()) fail("missing weight '%s'", k.c_str());
    return it->second.dims;
  };
  auto any = [&](const std::string& k) -> const HostTensor& {
    auto it = host_w_.find(k);
    if (it == host_w_.end()) fail("missing weight '%s'", k.c_str());
    return it->second;
  };
  // ---- content encoder
  const auto& d1 = shape_of("encoder.model.1.weight_v");
  if (d1.size() != 3 || d1[0] != C || d1[2] != 7) fail("encoder.model.1.weight_v must be (512, in_dim, 7)");
  front_.in_dim = (int)d1[1];
  front_.in_pad = round_up(front_.in_dim, 64);
  auto conv_taps = [&](const std::string& p, int O, int I, int Ip, int k) {   // (O, I, k) -> [O][k][Ip]
    const std::vector<float> w = fold_weight_norm(hw(p + ".weight_v", {O, I, k}), hw(p + ".weight_g", {O, 1, 1}));
    std::vector<float> pk((size_t)O * k * Ip, 0.f);
    for (int o = 0; o < O; ++o)
      for (int i = 0; i < I; ++i)
        for (int t = 0; t < k; ++t) pk[((size_t)o * k + t) * Ip + i] = w[((size_t)o * I + i) * k + t];
    return pk;
  };
  front_.w_in = up_w(conv_taps("encoder.model.1", C, front_.in_dim, front_.in_pad, 7), C, 7 * front_.in_pad);
  front_.b_in = up_f32(hw("encoder.model.1.bias", {C}).data);
  for (int n = 0; n < 2; ++n) {
    const std::string p = "encoder.model." + std::to_string(2 + n);
    front_.blk[n].w3 = up_w(conv_taps(p + ".block.2", C, C, C, 3), C, 3 * C);
    front_.blk[n].b3 = up_f32(hw(p + ".block.2.bias", {C}).data);
    const std::vector<float> w1 = conv_taps(p + ".block.4", C, C, C, 1), ws = conv_taps(p + ".shortcut", C, C, C, 1);
    std::vector<float> cat((size_t)C * 2 * C), bsum(C);
    for (int o = 0; o < C; ++o) {
      memcpy(&cat[(size_t)o * 2 * C], &w1[(size_t)o * C], C * sizeof(float));
      memcpy(&cat[(size_t)o * 2 * C + C], &ws[(size_t)o * C], C * sizeof(float));
      bsum[o] = hw(p + ".block.4.bias", {C}).data[o] + hw(p + ".shortcut.bias", {C}).data[o];
    }
    front_.blk[n].w1sc = up_w(cat, C, 2 * C);
    front_.blk[n].b1sc = up_f32(bsum);
  }
  const auto& d6 = shape_of("encoder.model.6.weight_v");
  front_.enc_out = (int)d6[0];
  if (front_.enc_out % 4 != 0 || front_.enc_out > 128) fail("encoder output dim must be a multiple of 4 and <= 128");
  front_.w_out = up_w(conv_taps("encoder.model.6", front_.enc_out, C, C, 7), front_.enc_out, 7 * C);
  front_.b_out = up_f32(hw("encoder.model.6.bias", {front_.enc_out}).data);
  // ---- GST reference encoder: Conv2d (O, I, 3, 3) + eval BatchNorm -> [O][(kh, kw, i)] * scale, bias = beta - mean * scale
  int cin = 1, f = 80;
  for (int l = 0; l < 6; ++l) {
    const std::string c = "gst.ref_enc.convs." + std::to_string(3 * l), bnk = "gst.ref_enc.convs." + std::to_string(3 * l + 1);
    const auto& d = shape_of(c + ".weight");
    if (d.size() != 4 || d[1] != cin || d[2] != 3 || d[3] != 3) fail("unexpected shape of %s.weight", c.c_str());
    const int O = (int)d[0];
    const HostTensor& w = hw(c + ".weight", {O, cin, 3, 3});
    const HostTensor &ga = hw(bnk + ".weight", {O}), &be = hw(bnk + ".bias", {O}), &rm = hw(bnk + ".running_mean", {O}),
                     &rv = hw(bnk + ".running_var", {O});
    const int Kp = round_up(9 * cin, 64);
    std::vector<float> pk((size_t)O * Kp, 0.f), bias(O);
    for (int o = 0; o < O; ++o) {
      const float sc = ga.data[o] / std::sqrt(rv.data[o] + 1e-5f);
      bias[o] = be.data[o] - rm.data[o] * sc;
      for (int i = 0; i < cin; ++i)
        for (int t = 0; t < 9; ++t) pk[(size_t)o * Kp + t * cin + i] = w.data[((size_t)o * cin + i) * 9 + t] * sc;
    }
    front_.gw[l] = up_w(pk, O, Kp);
    front_.gb[l] = up_f32(bias);
    front_.gc[l] = cin;
    front_.gkp[l] = Kp;
    cin = O;
    f = (f - 1) / 2 + 1;
  }
  front_.gc[6] = cin;
  front_.gst_f = f;
  const int G = f * cin, H = kGruH;
  const HostTensor& wih = hw("gst.ref_enc.gru.weight_ih_l0", {3 * H, G});
  const HostTensor& whh = hw("gst.ref_enc.gru.weight_hh_l0", {3 * H, H});
  {  // reference feature index = c * F' + f (NCHW view, style_encoder.py:185-186); ours = f * C + c (NHWC)
    std::vector<float> t((size_t)G * 3 * H);
    for (int j = 0; j < 3 * H; ++j)
      for (int c = 0; c < cin; ++c)
        for (int ff = 0; ff < f; ++ff) t[(size_t)(ff * cin + c) * 3 * H + j] = wih.data[(size_t)j * G + c * f + ff];
    front_.w_ihT = up_f32(t);
    std::vector<float> t2((size_t)H * 3 * H);
    for (int j = 0; j < 3 * H; ++j)
      for (int k = 0; k < H; ++k) t2[(size_t)k * 3 * H + j] = whh.data[(size_t)j * H + k];
    front_.w_hhT = up_f32(t2);
  }
  front_.b_ih = up_f32(hw("gst.ref_enc.gru.bias_ih_l0", {3 * H}).data);
  front_.b_hh = up_f32(hw("gst.ref_enc.gru.bias_hh_l0", {3 * H}).data);
  const auto& de = shape_of("gst.stl.gst_embs");
  const int ntok = (int)de[0], dk = (int)de[1];
  if (ntok > 64 || dk != 64) fail("style token layer: at most 64 tokens of dim 64 are supported (got %d x %d)", ntok, dk);
  front_.ntok = ntok;
  const HostTensor& emb = any("gst.stl.gst_embs");
  auto lin_tok = [&](const std::string& nm) {      // linear(tanh(gst_embs)): input independent
    const HostTensor &w = hw("gst.stl.mha." + nm + ".weight", {256, dk}), &bb = hw("gst.stl.mha." + nm + ".bias", {256});
    std::vector<float> o((size_t)ntok * 256);
    for (int n = 0; n < ntok; ++n)
      for (int j = 0; j < 256; ++j) {
        double a = bb.data[j];
        for (int d = 0; d < dk; ++d) a += (double)w.data[(size_t)j * dk + d] * std::tanh((double)emb.data[(size_t)n * dk + d]);
        o[(size_t)n * 256 + j] = (float)a;
      }
    return o;
  };
  front_.kmat = up_f32(lin_tok("linear_k"));
  front_.vmat = up_f32(lin_tok("linear_v"));
  front_.wq = up_f32(hw("gst.stl.mha.linear_q.weight", {256, H}).data);
  front_.bq = up_f32(hw("gst.stl.mha.linear_q.bias", {256}).data);
  front_.wo = up_f32(hw("gst.stl.mha.linear_out.weight", {256, 256}).data);
  front_.bo = up_f32(hw("gst.stl.mha.linear_out.bias", {256}).data);
  SRN_CUDA(cudaFuncSetAttribute(gru_stl_kernel<OT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)((kGruChunk * G + kGruH + 6 * kGruH + 768) * sizeof(float))));
  front_.present = true;
}

namespace {
inline void grow(DevBuf& b, size_t bytes) {
  if (b.bytes < bytes) b.alloc(bytes);
}
}  // namespace

// One GEMM of the front half: D[z][m][:] = sum_segs A(seg)[z][m + shift][:] . W^T, epilogue STORE_OT or STORE_F32.
template <typename OT>
void Engine<OT>::front_gemm(GemmParams& gp, const OT* W, int N, int K, const float* bias, int M, int nbatch,
                            bool f32_out, void* out, long long out_bstride, int ldo, float slope, cudaStream_t st) {
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
  if (f32_out) {
    gp.epi = EPI_STORE_F32;
    set_D<float>(gp, static_cast<float*>(out), N, M, out_bstride, ldo, nbatch);
  } else {
    gp.epi = EPI_STORE_OT;
    set_D<OT>(gp, static_cast<OT*>(out), N, M, out_bstride, ldo, nbatch);
  }
  launch_gemm(gp, bn, dim3(cdiv(N, bn), mtiles, nbatch), st);
}

template <typename OT>
void Engine<OT>::content_encode(const float* x, int B, int L, float* out, cudaStream_t st) {
  if (!front_.present) fail("content encoder weights (encoder.*) were not loaded");
  if (B < 1 || L < 4) fail("content_encode: need B >= 1 and L >= 4 (ReflectionPad1d(3))");
  const int C = 512, Ci = front_.in_pad, P = 3, Lp = L + 2 * P;
  if (front_.in_pad != front_.in_dim) fail("input_dim must be a multiple of 64");
  grow(fb_xpad_, (size_t)B * Lp * Ci * sizeof(OT));
  grow(fb_h_, (size_t)B * L * C * sizeof(OT));
  grow(fb_h2_, (size_t)B * L * C * sizeof(OT));
  grow(fb_apad_, (size_t)B * Lp * C * sizeof(OT));
  grow(fb_e_, (size_t)B * L * C * sizeof(OT));
  OT *xpad = fb_xpad_.template as<OT>(), *h = fb_h_.template as<OT>(), *h2 = fb_h2_.template as<OT>(),
     *apad = fb_apad_.template as<OT>(), *e = fb_e_.template as<OT>();
  auto pad_rows = [&](auto* src, long long sbs, int Cc, int pad, float slope, OT* dst) {
    using IT = std::remove_pointer_t<decltype(src)>;
    const long long n = (long long)(L + 2 * pad) * (Cc / 8);
    launch_k(reflect_pad_rows<std::remove_const_t<IT>, OT>, dim3((unsigned)((n + 255) / 256), B), dim3(256), st,
             (const std::remove_const_t<IT>*)src, sbs, L, Cc, pad, slope, dst, (long long)(L + 2 * pad) * Cc);
    ++launches;
  };
  auto base = [&]() {
    GemmParams gp;
    memset(&gp, 0, sizeof gp);
    gp.HB = 1;
    gp.alpha = 1.0f;
    gp.act_slope = 1.0f;
    gp.mask_step = 1;
    return gp;
  };
  // k7 conv on the reflection-padded input (serenade.py:289-292)
  pad_rows(x, (long long)L * front_.in_dim, front_.in_dim, P, 1.0f, xpad);
  {
    GemmParams gp = base();
    set_A<OT>(gp, 0, xpad, Ci, Lp, Lp, Ci, B);
    for (int k = 0; k < 7; ++k) add_seg(gp, 0, 0, k, Ci / BK);
    front_gemm(gp, front_.w_in, C, 7 * Ci, front_.b_in, L, B, false, h, (long long)L * C, C, 1.0f, st);
  }
  for (int n = 0; n < 2; ++n) {      // ResnetBlock(dilation 2^n) (serenade.py:361-376)
    const int d = 1 << n, La = L + 2 * d;
    pad_rows((const OT*)h, (long long)L * C, C, d, 0.2f, apad);
    {
      GemmParams gp = base();
      set_A<OT>(gp, 0, apad, C, La, La, C, B);
      for (int k = 0; k < 3; ++k) add_seg(gp, 0, 0, k * d, C / BK);
      front_gemm(gp, front_.blk[n].w3, C, 3 * C, front_.blk[n].b3, L, B, false, e, (long long)L * C, C, 0.2f, st);
    }
    {
      GemmParams gp = base();
      set_A<OT>(gp, 0, e, C, L, L, C, B);
      set_A<OT>(gp, 1, h, C, L, L, C, B);
      add_seg(gp, 0, 0, 0, C / BK);
      add_seg(gp, 1, 0, 0, C / BK);
      front_gemm(gp, front_.blk[n].w1sc, C, 2 * C, front_.blk[n].b1sc, L, B, false, h2, (long long)L * C, C, 1.0f, st);
    }
    std::swap(h, h2);
  }
  pad_rows((const OT*)h, (long long)L * C, C, P, 0.2f, apad);
  {
    GemmParams gp = base();
    set_A<OT>(gp, 0, apad, C, Lp, Lp, C, B);
    for (int k = 0; k < 7; ++k) add_seg(gp, 0, 0, k, C / BK);
    front_gemm(gp, front_.w_out, front_.enc_out, 7 * C, front_.b_out, L, B, true, out, (long long)L * front_.enc_out,
               front_.enc_out, 1.0f, st);
  }
}

template <typename OT>
void Engine<OT>::style_encode(const float* mel, int B, int L, float* out, cudaStream_t st) {
  if (!front_.present) fail("style encoder weights (gst.*) were not loaded");
  if (B < 1 || L < 1) fail("style_encode: empty input");
  int H = L, W = 80;
  const void* in = mel;
  bool in_f32 = true;
  DevBuf* act[2] = {&fb_act0_, &fb_act1_};
  for (int l = 0; l < 6; ++l) {
    const int Cin = front_.gc[l], Cout = front_.gc[l + 1], Kp = front_.gkp[l];
    const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
    const long long M = (long long)B * Ho * Wo;
    grow(fb_col_, (size_t)M * Kp * sizeof(OT));
    grow(*act[l & 1], (size_t)M * Cout * sizeof(OT));
    OT* col = fb_col_.template as<OT>();
    if (in_f32)
      launch_k(im2col_k3s2<float, OT>, dim3((unsigned)M), dim3(128), st, static_cast<const float*>(in), H, W, Cin, Ho, Wo,
               Kp, col);
    else
      launch_k(im2col_k3s2<OT, OT>, dim3((unsigned)M), dim3(128), st, static_cast<const OT*>(in), H, W, Cin, Ho, Wo, Kp,
               col);
    ++launches;
    GemmParams gp;
    memset(&gp, 0, sizeof gp);
    gp.HB = 1;
    gp.alpha = 1.0f;
    gp.mask_step = 1;
    set_A<OT>(gp, 0, col, Kp, (int)M, M, Kp, 1);
    add_seg(gp, 0, 0, 0, Kp / BK);
    OT* o = act[l & 1]->template as<OT>();
    front_gemm(gp, front_.gw[l], Cout, Kp, front_.gb[l], (int)M, 1, false, o, M * Cout, Cout, 0.0f, st);   // ReLU
    in = o;
    in_f32 = false;
    H = Ho;
    W = Wo;
  }
  const int G = W * front_.gc[6];
  const size_t smem = (size_t)(kGruChunk * G + kGruH + 6 * kGruH + 768) * sizeof(float);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof cfg);
  cfg.gridDim = dim3(B);
  cfg.blockDim = dim3(384);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  SRN_CUDA(cudaLaunchKernelEx(&cfg, gru_stl_kernel<OT>, static_cast<const OT*>(in), H, G, (const float*)front_.w_ihT,
                              (const float*)front_.b_ih, (const float*)front_.w_hhT, (const float*)front_.b_hh,
                              (const float*)front_.wq, (const float*)front_.bq, (const float*)front_.kmat,
                              (const float*)front_.vmat, front_.ntok, (const float*)front_.wo, (const float*)front_.bo,
                              out));
  ++launches;
}

}  // namespace srn
