// Engine<OT>: construction and weight repacking.  Included by engine.cu.
namespace srn {

template <typename OT>
inline OT to_ot(float v);
template <>
inline float to_ot<float>(float v) {
  // TF32 operands: round to nearest-even here; the tensor core would otherwise TRUNCATE the low 13 mantissa bits
  uint32_t u;
  memcpy(&u, &v, 4);
  if ((u & 0x7f800000u) != 0x7f800000u) u = (u + 0xfffu + ((u >> 13) & 1u)) & ~0x1fffu;
  memcpy(&v, &u, 4);
  return v;
}
template <>
inline __nv_bfloat16 to_ot<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

template <typename OT>
Engine<OT>::Engine(const SrnConfig& cfg) : cfg_(cfg) {
  if (cfg.channels != 512 || cfg.attention_head_dim != 512 || cfg.num_heads != 4)
    fail("this build is specialised for the ssc1 decoder geometry: channels=512, heads=4, head_dim=512 "
         "(got %d, %d, %d)", cfg.channels, cfg.num_heads, cfg.attention_head_dim);
  if (cfg.out_channels < 1 || cfg.out_channels > 128) fail("out_channels must be in [1,128]");
  if (cfg.in_channels <= cfg.out_channels || cfg.in_channels > 256) fail("in_channels must be in (out_channels, 256]");
  if (cfg.in_channels % 2 != 0) fail("in_channels must be even (SinusoidalPosEmb, decoder.py:52)");
  if (cfg.spk_embed_dim < 1 || cfg.num_mid_blocks < 0) fail("bad spk_embed_dim / num_mid_blocks");
  cmu_ = cfg.in_channels - cfg.out_channels;
  int dev = 0;
  SRN_CUDA(cudaGetDevice(&dev));
  SRN_CUDA(cudaDeviceGetAttribute(&num_sms_, cudaDevAttrMultiProcessorCount, dev));
}

template <typename OT>
Engine<OT>::~Engine() {
  for (auto& g : voc_graphs_)
    if (g.exec) cudaGraphExecDestroy(g.exec);
  voc_graphs_.clear();
  for (auto& e : pin_ev_)
    if (e) cudaEventDestroy(e);
  if (pin_) cudaFreeHost(pin_);
  plans_.clear();
  arena_.clear();
}

template <typename OT>
void Engine<OT>::load_weight(const std::string& key, const float* data, int ndim, const int64_t* dims) {
  if (ndim < 1 || ndim > 4) fail("weight %s: ndim %d not supported", key.c_str(), ndim);
  HostTensor t;
  size_t n = 1;
  for (int i = 0; i < ndim; ++i) {
    if (dims[i] < 1) fail("weight %s: bad dim", key.c_str());
    t.dims.push_back(dims[i]);
    n *= (size_t)dims[i];
  }
  t.data.assign(data, data + n);
  host_w_[key] = std::move(t);
  finalized_ = false;
}

template <typename OT>
const HostTensor& Engine<OT>::hw(const std::string& key, std::initializer_list<int64_t> dims) {
  const bool front = key.compare(0, 8, "encoder.") == 0 || key.compare(0, 4, "gst.") == 0;
  auto it = host_w_.find(front ? key : "estimator." + key);
  if (it == host_w_.end()) fail("missing weight '%s%s'", front ? "" : "estimator.", key.c_str());
  if (it->second.dims != std::vector<int64_t>(dims)) {
    std::string got, want;
    for (auto d : it->second.dims) got += std::to_string(d) + ",";
    for (auto d : dims) want += std::to_string(d) + ",";
    fail("weight 'estimator.%s' has shape (%s) but (%s) is expected", key.c_str(), got.c_str(), want.c_str());
  }
  return it->second;
}

template <typename OT>
float* Engine<OT>::up_f32(const std::vector<float>& v) {
  arena_.emplace_back();
  arena_.back().alloc(v.size() * sizeof(float), false);
  SRN_CUDA(cudaMemcpy(arena_.back().p, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice));
  return arena_.back().template as<float>();
}

template <typename OT>
OT* Engine<OT>::up_ot(const std::vector<float>& v) {
  std::vector<OT> tmp(v.size());
  for (size_t i = 0; i < v.size(); ++i) tmp[i] = to_ot<OT>(v[i]);
  arena_.emplace_back();
  arena_.back().alloc(tmp.size() * sizeof(OT), false);
  SRN_CUDA(cudaMemcpy(arena_.back().p, tmp.data(), tmp.size() * sizeof(OT), cudaMemcpyHostToDevice));
  return arena_.back().template as<OT>();
}

// GEMM weight [N][K], K-major.  bf16 mode: rounded to bf16.  TF32 mode: every row is stored as [hi | lo] with
// hi = tf32(w), lo = tf32(w - hi), and the GEMM runs its K loop twice (A.hi + A.lo, see set_B_weight): the weights
// then carry ~21 mantissa bits and only the activation operand is rounded to TF32.
template <typename OT>
OT* Engine<OT>::up_w(const std::vector<float>& v, int N, int K) {
  if ((size_t)N * K != v.size()) fail("up_w: %d x %d does not match %zu values", N, K, v.size());
  if constexpr (!OpTraits<OT>::kTF32) {
    return up_ot(v);
  } else {
    std::vector<float> sp((size_t)N * 2 * K);
    for (int n = 0; n < N; ++n)
      for (int k = 0; k < K; ++k) {
        const float w = v[(size_t)n * K + k];
        const float hi = to_ot<float>(w);
        sp[(size_t)n * 2 * K + k] = hi;
        sp[(size_t)n * 2 * K + K + k] = to_ot<float>(w - hi);
      }
    return up_f32(sp);
  }
}

// Conv1d weight (O, I, 3) -> [O][3][cin_pad]: K index = tap * cin_pad + i  (tap 0 multiplies x[t-1]).
template <typename OT>
OT* Engine<OT>::pack_conv3(const HostTensor& w, int cin, int cin_pad, float*) {
  const int O = (int)w.dims[0];
  std::vector<float> p((size_t)O * 3 * cin_pad, 0.f);
  for (int o = 0; o < O; ++o)
    for (int i = 0; i < cin; ++i)
      for (int k = 0; k < 3; ++k) p[((size_t)o * 3 + k) * cin_pad + i] = w.data[((size_t)o * cin + i) * 3 + k];
  return up_w(p, O, 3 * cin_pad);
}

template <typename OT>
ResnetW<OT> Engine<OT>::pack_resnet(const std::string& p, int cin) {
  const int C = 512, te = 2048, sd = cfg_.spk_embed_dim;
  ResnetW<OT> r;
  r.cin_pad = (cin + 63) / 64 * 64;
  r.w1 = pack_conv3(hw(p + ".block1.block.0.weight", {C, cin, 3}), cin, r.cin_pad);
  r.b1 = up_f32(hw(p + ".block1.block.0.bias", {C}).data);
  r.g1 = up_f32(hw(p + ".block1.block.1.weight", {C}).data);
  r.be1 = up_f32(hw(p + ".block1.block.1.bias", {C}).data);
  r.w2 = pack_conv3(hw(p + ".block2.block.0.weight", {C, C, 3}), C, C);
  r.b2 = up_f32(hw(p + ".block2.block.0.bias", {C}).data);
  r.g2 = up_f32(hw(p + ".block2.block.1.weight", {C}).data);
  r.be2 = up_f32(hw(p + ".block2.block.1.bias", {C}).data);
  {
    const HostTensor& w = hw(p + ".res_conv.weight", {C, cin, 1});
    std::vector<float> pk((size_t)C * r.cin_pad, 0.f);
    for (int o = 0; o < C; ++o)
      for (int i = 0; i < cin; ++i) pk[(size_t)o * r.cin_pad + i] = w.data[(size_t)o * cin + i];
    r.wr = up_w(pk, C, r.cin_pad);
  }
  r.br = up_f32(hw(p + ".res_conv.bias", {C}).data);
  r.mlp_w = up_f32(hw(p + ".mlp.1.weight", {C, te}).data);
  r.mlp_b = up_f32(hw(p + ".mlp.1.bias", {C}).data);
  r.ws_w = up_f32(hw(p + ".speaker_projection.W_scale.weight", {C, sd}).data);
  r.ws_b = up_f32(hw(p + ".speaker_projection.W_scale.bias", {C}).data);
  r.wb_w = up_f32(hw(p + ".speaker_projection.W_bias.weight", {C, sd}).data);
  r.wb_b = up_f32(hw(p + ".speaker_projection.W_bias.bias", {C}).data);
  return r;
}

template <typename OT>
TfmW<OT> Engine<OT>::pack_tfm(const std::string& p) {
  const int C = 512, inner = 2048, ffi = 2048;
  TfmW<OT> t;
  t.ln1g = up_f32(hw(p + ".norm1.weight", {C}).data);
  t.ln1b = up_f32(hw(p + ".norm1.bias", {C}).data);
  t.ln3g = up_f32(hw(p + ".norm3.weight", {C}).data);
  t.ln3b = up_f32(hw(p + ".norm3.bias", {C}).data);
  {
    std::vector<float> qkv;
    qkv.reserve((size_t)3 * inner * C);
    for (const char* n : {".attn1.to_q.weight", ".attn1.to_k.weight", ".attn1.to_v.weight"}) {
      const HostTensor& w = hw(p + n, {inner, C});
      qkv.insert(qkv.end(), w.data.begin(), w.data.end());
    }
    t.wqkv = up_w(qkv, 3 * inner, C);
  }
  t.wo = up_w(hw(p + ".attn1.to_out.0.weight", {C, inner}).data, C, inner);
  t.bo = up_f32(hw(p + ".attn1.to_out.0.bias", {C}).data);
  {
    // GEGLU: rows [0, ffi) are the value half, [ffi, 2 ffi) the gate half (chunk(2, dim=-1)).
    // Interleave so that value i and gate i are adjacent accumulator columns of one epilogue thread.
    const HostTensor& w = hw(p + ".ff.net.0.proj.weight", {2 * ffi, C});
    const HostTensor& bb = hw(p + ".ff.net.0.proj.bias", {2 * ffi});
    std::vector<float> wi((size_t)2 * ffi * C), bi((size_t)2 * ffi);
    for (int i = 0; i < ffi; ++i) {
      memcpy(&wi[(size_t)(2 * i) * C], &w.data[(size_t)i * C], C * sizeof(float));
      memcpy(&wi[(size_t)(2 * i + 1) * C], &w.data[(size_t)(ffi + i) * C], C * sizeof(float));
      bi[2 * i] = bb.data[i];
      bi[2 * i + 1] = bb.data[ffi + i];
    }
    t.wgl = up_w(wi, 2 * ffi, C);
    t.bgl = up_f32(bi);
  }
  t.wf = up_w(hw(p + ".ff.net.2.weight", {C, ffi}).data, C, ffi);
  t.bf = up_f32(hw(p + ".ff.net.2.bias", {C}).data);
  return t;
}

template <typename OT>
void Engine<OT>::finalize_weights() {
  for (auto& g : voc_graphs_)
    if (g.exec) cudaGraphExecDestroy(g.exec);
  voc_graphs_.clear();
  arena_.clear();
  plans_.clear();
  resnets_.clear();
  tfms_.clear();
  size_t n_est = 0;
  for (const auto& kv : host_w_) n_est += kv.first.compare(0, 10, "estimator.") == 0;
  decoder_present_ = n_est > 0;
  if (decoder_present_) pack_decoder(n_est);
  pack_frontend();      // optional encoder.* / gst.* weights of the conditioning front half
  pack_vocoder();       // optional vocoder.* weights
  if (!decoder_present_ && !front_.present && !voc_.present) fail("no weights were loaded");
  host_w_.clear();
  SRN_CUDA(cudaDeviceSynchronize());
  finalized_ = true;
}

template <typename OT>
void Engine<OT>::pack_decoder(size_t n_est) {
  const int C = 512, te = 2048;
  time_w1_ = up_f32(hw("time_mlp.linear_1.weight", {te, cfg_.in_channels}).data);
  time_b1_ = up_f32(hw("time_mlp.linear_1.bias", {te}).data);
  time_w2_ = up_f32(hw("time_mlp.linear_2.weight", {te, te}).data);
  time_b2_ = up_f32(hw("time_mlp.linear_2.bias", {te}).data);
  // resnets / transformers in execution order: down0, down1, mid*, up0, up1
  resnets_.push_back(pack_resnet("down_blocks.0.0", cfg_.in_channels));
  tfms_.push_back(pack_tfm("down_blocks.0.1.0"));
  resnets_.push_back(pack_resnet("down_blocks.1.0", C));
  tfms_.push_back(pack_tfm("down_blocks.1.1.0"));
  for (int i = 0; i < cfg_.num_mid_blocks; ++i) {
    resnets_.push_back(pack_resnet("mid_blocks." + std::to_string(i) + ".0", C));
    tfms_.push_back(pack_tfm("mid_blocks." + std::to_string(i) + ".1.0"));
  }
  resnets_.push_back(pack_resnet("up_blocks.0.0", 2 * C));
  tfms_.push_back(pack_tfm("up_blocks.0.1.0"));
  resnets_.push_back(pack_resnet("up_blocks.1.0", 2 * C));
  tfms_.push_back(pack_tfm("up_blocks.1.1.0"));
  // Downsample1D conv k3 s2 (decoder.py:107) and the plain k3 convs (decoder.py:271, :335)
  down_conv_[0].w = pack_conv3(hw("down_blocks.0.2.conv.weight", {C, C, 3}), C, C);
  down_conv_[0].b = up_f32(hw("down_blocks.0.2.conv.bias", {C}).data);
  down_conv_[1].w = pack_conv3(hw("down_blocks.1.2.weight", {C, C, 3}), C, C);
  down_conv_[1].b = up_f32(hw("down_blocks.1.2.bias", {C}).data);
  {
    // ConvTranspose1d(512, 512, 4, 2, 1), weight (in, out, k) (decoder.py:191):
    //   out[2m]   = W[:,:,3]^T x[m-1] + W[:,:,1]^T x[m]      -> even parity, K = [x[m-1] | x[m]]
    //   out[2m+1] = W[:,:,2]^T x[m]   + W[:,:,0]^T x[m+1]    -> odd parity,  K = [x[m] | x[m+1]]
    const HostTensor& w = hw("up_blocks.0.2.conv.weight", {C, C, 4});
    std::vector<float> ev((size_t)C * 2 * C), od((size_t)C * 2 * C);
    for (int o = 0; o < C; ++o)
      for (int i = 0; i < C; ++i) {
        const float* wk = &w.data[((size_t)i * C + o) * 4];
        ev[(size_t)o * 2 * C + i] = wk[3];
        ev[(size_t)o * 2 * C + C + i] = wk[1];
        od[(size_t)o * 2 * C + i] = wk[2];
        od[(size_t)o * 2 * C + C + i] = wk[0];
      }
    up_conv_[0].w = up_w(ev, C, 2 * C);
    up_conv_[0].w_odd = up_w(od, C, 2 * C);
    up_conv_[0].b = up_f32(hw("up_blocks.0.2.conv.bias", {C}).data);
  }
  up_conv_[1].w = pack_conv3(hw("up_blocks.1.2.weight", {C, C, 3}), C, C);
  up_conv_[1].b = up_f32(hw("up_blocks.1.2.bias", {C}).data);
  final_conv_.w = pack_conv3(hw("final_block.block.0.weight", {C, C, 3}), C, C);
  final_conv_.b = up_f32(hw("final_block.block.0.bias", {C}).data);
  final_g_ = up_f32(hw("final_block.block.1.weight", {C}).data);
  final_be_ = up_f32(hw("final_block.block.1.bias", {C}).data);
  proj_w_ = up_w(hw("final_proj.weight", {cfg_.out_channels, C, 1}).data, cfg_.out_channels, C);
  {
    std::vector<float> pb = hw("final_proj.bias", {cfg_.out_channels}).data;
    pb.resize(128, 0.f);          // the epilogue reads bias in whole 32-column chunks
    proj_b_ = up_f32(pb);
  }
  const size_t expected = 4 + resnets_.size() * 16 + tfms_.size() * 13 + 4 * 2 + 4 + 2;
  if (n_est != expected)
    fail("state dict has %zu estimator tensors, the decoder expects %zu (unexpected extra keys?)", n_est, expected);
}

}  // namespace srn
