// Engine<OT>: kernel launches, the estimator forward and the Euler solve loop.  Included by engine.cu.
namespace srn {

static inline float to_f_host(float v) { return v; }
static inline float to_f_host(__nv_bfloat16 v) { return __bfloat162float(v); }

// Every kernel is launched with programmatic stream serialization (see ptx.cuh: pdl_wait): the next kernel's
// launch latency and prologue hide under this kernel's tail, also inside the captured CUDA graph.
static bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("SRN_DISABLE_PDL");
    return !(e != nullptr && e[0] == '1');
  }();
  return on;
}

// Set before a launch_k call that must NOT start before its predecessor has completed (no programmatic edge): used once
// per solve, right after row_extent_kernel, so that every later kernel may read the extents before its dependency wait.
static thread_local bool g_next_launch_serialised = false;

static thread_local size_t g_next_launch_smem = 0;      // dynamic shared memory of the next launch_k call

template <typename... KArgs, typename... Args>
static void launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof cfg);
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.stream = st;
  cfg.dynamicSmemBytes = g_next_launch_smem;
  g_next_launch_smem = 0;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = (pdl_enabled() && !g_next_launch_serialised) ? 1 : 0;
  g_next_launch_serialised = false;
  SRN_CUDA(cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...));
}

template <typename OT, int BN, int EPI>
static void launch_gemm_impl(GemmParams gp, dim3 grid, int num_sms, cudaStream_t st) {
  gp.ntn = (int)grid.x;
  gp.ntm = (int)grid.y;
  gp.nz = (int)grid.z;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof cfg);
  cfg.blockDim = dim3(GemmCfg<BN>::kThreads);
  cfg.dynamicSmemBytes = GemmCfg<BN>::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  const long long tiles = (long long)grid.x * grid.y * grid.z;                 // 256 x BN tiles, one per CTA pair
  const unsigned pairs = (unsigned)std::min<long long>(tiles, num_sms / 2);    // persistent: one pair per SM pair
  cfg.gridDim = dim3(2 * pairs);
  SRN_CUDA(cudaLaunchKernelEx(&cfg, gemm_kernel<OT, BN, EPI>, gp));
}

template <typename OT, int BN>
static void launch_gemm_bn(const GemmParams& gp, dim3 grid, int num_sms, cudaStream_t st) {
  switch (gp.epi) {
    case EPI_STORE_F32: launch_gemm_impl<OT, BN, EPI_STORE_F32>(gp, grid, num_sms, st); break;
    case EPI_STORE_GN: launch_gemm_impl<OT, BN, EPI_STORE_GN>(gp, grid, num_sms, st); break;
    case EPI_STORE_OT: launch_gemm_impl<OT, BN, EPI_STORE_OT>(gp, grid, num_sms, st); break;
    case EPI_RESID_F32: launch_gemm_impl<OT, BN, EPI_RESID_F32>(gp, grid, num_sms, st); break;
    case EPI_RESID_OT: launch_gemm_impl<OT, BN, EPI_RESID_OT>(gp, grid, num_sms, st); break;
    case EPI_GEGLU: launch_gemm_impl<OT, BN, EPI_GEGLU>(gp, grid, num_sms, st); break;
    case EPI_QKV: launch_gemm_impl<OT, BN, EPI_QKV>(gp, grid, num_sms, st); break;
    case EPI_EULER: launch_gemm_impl<OT, BN, EPI_EULER>(gp, grid, num_sms, st); break;
    case EPI_ROWMAX: launch_gemm_impl<OT, BN, EPI_ROWMAX>(gp, grid, num_sms, st); break;
    case EPI_EXP: launch_gemm_impl<OT, BN, EPI_EXP>(gp, grid, num_sms, st); break;
    case EPI_PV: launch_gemm_impl<OT, BN, EPI_PV>(gp, grid, num_sms, st); break;
    case EPI_RESID_ACT: launch_gemm_impl<OT, BN, EPI_RESID_ACT>(gp, grid, num_sms, st); break;
    default: fail("unknown GEMM epilogue %d", gp.epi);
  }
}

template <typename OT>
void Engine<OT>::launch_gemm(const GemmParams& gp, int bn, dim3 grid, cudaStream_t st) {
  if (gp.nkb < 1 || gp.nseg < 1) fail("gemm with empty K");
  if (gp.epi == EPI_RESID_F32 && gp.resid != gp.out) fail("EPI_RESID_F32 updates its residual in place (TMA reduce-add)");
  if (gp.grp_ntn0 > 0 && (gp.win != 0 || gp.grp_nkb < 1 || (gp.epi != EPI_STORE_GN && gp.epi != EPI_STORE_OT)))
    fail("tile groups: plain operand path, store epilogues only");
  if (bn == 256) launch_gemm_bn<OT, 256>(gp, grid, num_sms_, st);
  else launch_gemm_bn<OT, 128>(gp, grid, num_sms_, st);
  ++launches;
}

template <typename OT>
void Engine<OT>::capture_tap(const Plan<OT>& pl, const Tap& tap, cudaStream_t st) {
  (void)pl;
  SRN_CUDA(cudaStreamSynchronize(st));
  DebugTensor d;
  d.B = tap.nb;
  d.rows = tap.rows;
  d.cols = tap.cols;
  d.data.resize((size_t)tap.nb * tap.rows * tap.cols);
  const size_t es = tap.is_ot ? sizeof(OT) : sizeof(float);
  std::vector<uint8_t> raw((size_t)tap.rows * tap.ld * es);
  for (int b = 0; b < tap.nb; ++b) {
    SRN_CUDA(cudaMemcpy(raw.data(), static_cast<const uint8_t*>(tap.ptr) + (size_t)b * tap.bstride * es, raw.size(),
                        cudaMemcpyDeviceToHost));
    for (int r = 0; r < tap.rows; ++r)
      for (int c = 0; c < tap.cols; ++c) {
        const size_t si = (size_t)r * tap.ld + c;
        float v;
        if (tap.is_ot) v = to_f_host(reinterpret_cast<const OT*>(raw.data())[si]);
        else v = reinterpret_cast<const float*>(raw.data())[si];
        d.data[((size_t)b * tap.rows + r) * tap.cols + c] = v;
      }
  }
  debug_tensors[tap.name] = std::move(d);
}

// Rows per block of a (row blocks, B) row-kernel grid: `dflt` when that grid fits one wave of `slots` resident blocks,
// else the smallest multiple of `gran` that makes it fit (SRN_ROW_ONEWAVE=0: always `dflt`).
static int row_rpb(int T, int B, int slots, int gran, int dflt) {
  static const bool onewave = [] {
    const char* e = getenv("SRN_ROW_ONEWAVE");
    return !(e != nullptr && e[0] == '0');
  }();
  if (!onewave || (long long)cdiv(T, dflt) * B <= slots || slots / B < 1) return dflt;
  const int per_b = slots / B;
  int rpb = cdiv(cdiv(T, per_b), gran) * gran;
  return rpb < dflt ? dflt : rpb;
}

template <typename OT>
void Engine<OT>::launch_op(Plan<OT>& pl, Op& op, int step, float* v_out, bool update_x, cudaStream_t st) {
  const Level& L = pl.lv[op.level];
  const uint8_t* mask = pl.mask.template as<uint8_t>();
  const int B = pl.B;
  const int* nvalid = gn_valid_only ? pl.nvalid.template as<int>() + (size_t)op.level * B : nullptr;
  switch (op.kind) {
    case OP_GEMM: {
      if (op.is_euler) {
        GemmParams gp = op.gp;
        gp.v_out = v_out;
        gp.x_out = update_x ? pl.x_state.template as<float>() : nullptr;
        gp.dt_ptr = pl.dts.template as<float>() + step;
        if (update_x && blend_active_) {
          gp.blend_mask = pl.blend_mask.template as<uint8_t>();
          gp.blend_src = pl.blend_src.template as<float>();
        }
        launch_gemm(gp, op.bn, op.grid, st);
      } else if (!noshift && (op.gp.noshift != nullptr || op.gp.qk_norm2 != nullptr)) {
        GemmParams gp = op.gp;      // srn_set_option("noshift", 0): no norms collected, every entry takes the row-max pass
        gp.noshift = nullptr;
        gp.qk_norm2 = nullptr;
        launch_gemm(gp, op.bn, op.grid, st);
      } else if (op.gp.gn_part != nullptr && gn_valid_only) {
        GemmParams gp = op.gp;
        gp.gn_mask = mask;
        gp.mask_bstride = pl.Tp;
        gp.mask_step = L.mask_step;
        launch_gemm(gp, op.bn, op.grid, st);
      } else {
        launch_gemm(op.gp, op.bn, op.grid, st);
      }
    } break;
    case OP_PACKX: {
      dim3 grid(cdiv(pl.T, 32), cdiv(cfg_.out_channels, 32), B);
      launch_k(pack_cf_to_cl<OT>, grid, dim3(32, 8), st, pl.x_state.template as<float>(), cfg_.out_channels, pl.T, mask,
               pl.Tp, pl.xin.template as<OT>(), (long long)pl.Tp * 256, 256, 0);
      ++launches;
    } break;
    case OP_GN_APPLY: {
      // rows per block: 16 (same-box A/B: 66 us per call against 76 us with 32 rows and 78 us with 8), or - SRN_ROW_ONEWAVE -
      // the smallest multiple of 4 that fits the launch in one wave of resident blocks (4 per SM)
      const int rpb = row_rpb(L.T, B, 4 * num_sms_, 4, 16);
      const float* tb = op.use_tbias ? pl.tb.template as<float>() + ((size_t)op.resnet_idx * pl.tb_steps + step) * 512
                                     : nullptr;
      dim3 grid(cdiv(L.T, rpb), B);
      launch_k(gn_apply_kernel<OT>, grid, dim3(256), st, static_cast<const OT*>(op.y), op.part, 2 * L.mtiles, L.T, (long long)L.R * 512, op.gamma,
               op.beta, tb, mask, pl.Tp, L.mask_step, static_cast<OT*>(op.out), rpb, nvalid, tb_bstride_);
      ++launches;
    } break;
    case OP_GN_RES_SPK_LN: {
      // rows per block: 32 (same-box A/B: 138 us per call, 16 rows 152 us, 8 rows 167 us); the per-block prologue
      // (GroupNorm statistics + folded constants) is about half of a 32-row block's time, so a launch that needs more
      // than one wave of resident blocks (2 per SM) gets the smallest multiple of 8 that fits it in one wave instead
      const int rpb = row_rpb(L.T, B, 2 * num_sms_, 8, 32);
      dim3 grid(cdiv(L.T, rpb), B);
      const float* sc = pl.spk_scale.template as<float>() + (size_t)op.resnet_idx * B * 512;
      const float* sb = pl.spk_bias.template as<float>() + (size_t)op.resnet_idx * B * 512;
      launch_k(gn_res_spk_ln<OT>, grid, dim3(256), st, static_cast<const OT*>(op.y), op.part, 2 * L.mtiles, L.T, (long long)L.R * 512, op.gamma,
               op.beta, static_cast<const OT*>(op.res), mask, pl.Tp, L.mask_step, sc, sb, op.ln_g, op.ln_b, op.hres, static_cast<OT*>(op.out), rpb,
               nvalid);
      ++launches;
    } break;
    case OP_LAYERNORM: {
      const long long nrows = (long long)B * L.R;
      launch_k(layernorm_kernel<OT>, dim3((unsigned)((nrows + 7) / 8)), dim3(256), st, pl.hres.template as<float>(), nrows,
               op.ln_g, op.ln_b, pl.ln.template as<OT>());
      ++launches;
    } break;
  }
  if (debug && op.tap.name != nullptr) {
    // debug key = "<tag>.<tap name>" (e.g. "down0.y1"), or the bare tap name for untagged ops
    const size_t dot = op.name.rfind('.');
    const std::string key = (dot == std::string::npos ? std::string() : op.name.substr(0, dot + 1)) + op.tap.name;
    Tap t = op.tap;
    t.name = key.c_str();
    capture_tap(pl, t, st);
  }
  if (debug && op.tap2.name != nullptr) {
    const size_t dot = op.name.rfind('.');
    const std::string key = (dot == std::string::npos ? std::string() : op.name.substr(0, dot + 1)) + op.tap2.name;
    Tap t = op.tap2;
    t.name = key.c_str();
    capture_tap(pl, t, st);
  }
}

// NVTX ranges (the reference has no tracing at all): one range per estimator call and one per
// UNet block (down0, down1, mid0, ... = the op names' prefixes), around the launches - in eager mode and while a graph is
// captured.  nvtx3 is header-only and costs a few nanoseconds per call when no tool is attached; SRN_NVTX=0 removes them.
static bool nvtx_enabled() {
  static const bool on = [] {
    const char* e = getenv("SRN_NVTX");
    return !(e != nullptr && e[0] == '0');
  }();
  return on;
}
struct NvtxRange {
  bool on;
  explicit NvtxRange(const char* name) : on(nvtx_enabled()) {
    if (on) nvtxRangePushA(name);
  }
  ~NvtxRange() {
    if (on) nvtxRangePop();
  }
};

template <typename OT>
void Engine<OT>::run_forward(Plan<OT>& pl, int step, float* v_out, bool update_x, cudaStream_t st) {
  char nm[48];
  snprintf(nm, sizeof nm, "srn.estimator step %d", step);
  NvtxRange whole(nm);
  std::string cur;
  bool open = false;
  for (Op& op : pl.ops) {
    if (nvtx_enabled()) {
      const size_t dot = op.name.find('.');
      const std::string blk = dot == std::string::npos ? op.name : op.name.substr(0, dot);
      if (blk != cur) {
        if (open) nvtxRangePop();
        nvtxRangePushA(blk.c_str());
        open = true;
        cur = blk;
      }
    }
    try {
      launch_op(pl, op, step, v_out, update_x, st);
    } catch (...) {
      if (open) nvtxRangePop();
      throw;
    }
  }
  if (open) nvtxRangePop();
}

// Reads (and clears) the device-side soft-abort flag of ptx.cuh: mbar_wait.  Call only after a synchronisation.
template <typename OT>
void Engine<OT>::check_abort() {
  int flag = 0;
  SRN_CUDA(cudaMemcpyFromSymbol(&flag, g_srn_abort, sizeof flag));
  if (flag != 0) {
    const int zero = 0;
    SRN_CUDA(cudaMemcpyToSymbol(g_srn_abort, &zero, sizeof zero));
    fail("a GEMM pipeline barrier timed out on the device (soft abort %d): results of the last calls are invalid", flag);
  }
}

template <typename OT>
void Engine<OT>::stage_inputs(Plan<OT>& pl, const float* x, const float* mu, const uint8_t* mask, const float* spk,
                              cudaMemcpyKind kind, cudaStream_t st) {
  const int B = pl.B, T = pl.T;
  SRN_CUDA(cudaMemcpyAsync(pl.x_state.p, x, (size_t)B * cfg_.out_channels * T * 4, kind, st));
  SRN_CUDA(cudaMemcpyAsync(pl.mu_in.p, mu, (size_t)B * cmu_ * T * 4, kind, st));
  SRN_CUDA(cudaMemcpy2DAsync(pl.mask.p, pl.Tp, mask, T, T, B, kind, st));
  SRN_CUDA(cudaMemcpyAsync(pl.spk_in.p, spk, (size_t)B * cfg_.spk_embed_dim * 4, kind, st));
}

// n floats from ordinary host memory to the device on `st` without blocking the host on the stream: the values are
// copied into a page-locked slot first and the slot is reused only once the copy out of it has completed (its event).
// Tables larger than a slot (never the case for the reference's shapes) take the synchronising path.
template <typename OT>
void Engine<OT>::upload_floats(void* dst_dev, const float* src_host, size_t n, cudaStream_t st) {
  if (n == 0) return;
  if (n > kPinSlotFloats) {
    SRN_CUDA(cudaMemcpyAsync(dst_dev, src_host, n * 4, cudaMemcpyHostToDevice, st));
    SRN_CUDA(cudaStreamSynchronize(st));
    return;
  }
  if (pin_ == nullptr) {
    SRN_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&pin_), kPinSlots * kPinSlotFloats * 4, cudaHostAllocDefault));
    for (auto& e : pin_ev_) SRN_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  const int slot = pin_next_;
  pin_next_ = (pin_next_ + 1) % kPinSlots;
  SRN_CUDA(cudaEventSynchronize(pin_ev_[slot]));      // (returns at once for an event never recorded)
  float* h = pin_ + (size_t)slot * kPinSlotFloats;
  std::memcpy(h, src_host, n * 4);
  SRN_CUDA(cudaMemcpyAsync(dst_dev, h, n * 4, cudaMemcpyHostToDevice, st));
  SRN_CUDA(cudaEventRecord(pin_ev_[slot], st));
}

// Step-invariant work hoisted out of the Euler loop (SURVEY.md finding 10): mu columns of the packed
// conv input, SpeakerAdapter scale / bias per resnet, and the time-embedding MLP + per-resnet time bias
// for every step (the time path depends only on t, which the whole batch shares: decoder.py:401-402, :97).
template <typename OT>
void Engine<OT>::prepass(Plan<OT>& pl, const std::vector<float>& ts, const std::vector<float>& dts, cudaStream_t st) {
  (void)dts;
  NvtxRange range("srn.prepass");
  const int B = pl.B, n = (int)ts.size();
  const int nres = (int)resnets_.size();
  const uint8_t* mask = pl.mask.template as<uint8_t>();
  {
    dim3 grid(cdiv(pl.T, 32), cdiv(cmu_, 32), B);
    launch_k(pack_cf_to_cl<OT>, grid, dim3(32, 8), st, pl.mu_in.template as<float>(), cmu_, pl.T, mask, pl.Tp,
             pl.xin.template as<OT>(), (long long)pl.Tp * 256, 256, cfg_.out_channels);
    ++launches;
  }
  if (gn_valid_only) {
    launch_k(count_valid_kernel, dim3(B), dim3(256), st, mask, pl.T, pl.Tp, B, pl.nvalid.template as<int>());
    ++launches;
  }
  SRN_CUDA(cudaMemsetAsync(pl.ns_count.p, 0, 16, st));
  launch_k(row_extent_kernel, dim3(B), dim3(256), st, mask, pl.T, pl.Tp, B, pl.rowext.template as<int>());
  ++launches;
  g_next_launch_serialised = true;      // (the next launch_k below: a full dependency, see gemm_mainloop.cuh: s_ext)
  const int sd = cfg_.spk_embed_dim;
  for (int r = 0; r < nres; ++r) {
    launch_k(gemv_kernel, dim3(512 / 8, B), dim3(256), st, resnets_[r].ws_w, resnets_[r].ws_b,
             pl.spk_in.template as<float>(), pl.spk_scale.template as<float>() + (size_t)r * B * 512, sd, 512,
             (int)ACT_NONE, (int)ACT_NONE);
    launch_k(gemv_kernel, dim3(512 / 8, B), dim3(256), st, resnets_[r].wb_w, resnets_[r].wb_b,
             pl.spk_in.template as<float>(), pl.spk_bias.template as<float>() + (size_t)r * B * 512, sd, 512,
             (int)ACT_NONE, (int)ACT_NONE);
    launches += 2;
  }
  SRN_CUDA(cudaGetLastError());
  const int dim = cfg_.in_channels;
  const float e = (float)(std::log(10000.0) / (double)(dim / 2 - 1));
  launch_k(sinusoid_kernel, dim3(n), dim3(128), st, pl.tsteps.template as<float>(), dim, e, pl.temb0.template as<float>());
  launch_k(gemv_kernel, dim3(2048 / 8, n), dim3(256), st, time_w1_, time_b1_, pl.temb0.template as<float>(),
           pl.temb1.template as<float>(), dim, 2048, (int)ACT_NONE, (int)ACT_SILU);
  launch_k(gemv_kernel, dim3(2048 / 8, n), dim3(256), st, time_w2_, time_b2_, pl.temb1.template as<float>(),
           pl.temb2.template as<float>(), 2048, 2048, (int)ACT_NONE, (int)ACT_NONE);
  launches += 3;
  for (int r = 0; r < nres; ++r) {
    launch_k(gemv_kernel, dim3(512 / 8, n), dim3(256), st, resnets_[r].mlp_w, resnets_[r].mlp_b,
             pl.temb2.template as<float>(), pl.tb.template as<float>() + (size_t)r * pl.tb_steps * 512, 2048, 512,
             (int)ACT_MISH, (int)ACT_NONE);
    ++launches;
  }
  SRN_CUDA(cudaGetLastError());
}

template <typename OT>
static void ensure_step_buffers(Plan<OT>& pl, int n, int in_channels, int nres) {
  if (n <= pl.tb_steps) return;
  const int cap = std::max(n, 16);
  pl.tsteps.alloc((size_t)cap * 4);
  pl.dts.alloc((size_t)cap * 4);
  pl.temb0.alloc((size_t)cap * in_channels * 4);
  pl.temb1.alloc((size_t)cap * 2048 * 4);
  pl.temb2.alloc((size_t)cap * 2048 * 4);
  pl.tb.alloc((size_t)nres * cap * 512 * 4);
  pl.tb_steps = cap;
  if (pl.graph) {
    cudaGraphExecDestroy(pl.graph);
    pl.graph = nullptr;
  }
}

template <typename OT>
void Engine<OT>::estimator_forward(const float* x, const float* mu, const uint8_t* mask, const float* spk, float t,
                                   int B, int T, float* v_out, cudaStream_t st) {
  Plan<OT>& pl = get_plan(B, T);
  last_plan_ = &pl;
  pl.staged_ts.clear();
  ensure_step_buffers(pl, 1, cfg_.in_channels, (int)resnets_.size());
  stage_inputs(pl, x, mu, mask, spk, cudaMemcpyDeviceToDevice, st);
  // t by value through a one-thread kernel: no staging buffer, so the call never blocks the host (a reference-style
  // Python Euler loop over `estimator(...)` stays asynchronous) and works under an outer stream capture
  launch_k(set_scalars_kernel, dim3(1), dim3(32), st, pl.tsteps.template as<float>(), t, pl.dts.template as<float>(), 0.f);
  ++launches;
  blend_active_ = false;
  prepass(pl, std::vector<float>{t}, std::vector<float>{0.f}, st);
  if (debug) {      // the hoisted time path (decoder.py:48-63, :113-157, :97): time embedding, per-resnet time biases
    capture_tap(pl, Tap{"time_emb", pl.temb2.p, 0, 1, 2048, 2048, 0, 1, 0}, st);
    capture_tap(pl, Tap{"time_bias", pl.tb.p, 0, (int)resnets_.size(), 512, pl.tb_steps * 512, 0, 1, 0}, st);
  }
  run_forward(pl, 0, v_out, false, st);
}

// Decoder.forward with one time per clip (decoder.py:401-402: `t` may be a vector; CFM.compute_loss draws one per
// sample).  The time path is the only place where t enters: the prepass computes B time biases per resnet instead of
// one per Euler step, and gn_apply picks its clip's.
template <typename OT>
void Engine<OT>::estimator_forward_t(const float* x, const float* mu, const uint8_t* mask, const float* spk,
                                     const float* t_host, int B, int T, float* v_out, cudaStream_t st) {
  estimator_forward_times(x, mu, mask, spk, t_host, false, B, B, T, v_out, st);
}

// The same with the time(s) in DEVICE memory - one value shared by the batch (n_t = 1: what the reference's solve_euler
// hands its estimator, a 0-dim CUDA tensor) or one per clip (n_t = B).  Nothing comes back to the host.
template <typename OT>
void Engine<OT>::estimator_forward_tdev(const float* x, const float* mu, const uint8_t* mask, const float* spk,
                                        const float* t_dev, int n_t, int B, int T, float* v_out, cudaStream_t st) {
  if (n_t != 1 && n_t != B) fail("estimator_forward_tdev: n_t must be 1 or B (%d), got %d", B, n_t);
  estimator_forward_times(x, mu, mask, spk, t_dev, true, n_t, B, T, v_out, st);
}

template <typename OT>
void Engine<OT>::estimator_forward_times(const float* x, const float* mu, const uint8_t* mask, const float* spk,
                                         const float* t_src, bool t_on_device, int n_t, int B, int T, float* v_out,
                                         cudaStream_t st) {
  Plan<OT>& pl = get_plan(B, T);
  last_plan_ = &pl;
  pl.staged_ts.clear();
  ensure_step_buffers(pl, n_t, cfg_.in_channels, (int)resnets_.size());
  stage_inputs(pl, x, mu, mask, spk, cudaMemcpyDeviceToDevice, st);
  if (t_on_device)
    SRN_CUDA(cudaMemcpyAsync(pl.tsteps.p, t_src, (size_t)n_t * 4, cudaMemcpyDeviceToDevice, st));
  else
    upload_floats(pl.tsteps.p, t_src, (size_t)n_t, st);        // (through pinned staging: no host sync)
  SRN_CUDA(cudaMemsetAsync(pl.dts.p, 0, (size_t)n_t * 4, st));
  blend_active_ = false;
  const std::vector<float> count(n_t, 0.f);                    // (the prepass reads the times on the device; it only
  prepass(pl, count, count, st);                               //  needs to know how many there are)
  tb_bstride_ = n_t == 1 ? 0 : 512;
  try {
    run_forward(pl, 0, v_out, false, st);
  } catch (...) {
    tb_bstride_ = 0;
    throw;
  }
  tb_bstride_ = 0;
}

// CFM.compute_loss, forward value only (flow_matching.py:95-133; the validation loss of trainers/ssc.py).  The caller
// supplies the noise z and the times t the reference would draw; y (the conditional flow) is returned like the
// reference does.  No gradients: the backward pass is not built.
template <typename OT>
void Engine<OT>::cfm_loss(const float* x1, const float* z, const float* t_host, const float* mu, const uint8_t* mask,
                          const float* spk, const uint8_t* mask_l, int B, int T, float sigma_min, float* y_out,
                          float* loss_out, cudaStream_t st, bool t_on_device) {
  Plan<OT>& pl = get_plan(B, T);
  pl.staged_ts.clear();
  const int C = cfg_.out_channels;
  const long long n = (long long)B * C * T;
  const float oms = (float)(1.0 - (double)sigma_min);
  ensure_step_buffers(pl, B, cfg_.in_channels, (int)resnets_.size());
  const int nblk = 1024;
  if (loss_v_.bytes < (size_t)n * 4) loss_v_.alloc((size_t)n * 4);
  if (loss_part_.bytes < (size_t)nblk * 16) loss_part_.alloc((size_t)nblk * 16);
  // t on the device first: the mix kernel reads it (estimator_forward_t stages the same values again)
  if (t_on_device)
    SRN_CUDA(cudaMemcpyAsync(pl.tsteps.p, t_host, (size_t)B * 4, cudaMemcpyDeviceToDevice, st));
  else
    upload_floats(pl.tsteps.p, t_host, (size_t)B, st);
  launch_k(cfm_mix_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), st, x1, z, (const float*)pl.tsteps.template as<float>(),
           (long long)C * T, n, oms, y_out);
  ++launches;
  estimator_forward_times(y_out, mu, mask, spk, t_host, t_on_device, B, B, T, loss_v_.template as<float>(), st);
  launch_k(cfm_loss_partial_kernel, dim3(nblk), dim3(256), st, (const float*)loss_v_.template as<float>(), x1, z, mask_l,
           mask_l != nullptr ? mask_l : mask, T, C, T, n, oms, loss_part_.template as<double>());
  launch_k(cfm_loss_final_kernel, dim3(1), dim3(32), st, (const double*)loss_part_.template as<double>(), nblk, C, loss_out);
  launches += 2;
  SRN_CUDA(cudaGetLastError());
}

template <typename OT>
void Engine<OT>::solve_staged(Plan<OT>& pl, const float* t_span, int n_steps, bool blend, cudaStream_t st,
                              bool t_span_on_device) {
  // flow_matching.py:79-91 in fp32: t += dt; dt = t_span[k+1] - t  (keeps the reference's last-ulp wobble)
  std::vector<float> ts(n_steps), dts(n_steps);      // (the prepass only needs their length: it reads the device tables)
  last_plan_ = &pl;
  const size_t tb_before = pl.tb_steps;
  ensure_step_buffers(pl, n_steps, cfg_.in_channels, (int)resnets_.size());
  if (t_span_on_device) {
    // the grid never visits the host: one thread builds the tables with the same additions (rowops.cuh)
    launch_k(tspan_tables_kernel, dim3(1), dim3(32), st, t_span, n_steps, pl.tsteps.template as<float>(),
             pl.dts.template as<float>());
    ++launches;
    pl.staged_ts.clear();
  } else {
    float t = t_span[0];
    float dt = t_span[1] - t_span[0];
    for (int s = 1; s <= n_steps; ++s) {
      ts[s - 1] = t;
      dts[s - 1] = dt;
      t = t + dt;
      if (s < n_steps) dt = t_span[s + 1] - t;
    }
    // The (t, dt) tables stay on the device between calls: a decode loop passes the same t_span every time, and then
    // nothing is uploaded; a new t_span goes through pinned staging (upload_floats), so the call never blocks the host.
    const std::vector<float> span(t_span, t_span + n_steps + 1);
    if (tb_before != (size_t)pl.tb_steps || pl.staged_ts != span) {
      upload_floats(pl.tsteps.p, ts.data(), (size_t)n_steps, st);
      upload_floats(pl.dts.p, dts.data(), (size_t)n_steps, st);
      pl.staged_ts = span;
    }
  }
  blend_active_ = blend;
  if (use_graph && !debug) {
    if (pl.graph == nullptr || pl.graph_steps != n_steps || pl.graph_blend != blend || pl.graph_gn_valid != gn_valid_only ||
        pl.graph_noshift != noshift) {
      if (pl.graph) {
        cudaGraphExecDestroy(pl.graph);
        pl.graph = nullptr;
      }
      if (capture_stream_ == nullptr) SRN_CUDA(cudaStreamCreateWithFlags(&capture_stream_, cudaStreamNonBlocking));
      const int64_t before = launches;
      cudaGraph_t g = nullptr;
      SRN_CUDA(cudaStreamBeginCapture(capture_stream_, cudaStreamCaptureModeThreadLocal));
      try {
        prepass(pl, ts, dts, capture_stream_);
        for (int s = 0; s < n_steps; ++s) run_forward(pl, s, nullptr, true, capture_stream_);
      } catch (...) {
        cudaStreamEndCapture(capture_stream_, &g);
        if (g) cudaGraphDestroy(g);
        launches = before;
        throw;
      }
      SRN_CUDA(cudaStreamEndCapture(capture_stream_, &g));
      pl.graph_nodes = launches - before;
      launches = before;
      cudaError_t e = cudaGraphInstantiate(&pl.graph, g, 0);
      cudaGraphDestroy(g);
      if (e != cudaSuccess) fail("cudaGraphInstantiate failed: %s", cudaGetErrorString(e));
      pl.graph_steps = n_steps;
      pl.graph_blend = blend;
      pl.graph_gn_valid = gn_valid_only;
      pl.graph_noshift = noshift;
    }
    SRN_CUDA(cudaGraphLaunch(pl.graph, st));
    launches += pl.graph_nodes;
  } else {
    prepass(pl, ts, dts, st);
    for (int s = 0; s < n_steps; ++s) run_forward(pl, s, nullptr, true, st);
  }
}

template <typename OT>
void Engine<OT>::softmax_stats(int64_t* out2) {
  out2[0] = out2[1] = 0;
  if (last_plan_ == nullptr) return;
  unsigned long long h[2] = {0, 0};
  SRN_CUDA(cudaDeviceSynchronize());
  SRN_CUDA(cudaMemcpy(h, last_plan_->ns_count.p, sizeof h, cudaMemcpyDeviceToHost));
  out2[0] = (int64_t)h[0];
  out2[1] = (int64_t)h[1];
}

template <typename OT>
void Engine<OT>::solve_euler(float* x, const float* mu, const uint8_t* mask, const float* spk, const float* t_span,
                             int n_steps, int B, int T, const uint8_t* blend_mask, const float* blend_src,
                             cudaStream_t st, bool t_span_on_device) {
  NvtxRange range("srn.solve_euler");
  Plan<OT>& pl = get_plan(B, T);
  stage_inputs(pl, x, mu, mask, spk, cudaMemcpyDeviceToDevice, st);
  const bool blend = blend_mask != nullptr;
  if (blend) {
    SRN_CUDA(cudaMemcpyAsync(pl.blend_mask.p, blend_mask, (size_t)B * T, cudaMemcpyDeviceToDevice, st));
    SRN_CUDA(cudaMemcpyAsync(pl.blend_src.p, blend_src, (size_t)B * cfg_.out_channels * T * 4, cudaMemcpyDeviceToDevice, st));
  }
  solve_staged(pl, t_span, n_steps, blend, st, t_span_on_device);
  SRN_CUDA(cudaMemcpyAsync(x, pl.x_state.p, (size_t)B * cfg_.out_channels * T * 4, cudaMemcpyDeviceToDevice, st));
}

template <typename OT>
void Engine<OT>::solve_euler_host(const float* z, const float* mu, const uint8_t* mask, const float* spk,
                                  const float* t_span, int n_steps, int B, int T, float* mel_out, cudaStream_t st) {
  NvtxRange range("srn.solve_euler_host");
  Plan<OT>& pl = get_plan(B, T);
  stage_inputs(pl, z, mu, mask, spk, cudaMemcpyHostToDevice, st);
  solve_staged(pl, t_span, n_steps, false, st);
  SRN_CUDA(cudaMemcpyAsync(mel_out, pl.x_state.p, (size_t)B * cfg_.out_channels * T * 4, cudaMemcpyDeviceToHost, st));
  SRN_CUDA(cudaStreamSynchronize(st));
  check_abort();
}

}  // namespace srn

namespace srn {

// Per-op timing of one estimator forward (CUDA events around `reps` back-to-back launches of each op).
// GEMM flops: algorithmic and as issued, see below; row-kernel bytes = the tensors they must read and
// write once.
#if SRN_TRACE
// One traced launch of a GEMM op: medians over CTAs of the phase lengths in SM clocks (stderr).
// stamps: 0 kernel entry, 1 after griddepcontrol.wait, 2 first operand stage landed (leader), 3 last MMA committed
// (leader), 4 first accumulator ready (epilogue warp 2), 5 last accumulator ready, 6 last chunk staged, 7 stores drained
template <typename OT>
void Engine<OT>::trace_op(Plan<OT>& pl, Op& op, cudaStream_t st) {
  const int ncta = 160;
  unsigned long long* d = nullptr;
  SRN_CUDA(cudaMalloc(&d, ncta * 8 * sizeof(unsigned long long)));
  SRN_CUDA(cudaMemset(d, 0, ncta * 8 * sizeof(unsigned long long)));
  op.gp.trace = d;
  launch_op(pl, op, 0, nullptr, true, st);
  op.gp.trace = nullptr;
  SRN_CUDA(cudaStreamSynchronize(st));
  std::vector<unsigned long long> h(ncta * 8);
  SRN_CUDA(cudaMemcpy(h.data(), d, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  SRN_CUDA(cudaFree(d));
  auto med = [&](int a, int b, bool leader_only, double* mx) {
    std::vector<double> v;
    for (int c = 0; c < ncta; ++c) {
      if (leader_only && (c & 1)) continue;
      const unsigned long long ta = h[c * 8 + a], tb = h[c * 8 + b];
      if (ta == 0 || tb == 0) continue;
      v.push_back((double)tb - (double)ta);
    }
    if (v.empty()) { *mx = 0; return 0.0; }
    std::sort(v.begin(), v.end());
    *mx = v.back();
    return v[v.size() / 2];
  };
  double m[8], x[8];
  m[0] = med(0, 1, false, &x[0]);   // launch overlap: entry -> dependencies resolved
  m[1] = med(1, 2, true, &x[1]);    // first stage latency
  m[2] = med(2, 3, true, &x[2]);    // main loops (all tiles)
  m[3] = med(1, 4, false, &x[3]);   // dependencies resolved -> first accumulator
  m[4] = med(5, 6, false, &x[4]);   // last tile's epilogue (TMEM -> staged)
  m[5] = med(6, 7, false, &x[5]);   // store drain
  m[6] = med(1, 7, false, &x[6]);   // body
  m[7] = med(0, 7, false, &x[7]);   // whole CTA
  fprintf(stderr, "[trace] %-18s pdl_wait %6.0f | first stage %6.0f | mainloop %7.0f | ->first acc %7.0f | last epilogue %6.0f (max %6.0f) | "
                  "drain %6.0f (max %6.0f) | body %7.0f (max %7.0f) | cta %7.0f clk\n",
          op.name.c_str(), m[0], m[1], m[2], m[3], m[4], x[4], m[5], x[5], m[6], x[6], m[7]);
}
#endif

#if SRN_TRACE
template <typename OT>
void Engine<OT>::trace_row_op(Plan<OT>& pl, Op& op, cudaStream_t st) {
  const int nblk = 4096;
  unsigned long long* d = nullptr;
  SRN_CUDA(cudaMalloc(&d, (size_t)nblk * 16 * sizeof(unsigned long long)));
  SRN_CUDA(cudaMemset(d, 0, (size_t)nblk * 16 * sizeof(unsigned long long)));
  SRN_CUDA(cudaMemcpyToSymbol(g_row_trace, &d, sizeof d));
  launch_op(pl, op, 0, nullptr, true, st);
  SRN_CUDA(cudaStreamSynchronize(st));
  unsigned long long* nul = nullptr;
  SRN_CUDA(cudaMemcpyToSymbol(g_row_trace, &nul, sizeof nul));
  std::vector<unsigned long long> h((size_t)nblk * 16);
  SRN_CUDA(cudaMemcpy(h.data(), d, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  SRN_CUDA(cudaFree(d));
  std::vector<double> stats_clk, row_clk, body_clk, dur_ns, ph[5];
  unsigned long long g0 = ~0ull, g1 = 0;
  int per_sm[256] = {0};
  int n = 0;
  for (int i = 0; i < nblk; ++i) {
    const unsigned long long* t = &h[(size_t)i * 16];
    if (t[0] == 0 || t[4] == 0) continue;
    ++n;
    stats_clk.push_back((double)(t[2] - t[1]));
    if (t[3] != 0) row_clk.push_back((double)(t[3] - t[2]));
    body_clk.push_back((double)(t[4] - t[1]));
    if (t[8] && t[12]) {
      ph[0].push_back((double)(t[8] - t[2]));
      ph[1].push_back((double)(t[9] - t[8]));
      ph[2].push_back((double)(t[10] - t[9]));
      ph[3].push_back((double)(t[11] - t[10]));
      ph[4].push_back((double)(t[12] - t[11]));
    }
    dur_ns.push_back((double)(t[7] - t[6]));
    g0 = std::min(g0, t[6]);
    g1 = std::max(g1, t[7]);
    ++per_sm[t[5] & 255];
  }
  auto med = [](std::vector<double>& v) { if (v.empty()) return 0.0; std::sort(v.begin(), v.end()); return v[v.size() / 2]; };
  double sum_ns = 0;
  for (double x : dur_ns) sum_ns += x;
  int late = 0;
  for (int i = 0; i < nblk; ++i) {
    const unsigned long long* t = &h[(size_t)i * 16];
    if (t[0] != 0 && t[6] > g0 + 3000) ++late;
  }
  int mx = 0, used = 0;
  for (int i = 0; i < 256; ++i) { mx = std::max(mx, per_sm[i]); used += per_sm[i] > 0; }
  fprintf(stderr, "[trace] %-18s blocks %d on %d SMs (max %d per SM) | stats+staging %5.0f clk | all rows but the last %6.0f clk | block body %7.0f clk | "
                  "span %6.1f us, mean block %5.1f us, mean concurrency %5.1f blocks, %d blocks start > 3 us late\n",
          op.name.c_str(), n, used, mx, med(stats_clk), med(row_clk), med(body_clk), (double)(g1 - g0) * 1e-3,
          n ? sum_ns / n * 1e-3 : 0.0, (g1 > g0) ? sum_ns / (double)(g1 - g0) : 0.0, late);
  fprintf(stderr, "[trace]   first row of warp 0: loads+mish half 0 %6.0f | half 1 %6.0f | speaker LN stats %6.0f | scale + hres store %6.0f | LN stats %6.0f clk\n",
          med(ph[0]), med(ph[1]), med(ph[2]), med(ph[3]), med(ph[4]));
}
#endif

template <typename OT>
std::string Engine<OT>::profile_ops(int B, int T, int reps) {
  Plan<OT>& pl = get_plan(B, T);
  cudaStream_t st = nullptr;
  ensure_step_buffers(pl, 1, cfg_.in_channels, (int)resnets_.size());
  SRN_CUDA(cudaMemset(pl.mask.p, 1, pl.mask.bytes));
  SRN_CUDA(cudaMemset(pl.tsteps.p, 0, 4));
  SRN_CUDA(cudaMemset(pl.dts.p, 0, 4));
  blend_active_ = false;
  const bool dbg = debug;
  debug = false;
  prepass(pl, std::vector<float>{0.f}, std::vector<float>{0.f}, st);
  run_forward(pl, 0, nullptr, true, st);
  SRN_CUDA(cudaStreamSynchronize(st));
  cudaEvent_t e0, e1;
  SRN_CUDA(cudaEventCreate(&e0));
  SRN_CUDA(cudaEventCreate(&e1));
  std::string js = "[";
  const size_t es = sizeof(OT);
  for (size_t i = 0; i < pl.ops.size(); ++i) {
    Op& op = pl.ops[i];
    SRN_CUDA(cudaEventRecord(e0, st));
    for (int r = 0; r < reps; ++r) launch_op(pl, op, 0, nullptr, true, st);
    SRN_CUDA(cudaEventRecord(e1, st));
    SRN_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    SRN_CUDA(cudaEventElapsedTime(&ms, e0, e1));
#if SRN_TRACE
    if (op.kind == OP_GEMM && getenv("SRN_TRACE") != nullptr) trace_op(pl, op, st);
    if ((op.kind == OP_GN_RES_SPK_LN || op.kind == OP_GN_APPLY) && getenv("SRN_TRACE") != nullptr) trace_row_op(pl, op, st);
#endif
    const Level& L = pl.lv[op.level];
    double flops = 0, bytes = 0, issued = 0;
    const char* kind = "";
    const double rows = (double)B * L.T;
    switch (op.kind) {
      case OP_GEMM:
        kind = op.bn == 256 ? "gemm_bn256" : "gemm_bn128";
        // issued: every K block the tensor cores run (incl. the 242 -> 256 channel padding, the hi + lo weight passes of
        // the TF32-class mode and the row-max pass); algorithmic: SURVEY 8d's F(L) share of this op - no padding, one
        // pass over the weights, nothing for the row-max pass (softmax is one Q K^T in F)
        issued = 2.0 * (double)op.gp.M * op.gp.N * ((double)(op.gp.nkb + (op.gp.grp_ntn0 > 0 ? op.gp.grp_nkb : 0)) * BK) *
                 (double)op.grid.z;
        flops = issued * op.alg_scale * ((OpTraits<OT>::kTF32 && op.gp.b_batch_mode == 0) ? 0.5 : 1.0);
        break;
      case OP_PACKX: kind = "pack_cf_to_cl"; bytes = (double)B * cfg_.out_channels * T * (4 + es); break;
      case OP_GN_APPLY: kind = "gn_apply"; bytes = rows * 512 * (es + es); break;
      case OP_GN_RES_SPK_LN: kind = "gn_res_spk_ln"; bytes = rows * 512 * (es + es + 4 + es); break;
      case OP_LAYERNORM: kind = "layernorm"; bytes = rows * 512 * (4 + es); break;
    }
    char buf[512];
    snprintf(buf, sizeof buf, "%s{\"op\": \"%s\", \"kernel\": \"%s\", \"ms\": %.6f, \"flops\": %.6e, \"flops_issued\": %.6e, \"bytes\": %.6e}",
             i ? ", " : "", op.name.c_str(), kind, ms / reps, flops, issued, bytes);
    js += buf;
  }
  js += "]";
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  debug = dbg;
  return js;
}

}  // namespace srn
