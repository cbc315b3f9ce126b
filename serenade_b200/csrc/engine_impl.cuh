// Engine<OT>: OT is the GEMM operand / activation storage type (bf16, or float for the TF32 mode).
#pragma once
#include <cuda_bf16.h>

#include <algorithm>
#include <type_traits>
#include <cmath>
#include <cstring>
#include <list>

#include <nvtx3/nvToolsExt.h>

#include "engine.h"
#include "rowops.cuh"
#include "frontops.cuh"
#include "vocops.cuh"

namespace srn {

template <typename OT>
struct ResnetW {
  OT* w1 = nullptr;  int cin_pad = 0;       // [512][3 * cin_pad]   conv k3 taps, K-major
  OT* w2 = nullptr;                         // [512][3 * 512]
  OT* wr = nullptr;                         // [512][cin_pad]       1x1 res_conv
  float *b1 = nullptr, *g1 = nullptr, *be1 = nullptr, *b2 = nullptr, *g2 = nullptr, *be2 = nullptr, *br = nullptr;
  float *mlp_w = nullptr, *mlp_b = nullptr;                  // Linear 2048 -> 512 (time bias)
  float *ws_w = nullptr, *ws_b = nullptr, *wb_w = nullptr, *wb_b = nullptr;   // SpeakerAdapter
};

template <typename OT>
struct TfmW {
  float *ln1g = nullptr, *ln1b = nullptr, *ln3g = nullptr, *ln3b = nullptr;
  OT* wqkv = nullptr;      // [6144][512]  = [Wq; Wk; Wv]
  OT* wo = nullptr;        // [512][2048]
  float* bo = nullptr;
  OT* wgl = nullptr;       // [4096][512]  GEGLU proj, value/gate rows interleaved
  float* bgl = nullptr;    // interleaved the same way
  OT* wf = nullptr;        // [512][2048]
  float* bf = nullptr;
};

template <typename OT>
struct ConvW {
  OT* w = nullptr;         // k3: [512][1536]; transposed conv: even / odd parity [512][1024] each
  OT* w_odd = nullptr;
  float* b = nullptr;
};

enum OpKind : int { OP_GEMM, OP_PACKX, OP_GN_APPLY, OP_GN_RES_SPK_LN, OP_LAYERNORM };

struct Tap {
  const char* name = nullptr;   // debug tensor name (null = none)
  const void* ptr = nullptr;
  int is_ot = 0;                // element type: OT or float
  int rows = 0, cols = 0, ld = 0;
  long long bstride = 0;
  int nb = 0;
  int cf = 0;                   // channels-first (B, cols, rows) source (only for v)
};

struct Op {
  OpKind kind = OP_GEMM;
  std::string name;
  // GEMM
  GemmParams gp;
  int bn = 128;
  dim3 grid;
  bool is_euler = false;
  double alg_scale = 1.0;       // algorithmic / issued FLOPs of this op (channel padding; 0 for the row-max pass)
  // row kernels
  const void* y = nullptr;      // conv output in operand width (EPI_STORE_GN)
  const float* part = nullptr;
  const float* gamma = nullptr;
  const float* beta = nullptr;
  const void* res = nullptr;    // res_conv output, operand width
  const float* ln_g = nullptr;
  const float* ln_b = nullptr;
  float* hres = nullptr;
  void* out = nullptr;
  int resnet_idx = -1;          // which time-bias / speaker table row (-1: none)
  int use_tbias = 0;
  int level = 0;
  Tap tap;
  Tap tap2;                     // grouped launches: the second output (debug taps)
};

struct Level {
  int T = 0;        // valid rows
  int R = 0;        // allocated rows per batch entry
  int Sp = 0;       // padded key length (row stride of scores / P / V^T)
  int mtiles = 0;
  int mask_step = 1;
};

template <typename OT>
struct Plan {
  int B = 0, T = 0, Tp = 0;
  Level lv[2];
  DevBuf xin, y, r, gnpart, a1, hres, ln, qk, vt, smax, ssum, P, ao, gg, hid0, hid1, hc, cv;
  DevBuf qk_norm2, noshift;   // shift-free softmax test: [B * 4][16] squared-norm maxima, [B * 4] flags (EPI_ROWMAX prologue)
  DevBuf ns_count;            // [2] int64: attention entries that ran shift-free / all entries, since the call began
  DevBuf x_state, mask, spk_in, mu_in, spk_scale, spk_bias, tsteps, dts, temb0, temb1, temb2, tb;
  DevBuf blend_mask, blend_src;
  DevBuf rowext;      // [2][B] last unmasked frame + 1 per batch entry at full / half rate (padded-frame skipping)
  DevBuf nvalid;      // [2][B] unmasked frames per batch entry at full / half rate (valid-frame GroupNorm mode)
  int tb_steps = 0;
  std::vector<Op> ops;
  int euler_op = -1;
  // graph cache
  cudaGraphExec_t graph = nullptr;
  int graph_steps = 0;
  bool graph_blend = false;
  bool graph_gn_valid = false;
  bool graph_noshift = true;
  std::vector<float> staged_ts;      // the t_span whose (t, dt) tables are on the device (no re-upload, no host sync)
  int64_t graph_nodes = 0;
  ~Plan() {
    if (graph) cudaGraphExecDestroy(graph);
  }
};

template <typename OT>
class Engine : public EngineBase {
 public:
  explicit Engine(const SrnConfig& cfg);
  ~Engine() override;
  void load_weight(const std::string& key, const float* data, int ndim, const int64_t* dims) override;
  void finalize_weights() override;
  void estimator_forward(const float* x, const float* mu, const uint8_t* mask, const float* spk, float t, int B,
                         int T, float* v_out, cudaStream_t st) override;
  void estimator_forward_t(const float* x, const float* mu, const uint8_t* mask, const float* spk, const float* t_host,
                           int B, int T, float* v_out, cudaStream_t st) override;
  void estimator_forward_tdev(const float* x, const float* mu, const uint8_t* mask, const float* spk, const float* t_dev,
                              int n_t, int B, int T, float* v_out, cudaStream_t st) override;
  void cfm_loss(const float* x1, const float* z, const float* t_host, const float* mu, const uint8_t* mask,
                const float* spk, const uint8_t* mask_l, int B, int T, float sigma_min, float* y_out, float* loss_out,
                cudaStream_t st, bool t_on_device = false) override;
  void solve_euler(float* x, const float* mu, const uint8_t* mask, const float* spk, const float* t_span,
                   int n_steps, int B, int T, const uint8_t* blend_mask, const float* blend_src,
                   cudaStream_t st, bool t_span_on_device = false) override;
  void solve_euler_host(const float* z, const float* mu, const uint8_t* mask, const float* spk, const float* t_span,
                        int n_steps, int B, int T, float* mel_out, cudaStream_t st) override;
  std::string profile_ops(int B, int T, int reps) override;
#if SRN_TRACE
  void trace_op(Plan<OT>& pl, Op& op, cudaStream_t st);
  void trace_row_op(Plan<OT>& pl, Op& op, cudaStream_t st);
#endif
  void content_encode(const float* x, int B, int L, float* out, cudaStream_t st) override;
  void style_encode(const float* mel, int B, int L, float* out, cudaStream_t st) override;
  void vocoder_decode(const float* mel, const float* scale, const float* shift, int B, int T, float* wave,
                      cudaStream_t st) override;
  int vocoder_hop() const override { return voc_.present ? voc_.hop : 0; }
  void softmax_stats(int64_t* out2) override;

 private:
  static constexpr int BK = OpTraits<OT>::kBK;
  SrnConfig cfg_;
  int cmu_ = 0;
  int num_sms_ = 148;
  bool finalized_ = false;
  std::map<std::string, HostTensor> host_w_;
  std::list<DevBuf> arena_;
  // packed weights
  float *time_w1_ = nullptr, *time_b1_ = nullptr, *time_w2_ = nullptr, *time_b2_ = nullptr;
  std::vector<ResnetW<OT>> resnets_;   // down0, down1, mid0.., up0, up1
  std::vector<TfmW<OT>> tfms_;
  ConvW<OT> down_conv_[2], up_conv_[2], final_conv_;
  float *final_g_ = nullptr, *final_be_ = nullptr;
  OT* proj_w_ = nullptr;
  float* proj_b_ = nullptr;
  // conditioning front half (engine_frontend.inl); optional: present only if encoder.* / gst.* weights were loaded
  struct FrontW {
    bool present = false;
    int in_dim = 0, in_pad = 0, enc_out = 0;
    OT* w_in = nullptr;   float* b_in = nullptr;          // k7 conv [512][7 * in_pad]
    struct { OT* w3 = nullptr; float* b3 = nullptr; OT* w1sc = nullptr; float* b1sc = nullptr; } blk[2];
    OT* w_out = nullptr;  float* b_out = nullptr;         // k7 conv [80][7 * 512]
    OT* gw[6] = {};       float* gb[6] = {};              // GST convs, BatchNorm folded: [Cout][Kp]
    int gc[7] = {}, gkp[6] = {}, gst_f = 0, ntok = 0;
    float *w_ihT = nullptr, *b_ih = nullptr, *w_hhT = nullptr, *b_hh = nullptr;
    float *wq = nullptr, *bq = nullptr, *kmat = nullptr, *vmat = nullptr, *wo = nullptr, *bo = nullptr;
  } front_;
  DevBuf fb_xpad_, fb_h_, fb_h2_, fb_apad_, fb_e_, fb_col_, fb_act0_, fb_act1_;
  // HiFi-GAN vocoder (engine_vocoder.inl); optional: present only if vocoder.* weights were loaded
  struct VocBlock {
    int k = 0;
    std::vector<OT*> w1, w2;          // per dilation: [n1][nseg * kc], [n2][nseg * kc]
    std::vector<float*> b1, b2;
    std::vector<std::vector<int>> sh1, sh2;     // per dilation: row shift of every K segment
    std::vector<int> g1, g2;                    // per dilation: time steps per GEMM row of convs1 / convs2
  };
  struct VocStage {
    int s = 0, c = 0, cp = 0;         // stride, channels, channels padded to a K block
    // Channel counts below one K block (C = 32 in bf16): g = BK / C consecutive time steps form one GEMM row
    // ([T][C] read as [T / g][g C], no padding); kc = K columns per segment, n1 / n2 = N of convs1 / convs2
    int g = 1;      // 1 = operand rows padded to cp (or C >= BK); > 1 = unpadded, every conv picks its own grouping
    bool padded = false;
    OT* w_up = nullptr;               // [s * C][3 * Cp_in]
    float* b_up = nullptr;
    std::vector<VocBlock> blocks;
  };
  struct VocW {
    bool present = false;
    int in_ch = 0, in_pad = 0, c0 = 0, k_in = 0, k_out = 0, n_blk = 0, n_dil = 0, hop = 1;
    OT *w_in = nullptr, *w_out = nullptr;
    float *b_in = nullptr, *b_out = nullptr;
    std::vector<int> sh_out;          // K segment shifts of the output conv
    int g_out = 1, kc_out = 0;
    std::vector<VocStage> stages;
  } voc_;
  DevBuf vb_ua_[2], vb_c_, vb_x_[4], vb_xa_, vb_ta_, vb_y4_, vb_mel_, vb_wave_, vb_aff_;
  struct VocGraph {
    int B = 0, T = 0;
    bool affine = false;
    cudaGraphExec_t exec = nullptr;
    int64_t nodes = 0;
  };
  std::list<VocGraph> voc_graphs_;
  void voc_body(const float* mel, const float* scale, const float* shift, int B, int T, float* wave, cudaStream_t st);
  bool decoder_present_ = false;
  void pack_vocoder();
  void voc_gemm(GemmParams& gp, const OT* W, int N, int K, const float* bias, int M, int nbatch, int mode, void* out,
                long long out_bstride, int ldo, float slope, cudaStream_t st, const float* resid = nullptr,
                OT* act_out = nullptr);
  void pack_frontend();
  void pack_decoder(size_t n_est);
  void front_gemm(GemmParams& gp, const OT* W, int N, int K, const float* bias, int M, int nbatch, bool f32_out,
                  void* out, long long out_bstride, int ldo, float slope, cudaStream_t st);
  static constexpr size_t kMaxPlans = 6;
  std::list<Plan<OT>> plans_;
  Plan<OT>* last_plan_ = nullptr;   // plan of the most recent estimator / solve call (softmax_stats)
  bool blend_active_ = false;
  int tb_bstride_ = 0;          // 512 while an estimator call with per-clip times runs (estimator_forward_t)
  DevBuf loss_v_, loss_part_, loss_lmask_;
  cudaStream_t capture_stream_ = nullptr;
  // Pinned staging for the small per-call tables (t, dt, per-clip t): the host values go through a page-locked slot so
  // that the upload needs no stream synchronisation; a slot is reused only after its own copy has finished.
  static constexpr int kPinSlots = 4;
  static constexpr size_t kPinSlotFloats = 4096;
  float* pin_ = nullptr;
  cudaEvent_t pin_ev_[kPinSlots] = {};
  int pin_next_ = 0;
  void upload_floats(void* dst_dev, const float* src_host, size_t n, cudaStream_t st);

  // weights
  const HostTensor& hw(const std::string& key, std::initializer_list<int64_t> dims);
  float* up_f32(const std::vector<float>& v);
  OT* up_ot(const std::vector<float>& v);
  OT* up_w(const std::vector<float>& v, int N, int K);
  ResnetW<OT> pack_resnet(const std::string& p, int cin);
  TfmW<OT> pack_tfm(const std::string& p);
  OT* pack_conv3(const HostTensor& w, int cin, int cin_pad, float* /*unused*/ = nullptr);

  // plan
  Plan<OT>& get_plan(int B, int T);
  void build_plan(Plan<OT>& pl);
  void add_resnet(Plan<OT>& pl, int ridx, int level, const OT* src0, int c0, int ld0, const OT* src1, int c1,
                  const std::string& tag, const TfmW<OT>& next_ln);
  void add_tfm(Plan<OT>& pl, int tidx, int level, OT* out_hc, const std::string& tag);
  GemmParams base_gemm(const Plan<OT>& pl, int level) const;

  // run
  void launch_op(Plan<OT>& pl, Op& op, int step, float* v_out, bool update_x, cudaStream_t st);
  void launch_gemm(const GemmParams& gp, int bn, dim3 grid, cudaStream_t st);
  void run_forward(Plan<OT>& pl, int step, float* v_out, bool update_x, cudaStream_t st);
  void stage_inputs(Plan<OT>& pl, const float* x, const float* mu, const uint8_t* mask, const float* spk,
                    cudaMemcpyKind kind, cudaStream_t st);
  void prepass(Plan<OT>& pl, const std::vector<float>& ts, const std::vector<float>& dts, cudaStream_t st);
  void solve_staged(Plan<OT>& pl, const float* t_span, int n_steps, bool blend, cudaStream_t st,
                    bool t_span_on_device = false);
  void capture_tap(const Plan<OT>& pl, const Tap& tap, cudaStream_t st);
  void check_abort();
  void estimator_forward_times(const float* x, const float* mu, const uint8_t* mask, const float* spk,
                               const float* t_src, bool t_on_device, int n_t, int B, int T, float* v_out, cudaStream_t st);
};

}  // namespace srn
