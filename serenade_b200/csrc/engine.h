// Host-side engine: weight repacking, per-(B,T) execution plans (buffers + TMA maps + op list),
// the estimator forward and the Euler solve loop (optionally replayed as one CUDA graph).
#pragma once
#include <memory>

#include "../../include/serenade_b200.h"
#include "common.h"
#include "gemm.cuh"

namespace srn {

struct HostTensor {
  std::vector<float> data;
  std::vector<int64_t> dims;
};

struct DebugTensor {
  std::vector<float> data;   // (B, rows, cols)
  int B = 0, rows = 0, cols = 0;
};

class EngineBase {
 public:
  virtual ~EngineBase() = default;
  virtual void load_weight(const std::string& key, const float* data, int ndim, const int64_t* dims) = 0;
  virtual void finalize_weights() = 0;
  virtual void estimator_forward(const float* x, const float* mu, const uint8_t* mask, const float* spk, float t,
                                 int B, int T, float* v_out, cudaStream_t st) = 0;
  // per-clip times t[B] (host): what Decoder.forward does when t is a vector (decoder.py:401-402)
  virtual void estimator_forward_t(const float* x, const float* mu, const uint8_t* mask, const float* spk,
                                   const float* t_host, int B, int T, float* v_out, cudaStream_t st) = 0;
  // the time(s) in device memory: n_t = 1 (shared by the batch) or B (one per clip); no host round trip
  virtual void estimator_forward_tdev(const float* x, const float* mu, const uint8_t* mask, const float* spk,
                                      const float* t_dev, int n_t, int B, int T, float* v_out, cudaStream_t st) = 0;
  // CFM.compute_loss forward value (flow_matching.py:95-133) for given noise z and times t; no gradients
  virtual void cfm_loss(const float* x1, const float* z, const float* t_host, const float* mu, const uint8_t* mask,
                        const float* spk, const uint8_t* mask_l, int B, int T, float sigma_min, float* y_out,
                        float* loss_out, cudaStream_t st, bool t_on_device = false) = 0;
  virtual void solve_euler(float* x, const float* mu, const uint8_t* mask, const float* spk, const float* t_span,
                           int n_steps, int B, int T, const uint8_t* blend_mask, const float* blend_src,
                           cudaStream_t st, bool t_span_on_device = false) = 0;
  virtual void solve_euler_host(const float* z, const float* mu, const uint8_t* mask, const float* spk,
                                const float* t_span, int n_steps, int B, int T, float* mel_out, cudaStream_t st) = 0;
  virtual std::string profile_ops(int B, int T, int reps) = 0;
  virtual void content_encode(const float* x, int B, int L, float* out, cudaStream_t st) = 0;
  virtual void style_encode(const float* mel, int B, int L, float* out, cudaStream_t st) = 0;
  virtual void vocoder_decode(const float* mel, const float* scale, const float* shift, int B, int T, float* wave,
                              cudaStream_t st) = 0;
  virtual int vocoder_hop() const = 0;
  virtual void softmax_stats(int64_t* out2) = 0;
  bool noshift = true;            // srn_set_option("noshift"): shift-free softmax entries allowed
  std::vector<int> voc_dilations = {1, 3, 5};     // HiFiGANResidualBlock dilations (not recoverable from the weights)
  int64_t launches = 0;
  bool use_graph = true;
  bool debug = false;
  // GroupNorm statistics over the unmasked frames of each utterance only (srn_set_gn_valid_only): a padded batch
  // then gives every utterance exactly what the reference gives it when decoded alone (batch 1, T = its length).
  // Off (default) = the reference's own padded-batch behaviour (statistics over all T_max frames, decoder.py:75).
  bool gn_valid_only = false;
  std::map<std::string, DebugTensor> debug_tensors;
};

std::unique_ptr<EngineBase> make_engine(const SrnConfig& cfg);

}  // namespace srn

struct SrnHandle {
  std::unique_ptr<srn::EngineBase> eng;
  int device = 0;
};
