/* serenade_b200 C ABI -- the drop-in boundary for Serenade's CFM mel-infilling decoder hot path.
 *
 * The reference (imulki/serenade) is pure Python/PyTorch and has no FFI of its own; the seam this
 * library replaces is the Python operator API of
 *   serenade/models/matcha_components/flow_matching.py:10-19  CFM.__init__      -> srn_create
 *   torch load_state_dict of `cfm_decoder.estimator.*` keys   (bin/ssc_decode.py:339) -> srn_load_weight / srn_finalize_weights
 *   decoder.py:384-467   Decoder.forward(x, mask, mu, t, spk) -> srn_estimator_forward
 *   flow_matching.py:65-93  CFM.solve_euler(x, t_span, mu, mask, spk) -> srn_solve_euler / srn_solve_euler_host
 * Plain pointers and sizes only; no torch types.  All functions return 0 on success, non-zero on
 * failure with a message available from srn_last_error().  A handle is bound to the CUDA device that
 * was current at srn_create and is not thread-safe.  There is no CPU fallback: every entry point
 * fails if no sm_100 device is present.
 *
 * Tensor layouts are the reference's: x, v (B, 80, T) f32 channels-first; mu (B, C_mu, T) f32;
 * mask (B, T) uint8 (non-zero = valid frame; the reference's bool (B,1,T)); spk (B, spk_dim) f32.
 */
#ifndef SERENADE_B200_H_
#define SERENADE_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct SrnHandle SrnHandle;

enum { SRN_PRECISION_BF16 = 0, SRN_PRECISION_TF32 = 1 };

typedef struct SrnConfig {
  int32_t in_channels;        /* x + mu channels: 242 (Serenade) or 244 (SerenadeNew); <= 256 */
  int32_t out_channels;       /* 80 */
  int32_t spk_embed_dim;      /* 256 */
  int32_t channels;           /* 512 (decoder_channels = (512, 512)) */
  int32_t num_heads;          /* 4 */
  int32_t attention_head_dim; /* 512 */
  int32_t num_mid_blocks;     /* 2 */
  int32_t precision;          /* SRN_PRECISION_* : GEMM operand type (accumulation is always f32) */
} SrnConfig;

const char* srn_last_error(void);
int srn_version(void);

int srn_create(const SrnConfig* cfg, SrnHandle** out);
int srn_destroy(SrnHandle* h);

/* Copy one parameter of the reference state dict (key relative to `cfm_decoder.`, i.e.
 * "estimator.down_blocks.0.0.block1.block.0.weight"), f32, HOST memory, dims in torch order (ndim <= 4). */
int srn_load_weight(SrnHandle* h, const char* key, const float* host_data, int32_t ndim, const int64_t* dims);
/* Validate that every expected key arrived with the expected shape, repack (K-major conv taps, fused QKV,
 * interleaved GEGLU, split transposed-conv parities, operand dtype) and upload. */
int srn_finalize_weights(SrnHandle* h);

/* v = Decoder.forward(x, mask, mu, t, spk).  DEVICE pointers.  `stream` is a cudaStream_t. */
int srn_estimator_forward(SrnHandle* h, const float* x, const float* mu, const uint8_t* mask, const float* spk,
                          float t, int32_t B, int32_t T, float* v_out, void* stream);

/* The same with one time per clip, t_host[B] (HOST): Decoder.forward accepts a vector t (decoder.py:401-402), which is
 * how CFM.compute_loss calls it (flow_matching.py:115,123). */
int srn_estimator_forward_t(SrnHandle* h, const float* x, const float* mu, const uint8_t* mask, const float* spk,
                            const float* t_host, int32_t B, int32_t T, float* v_out, void* stream);

/* The same with the time(s) in DEVICE memory: t_dev[n_t], n_t = 1 (one time for the whole batch - the 0-dim CUDA tensor
 * the reference's solve_euler hands its estimator, flow_matching.py:79-86) or n_t = B (one per clip).  The call enqueues
 * work on `stream` and never waits for it, so a reference-style Python Euler loop over estimator(...) stays asynchronous. */
int srn_estimator_forward_tdev(SrnHandle* h, const float* x, const float* mu, const uint8_t* mask, const float* spk,
                               const float* t_dev, int32_t n_t, int32_t B, int32_t T, float* v_out, void* stream);

/* CFM.compute_loss, FORWARD VALUE ONLY (flow_matching.py:95-133; what trainers/ssc.py evaluates on the dev set under
 * no_grad): given the target x1, the noise z and the times t_host[B] the reference would draw,
 *   y = (1 - (1 - sigma_min) t) z + t x1,  u = x1 - (1 - sigma_min) z,
 *   loss = sum(((estimator(y, mask, mu, t, spk) - u) * mask_l)^2) / (sum(mask_l) * n_feats); without mask_l the sum runs
 *   over every frame (padded ones too, as the reference's does) and the divisor is sum(mask) * n_feats.
 * x1, z, y_out (B, n_feats, T) f32 DEVICE; mask, mask_l (B, T) u8 DEVICE (mask_l nullable); loss_out one f32 on the
 * DEVICE.  There is no backward pass in this library. */
int srn_cfm_loss(SrnHandle* h, const float* x1, const float* z, const float* t_host, const float* mu, const uint8_t* mask,
                 const float* spk, const uint8_t* mask_l, int32_t B, int32_t T, float sigma_min, float* y_out,
                 float* loss_out, void* stream);

/* srn_cfm_loss with the times t_dev[B] in DEVICE memory, where the reference draws them (flow_matching.py:115:
 * torch.rand([b, 1, 1], device=mu.device)): no device-to-host read, the call never waits for the stream. */
int srn_cfm_loss_tdev(SrnHandle* h, const float* x1, const float* z, const float* t_dev, const float* mu,
                      const uint8_t* mask, const float* spk, const uint8_t* mask_l, int32_t B, int32_t T, float sigma_min,
                      float* y_out, float* loss_out, void* stream);

/* x <- solve_euler(x, t_span, mu, mask, spk) in place on DEVICE memory.  t_span is a HOST array of
 * n_steps + 1 f32 values.  blend_mask / blend_src (DEVICE, nullable): optional infill blend, after
 * every step x[b,:,t] = blend_src[b,:,t] where blend_mask[b,t] != 0.  The reference has no blend
 * (flow_matching.py:87); pass NULL for reference behaviour. */
int srn_solve_euler(SrnHandle* h, float* x_inout, const float* mu, const uint8_t* mask, const float* spk,
                    const float* t_span_host, int32_t n_steps, int32_t B, int32_t T, const uint8_t* blend_mask,
                    const float* blend_src, void* stream);

/* srn_solve_euler with t_span in DEVICE memory (n_steps + 1 f32) - what the reference's CFM.inference builds
 * (flow_matching.py:61: torch.linspace(..., device=mu.device)).  The (t, dt) tables are formed on the device with the same
 * fp32 additions as on the host path, so the result is bit-identical and the call never waits for the stream. */
int srn_solve_euler_tdev(SrnHandle* h, float* x_inout, const float* mu, const uint8_t* mask, const float* spk,
                         const float* t_span_dev, int32_t n_steps, int32_t B, int32_t T, const uint8_t* blend_mask,
                         const float* blend_src, void* stream);

/* Same with HOST buffers (pinned for best speed): uploads z/mu/mask/spk, solves, downloads the mel. */
int srn_solve_euler_host(SrnHandle* h, const float* z_host, const float* mu_host, const uint8_t* mask_host,
                         const float* spk_host, const float* t_span_host, int32_t n_steps, int32_t B, int32_t T,
                         float* mel_out_host, void* stream);

/* ---- Conditioning front half of Serenade.inference (reference serenade/models/serenade.py:181-206).  Available when
 * the state dict passed to srn_load_weight also held the `encoder.*` and `gst.*` keys of the Serenade model (keys as
 * in the reference state dict, e.g. "encoder.model.1.weight_v", "gst.ref_enc.convs.0.weight"; 0-d tensors such as
 * BatchNorm's num_batches_tracked are not sent).  DEVICE pointers, f32.
 *   srn_content_encode: enc = Serenade.encoder(x)      (Conv1dResnet, serenade.py:224-376)   x (B, L, in_dim) -> (B, L, 80)
 *   srn_style_encode:   spk = Serenade.gst(ref_logmel) (StyleEncoder, modules/gst/style_encoder.py) (B, L, 80) -> (B, 256) */
int srn_content_encode(SrnHandle* h, const float* x, int32_t B, int32_t L, float* enc_out, void* stream);
int srn_style_encode(SrnHandle* h, const float* logmel, int32_t B, int32_t L, float* style_out, void* stream);

/* HiFi-GAN vocoder, the stage after the CFM decoder in serenade-decode.  A handle may hold the
 * vocoder alone: load the generator's state dict with keys "vocoder." + the reference's
 * (serenade/vocoder/models/hifigan.py:21-161 after remove_weight_norm(): input_conv, upsamples.i.1, blocks.n.convs{1,2}.d.1,
 * output_conv.1), then srn_finalize_weights.  Structure (stages, strides = kernel / 2, residual blocks, kernel sizes) is
 * read off the weights; the residual blocks' dilations are not recoverable from them:
 *   srn_vocoder_set_dilations  before srn_finalize_weights; default {1, 3, 5} (hifigan.py:32)
 *   srn_vocoder_hop            samples per mel frame = product of the upsampling strides
 *   srn_vocoder_decode         = Vocoder.decode_batch (vocoder/vocoder.py:64-75): mel (B, T, in_channels) f32 DEVICE, as the
 *                              acoustic model returns it; scale / shift (DEVICE [in_channels], both or neither) fold
 *                              the de-normalisation with the target statistics and the normalisation with the
 *                              vocoder's (c * scale + shift); wave_out (B, T * hop) f32 DEVICE. */
int srn_vocoder_set_dilations(SrnHandle* h, const int32_t* dilations, int32_t n);
int srn_vocoder_hop(SrnHandle* h, int32_t* hop_out);
int srn_vocoder_decode(SrnHandle* h, const float* mel, const float* scale, const float* shift, int32_t B, int32_t T,
                       float* wave_out, void* stream);

/* CFM loss value from a velocity the caller already holds, WITHOUT any device allocation.  No counterpart in the
 * reference: added for callers that cannot call cudaMalloc (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid it).  Needs no handle; runs on the calling thread's current device.
 *   loss = sum(((v - u) * mask_l)^2) / (sum(mask_l) * C),  u = x1 - (1 - sigma_min) z   (the tail of srn_cfm_loss; with
 *   mask_l NULL the sum runs over every frame and the divisor is sum(mask) * C, like the reference's compute_loss)
 * v, x1, z (B, C, T) f32 DEVICE; mask, mask_l (B, T) u8 DEVICE (mask_l nullable); loss_out one f32 on the DEVICE.
 * The reduction's only scratch memory, 16 KB of per-block partial sums, is a file-scope `static __device__` array
 * (csrc/noalloc.cu; every GPU that loads the library holds it until exit), so the call makes no cudaMalloc / cudaFree and
 * calls no allocating routine.  Limits: B * C * T <= SRN_NOALLOC_MAX_ELEMS, larger calls fail with a message.  All calls
 * on one device share that array and it keeps the last call's contents (every slot read is rewritten first by the same
 * call, nothing needs clearing): calls that OVERLAP on the GPU - other streams or threads - race; serialise them. */
#define SRN_NOALLOC_MAX_ELEMS (1ll << 26)
int srn_cfm_loss_from_velocity_noalloc(const float* v, const float* x1, const float* z, const uint8_t* mask,
                                       const uint8_t* mask_l, int32_t B, int32_t C, int32_t T, float sigma_min,
                                       float* loss_out, void* stream);

/* Number of kernels of this library launched since the handle was created. */
int64_t srn_launch_count(const SrnHandle* h);
/* Use one CUDA graph per (B, T, n_steps) for srn_solve_euler (default 1). */
int srn_set_use_graph(SrnHandle* h, int32_t enable);
/* GroupNorm statistics of Block1D (decoder.py:75) over the unmasked frames of each utterance only (default 0).
 * 0 = the reference's padded-batch behaviour: statistics span all T_max frames, so an utterance's result depends on
 *     how far its batch was padded.
 * 1 = every utterance of a padded batch gets exactly what the reference gives it when decoded alone, as its CLI does
 *     (bin/ssc_decode.py:346-378, batch 1, T = its own length): what a length-bucketed batched decode needs. */
int srn_set_gn_valid_only(SrnHandle* h, int32_t enable);

/* Run-time options (take effect at the next call; a cached CUDA graph is re-captured when one changes):
 *   "noshift"  1 (default): attention entries whose scores are provably within +-64 (Cauchy-Schwarz on the Q / K row
 *              norms) skip the row-max pass and run softmax without a shift; 0: every entry takes the row-max pass.
 * Unknown names fail. */
int srn_set_option(SrnHandle* h, const char* name, int32_t value);
/* Which softmax path the attention blocks of the most recent srn_solve_euler* / srn_estimator_forward* call took:
 * out[0] = (batch, head, block, step) entries that ran shift-free, out[1] = all entries.  Synchronises the device. */
int srn_softmax_stats(SrnHandle* h, int64_t* out2);

/* Test hooks: when enabled, every named intermediate of the next forward is kept (as f32, channels-last
 * (B, rows, cols)) and can be fetched by name.  srn_debug_fetch returns the element count via *n_out
 * and copies min(capacity, n) values. */
int srn_debug_enable(SrnHandle* h, int32_t enable);
int srn_debug_fetch(SrnHandle* h, const char* name, float* host_out, int64_t capacity, int64_t* n_out,
                    int32_t* rows_out, int32_t* cols_out);

/* Per-op CUDA-event timing of one estimator forward at (B, T) on synthetic contents: writes a JSON array
 * [{"op": name, "kernel": kind, "ms": avg, "flops": algorithmic flops (no channel padding, nothing for the
 * row-max pass), "flops_issued": what the tensor cores ran, "bytes": algorithmic bytes}, ...]. */
int srn_profile_ops(SrnHandle* h, int32_t B, int32_t T, int32_t reps, char* json_out, int64_t capacity);

#ifdef __cplusplus
}
#endif
#endif /* SERENADE_B200_H_ */
