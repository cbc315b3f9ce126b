// extern "C" boundary (include/serenade_b200.h): status codes + srn_last_error, no exceptions cross it.
#include <cstring>

#include "engine.h"

namespace srn {

static thread_local std::string g_last_error;
void set_last_error(const std::string& msg) { g_last_error = msg; }

CUtensorMap make_tmap_3d(const void* base, int elem_bytes, uint64_t inner, uint64_t rows, uint64_t batch,
                         uint64_t ld_elems, uint64_t bstride_elems, uint32_t box_inner, uint32_t box_rows,
                         bool swizzle128) {
  using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = nullptr;
  if (encode == nullptr) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    SRN_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (fn == nullptr || qres != cudaDriverEntryPointSuccess) fail("cuTensorMapEncodeTiled not available from the driver");
    encode = reinterpret_cast<EncodeFn>(fn);
  }
  if (reinterpret_cast<uintptr_t>(base) % 16 != 0) fail("TMA base address not 16-byte aligned");
  if ((ld_elems * elem_bytes) % 16 != 0 || (bstride_elems * elem_bytes) % 16 != 0) fail("TMA stride not a multiple of 16 bytes");
  if (swizzle128 && box_inner * elem_bytes != 128) fail("TMA box inner extent must be 128 bytes (SWIZZLE_128B)");
  CUtensorMap m;
  memset(&m, 0, sizeof m);
  cuuint64_t dims[3] = {inner, rows, batch};
  cuuint64_t strides[2] = {ld_elems * (uint64_t)elem_bytes, bstride_elems * (uint64_t)elem_bytes};
  cuuint32_t box[3] = {box_inner, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  const CUtensorMapDataType dt = elem_bytes == 1   ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                 : elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                                   : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = encode(&m, dt, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) fail("cuTensorMapEncodeTiled failed with CUresult %d (inner=%llu rows=%llu batch=%llu ld=%llu)", (int)r,
                              (unsigned long long)inner, (unsigned long long)rows, (unsigned long long)batch,
                              (unsigned long long)ld_elems);
  return m;
}

}  // namespace srn

#define SRN_API_BEGIN try {
#define SRN_API_END                                \
  }                                                \
  catch (const std::exception& e) {                \
    srn::set_last_error(e.what());                 \
    return 1;                                      \
  }                                                \
  catch (...) {                                    \
    srn::set_last_error("unknown C++ exception");  \
    return 2;                                      \
  }                                                \
  return 0;

// Every entry point runs on the handle's device and leaves the caller's current device as it found it.
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) SRN_CUDA(cudaSetDevice(dev));
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};
#define check_handle(h)                                              \
  if ((h) == nullptr || !(h)->eng) srn::fail("null SrnHandle");      \
  DeviceGuard srn_device_guard((h)->device)

extern "C" {
#define SRN_EXPORT __attribute__((visibility("default")))

SRN_EXPORT const char* srn_last_error(void) { return srn::g_last_error.c_str(); }
SRN_EXPORT int srn_version(void) { return 100; }

SRN_EXPORT int srn_create(const SrnConfig* cfg, SrnHandle** out) {
  SRN_API_BEGIN
  if (cfg == nullptr || out == nullptr) srn::fail("srn_create: null argument");
  int dev = 0;
  SRN_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  SRN_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (prop.major != 10) srn::fail("serenade_b200 needs an sm_100 (B200) device; device %d is sm_%d%d and there is no fallback path", dev, prop.major, prop.minor);
  auto h = new SrnHandle;
  h->device = dev;
  try {
    h->eng = srn::make_engine(*cfg);
  } catch (...) {
    delete h;
    throw;
  }
  *out = h;
  SRN_API_END
}

SRN_EXPORT int srn_destroy(SrnHandle* h) {
  SRN_API_BEGIN
  if (h != nullptr) {
    DeviceGuard guard(h->device);
    delete h;
  }
  SRN_API_END
}

SRN_EXPORT int srn_load_weight(SrnHandle* h, const char* key, const float* host_data, int32_t ndim, const int64_t* dims) {
  SRN_API_BEGIN
  check_handle(h);
  if (key == nullptr || host_data == nullptr || dims == nullptr) srn::fail("srn_load_weight: null argument");
  h->eng->load_weight(key, host_data, ndim, dims);
  SRN_API_END
}

SRN_EXPORT int srn_finalize_weights(SrnHandle* h) {
  SRN_API_BEGIN
  check_handle(h);
  h->eng->finalize_weights();
  SRN_API_END
}

SRN_EXPORT int srn_estimator_forward(SrnHandle* h, const float* x, const float* mu, const uint8_t* mask, const float* spk,
                          float t, int32_t B, int32_t T, float* v_out, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (!x || !mu || !mask || !spk || !v_out) srn::fail("srn_estimator_forward: null argument");
  if (B < 1 || T < 2) srn::fail("srn_estimator_forward: need B >= 1 and T >= 2 (got B=%d T=%d)", B, T);
  h->eng->estimator_forward(x, mu, mask, spk, t, B, T, v_out, static_cast<cudaStream_t>(stream));
  SRN_API_END
}

SRN_EXPORT int srn_estimator_forward_t(SrnHandle* h, const float* x, const float* mu, const uint8_t* mask, const float* spk,
                                       const float* t_host, int32_t B, int32_t T, float* v_out, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (!x || !mu || !mask || !spk || !t_host || !v_out) srn::fail("srn_estimator_forward_t: null argument");
  if (B < 1 || T < 2) srn::fail("srn_estimator_forward_t: need B >= 1 and T >= 2 (got B=%d T=%d)", B, T);
  h->eng->estimator_forward_t(x, mu, mask, spk, t_host, B, T, v_out, static_cast<cudaStream_t>(stream));
  SRN_API_END
}

SRN_EXPORT int srn_estimator_forward_tdev(SrnHandle* h, const float* x, const float* mu, const uint8_t* mask,
                                          const float* spk, const float* t_dev, int32_t n_t, int32_t B, int32_t T,
                                          float* v_out, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (!x || !mu || !mask || !spk || !t_dev || !v_out) srn::fail("srn_estimator_forward_tdev: null argument");
  if (B < 1 || T < 2) srn::fail("srn_estimator_forward_tdev: need B >= 1 and T >= 2 (got B=%d T=%d)", B, T);
  if (n_t != 1 && n_t != B) srn::fail("srn_estimator_forward_tdev: n_t must be 1 or B = %d (got %d)", B, n_t);
  h->eng->estimator_forward_tdev(x, mu, mask, spk, t_dev, n_t, B, T, v_out, static_cast<cudaStream_t>(stream));
  SRN_API_END
}

SRN_EXPORT int srn_cfm_loss(SrnHandle* h, const float* x1, const float* z, const float* t_host, const float* mu,
                            const uint8_t* mask, const float* spk, const uint8_t* mask_l, int32_t B, int32_t T,
                            float sigma_min, float* y_out, float* loss_out, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (!x1 || !z || !t_host || !mu || !mask || !spk || !y_out || !loss_out) srn::fail("srn_cfm_loss: null argument");
  if (B < 1 || T < 2) srn::fail("srn_cfm_loss: need B >= 1 and T >= 2 (got B=%d T=%d)", B, T);
  h->eng->cfm_loss(x1, z, t_host, mu, mask, spk, mask_l, B, T, sigma_min, y_out, loss_out, static_cast<cudaStream_t>(stream));
  SRN_API_END
}

SRN_EXPORT int srn_cfm_loss_tdev(SrnHandle* h, const float* x1, const float* z, const float* t_dev, const float* mu,
                                 const uint8_t* mask, const float* spk, const uint8_t* mask_l, int32_t B, int32_t T,
                                 float sigma_min, float* y_out, float* loss_out, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (!x1 || !z || !t_dev || !mu || !mask || !spk || !y_out || !loss_out) srn::fail("srn_cfm_loss_tdev: null argument");
  if (B < 1 || T < 2) srn::fail("srn_cfm_loss_tdev: need B >= 1 and T >= 2 (got B=%d T=%d)", B, T);
  h->eng->cfm_loss(x1, z, t_dev, mu, mask, spk, mask_l, B, T, sigma_min, y_out, loss_out, static_cast<cudaStream_t>(stream),
                   true);
  SRN_API_END
}

SRN_EXPORT int srn_solve_euler(SrnHandle* h, float* x_inout, const float* mu, const uint8_t* mask, const float* spk,
                    const float* t_span_host, int32_t n_steps, int32_t B, int32_t T, const uint8_t* blend_mask,
                    const float* blend_src, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (!x_inout || !mu || !mask || !spk || !t_span_host) srn::fail("srn_solve_euler: null argument");
  if (B < 1 || T < 2 || n_steps < 1) srn::fail("srn_solve_euler: need B >= 1, T >= 2, n_steps >= 1 (got %d, %d, %d)", B, T, n_steps);
  if ((blend_mask == nullptr) != (blend_src == nullptr)) srn::fail("srn_solve_euler: blend_mask and blend_src must be given together");
  h->eng->solve_euler(x_inout, mu, mask, spk, t_span_host, n_steps, B, T, blend_mask, blend_src,
                      static_cast<cudaStream_t>(stream));
  SRN_API_END
}

SRN_EXPORT int srn_solve_euler_tdev(SrnHandle* h, float* x_inout, const float* mu, const uint8_t* mask, const float* spk,
                                    const float* t_span_dev, int32_t n_steps, int32_t B, int32_t T,
                                    const uint8_t* blend_mask, const float* blend_src, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (!x_inout || !mu || !mask || !spk || !t_span_dev) srn::fail("srn_solve_euler_tdev: null argument");
  if (B < 1 || T < 2 || n_steps < 1)
    srn::fail("srn_solve_euler_tdev: need B >= 1, T >= 2, n_steps >= 1 (got %d, %d, %d)", B, T, n_steps);
  if ((blend_mask == nullptr) != (blend_src == nullptr))
    srn::fail("srn_solve_euler_tdev: blend_mask and blend_src must be given together");
  h->eng->solve_euler(x_inout, mu, mask, spk, t_span_dev, n_steps, B, T, blend_mask, blend_src,
                      static_cast<cudaStream_t>(stream), true);
  SRN_API_END
}

SRN_EXPORT int srn_solve_euler_host(SrnHandle* h, const float* z_host, const float* mu_host, const uint8_t* mask_host,
                         const float* spk_host, const float* t_span_host, int32_t n_steps, int32_t B, int32_t T,
                         float* mel_out_host, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (!z_host || !mu_host || !mask_host || !spk_host || !t_span_host || !mel_out_host) srn::fail("srn_solve_euler_host: null argument");
  if (B < 1 || T < 2 || n_steps < 1) srn::fail("srn_solve_euler_host: need B >= 1, T >= 2, n_steps >= 1");
  h->eng->solve_euler_host(z_host, mu_host, mask_host, spk_host, t_span_host, n_steps, B, T, mel_out_host,
                           static_cast<cudaStream_t>(stream));
  SRN_API_END
}

SRN_EXPORT int srn_content_encode(SrnHandle* h, const float* x, int32_t B, int32_t L, float* out, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (x == nullptr || out == nullptr) srn::fail("srn_content_encode: null pointer");
  h->eng->content_encode(x, B, L, out, static_cast<cudaStream_t>(stream));
  SRN_API_END
}

SRN_EXPORT int srn_style_encode(SrnHandle* h, const float* logmel, int32_t B, int32_t L, float* out, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (logmel == nullptr || out == nullptr) srn::fail("srn_style_encode: null pointer");
  h->eng->style_encode(logmel, B, L, out, static_cast<cudaStream_t>(stream));
  SRN_API_END
}

SRN_EXPORT int srn_vocoder_set_dilations(SrnHandle* h, const int32_t* dilations, int32_t n) {
  SRN_API_BEGIN
  check_handle(h);
  if (dilations == nullptr || n < 1 || n > 8) srn::fail("srn_vocoder_set_dilations: 1..8 dilations expected");
  h->eng->voc_dilations.assign(dilations, dilations + n);
  for (int d : h->eng->voc_dilations)
    if (d < 1) srn::fail("srn_vocoder_set_dilations: dilations must be >= 1");
  SRN_API_END
}

SRN_EXPORT int srn_vocoder_hop(SrnHandle* h, int32_t* hop_out) {
  SRN_API_BEGIN
  check_handle(h);
  if (hop_out == nullptr) srn::fail("srn_vocoder_hop: null pointer");
  *hop_out = h->eng->vocoder_hop();
  if (*hop_out == 0) srn::fail("vocoder weights (vocoder.*) were not loaded");
  SRN_API_END
}

SRN_EXPORT int srn_vocoder_decode(SrnHandle* h, const float* mel, const float* scale, const float* shift, int32_t B,
                                  int32_t T, float* wave_out, void* stream) {
  SRN_API_BEGIN
  check_handle(h);
  if (mel == nullptr || wave_out == nullptr) srn::fail("srn_vocoder_decode: null pointer");
  h->eng->vocoder_decode(mel, scale, shift, B, T, wave_out, static_cast<cudaStream_t>(stream));
  SRN_API_END
}

SRN_EXPORT int64_t srn_launch_count(const SrnHandle* h) { return (h && h->eng) ? h->eng->launches : -1; }

SRN_EXPORT int srn_set_use_graph(SrnHandle* h, int32_t enable) {
  SRN_API_BEGIN
  check_handle(h);
  h->eng->use_graph = enable != 0;
  SRN_API_END
}

SRN_EXPORT int srn_set_gn_valid_only(SrnHandle* h, int32_t enable) {
  SRN_API_BEGIN
  check_handle(h);
  h->eng->gn_valid_only = enable != 0;
  SRN_API_END
}

SRN_EXPORT int srn_set_option(SrnHandle* h, const char* name, int32_t value) {
  SRN_API_BEGIN
  check_handle(h);
  if (name == nullptr) srn::fail("srn_set_option: null name");
  if (strcmp(name, "noshift") == 0) h->eng->noshift = value != 0;
  else srn::fail("srn_set_option: unknown option '%s'", name);
  SRN_API_END
}

SRN_EXPORT int srn_softmax_stats(SrnHandle* h, int64_t* out2) {
  SRN_API_BEGIN
  check_handle(h);
  if (out2 == nullptr) srn::fail("srn_softmax_stats: null pointer");
  h->eng->softmax_stats(out2);
  SRN_API_END
}

SRN_EXPORT int srn_debug_enable(SrnHandle* h, int32_t enable) {
  SRN_API_BEGIN
  check_handle(h);
  h->eng->debug = enable != 0;
  h->eng->debug_tensors.clear();
  SRN_API_END
}

SRN_EXPORT int srn_debug_fetch(SrnHandle* h, const char* name, float* host_out, int64_t capacity, int64_t* n_out,
                    int32_t* rows_out, int32_t* cols_out) {
  SRN_API_BEGIN
  check_handle(h);
  auto it = h->eng->debug_tensors.find(name ? name : "");
  if (it == h->eng->debug_tensors.end()) srn::fail("srn_debug_fetch: no debug tensor named '%s'", name ? name : "(null)");
  const auto& d = it->second;
  if (n_out) *n_out = (int64_t)d.data.size();
  if (rows_out) *rows_out = d.rows;
  if (cols_out) *cols_out = d.cols;
  if (host_out && capacity > 0) memcpy(host_out, d.data.data(), sizeof(float) * std::min<int64_t>(capacity, d.data.size()));
  SRN_API_END
}

SRN_EXPORT int srn_profile_ops(SrnHandle* h, int32_t B, int32_t T, int32_t reps, char* json_out, int64_t capacity) {
  SRN_API_BEGIN
  check_handle(h);
  if (json_out == nullptr || capacity < 2 || reps < 1) srn::fail("srn_profile_ops: bad arguments");
  const std::string js = h->eng->profile_ops(B, T, reps);
  if ((int64_t)js.size() + 1 > capacity) srn::fail("srn_profile_ops: output buffer too small (%zu needed)", js.size() + 1);
  memcpy(json_out, js.c_str(), js.size() + 1);
  SRN_API_END
}

}  // extern "C"
