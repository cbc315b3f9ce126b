// srn_cfm_loss_from_velocity_noalloc: the CFM loss value (flow_matching.py:112-133) from a velocity tensor the caller
// already holds, WITHOUT any device allocation - for callers that cannot call cudaMalloc (code under CUDA-graph capture,
// tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// Nothing else in the library goes through this file.
//
// It is the two-stage deterministic reduction of srn_cfm_loss (rowops.cuh: cfm_loss_partial_kernel / cfm_loss_final_kernel,
// same block count, same order of additions, doubles) with the per-block partial sums - the one piece of scratch memory
// that reduction needs, which Engine::cfm_loss keeps in a cudaMalloc'ed buffer - in a file-scope __device__ array:
//
//   * g_noalloc_partial is 16 KB (kNoAllocBlocks x {sum, count} doubles).  EVERY GPU that loads the library holds it
//     until the process exits.
//   * In host code the array's NAME is a host-side stand-in, not a device address.  The device address is taken once per
//     device, with that device current, by cudaGetSymbolAddress (below) and passed to the kernels.
//   * All calls on a device share the array and it keeps the last call's contents.  Nothing has to start at zero: every
//     slot the second kernel reads is written by the first kernel of the same call.  Calls that OVERLAP on the GPU (two
//     streams, two threads) race on it - serialise them.
//   * Sizes are bounded (kNoAllocMaxElems elements per tensor); larger calls are rejected, not truncated.
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>

#include "../../include/serenade_b200.h"
#include "common.h"

namespace {

constexpr int kNoAllocBlocks = 1024;                      // blocks of the first stage = partial-sum slots
constexpr long long kNoAllocMaxElems = SRN_NOALLOC_MAX_ELEMS;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) double g_noalloc_partial[2 * kNoAllocBlocks];

// partial[2 * block] = sum over the block's elements of (v - u)^2 with u = x1 - (1 - sigma) z over the frames `sel` selects
// (null: every frame - without mask_l the reference sums over padded frames too); partial[2 * block + 1] = frames with
// cntm != 0 the block saw in channel 0 (the divisor).  Same arithmetic and order as rowops.cuh: cfm_loss_partial_kernel.
__global__ void __launch_bounds__(256) noalloc_loss_partial(const float* __restrict__ v, const float* __restrict__ x1,
                                                            const float* __restrict__ z, const uint8_t* __restrict__ sel,
                                                            const uint8_t* __restrict__ cntm, int C, int T, long long n,
                                                            float one_minus_sigma, double* __restrict__ partial) {
  __shared__ double red[2][256];
  double acc = 0.0, cnt = 0.0;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const int tt = (int)(i % T);
    const long long bc = i / T;
    const int c = (int)(bc % C), b = (int)(bc / C);
    const long long mi = (long long)b * T + tt;
    if (sel == nullptr || sel[mi] != 0) {
      const float u = __fsub_rn(x1[i], __fmul_rn(one_minus_sigma, z[i]));
      const double d = (double)v[i] - (double)u;
      acc += d * d;
    }
    if (c == 0 && cntm[mi] != 0) cnt += 1.0;
  }
  red[0][threadIdx.x] = acc;
  red[1][threadIdx.x] = cnt;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      red[0][threadIdx.x] += red[0][threadIdx.x + s];
      red[1][threadIdx.x] += red[1][threadIdx.x + s];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    partial[2 * blockIdx.x] = red[0][0];
    partial[2 * blockIdx.x + 1] = red[1][0];
  }
}

// loss = sum / (frames * C)   (flow_matching.py:128-132)
__global__ void noalloc_loss_final(const double* __restrict__ partial, int nblocks, int C, float* __restrict__ loss) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double s = 0.0, c = 0.0;
  for (int i = 0; i < nblocks; ++i) { s += partial[2 * i]; c += partial[2 * i + 1]; }
  *loss = (float)(s / (c * (double)C));
}

// Device address of g_noalloc_partial on the current device: cudaGetSymbolAddress once per device, with that device
// current (this function runs on the calling thread's current device), return code checked.
double* scratch_for_current_device() {
  static std::mutex mu;
  static double* addr[kMaxDevices] = {};
  int dev = -1;
  SRN_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) srn::fail("srn_cfm_loss_from_velocity_noalloc: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    SRN_CUDA(cudaGetSymbolAddress(&p, g_noalloc_partial));
    if (p == nullptr) srn::fail("cudaGetSymbolAddress returned a null address for the no-alloc scratch array");
    addr[dev] = static_cast<double*>(p);
  }
  return addr[dev];
}

}  // namespace

extern "C" __attribute__((visibility("default"))) int srn_cfm_loss_from_velocity_noalloc(
    const float* v, const float* x1, const float* z, const uint8_t* mask, const uint8_t* mask_l, int32_t B, int32_t C,
    int32_t T, float sigma_min, float* loss_out, void* stream) {
  try {
    if (!v || !x1 || !z || !mask || !loss_out) srn::fail("srn_cfm_loss_from_velocity_noalloc: null argument");
    if (B < 1 || C < 1 || T < 1) srn::fail("srn_cfm_loss_from_velocity_noalloc: need B, C, T >= 1 (got %d, %d, %d)", B, C, T);
    const long long n = (long long)B * C * T;
    if (n > kNoAllocMaxElems)
      srn::fail("srn_cfm_loss_from_velocity_noalloc: B * C * T = %lld elements exceeds the limit of %lld of the no-allocation "
                "entry point (SRN_NOALLOC_MAX_ELEMS); use srn_cfm_loss", n, kNoAllocMaxElems);
    double* scratch = scratch_for_current_device();
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const float oms = (float)(1.0 - (double)sigma_min);
    noalloc_loss_partial<<<kNoAllocBlocks, 256, 0, st>>>(v, x1, z, mask_l, mask_l != nullptr ? mask_l : mask, C, T, n, oms,
                                                         scratch);
    noalloc_loss_final<<<1, 32, 0, st>>>(scratch, kNoAllocBlocks, C, loss_out);
    SRN_CUDA(cudaGetLastError());
  } catch (const std::exception& e) {
    srn::set_last_error(e.what());
    return 1;
  } catch (...) {
    srn::set_last_error("unknown C++ exception");
    return 2;
  }
  return 0;
}
