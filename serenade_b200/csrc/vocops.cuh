// Element-wise kernels of the HiFi-GAN vocoder (reference serenade/vocoder/models/hifigan.py:167-186,
// serenade/vocoder/layers/residual_block.py:247-258).  The convolutions and transposed convolutions run on the tcgen05
// GEMM core; what is left between them is LeakyReLU + conversion to the operand type, the residual-stream copies of a
// stage, the mean over its residual blocks and the final tanh:
//   voc_prepare_mel    (B, T, 80) f32 -> normalised (x * scale + shift), zero-padded channels-last operand
//   voc_lrelu_cast     f32 [rows][C] -> LeakyReLU -> operand [rows][Cp] (pad columns zero) and up to 3 f32 copies
//   voc_mean_lrelu     mean of up to 4 f32 tensors -> LeakyReLU -> operand [rows][Cp]
//   voc_tanh_out       column 0 of the f32 [rows][4] output-conv result -> tanh -> wave
// All are single-pass, 8 channels (16 / 32 bytes) per thread, HBM-bound.
#pragma once
#include <cuda_bf16.h>
#include <cstdint>

#include "frontops.cuh"

namespace srn {

// grid = ceil(rows * Cp / 8 / 256); Cin <= Cp, both % 8 == 0.  scale / shift may be null (identity).
template <typename OT>
__global__ void __launch_bounds__(256) voc_prepare_mel(const float* __restrict__ mel, long long rows, int Cin, int Cp,
                                                       const float* __restrict__ scale, const float* __restrict__ shift,
                                                       OT* __restrict__ dst) {
  pdl_launch_dependents();
  pdl_wait();
  const int c8 = Cp / 8;
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= rows * c8) return;
  const long long r = idx / c8;
  const int c = (int)(idx % c8) * 8;
  float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c < Cin) {
    load8(mel + r * Cin + c, v);
    if (scale != nullptr) {
      float sc[8], sh[8];
      load8(scale + c, sc);
      load8(shift + c, sh);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = fmaf(v[i], sc[i], sh[i]);
    }
  }
  store8_op(dst + r * Cp + c, v);
}

template <typename OT>
__global__ void __launch_bounds__(256) voc_lrelu_cast(const float* __restrict__ src, long long rows, int C, int Cp,
                                                      float slope, OT* __restrict__ dst, float* __restrict__ copy0,
                                                      float* __restrict__ copy1, float* __restrict__ copy2) {
  pdl_launch_dependents();
  pdl_wait();
  const int c8 = Cp / 8;
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= rows * c8) return;
  const long long r = idx / c8;
  const int c = (int)(idx % c8) * 8;
  float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c < C) {
    load8(src + r * C + c, v);
    if (copy0 != nullptr) store8(copy0 + r * C + c, v);
    if (copy1 != nullptr) store8(copy1 + r * C + c, v);
    if (copy2 != nullptr) store8(copy2 + r * C + c, v);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * slope;
  }
  store8_op(dst + r * Cp + c, v);
}

template <typename OT>
__global__ void __launch_bounds__(256) voc_mean_lrelu(const float* __restrict__ x0, const float* __restrict__ x1,
                                                      const float* __restrict__ x2, const float* __restrict__ x3, int n,
                                                      long long rows, int C, int Cp, float slope, OT* __restrict__ dst) {
  pdl_launch_dependents();
  pdl_wait();
  const int c8 = Cp / 8;
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= rows * c8) return;
  const long long r = idx / c8;
  const int c = (int)(idx % c8) * 8;
  float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c < C) {
    load8(x0 + r * C + c, v);
    auto add = [&](const float* x) {       // same order as `cs += block_j(c)` (hifigan.py:181-182)
      float w[8];
      load8(x + r * C + c, w);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] += w[i];
    };
    if (n > 1) add(x1);
    if (n > 2) add(x2);
    if (n > 3) add(x3);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      v[i] = v[i] / (float)n;
      v[i] = v[i] > 0.f ? v[i] : v[i] * slope;
    }
  }
  store8_op(dst + r * Cp + c, v);
}

// y: [rows / g][ld] f32, sample r = column r % g of row r / g
__global__ void __launch_bounds__(256) voc_tanh_out(const float* __restrict__ y, long long rows, int g, int ld,
                                                    float* __restrict__ wave) {
  pdl_launch_dependents();
  pdl_wait();
  const long long r = (long long)blockIdx.x * 256 + threadIdx.x;
  if (r < rows) wave[r] = tanhf(y[(r / g) * ld + (r % g)]);
}

}  // namespace srn
