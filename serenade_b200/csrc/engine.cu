// Template instantiation of the engine for the two operand precisions.
#include "engine_impl.cuh"

#include "engine_weights.inl"
#include "engine_plan.inl"
#include "engine_run.inl"

namespace srn {

template <typename OT>
static void configure_kernels() {
  SRN_CUDA(cudaFuncSetAttribute(gemm_kernel<OT, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                GemmCfg<128>::kSmemBytes));
  SRN_CUDA(cudaFuncSetAttribute(gemm_kernel<OT, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                GemmCfg<256>::kSmemBytes));
}

std::unique_ptr<EngineBase> make_engine(const SrnConfig& cfg) {
  if (cfg.precision == SRN_PRECISION_BF16) {
    configure_kernels<__nv_bfloat16>();
    return std::unique_ptr<EngineBase>(new Engine<__nv_bfloat16>(cfg));
  }
  if (cfg.precision == SRN_PRECISION_TF32) {
    configure_kernels<float>();
    return std::unique_ptr<EngineBase>(new Engine<float>(cfg));
  }
  fail("unknown precision %d", cfg.precision);
}

}  // namespace srn
