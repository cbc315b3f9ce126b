// Template instantiation of the engine for the two operand precisions.
#include "engine_impl.cuh"

#include "engine_weights.inl"
#include "engine_plan.inl"
#include "engine_run.inl"
#include "engine_frontend.inl"
#include "engine_vocoder.inl"

namespace srn {

template <typename OT, int BN, int EPI>
static void configure_one() {
  SRN_CUDA(cudaFuncSetAttribute(gemm_kernel<OT, BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                GemmCfg<BN>::kSmemBytes));
  SRN_CUDA(cudaFuncSetAttribute(gemm_kernel<OT, BN, EPI>, cudaFuncAttributeNonPortableClusterSizeAllowed, 0));
}

template <typename OT, int BN>
static void configure_bn() {
  configure_one<OT, BN, EPI_STORE_F32>();
  configure_one<OT, BN, EPI_STORE_GN>();
  configure_one<OT, BN, EPI_STORE_OT>();
  configure_one<OT, BN, EPI_RESID_F32>();
  configure_one<OT, BN, EPI_RESID_OT>();
  configure_one<OT, BN, EPI_GEGLU>();
  configure_one<OT, BN, EPI_QKV>();
  configure_one<OT, BN, EPI_EULER>();
  configure_one<OT, BN, EPI_ROWMAX>();
  configure_one<OT, BN, EPI_EXP>();
  configure_one<OT, BN, EPI_PV>();
  configure_one<OT, BN, EPI_RESID_ACT>();
}

template <typename OT>
static void configure_kernels() {
  configure_bn<OT, 128>();
  configure_bn<OT, 256>();
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
