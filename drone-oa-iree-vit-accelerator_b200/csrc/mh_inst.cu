// mh_inst.cu -- instantiations of the multi-head S = 256 kernels (mh_kernel.cuh) for (E, P) = (128, 64).
#include "mh_kernel.cuh"

namespace ita {

#define ITA_MH_PICK(M)                                                                                                  \
  if (mode == M && softmax == 0) return dump ? k_ita_mh<128, 64, M, 0, true> : k_ita_mh<128, 64, M, 0, false>;          \
  if (mode == M && softmax == 1) return dump ? k_ita_mh<128, 64, M, 1, true> : k_ita_mh<128, 64, M, 1, false>;

MhKernelFn mh_pick(int E, int P, int mode, int softmax, bool dump) {
  if (E != 128 || P != 64) return nullptr;
  ITA_MH_PICK(RQM_FP)
  ITA_MH_PICK(RQM_TORCH_FP)
  ITA_MH_PICK(RQM_TORCH_FP2)
  ITA_MH_PICK(RQM_INT_GEN)
  ITA_MH_PICK(RQM_TORCH)
  return nullptr;
}

#define ITA_MO_PICK(M)                                                                                                  \
  if (mode == M && per_head) return dump ? k_ita_mh_out<128, 64, M, true, true> : k_ita_mh_out<128, 64, M, true, false>; \
  if (mode == M && !per_head) return k_ita_mh_out<128, 64, M, false, false>;

MoKernelFn mo_pick(int E, int P, int mode, bool per_head, bool dump) {
  if (E != 128 || P != 64) return nullptr;
  ITA_MO_PICK(RQM_FP)
  ITA_MO_PICK(RQM_TORCH_FP)
  ITA_MO_PICK(RQM_TORCH_FP2)
  ITA_MO_PICK(RQM_INT_GEN)
  ITA_MO_PICK(RQM_TORCH)
  return nullptr;
}

int mh_smem_bytes(int E, int P) { return (E == 128 && P == 64) ? MhCfg<128, 64>::SMEM_BYTES : 0; }

}  // namespace ita
