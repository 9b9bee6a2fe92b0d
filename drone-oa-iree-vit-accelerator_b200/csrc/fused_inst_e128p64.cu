// fused_inst_e128p64.cu -- instantiates the fused attention kernel for E=128, P=64 (see fused_kernel.cuh).
#include "fused_kernel.cuh"
namespace ita {
ITA_FUSED_DEFINE_SHAPE(128, 64)
}
