// fused_inst_e128p128.cu -- instantiates the fused attention kernel for E=128, P=128 (see fused_kernel.cuh).
#include "fused_kernel.cuh"
namespace ita {
ITA_FUSED_DEFINE_SHAPE(128, 128)
}
