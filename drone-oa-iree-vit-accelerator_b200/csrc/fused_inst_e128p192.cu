// fused_inst_e128p192.cu -- instantiates the fused attention kernel for E=128, P=192 (see fused_kernel.cuh).
#include "fused_kernel.cuh"
namespace ita {
ITA_FUSED_DEFINE_SHAPE(128, 192)
}
