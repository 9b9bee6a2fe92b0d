/* ita_dispatch_shim.c -- the C symbol the IREE custom dispatch of the reference calls, backed by libita_b200.so.
 *
 * The reference's compiled model reaches its attention block through a statically imported C function
 * (hal.import.static, samples/inference_udp_FPGA_custom_dispatch/plugin/ITA_spec.mlir:30-33,47), which in the reference is
 * a memcpy stub in two spellings:
 *     void ITASelfAttention_workgroup(const T* input, T* output)                      plugin/ITA_dispatch.c:17-27
 *     void ITASelfAttention_workgroup_expanded(<two expanded memrefs, 10 arguments>)  plugin/ITA_dispatch.c:31-56
 * with T = f32 in the MLIR declaration (memref<1x128x128xf32>) and float16_t in the C file.  This shim exports both
 * spellings with the same calling convention and makes them compute the real operator: quant stub -> int8 ITA attention
 * (tcgen05 kernel) -> dequant stub, on caller-owned HOST buffers of shape [1,128,E], synchronously, as the local-sync HAL
 * expects.  T is float by default; build with -DITA_DISPATCH_F16 for the reference C file's float16_t.
 *
 * The dispatch carries no weights or quantisation parameters, so the layer it runs is bound
 * beforehand with ita_dispatch_bind().  The reference symbol returns void; the status of the last call is kept in
 * ita_dispatch_last_status() (0 = OK; the enclosing IREE dispatch function would return it,
 * third_party/iree/runtime/src/iree/hal/local/executable_library.h:362-371), and on error the output is zero-filled.
 * A freestanding object cannot launch CUDA work, so unlike the reference stub this is a hosted shared object.
 */
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/ita_b200.h"

#ifdef ITA_DISPATCH_F16
typedef uint16_t ita_dispatch_elem_t; /* IEEE binary16 bit pattern, as `typedef uint16_t float16_t` in the reference */
#else
typedef float ita_dispatch_elem_t;
#endif

static ita_mha_handle *g_handle = NULL;
static float g_s_in = 1.0f, g_s_out = 1.0f;
static int32_t g_elems = 0; /* S * E of the bound layer */
static int g_status = 0;

#define ITA_SHIM_API __attribute__((visibility("default")))

/* Install the layer the dispatch computes: an ita_mha_handle plus the scales of its quant / dequant stubs
 * (attention_blocks[i].quant.scale, attention_blocks[i].out_proj.scale).  seq_len * embed_dim = elements per call. */
ITA_SHIM_API int ita_dispatch_bind(ita_mha_handle *h, float s_in, float s_out, int32_t seq_len, int32_t embed_dim) {
  if (!h || !(s_in > 0.f) || !(s_out > 0.f) || seq_len <= 0 || embed_dim <= 0) return ITA_ERR_INVALID;
  g_handle = h;
  g_s_in = s_in;
  g_s_out = s_out;
  g_elems = seq_len * embed_dim;
  g_status = 0;
  return ITA_OK;
}
ITA_SHIM_API int ita_dispatch_last_status(void) { return g_status; }
ITA_SHIM_API int ita_dispatch_elem_size(void) { return (int)sizeof(ita_dispatch_elem_t); }

#ifdef ITA_DISPATCH_F16
static float half_to_float(uint16_t h) {
  const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1Fu, man = h & 0x3FFu, bits;
  if (exp == 0) {
    if (man == 0) {
      bits = sign;
    } else { /* subnormal: normalise */
      int e = -1;
      do { ++e; man <<= 1; } while (!(man & 0x400u));
      bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3FFu) << 13);
    }
  } else if (exp == 31) {
    bits = sign | 0x7F800000u | (man << 13);
  } else {
    bits = sign | ((exp + 112u) << 23) | (man << 13);
  }
  float f;
  memcpy(&f, &bits, 4);
  return f;
}
static uint16_t float_to_half(float f) { /* round to nearest even */
  uint32_t x;
  memcpy(&x, &f, 4);
  const uint32_t sign = (x >> 16) & 0x8000u;
  const int32_t exp = (int32_t)((x >> 23) & 0xFF) - 127 + 15;
  uint32_t man = x & 0x7FFFFFu;
  if (((x >> 23) & 0xFF) == 0xFF) return (uint16_t)(sign | 0x7C00u | (man ? 0x200u : 0));
  if (exp >= 31) return (uint16_t)(sign | 0x7C00u);
  if (exp <= 0) {
    if (exp < -10) return (uint16_t)sign;
    man |= 0x800000u;
    const int shift = 14 - exp;
    uint32_t h = man >> shift;
    const uint32_t rem = man & ((1u << shift) - 1), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (h & 1))) ++h;
    return (uint16_t)(sign | h);
  }
  uint32_t h = ((uint32_t)exp << 10) | (man >> 13);
  const uint32_t rem = man & 0x1FFFu;
  if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) ++h;
  return (uint16_t)(sign | h);
}
#endif

static void run(const ita_dispatch_elem_t *input, ita_dispatch_elem_t *output) {
  if (!input || !output) {
    g_status = ITA_ERR_INVALID;
    return;
  }
  if (!g_handle) {
    g_status = ITA_ERR_INVALID;
    memset(output, 0, sizeof(ita_dispatch_elem_t) * (size_t)(g_elems > 0 ? g_elems : 0));
    return;
  }
#ifdef ITA_DISPATCH_F16
  float *tmp = (float *)malloc(sizeof(float) * 2 * (size_t)g_elems);
  if (!tmp) {
    g_status = ITA_ERR_NOMEM;
    return;
  }
  for (int32_t i = 0; i < g_elems; ++i) tmp[i] = half_to_float(input[i]);
  g_status = ita_mha_forward_host_f32(g_handle, tmp, tmp + g_elems, 1, g_s_in, g_s_out, ITA_PATH_AUTO, NULL);
  if (g_status) memset(output, 0, sizeof(ita_dispatch_elem_t) * (size_t)g_elems);
  else
    for (int32_t i = 0; i < g_elems; ++i) output[i] = float_to_half(tmp[g_elems + i]);
  free(tmp);
#else
  g_status = ita_mha_forward_host_f32(g_handle, input, output, 1, g_s_in, g_s_out, ITA_PATH_AUTO, NULL);
  if (g_status) memset(output, 0, sizeof(ita_dispatch_elem_t) * (size_t)g_elems);
#endif
}

/* reference plugin/ITA_dispatch.c:17-27 */
ITA_SHIM_API void ITASelfAttention_workgroup(const ita_dispatch_elem_t *input, ita_dispatch_elem_t *output) { run(input, output); }

/* reference plugin/ITA_dispatch.c:31-56: each memref expanded to (base, aligned, offset, size, stride); the data starts
 * at aligned + offset (the reference ignores the offset; the subspan it is called with has offset 0, ITA_spec.mlir:40-43) */
ITA_SHIM_API void ITASelfAttention_workgroup_expanded(const ita_dispatch_elem_t *binding0, const ita_dispatch_elem_t *binding0_aligned,
                                                      size_t binding0_offset, size_t binding0_size, size_t binding0_stride,
                                                      ita_dispatch_elem_t *binding1, ita_dispatch_elem_t *binding1_aligned,
                                                      size_t binding1_offset, size_t binding1_size, size_t binding1_stride) {
  (void)binding0_size; (void)binding0_stride; (void)binding1_size; (void)binding1_stride;
  const ita_dispatch_elem_t *in = binding0_aligned ? binding0_aligned : binding0;
  ita_dispatch_elem_t *out = binding1_aligned ? binding1_aligned : binding1;
  if (!in || !out) {
    g_status = ITA_ERR_INVALID;
    return;
  }
  run(in + binding0_offset, out + binding1_offset);
}
