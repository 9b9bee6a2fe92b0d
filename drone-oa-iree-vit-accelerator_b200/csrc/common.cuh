// common.cuh -- shared definitions of libita_b200: handle layout, error plumbing, the requantisation arithmetic
//  in the exact forms the kernels use.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/ita_b200.h"

namespace ita {

// ---- error plumbing --------------------------------------------------------------------------------------------------
void set_error(const char *fmt, ...);
extern thread_local char g_err[512];
void count_launch(int n = 1);

#define ITA_CUDA_TRY(expr)                                                                              \
  do {                                                                                                  \
    cudaError_t _e = (expr);                                                                            \
    if (_e != cudaSuccess) {                                                                            \
      ita::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));             \
      return ITA_ERR_CUDA;                                                                              \
    }                                                                                                   \
  } while (0)

// ---- requantisation --------------------------------------------------------------------------------------------------
// One stage's parameters for one head, in the form the device code consumes.
struct Rq {
  int32_t mult, shift, add, mode;
  float eff;
  long long rnd;   // 1 << (shift-1) for ITA_ROUND_HALF_UP with shift > 0, else 0
};

__host__ __device__ __forceinline__ int32_t sat8(long long v) { return (int32_t)(v < -128 ? -128 : (v > 127 ? 127 : v)); }

// General, exact for every parameter value (64-bit intermediate, clamp in 64 bit).
__device__ __forceinline__ int32_t requant_general(int32_t acc, const Rq &r) {
  if (r.mode == ITA_ROUND_TORCH_RNE_F32) {
    float f = __fmul_rn(__int2float_rn(acc), r.eff);   // one IEEE multiply, no contraction
    long long y = (long long)__float2ll_rn(f) + r.add;
    return sat8(y);
  }
  long long p = (long long)acc * (long long)r.mult + r.rnd;
  int sh = r.shift;
  long long y = (sh >= 63 ? (p < 0 ? -1LL : 0LL) : (p >> sh)) + r.add;
  return sat8(y);
}

// ---- handle ----------------------------------------------------------------------------------------------------------
struct FusedPlan;   // fused.cu
struct FfnPlan;     // ffn.cu
struct MhPlan;      // mh.cu

}  // namespace ita

struct ita_mha_handle {
  int32_t S, E, P, H, F;
  int32_t round_mode, softmax_mode, head_mode;
  int32_t device;
  int32_t sm_count;
  // device copies, golden-model layouts (staged kernels)
  int8_t *Wq = nullptr, *Wk = nullptr, *Wv = nullptr, *Wo = nullptr, *Wff = nullptr, *Wff2 = nullptr;
  int32_t *Bq = nullptr, *Bk = nullptr, *Bv = nullptr, *Bo = nullptr, *Bff = nullptr, *Bff2 = nullptr;
  ita::Rq *rq_dev = nullptr;            // [ITA_NUM_STAGES * H]
  ita::Rq rq_host[ITA_NUM_STAGES * 16];  // H <= 16
  // scratch for forward paths that need intermediates in HBM (grown on demand)
  int8_t *scratch = nullptr;
  size_t scratch_bytes = 0;
  int8_t *io_dev = nullptr;             // device staging for ita_mha_forward_host
  size_t io_bytes = 0;
  int *abort_flag = nullptr;            // device int: set by a kernel whose bounded wait timed out
  // ita_mha_forward_host pipeline (created on first use): aux[0] H2D copies, aux[1] D2H copies
  static constexpr int kHostChunks = 8;
  cudaStream_t aux[2] = {nullptr, nullptr};
  cudaEvent_t ev_in[kHostChunks] = {}, ev_k[kHostChunks] = {}, ev_start = nullptr, ev_done[2] = {nullptr, nullptr};
  bool pipe_ready = false;
  ita::FusedPlan *fused = nullptr;      // non-null when the fused tcgen05 kernel supports this handle
  ita::MhPlan *mh = nullptr;            // non-null when the multi-head S = 256 tcgen05 kernels support this handle
  std::string fused_reason;             // why neither, otherwise
  bool warned_staged = false;           // ITA_PATH_AUTO fallback notice printed
  ita::FfnPlan *ffn = nullptr;          // non-null when the fused tcgen05 FFN kernel supports this handle
  std::string ffn_reason;
  int8_t *layer_scratch = nullptr;      // ita_encoder_layer_forward intermediates
  size_t layer_scratch_bytes = 0;
};

namespace ita {
int ensure_scratch(ita_mha_handle *h, size_t bytes);
// staged.cu
int staged_forward(ita_mha_handle *h, const int8_t *x, int8_t *y, int64_t B, const ita_mha_stages *st, cudaStream_t s);
int staged_step(ita_mha_handle *h, int step, const void *in0, const void *in1, void *out, int64_t B, cudaStream_t s);
int staged_ffn(ita_mha_handle *h, const int8_t *x, int8_t *f1, int8_t *f2, int64_t B, cudaStream_t s);
int launch_quantize(const float *x, int8_t *q, int64_t n, float scale, cudaStream_t s);
int launch_dequantize(const int8_t *q, float *y, int64_t n, float scale, cudaStream_t s);
// fused.cu
int fused_plan_create(ita_mha_handle *h, const ita_mha_desc *d);   // sets h->fused or h->fused_reason; returns ITA_OK unless CUDA fails
void fused_plan_destroy(ita_mha_handle *h);
int fused_forward(ita_mha_handle *h, const int8_t *x, int8_t *y, int64_t B, const ita_mha_stages *dump, cudaStream_t s);
// mh.cu
int mh_plan_create(ita_mha_handle *h, const ita_mha_desc *d);      // sets h->mh or h->fused_reason
void mh_plan_destroy(ita_mha_handle *h);
int mh_forward(ita_mha_handle *h, const int8_t *x, int8_t *y, int64_t B, const ita_mha_stages *dump, cudaStream_t s);
std::string describe_mh(const ita_mha_handle *h);
// ffn.cu
int ffn_plan_create(ita_mha_handle *h, const ita_mha_desc *d);     // sets h->ffn or h->ffn_reason
void ffn_plan_destroy(ita_mha_handle *h);
int ffn_fused_forward(ita_mha_handle *h, const int8_t *x, int8_t *f1, int8_t *f2, int64_t B, cudaStream_t s);
std::string describe_fused(const ita_mha_handle *h);   // fused.cu
std::string describe_ffn(const ita_mha_handle *h);     // ffn.cu
// layer.cu
int launch_resid_ln(const float *x, const int8_t *q, float sq, const float *gamma, const float *beta, float eps, float *y, int8_t *yq,
                    float s_next, int64_t rows, int E, cudaStream_t s);
int layer_forward(ita_mha_handle *h, const ita_layer_params *lp, const float *x, float *y, int64_t B, cudaStream_t s);
// tokenizer.cu
int launch_tokenizer(const void *img, int is_u8, int64_t B, int E, const float *w, const float *cb, const float *gamma, const float *beta,
                     float eps, float *tok, int8_t *tokq, float q_scale, cudaStream_t s);
int launch_lstm_cell(const float *gates, const float *c_prev, float *h, float *c, int64_t B, int Hd, cudaStream_t s);
int launch_fuse_conv(const float *x, int64_t B, int E, const float *w, const float *cb, float *out, cudaStream_t s);
// selftest.cu
int selftest_run(int device, std::string &report);
int microbench_run(int device, int kind, double *out, int n_out);
}  // namespace ita
