// layer.cu -- what sits between the two int8 blocks of an encoder layer of ITAViTLSTM, and the layer as one call:
//
//     x (f32) --quant--> int8 attention --dequant--> + x --> LayerNorm --quant--> int8 FFN --dequant--> + . --> LayerNorm --> f32
//
// Reference: the loop body of ITALSTMNetVIT_QAT.forward (models/ITA_single_layer_upsample_shuffle/QAT/model.py:100-113;
// the residual `self.add` and the LayerNorms carry no qconfig, so they are float ops on the dequantised block outputs).
// The two int8 blocks are the tcgen05 kernels (fused.cu, ffn.cu); here are the HBM-bound float stages between them, each
// ONE pass over the tokens: dequantise + residual add + LayerNorm (+ the next block's quant stub) in a single kernel, one
// warp per token row, coalesced 16-byte accesses.
#include "common.cuh"

namespace ita {

// y = LayerNorm(x + sq * q) * gamma + beta over the last dim (E = 32 * VEC); optionally yq = sat8(rne(y / s_next)).
// The dequantise (one fp32 multiply) and the residual add (one fp32 add) are kept as two roundings, as in the reference.
// Each warp takes kRowsPerWarp consecutive token rows and issues all of their loads before the first reduction, so that
// ~2 KB per warp are in flight (one row alone is 256 - 512 B: latency-bound at 40 % of the HBM rate, measured).
constexpr int kRowsPerWarp = 4;
template <int VEC>
__global__ void __launch_bounds__(256) k_resid_ln(const float *__restrict__ x, const int8_t *__restrict__ q, float sq,
                                                  const float *__restrict__ gamma, const float *__restrict__ beta, float eps,
                                                  float *__restrict__ y, int8_t *__restrict__ yq, float inv_next, long long rows) {
  constexpr int E = 32 * VEC, R = kRowsPerWarp;
  const int lane = threadIdx.x & 31;
  const long long row0 = ((long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * R;
  if (row0 >= rows) return;
  float v[R][VEC];
  float g[VEC], bt[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    g[i] = gamma[lane * VEC + i];
    bt[i] = beta[lane * VEC + i];
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const long long row = row0 + r < rows ? row0 + r : rows - 1;       // tail rows re-read the last row (not stored)
    const float *xr = x + row * E + lane * VEC;
    const int8_t *qr = q + row * E + lane * VEC;
    if constexpr (VEC == 2) {
      const float2 a = *reinterpret_cast<const float2 *>(xr);
      const char2 b = *reinterpret_cast<const char2 *>(qr);
      v[r][0] = __fadd_rn(a.x, __fmul_rn(sq, (float)b.x));
      v[r][1] = __fadd_rn(a.y, __fmul_rn(sq, (float)b.y));
    } else {
      const float4 a = *reinterpret_cast<const float4 *>(xr);
      const char4 b = *reinterpret_cast<const char4 *>(qr);
      v[r][0] = __fadd_rn(a.x, __fmul_rn(sq, (float)b.x));
      v[r][1] = __fadd_rn(a.y, __fmul_rn(sq, (float)b.y));
      v[r][2] = __fadd_rn(a.z, __fmul_rn(sq, (float)b.z));
      v[r][3] = __fadd_rn(a.w, __fmul_rn(sq, (float)b.w));
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VEC; ++i) s += v[r][i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s * (1.0f / E);
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VEC; ++i) ss += (v[r][i] - mean) * (v[r][i] - mean);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float rstd = 1.0f / sqrtf(ss * (1.0f / E) + eps);
    if (row0 + r >= rows) continue;                                     // warp-uniform
    float o[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) o[i] = (v[r][i] - mean) * rstd * g[i] + bt[i];
    float *yr = y + (row0 + r) * E + lane * VEC;
    if constexpr (VEC == 2) *reinterpret_cast<float2 *>(yr) = make_float2(o[0], o[1]);
    else *reinterpret_cast<float4 *>(yr) = make_float4(o[0], o[1], o[2], o[3]);
    if (yq) {
      int8_t t[VEC];
#pragma unroll
      for (int i = 0; i < VEC; ++i) t[i] = (int8_t)(int)fminf(fmaxf(nearbyintf(o[i] * inv_next), -128.f), 127.f);
      int8_t *qo = yq + (row0 + r) * E + lane * VEC;
      if constexpr (VEC == 2) *reinterpret_cast<char2 *>(qo) = make_char2(t[0], t[1]);
      else *reinterpret_cast<char4 *>(qo) = make_char4(t[0], t[1], t[2], t[3]);
    }
  }
}

int launch_resid_ln(const float *x, const int8_t *q, float sq, const float *gamma, const float *beta, float eps, float *y, int8_t *yq,
                    float s_next, int64_t rows, int E, cudaStream_t s) {
  if (rows <= 0) return ITA_OK;
  if (E != 64 && E != 128) {
    set_error("residual + LayerNorm kernel is instantiated for E in {64, 128}, got %d", E);
    return ITA_ERR_UNSUPPORTED;
  }
  if (yq && !(s_next > 0.f)) {
    set_error("residual + LayerNorm: the quantised output needs a positive scale");
    return ITA_ERR_INVALID;
  }
  const float inv = yq ? 1.0f / s_next : 0.f;      // formed in fp32 like ATen's quantize_per_tensor
  const long long per_block = 8 * kRowsPerWarp;
  const unsigned blocks = (unsigned)((rows + per_block - 1) / per_block);
  if (E == 64) k_resid_ln<2><<<blocks, 256, 0, s>>>(x, q, sq, gamma, beta, eps, y, yq, inv, rows);
  else k_resid_ln<4><<<blocks, 256, 0, s>>>(x, q, sq, gamma, beta, eps, y, yq, inv, rows);
  count_launch();
  ITA_CUDA_TRY(cudaGetLastError());
  return ITA_OK;
}

// One encoder layer, f32 tokens in -> f32 tokens out (QAT/model.py:100-113).
int layer_forward(ita_mha_handle *h, const ita_layer_params *lp, const float *x, float *y, int64_t B, cudaStream_t s) {
  if (!h->F) {
    set_error("ita_encoder_layer_forward: the handle was created without FFN parameters (F == 0)");
    return ITA_ERR_INVALID;
  }
  if (h->H != 1 && h->head_mode != ITA_HEADS_TORCH) {
    set_error("ita_encoder_layer_forward: multi-head handles must use the torch head convention");
    return ITA_ERR_INVALID;
  }
  if (B <= 0) return ITA_OK;
  const size_t SE = (size_t)h->S * h->E;
  // per frame: xq, yq, x1q, f2q (int8) + x1 (f32)
  const size_t per_frame = 8 * SE;
  int64_t chunk = (int64_t)((size_t(1) << 30) / per_frame);
  chunk = chunk < 1 ? 1 : (chunk > B ? B : chunk);
  if (h->layer_scratch_bytes < per_frame * (size_t)chunk) {
    if (h->layer_scratch) cudaFree(h->layer_scratch);
    h->layer_scratch = nullptr;
    h->layer_scratch_bytes = 0;
    if (cudaMalloc(&h->layer_scratch, per_frame * (size_t)chunk) != cudaSuccess) {
      cudaGetLastError();
      set_error("ita_encoder_layer_forward: cudaMalloc(%zu) failed", per_frame * (size_t)chunk);
      return ITA_ERR_NOMEM;
    }
    h->layer_scratch_bytes = per_frame * (size_t)chunk;
  }
  for (int64_t b0 = 0; b0 < B; b0 += chunk) {
    const int64_t nb = B - b0 < chunk ? B - b0 : chunk;
    const size_t n = (size_t)nb * SE;
    float *x1 = reinterpret_cast<float *>(h->layer_scratch);
    int8_t *xq = h->layer_scratch + 4 * SE * (size_t)chunk, *yq = xq + SE * (size_t)chunk, *x1q = yq + SE * (size_t)chunk,
           *f2q = x1q + SE * (size_t)chunk;
    const float *xb = x + (size_t)b0 * SE;
    int rc = launch_quantize(xb, xq, (int64_t)n, lp->s_attn_in, s);
    if (rc) return rc;
    rc = (h->fused || h->mh) ? fused_forward(h, xq, yq, nb, nullptr, s) : staged_forward(h, xq, yq, nb, nullptr, s);
    if (rc) return rc;
    rc = launch_resid_ln(xb, yq, lp->s_attn_out, lp->ln1_w, lp->ln1_b, lp->ln_eps, x1, x1q, lp->s_ffn_in, nb * h->S, h->E, s);
    if (rc) return rc;
    rc = h->ffn ? ffn_fused_forward(h, x1q, nullptr, f2q, nb, s) : staged_ffn(h, x1q, nullptr, f2q, nb, s);
    if (rc) return rc;
    rc = launch_resid_ln(x1, f2q, lp->s_ffn_out, lp->ln2_w, lp->ln2_b, lp->ln_eps, y + (size_t)b0 * SE, nullptr, 0.f, nb * h->S, h->E, s);
    if (rc) return rc;
  }
  return ITA_OK;
}

}  // namespace ita
