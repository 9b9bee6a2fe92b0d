// tokenizer.cu -- the producer in front of the hot path: OverlapPatchMerging of ITAViTLSTM (reference
// models/ITA/QAT/layers.py:25-45: Conv2d(1, E, 7, stride 2, padding 3) on a 60x90 depth frame -> bilinear resize of the
// 30x45 map to 8x16 (align_corners=False) -> LayerNorm(E) over the 128 tokens), fused with the first encoder layer's quant
// stub, so that a frame enters the GPU as 5400 bytes (samples/inference_udp_FPGA_custom_dispatch/main.cpp:168 scales the
// uint8 image by 1/255) and the int8 token tile [128, E] the attention kernel consumes is produced without a float
// round trip through Python.
//
// One CTA per frame, one thread per token.  The resize is linear and so is the convolution, so a token is
//     bias + W[E x 49] . patch,   patch[ky,kx] = sum over the 4 bilinear taps of weight_tap * image[2*ty + ky - 3, 2*tx + kx - 3]
// (zero padding): 4*49 blends + 49*E FMAs per token instead of 4*49*E (only 512 of the 1350 conv pixels are ever sampled and
// none is shared between tokens).  The summation order differs from conv-then-resize by float rounding only (1e-6 relative;
// tests state the tolerance).  HBM: 5.4 KB in (u8) / 21.6 KB (f32), 128*E*5 bytes out per frame.
#include "common.cuh"

namespace ita {

constexpr int kImgH = 60, kImgW = 90, kConvH = 30, kConvW = 45, kTokH = 8, kTokW = 16, kTaps = 49;

template <int E, bool U8>
__global__ void __launch_bounds__(128) k_tokenize(const void *__restrict__ img, const float *__restrict__ w, const float *__restrict__ cb,
                                                  const float *__restrict__ gamma, const float *__restrict__ beta, float eps,
                                                  float *__restrict__ tok, int8_t *__restrict__ tokq, float inv_q) {
  __shared__ __align__(16) float s_img[kImgH * kImgW];
  __shared__ __align__(16) float s_w[kTaps * E];          // transposed: [tap][channel]
  const long long f = blockIdx.x;
  const int t = threadIdx.x;
  if constexpr (U8) {
    // u8 / 255.0f (a true division, as main.cpp:168) has 256 possible results: a table per CTA instead of 5400 divisions;
    // the frame is read as 16-byte words (5400 = 337 * 16 + 8)
    __shared__ float s_lut[256];
    for (int i = t; i < 256; i += 128) s_lut[i] = __fdiv_rn((float)i, 255.0f);
    __syncthreads();
    const uint8_t *src = reinterpret_cast<const uint8_t *>(img) + f * (kImgH * kImgW);
    if ((reinterpret_cast<uintptr_t>(src) & 7) == 0) {
      const uint2 *s8 = reinterpret_cast<const uint2 *>(src);
      for (int i = t; i < kImgH * kImgW / 8; i += 128) {
        const uint2 v = s8[i];
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          s_img[i * 8 + b] = s_lut[(v.x >> (8 * b)) & 0xFF];
          s_img[i * 8 + 4 + b] = s_lut[(v.y >> (8 * b)) & 0xFF];
        }
      }
    } else {
      for (int i = t; i < kImgH * kImgW; i += 128) s_img[i] = s_lut[src[i]];
    }
  } else {
    const float *src = reinterpret_cast<const float *>(img) + f * (kImgH * kImgW);
    for (int i = t; i < kImgH * kImgW; i += 128) s_img[i] = src[i];
  }
  for (int i = t; i < kTaps * E; i += 128) s_w[i] = w[(i % E) * kTaps + i / E];          // w is [E][49]; s_w [tap][channel]
  __syncthreads();

  // ---- bilinear source coordinates (ATen area_pixel_compute_source_index, align_corners=False)
  const int oy = t / kTokW, ox = t % kTokW;
  const float sy = fmaxf(((float)oy + 0.5f) * ((float)kConvH / kTokH) - 0.5f, 0.f), sx = fmaxf(((float)ox + 0.5f) * ((float)kConvW / kTokW) - 0.5f, 0.f);
  const int y0 = (int)sy, x0 = (int)sx;
  const int y1 = y0 + (y0 < kConvH - 1 ? 1 : 0), x1 = x0 + (x0 < kConvW - 1 ? 1 : 0);
  const float ly1 = sy - (float)y0, lx1 = sx - (float)x0, ly0 = 1.f - ly1, lx0 = 1.f - lx1;
  const float wt[4] = {ly0 * lx0, ly0 * lx1, ly1 * lx0, ly1 * lx1};
  const int ty[4] = {y0, y0, y1, y1}, tx[4] = {x0, x1, x0, x1};
  // ---- the blended 7x7 patch of this token
  float patch[kTaps];
#pragma unroll
  for (int k = 0; k < kTaps; ++k) patch[k] = 0.f;
#pragma unroll
  for (int tp = 0; tp < 4; ++tp) {
    const int by = 2 * ty[tp] - 3, bx = 2 * tx[tp] - 3;
#pragma unroll
    for (int ky = 0; ky < 7; ++ky) {
      const int iy = by + ky;
      const bool rowok = iy >= 0 && iy < kImgH;
#pragma unroll
      for (int kx = 0; kx < 7; ++kx) {
        const int ix = bx + kx;
        const float v = (rowok && ix >= 0 && ix < kImgW) ? s_img[iy * kImgW + ix] : 0.f;
        patch[ky * 7 + kx] = fmaf(wt[tp], v, patch[ky * 7 + kx]);
      }
    }
  }
  // ---- W . patch + bias
  float acc[E];
#pragma unroll
  for (int c = 0; c < E; ++c) acc[c] = cb[c];
#pragma unroll
  for (int k = 0; k < kTaps; ++k) {
    const float4 *wr = reinterpret_cast<const float4 *>(s_w + k * E);
#pragma unroll
    for (int c4 = 0; c4 < E / 4; ++c4) {
      const float4 ww = wr[c4];                    // same address in every thread: a shared-memory broadcast
      acc[c4 * 4 + 0] = fmaf(ww.x, patch[k], acc[c4 * 4 + 0]);
      acc[c4 * 4 + 1] = fmaf(ww.y, patch[k], acc[c4 * 4 + 1]);
      acc[c4 * 4 + 2] = fmaf(ww.z, patch[k], acc[c4 * 4 + 2]);
      acc[c4 * 4 + 3] = fmaf(ww.w, patch[k], acc[c4 * 4 + 3]);
    }
  }
  // ---- LayerNorm over the E channels of the token (biased variance), then the quant stub
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < E; ++c) s += acc[c];
  const float mean = s * (1.0f / E);
  float ss = 0.f;
#pragma unroll
  for (int c = 0; c < E; ++c) ss += (acc[c] - mean) * (acc[c] - mean);
  const float rstd = 1.0f / sqrtf(ss * (1.0f / E) + eps);
  float *to = tok ? tok + (f * (kTokH * kTokW) + t) * (long long)E : nullptr;
  int8_t *tq = tokq ? tokq + (f * (kTokH * kTokW) + t) * (long long)E : nullptr;
#pragma unroll
  for (int c4 = 0; c4 < E / 4; ++c4) {
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = (acc[c4 * 4 + i] - mean) * rstd * gamma[c4 * 4 + i] + beta[c4 * 4 + i];
    if (to) reinterpret_cast<float4 *>(to)[c4] = make_float4(o[0], o[1], o[2], o[3]);
    if (tq) {
      char4 q;
      q.x = (int8_t)(int)fminf(fmaxf(nearbyintf(o[0] * inv_q), -128.f), 127.f);
      q.y = (int8_t)(int)fminf(fmaxf(nearbyintf(o[1] * inv_q), -128.f), 127.f);
      q.z = (int8_t)(int)fminf(fmaxf(nearbyintf(o[2] * inv_q), -128.f), 127.f);
      q.w = (int8_t)(int)fminf(fmaxf(nearbyintf(o[3] * inv_q), -128.f), 127.f);
      reinterpret_cast<char4 *>(tq)[c4] = q;
    }
  }
}

int launch_tokenizer(const void *img, int is_u8, int64_t B, int E, const float *w, const float *cb, const float *gamma, const float *beta,
                     float eps, float *tok, int8_t *tokq, float q_scale, cudaStream_t s) {
  if (B <= 0) return ITA_OK;
  if (E != 64 && E != 128) {
    set_error("tokenizer kernel is instantiated for E in {64, 128}, got %d", E);
    return ITA_ERR_UNSUPPORTED;
  }
  if (tokq && !(q_scale > 0.f)) {
    set_error("tokenizer: the quantised output needs a positive scale");
    return ITA_ERR_INVALID;
  }
  if (B > 0x7FFFFFFFLL) {
    set_error("tokenizer: batch too large for one launch");
    return ITA_ERR_INVALID;
  }
  const float inv = tokq ? 1.0f / q_scale : 0.f;          // formed in fp32 like ATen's quantize_per_tensor
  const unsigned grid = (unsigned)B;
  if (E == 64) {
    if (is_u8) k_tokenize<64, true><<<grid, 128, 0, s>>>(img, w, cb, gamma, beta, eps, tok, tokq, inv);
    else k_tokenize<64, false><<<grid, 128, 0, s>>>(img, w, cb, gamma, beta, eps, tok, tokq, inv);
  } else {
    if (is_u8) k_tokenize<128, true><<<grid, 128, 0, s>>>(img, w, cb, gamma, beta, eps, tok, tokq, inv);
    else k_tokenize<128, false><<<grid, 128, 0, s>>>(img, w, cb, gamma, beta, eps, tok, tokq, inv);
  }
  count_launch();
  ITA_CUDA_TRY(cudaGetLastError());
  return ITA_OK;
}


// ---------------------------------------------------------------------------------------------------- feature fusion
// The consumer behind the encoder: ITALSTMNetVIT_QAT.forward's fusion stage (reference
// models/ITA_single_layer_upsample_shuffle/QAT/model.py:116-121): tokens [128, E] viewed as a [E, 8, 16] map ->
//   PixelShuffle(2)                                    [E/4, 16, 32]
//   Upsample(16x32, bilinear, align_corners=True)      [E,   16, 32]
//   cat -> Conv2d(E/4 + E -> 9, 3x3, padding 1) -> flatten -> [9 * 16 * 32] (the decoder's input)
// as one kernel: one CTA per frame, the token tile staged in shared memory, the fused map built 16 channels at a time in
// shared memory (never in HBM: torch writes and re-reads 160 KB per frame for the cat alone), each thread owning two
// horizontally adjacent output pixels x 9 output channels in registers.  fp32 throughout (cuDNN's default for this conv is
// TF32); summation order differs from the reference by float rounding only.
constexpr int kFuH = 16, kFuW = 32, kFuOut = 9, kFuChunk = 16, kFuWPad = 12, kFuTokPad = 129;

template <int E>
__global__ void __launch_bounds__(256, 2) k_fuse_conv(const float *__restrict__ x, const float *__restrict__ w, const float *__restrict__ cb,
                                                      float *__restrict__ out) {
  constexpr int CIN = E / 4 + E, NCHUNK = CIN / kFuChunk;
  static_assert(CIN % kFuChunk == 0 && (E / 4) % kFuChunk == 0, "channel chunks must not straddle the shuffle / upsample boundary");
  extern __shared__ __align__(16) float sm[];
  float *s_tok = sm;                                        // [E][kFuTokPad]: channel-major (the [E, 8, 16] map), padded against bank conflicts
  float *s_map = s_tok + E * kFuTokPad;                     // [16 channels][18 rows][34 cols] with a zero halo
  float *s_w = s_map + kFuChunk * 18 * 34;                  // [CIN][9 taps][12] (9 outputs, padded)
  const long long f = blockIdx.x;
  const int t = threadIdx.x;
  {
    const float *src = x + f * 128 * E;                     // [token][channel] -> [channel][token]
    for (int i = t; i < 128 * E; i += 256) s_tok[(i % E) * kFuTokPad + i / E] = src[i];
    for (int i = t; i < CIN * 9 * kFuWPad; i += 256) {
      const int o = i % kFuWPad, tap = (i / kFuWPad) % 9, c = i / (kFuWPad * 9);
      s_w[i] = o < kFuOut ? w[(o * CIN + c) * 9 + tap] : 0.f;            // w is [9][CIN][3][3]
    }
    for (int i = t; i < kFuChunk * 18 * 34; i += 256) s_map[i] = 0.f;    // the halo stays zero, the interior is overwritten per chunk
  }
  // this thread's two output pixels
  const int Y = t / 16, X0 = (t % 16) * 2;
  float acc[2][kFuOut];
#pragma unroll
  for (int o = 0; o < kFuOut; ++o) acc[0][o] = acc[1][o] = cb[o];
  // bilinear source coordinates of the upsample branch (ATen: scale = (in - 1) / (out - 1), align_corners=True)
  const float sc_y = (float)(8 - 1) / (float)(kFuH - 1), sc_x = (float)(16 - 1) / (float)(kFuW - 1);
  for (int ch = 0; ch < NCHUNK; ++ch) {
    __syncthreads();                                        // staging done / the previous chunk's readers are done
    // ---- build 16 channels of the fused map: 16 * 512 elements, 32 per thread (consecutive threads -> consecutive X)
    for (int i = t; i < kFuChunk * kFuH * kFuW; i += 256) {
      const int px = i % kFuW, py = (i / kFuW) % kFuH, cl = i / (kFuW * kFuH);
      const int c = ch * kFuChunk + cl;
      float v;
      if (c < E / 4) {                                      // PixelShuffle: out[c, 2h+i, 2w+j] = in[4c + 2i + j, h, w]
        const int cs = c * 4 + (py & 1) * 2 + (px & 1);
        v = s_tok[cs * kFuTokPad + (py >> 1) * 16 + (px >> 1)];
      } else {
        const int cu = c - E / 4;
        const float sy = sc_y * (float)py, sx = sc_x * (float)px;
        const int y0 = (int)sy, x0 = (int)sx;
        const int y1 = y0 + (y0 < 7 ? 1 : 0), x1 = x0 + (x0 < 15 ? 1 : 0);
        const float ly1 = sy - (float)y0, lx1 = sx - (float)x0, ly0 = 1.f - ly1, lx0 = 1.f - lx1;
        const float *tc = s_tok + cu * kFuTokPad;
        const float v00 = tc[y0 * 16 + x0], v01 = tc[y0 * 16 + x1], v10 = tc[y1 * 16 + x0], v11 = tc[y1 * 16 + x1];
        v = ly0 * (lx0 * v00 + lx1 * v01) + ly1 * (lx0 * v10 + lx1 * v11);
      }
      s_map[(cl * 18 + py + 1) * 34 + px + 1] = v;
    }
    __syncthreads();
    // ---- 3x3 convolution over the chunk's channels
#pragma unroll 2
    for (int cl = 0; cl < kFuChunk; ++cl) {
      const float *mp = s_map + (cl * 18 + Y) * 34 + X0;    // top-left of the 3 x 4 input window (halo offset included)
      float in[3][4];
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) in[r][q] = mp[r * 34 + q];
      const float4 *wp = reinterpret_cast<const float4 *>(s_w + (ch * kFuChunk + cl) * 9 * kFuWPad);
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const float4 w0 = wp[tap * 3], w1 = wp[tap * 3 + 1], w2 = wp[tap * 3 + 2];
        const float wv[kFuWPad] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y, w2.z, w2.w};
        const float a0 = in[tap / 3][tap % 3], a1 = in[tap / 3][tap % 3 + 1];
#pragma unroll
        for (int o = 0; o < kFuOut; ++o) {
          acc[0][o] = fmaf(wv[o], a0, acc[0][o]);
          acc[1][o] = fmaf(wv[o], a1, acc[1][o]);
        }
      }
    }
  }
  float *op = out + f * (kFuOut * kFuH * kFuW) + Y * kFuW + X0;
#pragma unroll
  for (int o = 0; o < kFuOut; ++o) *reinterpret_cast<float2 *>(op + o * kFuH * kFuW) = make_float2(acc[0][o], acc[1][o]);
}

int launch_fuse_conv(const float *x, int64_t B, int E, const float *w, const float *cb, float *out, cudaStream_t s) {
  if (B <= 0) return ITA_OK;
  if (E != 64 && E != 128) {
    set_error("fusion kernel is instantiated for E in {64, 128}, got %d", E);
    return ITA_ERR_UNSUPPORTED;
  }
  if (B > 0x7FFFFFFFLL) {
    set_error("fusion kernel: batch too large for one launch");
    return ITA_ERR_INVALID;
  }
  const int cin = E / 4 + E;
  const size_t smem = sizeof(float) * ((size_t)E * kFuTokPad + (size_t)kFuChunk * 18 * 34 + (size_t)cin * 9 * kFuWPad);
  static bool attr_set[2] = {false, false};
  const int which = E == 64 ? 0 : 1;
  if (!attr_set[which]) {
    cudaError_t e = E == 64 ? cudaFuncSetAttribute((const void *)k_fuse_conv<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                            : cudaFuncSetAttribute((const void *)k_fuse_conv<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
      cudaGetLastError();
      set_error("fusion kernel: cudaFuncSetAttribute(%zu B shared memory) failed: %s", smem, cudaGetErrorString(e));
      return ITA_ERR_CUDA;
    }
    attr_set[which] = true;
  }
  if (E == 64) k_fuse_conv<64><<<(unsigned)B, 256, smem, s>>>(x, w, cb, out);
  else k_fuse_conv<128><<<(unsigned)B, 256, smem, s>>>(x, w, cb, out);
  count_launch();
  ITA_CUDA_TRY(cudaGetLastError());
  return ITA_OK;
}

}  // namespace ita
