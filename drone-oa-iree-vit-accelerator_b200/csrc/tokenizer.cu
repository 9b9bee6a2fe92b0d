// tokenizer.cu -- the producer in front of the hot path: OverlapPatchMerging of ITAViTLSTM (reference
// models/ITA/QAT/layers.py:25-45: Conv2d(1, E, 7, stride 2, padding 3) on a 60x90 depth frame -> bilinear resize of the
// 30x45 map to 8x16 (align_corners=False) -> LayerNorm(E) over the 128 tokens), fused with the first encoder layer's quant
// stub, so that a frame enters the GPU as 5400 bytes (samples/inference_udp_FPGA_custom_dispatch/main.cpp:168 scales the
// uint8 image by 1/255) and the int8 token tile [128, E] the attention kernel consumes is produced without a float
// round trip through Python.
//
// One CTA per frame; a thread blends one token's patch, then owns 4 tokens x E/4 channels of the product.  The resize is linear
// and so is the convolution, so a token is
//     bias + W[E x 49] . patch,   patch[ky,kx] = sum over the 4 bilinear taps of weight_tap * image[2*ty + ky - 3, 2*tx + kx - 3]
// (zero padding): 4*49 blends + 49*E FMAs per token instead of 4*49*E (only 512 of the 1350 conv pixels are ever sampled and
// none is shared between tokens).  The summation order differs from conv-then-resize by float rounding only (1e-6 relative;
// tests state the tolerance).  HBM: 5.4 KB in (u8) / 21.6 KB (f32), 128*E*5 bytes out per frame.
#include "common.cuh"

namespace ita {

constexpr int kImgH = 60, kImgW = 90, kConvH = 30, kConvW = 45, kTokH = 8, kTokW = 16, kTaps = 49;
constexpr int kTokBuf = kImgH * kImgW > kTaps * 128 ? kImgH * kImgW : kTaps * 128;   // floats

template <int E, bool U8>
__global__ void __launch_bounds__(128) k_tokenize(const void *__restrict__ img, const float *__restrict__ w, const float *__restrict__ cb,
                                                  const float *__restrict__ gamma, const float *__restrict__ beta, float eps,
                                                  float *__restrict__ tok, int8_t *__restrict__ tokq, float inv_q) {
  extern __shared__ __align__(16) float tk_smem[];
  float *s_img = tk_smem;                                 // the frame (5400 floats), later the blended patches [49][128]
  float *s_w = tk_smem + kTokBuf;                         // transposed: [tap][channel]
  float *s_lut = s_w + kTaps * E;                         // u8 -> f32(u8) / 255 (U8 only)
  const long long f = blockIdx.x;
  const int t = threadIdx.x;
  if constexpr (U8) {
    // u8 / 255.0f (a true division, as main.cpp:168) has 256 possible results: a table per CTA instead of 5400 divisions;
    // the frame is read as 16-byte words (5400 = 337 * 16 + 8)
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
  // ---- the blended patches go through shared memory ([tap][token]), so that the [128 x 49] . [49 x E] product can be register
  //      tiled: a thread owns 4 tokens x E/4 channels and every shared-memory load feeds 8-16 FMAs (one token x all channels per
  //      thread, the first version, was bound by the weight broadcasts: 4 FMAs per LDS.128)
  float *s_patch = s_img;                          // the image is dead once every thread has blended its patch
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kTaps; ++k) s_patch[k * 128 + t] = patch[k];
  __syncthreads();
  constexpr int CG = E / 4;                        // channels per thread; lane = 4 * token group + channel group
  const int tg = t >> 2, cg = t & 3;               // tokens 4 tg .. 4 tg + 3, channels cg * CG .. + CG
  float acc[4][CG];
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int c = 0; c < CG; ++c) acc[q][c] = cb[cg * CG + c];
#pragma unroll 7
  for (int k = 0; k < kTaps; ++k) {
    const float4 pv = *reinterpret_cast<const float4 *>(s_patch + k * 128 + tg * 4);
    const float pq[4] = {pv.x, pv.y, pv.z, pv.w};
    const float4 *wr = reinterpret_cast<const float4 *>(s_w + k * E + cg * CG);
#pragma unroll
    for (int c4 = 0; c4 < CG / 4; ++c4) {
      const float4 ww = wr[c4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        acc[q][c4 * 4 + 0] = fmaf(ww.x, pq[q], acc[q][c4 * 4 + 0]);
        acc[q][c4 * 4 + 1] = fmaf(ww.y, pq[q], acc[q][c4 * 4 + 1]);
        acc[q][c4 * 4 + 2] = fmaf(ww.z, pq[q], acc[q][c4 * 4 + 2]);
        acc[q][c4 * 4 + 3] = fmaf(ww.w, pq[q], acc[q][c4 * 4 + 3]);
      }
    }
  }
  // ---- LayerNorm over the E channels of a token (biased variance; the four channel groups of a token are adjacent lanes),
  //      then the quant stub
  float gm[CG], bt[CG];
#pragma unroll
  for (int c = 0; c < CG; ++c) {
    gm[c] = gamma[cg * CG + c];
    bt[c] = beta[cg * CG + c];
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float sm_ = 0.f;
#pragma unroll
    for (int c = 0; c < CG; ++c) sm_ += acc[q][c];
    sm_ += __shfl_xor_sync(0xffffffffu, sm_, 1);
    sm_ += __shfl_xor_sync(0xffffffffu, sm_, 2);
    const float mean = sm_ * (1.0f / E);
    float ss = 0.f;
#pragma unroll
    for (int c = 0; c < CG; ++c) ss += (acc[q][c] - mean) * (acc[q][c] - mean);
    ss += __shfl_xor_sync(0xffffffffu, ss, 1);
    ss += __shfl_xor_sync(0xffffffffu, ss, 2);
    const float rstd = 1.0f / sqrtf(ss * (1.0f / E) + eps);
    const long long row = f * (kTokH * kTokW) + tg * 4 + q;
    float *to = tok ? tok + row * (long long)E + cg * CG : nullptr;
    int8_t *tq = tokq ? tokq + row * (long long)E + cg * CG : nullptr;
#pragma unroll
    for (int c4 = 0; c4 < CG / 4; ++c4) {
      float o[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) o[i] = (acc[q][c4 * 4 + i] - mean) * rstd * gm[c4 * 4 + i] + bt[c4 * 4 + i];
      if (to) reinterpret_cast<float4 *>(to)[c4] = make_float4(o[0], o[1], o[2], o[3]);
      if (tq) {
        char4 qv;
        qv.x = (int8_t)(int)fminf(fmaxf(nearbyintf(o[0] * inv_q), -128.f), 127.f);
        qv.y = (int8_t)(int)fminf(fmaxf(nearbyintf(o[1] * inv_q), -128.f), 127.f);
        qv.z = (int8_t)(int)fminf(fmaxf(nearbyintf(o[2] * inv_q), -128.f), 127.f);
        qv.w = (int8_t)(int)fminf(fmaxf(nearbyintf(o[3] * inv_q), -128.f), 127.f);
        reinterpret_cast<char4 *>(tq)[c4] = qv;
      }
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
  const size_t smem = sizeof(float) * ((size_t)kTokBuf + (size_t)kTaps * E + 256);
  if (smem > 48 * 1024) {
    static bool attr_done[64] = {};                      // function attributes are per device
    int dev = 0;
    ITA_CUDA_TRY(cudaGetDevice(&dev));
    bool &attr_set = attr_done[dev & 63];
    if (!attr_set) {
      cudaError_t e1 = cudaFuncSetAttribute((const void *)k_tokenize<128, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaError_t e2 = cudaFuncSetAttribute((const void *)k_tokenize<128, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e1 != cudaSuccess || e2 != cudaSuccess) {
        cudaGetLastError();
        set_error("tokenizer: cudaFuncSetAttribute(%zu B shared memory) failed", smem);
        return ITA_ERR_CUDA;
      }
      attr_set = true;
    }
  }
  if (E == 64) {
    if (is_u8) k_tokenize<64, true><<<grid, 128, smem, s>>>(img, w, cb, gamma, beta, eps, tok, tokq, inv);
    else k_tokenize<64, false><<<grid, 128, smem, s>>>(img, w, cb, gamma, beta, eps, tok, tokq, inv);
  } else {
    if (is_u8) k_tokenize<128, true><<<grid, 128, smem, s>>>(img, w, cb, gamma, beta, eps, tok, tokq, inv);
    else k_tokenize<128, false><<<grid, 128, smem, s>>>(img, w, cb, gamma, beta, eps, tok, tokq, inv);
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
// shared memory (never in HBM: torch writes and re-reads 160 KB per frame for the cat alone), each thread owning four
// horizontally adjacent output pixels x 9 output channels in registers for half of the channels (the first version, two pixels
// per thread over all channels, was bound by the shared-memory weight broadcasts, not by the FMA pipe).  fp32 throughout (cuDNN's default for this conv is
// TF32); summation order differs from the reference by float rounding only.
constexpr int kFuH = 16, kFuW = 32, kFuOut = 9, kFuChunk = 16, kFuWPad = 12, kFuTokPad = 129, kFuMapW = 36, kFuMapCh = 18 * kFuMapW;

template <int E>
__global__ void __launch_bounds__(256, 2) k_fuse_conv(const float *__restrict__ x, const float *__restrict__ w, const float *__restrict__ cb,
                                                      float *__restrict__ out) {
  constexpr int CIN = E / 4 + E, NCHUNK = CIN / kFuChunk;
  static_assert(CIN % kFuChunk == 0 && (E / 4) % kFuChunk == 0, "channel chunks must not straddle the shuffle / upsample boundary");
  extern __shared__ __align__(16) float sm[];
  float *s_tok = sm;                                        // [E][kFuTokPad]: channel-major (the [E, 8, 16] map), padded against bank conflicts
  float *s_map = s_tok + E * kFuTokPad;                     // [16 channels][18 rows][36 cols]: zero halo (row 0 / 17, col 0 / 33), rows 16-byte aligned
  float *s_w = s_map + kFuChunk * kFuMapCh;                  // [CIN][9 taps][12] (9 outputs, padded)
  const long long f = blockIdx.x;
  const int t = threadIdx.x;
  {
    const float *src = x + f * 128 * E;                     // [token][channel] -> [channel][token]
    for (int i = t; i < 128 * E; i += 256) s_tok[(i % E) * kFuTokPad + i / E] = src[i];
    for (int i = t; i < CIN * 9 * kFuWPad; i += 256) {
      const int o = i % kFuWPad, tap = (i / kFuWPad) % 9, c = i / (kFuWPad * 9);
      s_w[i] = o < kFuOut ? w[(o * CIN + c) * 9 + tap] : 0.f;            // w is [9][CIN][3][3]
    }
    for (int i = t; i < kFuChunk * kFuMapCh; i += 256) s_map[i] = 0.f;    // the halo stays zero, the interior is overwritten per chunk
  }
  // Convolution ownership: the two halves of the CTA split the channels of every chunk (half 0: channels 0-7, half 1: 8-15);
  // inside a half, thread tt owns the four horizontally adjacent output pixels (Y, X0 .. X0 + 3) x 9 output channels, so one
  // weight vector (3 x LDS.128, a broadcast) feeds 36 FMAs and the 3 x 6 input window of a channel is 6 vector loads.
  const int half = t >> 7, tt = t & 127;
  const int Y = tt >> 3, X0 = (tt & 7) * 4;
  float acc[4][kFuOut];
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int o = 0; o < kFuOut; ++o) acc[q][o] = 0.f;
  // ---- build-phase ownership: thread t fills pixels (t / 32, t % 32) and (t / 32 + 8, t % 32) of every channel of a chunk; their
  //      pixel-shuffle sources and bilinear taps (ATen: scale = (in - 1) / (out - 1), align_corners=True) do not depend on the
  //      channel, so indices and the four tap weights are formed once
  const float sc_y = (float)(8 - 1) / (float)(kFuH - 1), sc_x = (float)(16 - 1) / (float)(kFuW - 1);
  int b_map[2], b_sh_tok[2], b_sh_sub[2], b_o00[2], b_o01[2], b_o10[2], b_o11[2];
  float b_w00[2], b_w01[2], b_w10[2], b_w11[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int py = t / kFuW + 8 * k, px = t % kFuW;
    b_map[k] = (py + 1) * kFuMapW + px + 1;
    b_sh_tok[k] = (py >> 1) * 16 + (px >> 1);
    b_sh_sub[k] = (py & 1) * 2 + (px & 1);
    const float sy = sc_y * (float)py, sx = sc_x * (float)px;
    const int y0 = (int)sy, x0 = (int)sx;
    const int y1 = y0 + (y0 < 7 ? 1 : 0), x1 = x0 + (x0 < 15 ? 1 : 0);
    const float ly1 = sy - (float)y0, lx1 = sx - (float)x0, ly0 = 1.f - ly1, lx0 = 1.f - lx1;
    b_o00[k] = y0 * 16 + x0; b_o01[k] = y0 * 16 + x1; b_o10[k] = y1 * 16 + x0; b_o11[k] = y1 * 16 + x1;
    b_w00[k] = ly0 * lx0; b_w01[k] = ly0 * lx1; b_w10[k] = ly1 * lx0; b_w11[k] = ly1 * lx1;
  }
  for (int ch = 0; ch < NCHUNK; ++ch) {
    __syncthreads();                                        // staging done / the previous chunk's readers are done
    // ---- build 16 channels of the fused map (consecutive threads -> consecutive X: conflict-free stores)
#pragma unroll 4
    for (int cl = 0; cl < kFuChunk; ++cl) {
      const int c = ch * kFuChunk + cl;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        float v;
        if (c < E / 4) {                                    // PixelShuffle: out[c, 2h+i, 2w+j] = in[4c + 2i + j, h, w]
          v = s_tok[(c * 4 + b_sh_sub[k]) * kFuTokPad + b_sh_tok[k]];
        } else {
          const float *tc = s_tok + (c - E / 4) * kFuTokPad;
          v = b_w00[k] * tc[b_o00[k]] + b_w01[k] * tc[b_o01[k]] + b_w10[k] * tc[b_o10[k]] + b_w11[k] * tc[b_o11[k]];
        }
        s_map[cl * kFuMapCh + b_map[k]] = v;
      }
    }
    __syncthreads();
    // ---- 3x3 convolution over this half's 8 channels of the chunk
#pragma unroll 2
    for (int ci = 0; ci < kFuChunk / 2; ++ci) {
      const int cl = half * (kFuChunk / 2) + ci;
      const float *mp = s_map + cl * kFuMapCh + Y * kFuMapW + X0;     // top-left of the 3 x 6 input window (halo offset included)
      float in[3][6];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float4 a = *reinterpret_cast<const float4 *>(mp + r * kFuMapW);
        const float2 b2 = *reinterpret_cast<const float2 *>(mp + r * kFuMapW + 4);
        in[r][0] = a.x; in[r][1] = a.y; in[r][2] = a.z; in[r][3] = a.w; in[r][4] = b2.x; in[r][5] = b2.y;
      }
      const float4 *wp = reinterpret_cast<const float4 *>(s_w + (ch * kFuChunk + cl) * 9 * kFuWPad);
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const float4 w0 = wp[tap * 3], w1 = wp[tap * 3 + 1], w2 = wp[tap * 3 + 2];
        const float wv[kFuWPad] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y, w2.z, w2.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float a = in[tap / 3][tap % 3 + q];
#pragma unroll
          for (int o = 0; o < kFuOut; ++o) acc[q][o] = fmaf(wv[o], a, acc[q][o]);
        }
      }
    }
  }
  // ---- add the two channel halves (through the map buffer, dead now) and the bias, store
  __syncthreads();
  float *red = s_map;                                       // [128 threads][36]
  if (half == 1) {
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int o = 0; o < kFuOut; ++o) red[(q * kFuOut + o) * 128 + tt] = acc[q][o];
  }
  __syncthreads();
  if (half == 0) {
    float *op = out + f * (kFuOut * kFuH * kFuW) + Y * kFuW + X0;
#pragma unroll
    for (int o = 0; o < kFuOut; ++o) {
      const float bo = cb[o];
      float4 v;
      v.x = acc[0][o] + red[(0 * kFuOut + o) * 128 + tt] + bo;
      v.y = acc[1][o] + red[(1 * kFuOut + o) * 128 + tt] + bo;
      v.z = acc[2][o] + red[(2 * kFuOut + o) * 128 + tt] + bo;
      v.w = acc[3][o] + red[(3 * kFuOut + o) * 128 + tt] + bo;
      *reinterpret_cast<float4 *>(op + o * kFuH * kFuW) = v;
    }
  }
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
  const size_t smem = sizeof(float) * ((size_t)E * kFuTokPad + (size_t)kFuChunk * kFuMapCh + (size_t)cin * 9 * kFuWPad);
  static bool attr_done[64][2] = {};                     // function attributes are per device
  int dev = 0;
  ITA_CUDA_TRY(cudaGetDevice(&dev));
  bool *attr_set = attr_done[dev & 63];
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

// ---------------------------------------------------------------------------------------------------- LSTM cell
// The pointwise half of one LSTM time step (nn.LSTM gate order i, f, g, o; reference model.py:68,128): the gate pre-activations
// [B, 4*Hd] come from two cuBLAS GEMMs, this kernel turns them and c_prev into (h, c) in one pass instead of ~8 elementwise
// launches per layer.
__global__ void __launch_bounds__(256) k_lstm_cell(const float *__restrict__ gates, const float *__restrict__ c_prev, float *__restrict__ h,
                                                   float *__restrict__ c, long long n, int Hd) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long b = i / Hd;
  const int j = (int)(i - b * Hd);
  const float *g = gates + b * 4 * Hd;
  const float gi = 1.0f / (1.0f + expf(-g[j])), gf = 1.0f / (1.0f + expf(-g[Hd + j])), gg = tanhf(g[2 * Hd + j]),
              go = 1.0f / (1.0f + expf(-g[3 * Hd + j]));
  const float cn = gf * c_prev[i] + gi * gg;
  c[i] = cn;
  h[i] = go * tanhf(cn);
}

int launch_lstm_cell(const float *gates, const float *c_prev, float *h, float *c, int64_t B, int Hd, cudaStream_t s) {
  if (B <= 0) return ITA_OK;
  if (Hd <= 0) {
    set_error("LSTM cell: bad hidden size %d", Hd);
    return ITA_ERR_INVALID;
  }
  const long long n = (long long)B * Hd;
  k_lstm_cell<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(gates, c_prev, h, c, n, Hd);
  count_launch();
  ITA_CUDA_TRY(cudaGetLastError());
  return ITA_OK;
}

}  // namespace ita
