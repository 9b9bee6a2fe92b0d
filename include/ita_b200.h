/* ita_b200.h -- C-ABI of libita_b200.so: the B200 (sm_100a) implementation of the ITA int8 multi-head attention.
 *
 * This is the drop-in boundary for ONE hot path of OpenHardware-Initiative/Drone-OA-IREE-ViT-Accelerator: the ITA
 * attention block that ITAViTLSTM offloads.  Plain C types only (no torch / CUDA types in any signature); every
 * function returns 0 on success and non-zero on error (the IREE dispatch convention: a non-zero return from the
 * enclosing dispatch function means failure, third_party/iree/runtime/src/iree/hal/local/executable_library.h:362-371).
 * ita_last_error() returns a thread-local message for the last failing call.
 *
 * Which reference interface each entry point replaces (paths relative to the reference repo):
 *
 *   ita_mha_create / ita_mha_destroy
 *       the golden-model constructor  Transformer(ITA_N, path, activation, S,P,E,F,H, Q,K,V, Wq,Wk,Wv,Wo, Bq,Bk,Bv,Bo,
 *       FF_in,Wff,Wff2,Bff,Bff2, quant_params=...)      tests/export_and_validation_W_B.py:518 (tensor layouts :479-491,
 *       quant_params dict :247-305) and the parameter-carrying half of ITASelfAttention_QAT.__init__
 *       models/ITA/QAT/layers.py:79-99.
 *   ita_mha_forward
 *       ITASelfAttention_QAT.forward between the quant stub and the dequant stub   models/ITA/QAT/layers.py:103-126,
 *       == step1_Qp..step6_O(/step7_Osum) of the golden model  tests/export_and_validation_W_B.py:520-542,
 *       == what the IREE custom dispatch  void ITASelfAttention_workgroup(const T* input, T* output)  is meant to
 *       compute   samples/inference_udp_FPGA_custom_dispatch/plugin/ITA_dispatch.c:17-27 (decl ITA_spec.mlir:30-33):
 *       caller-owned in/out buffers of shape [B,S,E], no aliasing.
 *   ita_mha_forward_host, ita_mha_forward_host_async
 *       the same call for a caller that owns HOST buffers (what main.cpp hands to iree_runtime_call_invoke,
 *       samples/inference_udp_FPGA_custom_dispatch/main.cpp:168-195): H2D copy, kernels, D2H copy, synchronous.
 *   ita_mha_forward_stages
 *       the per-stage attributes the reference harness reads from the golden model: Qp_requant, Kp_requant, Vp_requant,
 *       A_requant, A_partial_softmax, O_soft_requant, Out_soft_requant   tests/export_and_validation_W_B.py:520-542.
 *   ita_mha_step
 *       one golden-model step (step1_Qp ... step6_O) on caller-provided stage inputs, so that the harness' pattern
 *       "overwrite A_partial_softmax, then call step5_AV" (tests/export_and_validation_W_B.py:531-540) keeps working.
 *   ita_ffn_forward
 *       ITAFeedForward_QAT.forward  models/ITA/QAT/layers.py:61-75 == golden feedforward_layer
 *       tests/export_and_validation_W_B.py:544-553.
 *   ita_residual_layernorm, ita_encoder_layer_forward
 *       the float ops between the int8 blocks and the encoder-layer loop body
 *       models/ITA_single_layer_upsample_shuffle/QAT/model.py:100-113.
 *   ita_tokenizer_forward
 *       OverlapPatchMerging.forward  models/ITA/QAT/layers.py:25-45 (conv 7x7 s2 + bilinear to 8x16 + LayerNorm) fused with
 *       the 1/255 image scaling of samples/inference_udp_FPGA_custom_dispatch/main.cpp:168 and the first quant stub.
 *   ita_fusion_conv_forward
 *       pixel-shuffle | upsample -> cat -> conv 3x3   models/ITA_single_layer_upsample_shuffle/QAT/model.py:116-121.
 *   ita_quantize_f32 / ita_dequantize_s8
 *       the QuantStub / DeQuantStub at the float boundary of the operator   models/ITA/QAT/layers.py:103,126.
 *
 * Threading: a handle may be used from one host thread at a time; distinct handles are independent.  The *_host entry
 * points stage through ONE per-handle device buffer: at most one ita_mha_forward_host[_async] call per handle may be in
 * flight (use one handle per stream to overlap calls, as bench.py does); the first call of a larger batch reallocates
 * that buffer, which synchronises the device.  Device-side time-outs of the fused kernels (every in-kernel wait is
 * bounded) are reported by ita_mha_sync_check and by the synchronous entry points, not by the asynchronous ones.  All device work
 * of a call is enqueued on the CUDA stream passed in (a cudaStream_t cast to void*; NULL = the legacy default stream)
 * and, except for the *_host entry points, the call returns without synchronising.
 */
#ifndef ITA_B200_H_
#define ITA_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ITA_B200_ABI_VERSION 1

#if defined(__GNUC__)
#define ITA_API __attribute__((visibility("default")))
#else
#define ITA_API
#endif

/* requant rounding */
enum { ITA_ROUND_HALF_UP = 0, ITA_ROUND_FLOOR = 1, ITA_ROUND_TORCH_RNE_F32 = 2 };
/* ITAmax variant */
enum { ITA_SOFTMAX_REPO_GLOBAL = 0, ITA_SOFTMAX_STREAM16 = 1 };
/* multi-head convention */
enum { ITA_HEADS_PYITA = 0, ITA_HEADS_TORCH = 1 };
/* requant stages: index into mult/shift/add/eff[stage*H + head]; 0..6 is the golden model's order */
enum { ITA_ST_Q = 0, ITA_ST_K, ITA_ST_V, ITA_ST_QK, ITA_ST_AV, ITA_ST_O, ITA_ST_SUM, ITA_ST_FC1, ITA_ST_RELU, ITA_ST_FC2,
       ITA_NUM_STAGES };
/* golden-model steps for ita_mha_step */
enum { ITA_STEP1_QP = 1, ITA_STEP2_KP = 2, ITA_STEP3_VP = 3, ITA_STEP4_QK = 4, ITA_STEP4_SOFTMAX = 40, ITA_STEP5_AV = 5,
       ITA_STEP6_O = 6, ITA_STEP7_OSUM = 7 };
/* which kernels a forward call uses */
enum { ITA_PATH_AUTO = 0,     /* fused tcgen05 kernel when the shape/params are supported, else the staged kernels */
       ITA_PATH_FUSED = 1,    /* fused tcgen05 kernel; error if unsupported                                        */
       ITA_PATH_STAGED = 2 }; /* one kernel per stage, every intermediate written to HBM scratch                    */

/* error codes */
enum { ITA_OK = 0, ITA_ERR_INVALID = 1, ITA_ERR_UNSUPPORTED = 2, ITA_ERR_CUDA = 3, ITA_ERR_NOMEM = 4 };

/* All pointers in the descriptor are HOST pointers; the handle keeps device copies.  Layouts are the golden model's
 * (tests/export_and_validation_W_B.py:479-491): Wq/Wk/Wv [H,E,P] s8, Wo [H,P,E] s8, Bq/Bk/Bv [H,P] s32, Bo [H,E] s32,
 * Wff [1,E,F], Wff2 [1,F,E], Bff [1,F], Bff2 [1,E].  P is the per-head projection width. */
typedef struct ita_mha_desc {
  int32_t abi_version;          /* ITA_B200_ABI_VERSION */
  int32_t S, E, P, H, F;        /* F == 0: no FFN parameters */
  int32_t round_mode, softmax_mode, head_mode;
  int32_t device;               /* CUDA device ordinal, -1 = current device */
  const int8_t *Wq, *Wk, *Wv, *Wo;
  const int32_t *Bq, *Bk, *Bv, *Bo;
  const int8_t *Wff, *Wff2;     /* may be NULL when F == 0 */
  const int32_t *Bff, *Bff2;
  const int32_t *mult, *shift, *add;  /* [ITA_NUM_STAGES, H] */
  const float *eff;                   /* [ITA_NUM_STAGES, H]; required for ITA_ROUND_TORCH_RNE_F32, else may be NULL */
} ita_mha_desc;

typedef struct ita_mha_handle ita_mha_handle;

/* Device pointers to per-stage results (any may be NULL = not wanted).  Shapes: Qp/Kp/Vp/O [B,H,S,P] s8,
 * A [B,H,S,S] s8, Asm [B,H,S,S] u8, Out [B,H,S,E] s8 (ITA_HEADS_TORCH: [B,1,S,E]). */
typedef struct ita_mha_stages {
  int8_t *Qp, *Kp, *Vp, *A;
  uint8_t *Asm;
  int8_t *O, *Out;
} ita_mha_stages;

ITA_API int ita_b200_abi_version(void);
ITA_API const char *ita_last_error(void);
/* number of CUDA kernels this library has launched in the calling process since load (for bench.py's gpu_launches) */
ITA_API int64_t ita_kernel_launch_count(void);

ITA_API int ita_mha_create(const ita_mha_desc *desc, ita_mha_handle **out);
ITA_API void ita_mha_destroy(ita_mha_handle *h);
/* 1 if the fused tcgen05 kernel can run this handle's shape and parameters, else 0 (reason in ita_last_error()) */
ITA_API int ita_mha_fused_supported(const ita_mha_handle *h);

/* x_dev [batch,S,E] s8 -> y_dev [batch,S,E] s8, both DEVICE pointers; asynchronous on `cuda_stream`. */
ITA_API int ita_mha_forward(ita_mha_handle *h, const int8_t *x_dev, int8_t *y_dev, int64_t batch, int32_t path, void *cuda_stream);
/* same with HOST pointers (pinned or pageable): H2D, kernels, D2H, then synchronises the stream. */
ITA_API int ita_mha_forward_host(ita_mha_handle *h, const int8_t *x_host, int8_t *y_host, int64_t batch, int32_t path,
                         void *cuda_stream);
/* same without the final synchronisation: every copy and kernel is ordered on `cuda_stream` (helper streams are joined
 * back into it), y_host is valid once the stream has been synchronised (ita_mha_sync_check also reports device-side
 * time-outs).  With two handles on two streams the H2D copies of one call overlap the D2H copies of the other. */
ITA_API int ita_mha_forward_host_async(ita_mha_handle *h, const int8_t *x_host, int8_t *y_host, int64_t batch, int32_t path,
                               void *cuda_stream);
/* The whole operator on HOST float buffers: y = s_out * attention(sat8(rne(x / s_in))), x/y [batch,S,E] f32
 * (ITASelfAttention_QAT.forward with its quant / dequant stubs, models/ITA/QAT/layers.py:101-127); synchronous.  This is
 * what the IREE dispatch symbol ITASelfAttention_workgroup(in, out) computes (csrc/ita_dispatch_shim.c). */
ITA_API int ita_mha_forward_host_f32(ita_mha_handle *h, const float *x_host, float *y_host, int64_t batch, float s_in, float s_out,
                                     int32_t path, void *cuda_stream);
/* which kernels / requantisation forms this handle runs, as text (diagnostics; written into buf, NUL-terminated) */
ITA_API int ita_mha_describe(const ita_mha_handle *h, char *buf, int64_t len);
/* staged kernels; writes every requested intermediate (DEVICE pointers) and y_dev (may be NULL). */
ITA_API int ita_mha_forward_stages(ita_mha_handle *h, const int8_t *x_dev, int8_t *y_dev, int64_t batch,
                           const ita_mha_stages *stages, void *cuda_stream);
/* Diagnostic twin of ita_mha_forward_stages: runs the FUSED kernel and additionally dumps its on-chip intermediates
 * (Qp, Kp, Vp, A, Asm, O; `Out` is ignored) in the natural layouts.  Synchronises the stream. */
ITA_API int ita_mha_forward_fused_debug(ita_mha_handle *h, const int8_t *x_dev, int8_t *y_dev, int64_t batch,
                                const ita_mha_stages *stages, void *cuda_stream);
/* Synchronise `cuda_stream` and report (and clear) a device-side pipeline time-out raised by earlier asynchronous
 * forward calls on this handle (every in-kernel wait is bounded; a protocol error surfaces here, not as a hang). */
ITA_API int ita_mha_sync_check(ita_mha_handle *h, void *cuda_stream);
/* One golden-model step on DEVICE buffers laid out as in ita_mha_stages; `in0`/`in1` are the step's inputs:
 *   STEP1/2/3: in0 = X [B,S,E]                       -> out [B,H,S,P]
 *   STEP4_QK : in0 = Qp, in1 = Kp                    -> out = A [B,H,S,S] s8
 *   STEP4_SOFTMAX: in0 = A                           -> out = Asm (u8)
 *   STEP5_AV : in0 = Asm (u8), in1 = Vp              -> out = O
 *   STEP6_O  : in0 = O                               -> out = Out [B,H,S,E] (HEADS_TORCH: Y [B,S,E])
 *   STEP7_OSUM: in0 = Out                            -> out = Y [B,S,E]                                              */
ITA_API int ita_mha_step(ita_mha_handle *h, int32_t step, const void *in0, const void *in1, void *out, int64_t batch,
                 void *cuda_stream);

/* FFN block: x_dev [batch,S,E] s8 -> f2_dev [batch,S,E] s8; f1_dev ([batch,S,F] s8, post-ReLU) may be NULL. */
ITA_API int ita_ffn_forward(ita_mha_handle *h, const int8_t *x_dev, int8_t *f1_dev, int8_t *f2_dev, int64_t batch, void *cuda_stream);

/* same with an explicit path: ITA_PATH_FUSED = the tcgen05 FFN kernel (F1 stays in shared memory; f1_dev, when given, is an
 * extra copy), ITA_PATH_STAGED = two CUDA-core GEMM kernels, ITA_PATH_AUTO = fused when ita_ffn_fused_supported(). */
ITA_API int ita_ffn_forward_path(ita_mha_handle *h, const int8_t *x_dev, int8_t *f1_dev, int8_t *f2_dev, int64_t batch, int32_t path,
                                 void *cuda_stream);
ITA_API int ita_ffn_fused_supported(const ita_mha_handle *h);

/* The float stages between the two int8 blocks of an encoder layer (reference
 * models/ITA_single_layer_upsample_shuffle/QAT/model.py:104-106,111-113: dequant stub, `self.add.add`, LayerNorm, and the
 * next block's quant stub), one pass:  y = LayerNorm(x + q_scale * q) * ln_w + ln_b  over the last dimension E (64 or 128);
 * yq = sat8(rne(y / yq_scale)) when yq_dev is not NULL.  All DEVICE pointers; x/y [rows,E] f32, q/yq [rows,E] s8. */
ITA_API int ita_residual_layernorm(const float *x_dev, const int8_t *q_dev, float q_scale, const float *ln_w_dev, const float *ln_b_dev,
                                   float eps, float *y_dev, int8_t *yq_dev, float yq_scale, int64_t rows, int32_t E, void *cuda_stream);

/* One encoder layer of ITAViTLSTM, f32 tokens in -> f32 tokens out: the loop body of ITALSTMNetVIT_QAT.forward
 * (models/ITA_single_layer_upsample_shuffle/QAT/model.py:100-113) -- quant stub, int8 attention, dequant + residual +
 * LayerNorm, quant stub, int8 FFN, dequant + residual + LayerNorm.  The handle must carry FFN parameters (F > 0). */
typedef struct ita_layer_params {
  float s_attn_in;    /* attention_blocks[i].quant.scale        */
  float s_attn_out;   /* attention_blocks[i].out_proj.scale     */
  float s_ffn_in;     /* ffn_blocks[i].quant.scale              */
  float s_ffn_out;    /* ffn_blocks[i].fc2.scale                */
  float ln_eps;       /* nn.LayerNorm eps (1e-5)                */
  const float *ln1_w, *ln1_b, *ln2_w, *ln2_b;   /* DEVICE pointers, [E]: norms1[i], norms2[i] */
} ita_layer_params;
ITA_API int ita_encoder_layer_forward(ita_mha_handle *h, const ita_layer_params *lp, const float *x_dev, float *y_dev, int64_t batch,
                                      void *cuda_stream);

/* ita_encoder_layer_forward for callers that must not allocate.  It has NO counterpart in the reference (the reference's
 * layer runs inside torch / IREE, which own their memory); it exists for hosts of this library that cannot let a dispatch
 * call cudaMalloc: the drone's per-packet inference loop (cudaMalloc / cudaFree stall the whole device and would land inside
 * the control period), a runtime that has already handed out all device memory from its own pool, code running inside
 * another component's stream-ordered region.  Same arguments, same five kernels and bit-identical results as
 * ita_encoder_layer_forward, but the per-frame intermediates (8 * S * E bytes per frame) are kept in a fixed array inside
 * the library's device image instead of an allocation, and the call itself performs no cudaMalloc / cudaFree.
 *   - Limit: batch * 8 * S * E <= ITA_LAYER_STATIC_SCRATCH_BYTES (256 frames at S=128, E=64; 128 frames at E=128).  A larger
 *     batch returns ITA_ERR_INVALID with the limit in ita_last_error(); it is never split or spilled.
 *   - Cost: every GPU that loads libita_b200.so carries those 16 MiB until the process exits, used or not.
 *   - The handle must run the single-head tcgen05 kernels for both blocks (ita_mha_fused_supported and
 *     ita_ffn_fused_supported, H == 1); other handles get ITA_ERR_UNSUPPORTED because their kernels keep allocated scratch.
 *     (ita_mha_create itself allocates the parameter copies: create handles at start-up.)
 *   - The first call on each device resolves the array's device address; make it at start-up, before any stream capture.
 *   - One array per GPU per process, shared by all handles and streams, never cleared (every intermediate is written in
 *     full before it is read, so stale contents are harmless): calls that may overlap on the GPU race on it -- keep them
 *     on one stream or order them with events. */
#define ITA_LAYER_STATIC_SCRATCH_BYTES (16 * 1024 * 1024)
ITA_API int ita_encoder_layer_forward_static(ita_mha_handle *h, const ita_layer_params *lp, const float *x_dev, float *y_dev,
                                             int64_t batch, void *cuda_stream);

/* The producer in front of the hot path: 60x90 depth frames -> 128 tokens of E channels (OverlapPatchMerging,
 * models/ITA/QAT/layers.py:25-45), all DEVICE pointers.  img_dev: [batch, 60, 90] uint8 (img_is_u8 != 0; scaled by 1/255 as
 * main.cpp:168 does) or float32.  conv_w_dev [E,1,7,7], conv_b_dev [E], ln_w_dev / ln_b_dev [E] as in the module's
 * state_dict.  Outputs (either may be NULL): tokens_dev f32 [batch,128,E]; tokens_q_dev int8 [batch,128,E] =
 * sat8(rne(tokens / q_scale)) (the first encoder layer's quant stub).  E in {64, 128}. */
ITA_API int ita_tokenizer_forward(const void *img_dev, int32_t img_is_u8, int64_t batch, int32_t E, const float *conv_w_dev,
                                  const float *conv_b_dev, const float *ln_w_dev, const float *ln_b_dev, float ln_eps, float *tokens_dev,
                                  int8_t *tokens_q_dev, float q_scale, void *cuda_stream);

/* The consumer behind the encoder: the fusion stage of ITALSTMNetVIT_QAT.forward (models/ITA_single_layer_upsample_shuffle/
 * QAT/model.py:116-121), all DEVICE pointers: tokens x_dev f32 [batch,128,E] viewed as [E,8,16] -> PixelShuffle(2) |
 * bilinear Upsample(16x32, align_corners=True) -> cat -> Conv2d(E/4+E -> 9, 3x3, padding 1) -> out_dev f32 [batch, 9*16*32]
 * (the decoder's input).  conv_w_dev [9, E/4+E, 3, 3], conv_b_dev [9] as in the module's state_dict.  E in {64, 128}. */
ITA_API int ita_fusion_conv_forward(const float *x_dev, int64_t batch, int32_t E, const float *conv_w_dev, const float *conv_b_dev,
                                    float *out_dev, void *cuda_stream);

/* Pointwise half of one LSTM time step (nn.LSTM gate order i, f, g, o; the model's 3-layer LSTM, QAT/model.py:68,128), DEVICE
 * pointers: gates_dev f32 [batch, 4*hidden] (= x W_ih^T + b_ih + h W_hh^T + b_hh), c_prev_dev [batch, hidden] ->
 * c = sigmoid(f) * c_prev + sigmoid(i) * tanh(g), h = sigmoid(o) * tanh(c).  h_dev / c_dev may alias nothing else. */
ITA_API int ita_lstm_cell_forward(const float *gates_dev, const float *c_prev_dev, float *h_dev, float *c_dev, int64_t batch, int32_t hidden,
                                  void *cuda_stream);

/* float boundary: q = sat8(rne(x / scale)),  y = scale * q   (n elements, DEVICE pointers) */
ITA_API int ita_quantize_f32(const float *x_dev, int8_t *q_dev, int64_t n, float scale, void *cuda_stream);
ITA_API int ita_dequantize_s8(const int8_t *q_dev, float *y_dev, int64_t n, float scale, void *cuda_stream);

/* GPU self-tests of the tcgen05 building blocks (descriptor layouts, TMEM load, TMA).  Returns 0 when every check
 * passes; a human-readable report is written into `report` (NUL-terminated, at most report_len bytes). */
ITA_API int ita_selftest(int32_t device, char *report, int64_t report_len);

/* Micro-benchmarks used by bench.py for the roofline denominators (results in `out`):
 *   kind 0: dense tcgen05 kind::i8 MMA issue rate -> out[0] = int8 TOP/s over the whole GPU
 *   kind 1: device-to-device copy                 -> out[0] = GB/s (read+write)
 *   kind 2: TMEM read bandwidth (developer)       -> out[0..7] = bytes per SM clock per SM
 *   kind 3: kind 0 back to back for out[0] seconds (input; default 2) -> out[0] = sustained int8 TOP/s, out[1] = seconds  */
ITA_API int ita_microbench(int32_t device, int32_t kind, double *out, int32_t n_out);

#ifdef __cplusplus
}
#endif
#endif /* ITA_B200_H_ */
