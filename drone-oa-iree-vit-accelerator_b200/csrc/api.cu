// api.cu -- the extern "C" surface declared in include/ita_b200.h.
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <new>

#include "common.cuh"

namespace ita {

thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int ensure_scratch(ita_mha_handle *h, size_t bytes) {
  if (h->scratch_bytes >= bytes) return ITA_OK;
  if (h->scratch) cudaFree(h->scratch);
  h->scratch = nullptr;
  h->scratch_bytes = 0;
  if (cudaMalloc(&h->scratch, bytes) != cudaSuccess) {
    cudaGetLastError();
    set_error("cudaMalloc(%zu) for intermediate scratch failed", bytes);
    return ITA_ERR_NOMEM;
  }
  h->scratch_bytes = bytes;
  return ITA_OK;
}

template <class T>
static int upload(T **dst, const T *src, size_t n) {
  *dst = nullptr;
  if (!src || !n) return ITA_OK;
  if (cudaMalloc(dst, n * sizeof(T)) != cudaSuccess) {
    cudaGetLastError();
    set_error("cudaMalloc(%zu) failed", n * sizeof(T));
    return ITA_ERR_NOMEM;
  }
  ITA_CUDA_TRY(cudaMemcpy(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice));
  return ITA_OK;
}

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (dev != prev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

static int check_abort(ita_mha_handle *h, cudaStream_t s, const char *what) {
  int flag = 0;
  ITA_CUDA_TRY(cudaMemcpyAsync(&flag, h->abort_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
  ITA_CUDA_TRY(cudaStreamSynchronize(s));
  if (flag) {
    cudaMemsetAsync(h->abort_flag, 0, sizeof(int), s);
    set_error("%s: a device-side bounded wait timed out (pipeline protocol error)", what);
    return ITA_ERR_CUDA;
  }
  return ITA_OK;
}

static int forward_dev(ita_mha_handle *h, const int8_t *x, int8_t *y, int64_t B, int path, cudaStream_t s) {
  if (path == ITA_PATH_AUTO) {
    path = (h->fused || h->mh) ? ITA_PATH_FUSED : ITA_PATH_STAGED;
    if (!h->fused && !h->mh && !h->warned_staged) {     // say it once: the staged kernels are the correctness path, not the fast one
      h->warned_staged = true;
      fprintf(stderr, "libita_b200: ITA_PATH_AUTO falls back to the staged (CUDA-core) kernels for S=%d E=%d P=%d H=%d: %s\n", h->S, h->E,
              h->P, h->H, h->fused_reason.c_str());
    }
  }
  if (path == ITA_PATH_FUSED) return fused_forward(h, x, y, B, nullptr, s);
  if (path == ITA_PATH_STAGED) return staged_forward(h, x, y, B, nullptr, s);
  set_error("unknown path %d", path);
  return ITA_ERR_INVALID;
}

}  // namespace ita

using namespace ita;

extern "C" {

int ita_b200_abi_version(void) { return ITA_B200_ABI_VERSION; }
const char *ita_last_error(void) { return g_err; }
int64_t ita_kernel_launch_count(void) { return (int64_t)g_launches.load(std::memory_order_relaxed); }

int ita_mha_create(const ita_mha_desc *d, ita_mha_handle **out) {
  if (!d || !out) {
    set_error("ita_mha_create: null argument");
    return ITA_ERR_INVALID;
  }
  *out = nullptr;
  if (d->abi_version != ITA_B200_ABI_VERSION) {
    set_error("ita_mha_create: abi_version %d, library has %d", d->abi_version, ITA_B200_ABI_VERSION);
    return ITA_ERR_INVALID;
  }
  if (d->S <= 0 || d->E <= 0 || d->P <= 0 || d->H <= 0 || d->H > 16 || d->F < 0) {
    set_error("ita_mha_create: bad dims S=%d E=%d P=%d H=%d F=%d (need >0, H<=16)", d->S, d->E, d->P, d->H, d->F);
    return ITA_ERR_INVALID;
  }
  if (d->round_mode < 0 || d->round_mode > 2 || d->softmax_mode < 0 || d->softmax_mode > 1 || d->head_mode < 0 || d->head_mode > 1) {
    set_error("ita_mha_create: bad mode (round %d, softmax %d, heads %d)", d->round_mode, d->softmax_mode, d->head_mode);
    return ITA_ERR_INVALID;
  }
  if (d->softmax_mode == ITA_SOFTMAX_STREAM16 && d->S % 16) {
    set_error("ita_mha_create: streaming ITAmax needs S %% 16 == 0");
    return ITA_ERR_INVALID;
  }
  if (!d->Wq || !d->Wk || !d->Wv || !d->Wo || !d->Bq || !d->Bk || !d->Bv || !d->Bo || !d->mult || !d->shift || !d->add) {
    set_error("ita_mha_create: null weight / bias / requant pointer");
    return ITA_ERR_INVALID;
  }
  if (d->round_mode == ITA_ROUND_TORCH_RNE_F32 && !d->eff) {
    set_error("ita_mha_create: ITA_ROUND_TORCH_RNE_F32 needs eff[]");
    return ITA_ERR_INVALID;
  }
  if (d->F > 0 && (!d->Wff || !d->Wff2 || !d->Bff || !d->Bff2)) {
    set_error("ita_mha_create: F > 0 needs Wff/Wff2/Bff/Bff2");
    return ITA_ERR_INVALID;
  }
  // (mult, shift) are unused by the fp32 requant form: only eff[] matters there
  for (int i = 0; i < ITA_NUM_STAGES * d->H && d->round_mode != ITA_ROUND_TORCH_RNE_F32; ++i) {
    if (d->shift[i] < 0 || d->shift[i] > 62 || d->mult[i] < 0 || d->mult[i] > 65535) {
      set_error("ita_mha_create: stage %d head %d: mult %d / shift %d out of range (0..65535 / 0..62)", i / d->H, i % d->H,
                d->mult[i], d->shift[i]);
      return ITA_ERR_INVALID;
    }
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    set_error("ita_mha_create: no CUDA device (this library has no CPU fallback)");
    return ITA_ERR_CUDA;
  }
  int dev = d->device;
  if (dev < 0) ITA_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= ndev) {
    set_error("ita_mha_create: device %d out of range (%d devices)", dev, ndev);
    return ITA_ERR_INVALID;
  }
  DeviceGuard guard(dev);
  if (!guard.ok) {
    set_error("ita_mha_create: cudaSetDevice(%d) failed", dev);
    return ITA_ERR_CUDA;
  }
  ita_mha_handle *h = new (std::nothrow) ita_mha_handle();
  if (!h) return ITA_ERR_NOMEM;
  h->S = d->S; h->E = d->E; h->P = d->P; h->H = d->H; h->F = d->F;
  h->round_mode = d->round_mode; h->softmax_mode = d->softmax_mode; h->head_mode = d->head_mode;
  h->device = dev;
  cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, dev);
  const size_t S = d->S, E = d->E, P = d->P, H = d->H, F = d->F;
  (void)S;
  int rc = ITA_OK;
  auto up = [&](int r) { if (!rc) rc = r; };
  up(upload(&h->Wq, d->Wq, H * E * P)); up(upload(&h->Wk, d->Wk, H * E * P)); up(upload(&h->Wv, d->Wv, H * E * P));
  up(upload(&h->Wo, d->Wo, H * P * E));
  up(upload(&h->Bq, d->Bq, H * P)); up(upload(&h->Bk, d->Bk, H * P)); up(upload(&h->Bv, d->Bv, H * P));
  up(upload(&h->Bo, d->Bo, H * E));
  if (F) {
    up(upload(&h->Wff, d->Wff, E * F)); up(upload(&h->Wff2, d->Wff2, F * E));
    up(upload(&h->Bff, d->Bff, F)); up(upload(&h->Bff2, d->Bff2, E));
  }
  for (int i = 0; i < ITA_NUM_STAGES * d->H; ++i) {
    Rq r;
    r.mult = d->mult[i]; r.shift = d->shift[i]; r.add = d->add[i]; r.mode = d->round_mode;
    if (d->round_mode == ITA_ROUND_TORCH_RNE_F32 && (r.shift < 0 || r.shift > 62)) r.shift = 0;
    r.eff = d->eff ? d->eff[i] : 0.f;
    r.rnd = (d->round_mode == ITA_ROUND_HALF_UP && r.shift > 0) ? (1LL << (r.shift - 1)) : 0LL;
    h->rq_host[i] = r;
  }
  up(upload(&h->rq_dev, h->rq_host, (size_t)ITA_NUM_STAGES * d->H));
  if (!rc && cudaMalloc(&h->abort_flag, sizeof(int)) != cudaSuccess) {
    cudaGetLastError();
    set_error("cudaMalloc(abort flag) failed");
    rc = ITA_ERR_NOMEM;
  }
  if (!rc && cudaMemset(h->abort_flag, 0, sizeof(int)) != cudaSuccess) rc = ITA_ERR_CUDA;
  if (!rc) rc = fused_plan_create(h, d);
  if (!rc) rc = ffn_plan_create(h, d);
  if (rc) {
    ita_mha_destroy(h);
    return rc;
  }
  *out = h;
  return ITA_OK;
}

void ita_mha_destroy(ita_mha_handle *h) {
  if (!h) return;
  DeviceGuard guard(h->device);
  fused_plan_destroy(h);
  ffn_plan_destroy(h);
  cudaFree(h->layer_scratch);
  cudaFree(h->Wq); cudaFree(h->Wk); cudaFree(h->Wv); cudaFree(h->Wo); cudaFree(h->Wff); cudaFree(h->Wff2);
  cudaFree(h->Bq); cudaFree(h->Bk); cudaFree(h->Bv); cudaFree(h->Bo); cudaFree(h->Bff); cudaFree(h->Bff2);
  cudaFree(h->rq_dev); cudaFree(h->scratch); cudaFree(h->io_dev); cudaFree(h->abort_flag);
  if (h->pipe_ready) {
    for (int i = 0; i < 2; ++i) { cudaStreamDestroy(h->aux[i]); cudaEventDestroy(h->ev_done[i]); }
    for (int i = 0; i < ita_mha_handle::kHostChunks; ++i) { cudaEventDestroy(h->ev_in[i]); cudaEventDestroy(h->ev_k[i]); }
    cudaEventDestroy(h->ev_start);
  }
  delete h;
}

int ita_mha_fused_supported(const ita_mha_handle *h) {
  if (!h) return 0;
  if (h->fused || h->mh) return 1;
  set_error("%s", h->fused_reason.c_str());
  return 0;
}

static int check_fwd_args(const char *fn, const ita_mha_handle *h, const void *x, const void *y, int64_t batch) {
  if (!h || (batch > 0 && (!x || !y)) || batch < 0) {
    set_error("%s: null handle/buffer or negative batch", fn);
    return ITA_ERR_INVALID;
  }
  return ITA_OK;
}

int ita_mha_forward(ita_mha_handle *h, const int8_t *x_dev, int8_t *y_dev, int64_t batch, int32_t path, void *cuda_stream) {
  int rc = check_fwd_args("ita_mha_forward", h, x_dev, y_dev, batch);
  if (rc) return rc;
  DeviceGuard guard(h->device);
  return forward_dev(h, x_dev, y_dev, batch, path, (cudaStream_t)cuda_stream);
}

static int forward_host_impl(ita_mha_handle *h, const int8_t *x_host, int8_t *y_host, int64_t batch, int32_t path, void *cuda_stream,
                             bool synchronous);

int ita_mha_forward_host(ita_mha_handle *h, const int8_t *x_host, int8_t *y_host, int64_t batch, int32_t path, void *cuda_stream) {
  int rc = forward_host_impl(h, x_host, y_host, batch, path, cuda_stream, true);
  if (rc || batch == 0) return rc;
  DeviceGuard guard(h->device);
  return check_abort(h, (cudaStream_t)cuda_stream, "ita_mha_forward_host");
}

int ita_mha_forward_host_async(ita_mha_handle *h, const int8_t *x_host, int8_t *y_host, int64_t batch, int32_t path, void *cuda_stream) {
  return forward_host_impl(h, x_host, y_host, batch, path, cuda_stream, false);
}

static int forward_host_impl(ita_mha_handle *h, const int8_t *x_host, int8_t *y_host, int64_t batch, int32_t path, void *cuda_stream,
                             bool synchronous) {
  int rc = check_fwd_args("ita_mha_forward_host", h, x_host, y_host, batch);
  if (rc) return rc;
  if (batch == 0) return ITA_OK;
  DeviceGuard guard(h->device);
  cudaStream_t s = (cudaStream_t)cuda_stream;
  const size_t frame = (size_t)h->S * h->E;
  const size_t need = 2 * frame * (size_t)batch;
  if (h->io_bytes < need) {
    if (h->io_dev) cudaFree(h->io_dev);
    h->io_dev = nullptr;
    h->io_bytes = 0;
    if (cudaMalloc(&h->io_dev, need) != cudaSuccess) {
      cudaGetLastError();
      set_error("ita_mha_forward_host: cudaMalloc(%zu) failed", need);
      return ITA_ERR_NOMEM;
    }
    h->io_bytes = need;
  }
  int8_t *xd = h->io_dev, *yd = h->io_dev + frame * (size_t)batch;
  // Pipeline the batch in chunks: the H2D copy of chunk i+1 and the D2H copy of chunk i-1 overlap the kernel of chunk i
  // when the caller's buffers are pinned (PCIe is full duplex); with pageable memory the copies serialise, still correct.
  // The copies are the cost (PCIe 5 x16: ~48 GB/s each way at once, the kernel is ~40x faster), so the chunks are small
  // enough to keep the un-overlapped first H2D / last D2H short and large enough (>= 4 MB) to run at link speed.
  // The synchronous call stands alone, so more chunks shorten its un-overlapped ends; the asynchronous call is meant to be
  // overlapped with another handle's, where the host-side cost of queueing the chunks is what limits (measured: 8 chunks
  // 4.4 M frames/s, 2 chunks 5.2 M).
  int64_t nchunk = synchronous ? (batch >= 4096 ? ita_mha_handle::kHostChunks : batch >= 1024 ? 4 : 1) : (batch >= 1024 ? 2 : 1);
  if (const char *e = getenv("ITA_HOST_CHUNKS")) {   // tuning knob
    const int want = atoi(e);
    if (want >= 1 && want <= ita_mha_handle::kHostChunks && batch >= want) nchunk = want;
  }
  if (path == ITA_PATH_AUTO) path = (h->fused || h->mh) ? ITA_PATH_FUSED : ITA_PATH_STAGED;
  if (path == ITA_PATH_STAGED) nchunk = 1;   // the staged path shares one scratch area
  const int64_t per = (batch + nchunk - 1) / nchunk;
  if (nchunk > 1 && !h->pipe_ready) {
    for (int i = 0; i < 2; ++i) ITA_CUDA_TRY(cudaStreamCreateWithFlags(&h->aux[i], cudaStreamNonBlocking));
    for (int i = 0; i < ita_mha_handle::kHostChunks; ++i) {
      ITA_CUDA_TRY(cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming));
      ITA_CUDA_TRY(cudaEventCreateWithFlags(&h->ev_k[i], cudaEventDisableTiming));
    }
    ITA_CUDA_TRY(cudaEventCreateWithFlags(&h->ev_start, cudaEventDisableTiming));
    for (int i = 0; i < 2; ++i) ITA_CUDA_TRY(cudaEventCreateWithFlags(&h->ev_done[i], cudaEventDisableTiming));
    h->pipe_ready = true;
  }
  if (nchunk > 1) {
    ITA_CUDA_TRY(cudaEventRecord(h->ev_start, s));
    ITA_CUDA_TRY(cudaStreamWaitEvent(h->aux[0], h->ev_start, 0));
    ITA_CUDA_TRY(cudaStreamWaitEvent(h->aux[1], h->ev_start, 0));
  }
  rc = ITA_OK;
  // inside the loop a CUDA error must not return before the helper streams are joined back into `s` below
#define ITA_CUDA_STEP(expr)                                                                                \
  {                                                                                                        \
    cudaError_t _e = (expr);                                                                               \
    if (_e != cudaSuccess) {                                                                               \
      set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));                     \
      rc = ITA_ERR_CUDA;                                                                                   \
      break;                                                                                               \
    }                                                                                                      \
  }
  for (int64_t c = 0; c < nchunk && !rc; ++c) {
    const int64_t b0 = c * per, nb = (b0 + per <= batch) ? per : batch - b0;
    if (nb <= 0) break;
    if (nchunk == 1) {
      ITA_CUDA_STEP(cudaMemcpyAsync(xd, x_host, frame * nb, cudaMemcpyHostToDevice, s));
      rc = forward_dev(h, xd, yd, nb, path, s);
      if (rc) break;
      ITA_CUDA_STEP(cudaMemcpyAsync(y_host, yd, frame * nb, cudaMemcpyDeviceToHost, s));
    } else {
      // aux[0]: H2D copies, s: kernels, aux[1]: D2H copies
      ITA_CUDA_STEP(cudaMemcpyAsync(xd + frame * b0, x_host + frame * b0, frame * nb, cudaMemcpyHostToDevice, h->aux[0]));
      ITA_CUDA_STEP(cudaEventRecord(h->ev_in[c], h->aux[0]));
      ITA_CUDA_STEP(cudaStreamWaitEvent(s, h->ev_in[c], 0));
      rc = forward_dev(h, xd + frame * b0, yd + frame * b0, nb, path, s);
      if (rc) break;
      ITA_CUDA_STEP(cudaEventRecord(h->ev_k[c], s));
      ITA_CUDA_STEP(cudaStreamWaitEvent(h->aux[1], h->ev_k[c], 0));
      ITA_CUDA_STEP(cudaMemcpyAsync(y_host + frame * b0, yd + frame * b0, frame * nb, cudaMemcpyDeviceToHost, h->aux[1]));
    }
  }
#undef ITA_CUDA_STEP
  if (nchunk > 1) {
    // join the helper streams back into `s`
    cudaEventRecord(h->ev_done[0], h->aux[0]);
    cudaEventRecord(h->ev_done[1], h->aux[1]);
    cudaStreamWaitEvent(s, h->ev_done[0], 0);
    cudaStreamWaitEvent(s, h->ev_done[1], 0);
  }
  return rc;
}

/* f32 host tokens in -> f32 host tokens out: quant stub, attention, dequant stub (ITASelfAttention_QAT.forward,
 * models/ITA/QAT/layers.py:101-127) on caller-owned HOST buffers, synchronous -- the contract of the IREE dispatch symbol
 * (csrc/ita_dispatch_shim.c). */
int ita_mha_forward_host_f32(ita_mha_handle *h, const float *x_host, float *y_host, int64_t batch, float s_in, float s_out, int32_t path,
                             void *cuda_stream) {
  int rc = check_fwd_args("ita_mha_forward_host_f32", h, x_host, y_host, batch);
  if (rc) return rc;
  if (!(s_in > 0.f)) {
    set_error("ita_mha_forward_host_f32: the input scale must be positive");
    return ITA_ERR_INVALID;
  }
  if (batch == 0) return ITA_OK;
  DeviceGuard guard(h->device);
  cudaStream_t s = (cudaStream_t)cuda_stream;
  const size_t n = (size_t)h->S * h->E * (size_t)batch, need = 6 * n;
  if (h->io_bytes < need) {
    if (h->io_dev) cudaFree(h->io_dev);
    h->io_dev = nullptr;
    h->io_bytes = 0;
    if (cudaMalloc(&h->io_dev, need) != cudaSuccess) {
      cudaGetLastError();
      set_error("ita_mha_forward_host_f32: cudaMalloc(%zu) failed", need);
      return ITA_ERR_NOMEM;
    }
    h->io_bytes = need;
  }
  float *fd = reinterpret_cast<float *>(h->io_dev);
  int8_t *xq = h->io_dev + 4 * n, *yq = xq + n;
  ITA_CUDA_TRY(cudaMemcpyAsync(fd, x_host, 4 * n, cudaMemcpyHostToDevice, s));
  if ((rc = launch_quantize(fd, xq, (int64_t)n, s_in, s))) return rc;
  if ((rc = forward_dev(h, xq, yq, batch, path, s))) return rc;
  if ((rc = launch_dequantize(yq, fd, (int64_t)n, s_out, s))) return rc;
  ITA_CUDA_TRY(cudaMemcpyAsync(y_host, fd, 4 * n, cudaMemcpyDeviceToHost, s));
  return check_abort(h, s, "ita_mha_forward_host_f32");
}

/* Human-readable description of what the handle will run (kernel family, frames in flight, requant form). */
int ita_mha_describe(const ita_mha_handle *h, char *buf, int64_t len) {
  if (!h || !buf || len <= 0) {
    set_error("ita_mha_describe: bad arguments");
    return ITA_ERR_INVALID;
  }
  std::string d = describe_fused(h) + "; " + describe_ffn(h);
  strncpy(buf, d.c_str(), (size_t)len - 1);
  buf[len - 1] = 0;
  return ITA_OK;
}

int ita_mha_forward_stages(ita_mha_handle *h, const int8_t *x_dev, int8_t *y_dev, int64_t batch, const ita_mha_stages *stages,
                           void *cuda_stream) {
  if (!h || (batch > 0 && !x_dev) || batch < 0) {
    set_error("ita_mha_forward_stages: null handle/input or negative batch");
    return ITA_ERR_INVALID;
  }
  DeviceGuard guard(h->device);
  return staged_forward(h, x_dev, y_dev, batch, stages, (cudaStream_t)cuda_stream);
}

/* Diagnostic twin of ita_mha_forward_stages that runs the FUSED kernel and additionally dumps its shared-memory
 * intermediates (Qp, Kp, Vp, A, Asm, O) in the natural layouts; used by the parity tests to localise a mismatch.
 * Synchronises and reports device-side pipeline time-outs. */
int ita_mha_forward_fused_debug(ita_mha_handle *h, const int8_t *x_dev, int8_t *y_dev, int64_t batch, const ita_mha_stages *stages,
                                void *cuda_stream) {
  int rc = check_fwd_args("ita_mha_forward_fused_debug", h, x_dev, y_dev, batch);
  if (rc) return rc;
  DeviceGuard guard(h->device);
  ita_mha_stages none{};
  rc = fused_forward(h, x_dev, y_dev, batch, stages ? stages : &none, (cudaStream_t)cuda_stream);
  if (rc) return rc;
  return check_abort(h, (cudaStream_t)cuda_stream, "ita_mha_forward_fused_debug");
}

/* Synchronise the stream and report (and clear) a device-side pipeline time-out of earlier asynchronous forward calls. */
int ita_mha_sync_check(ita_mha_handle *h, void *cuda_stream) {
  if (!h) {
    set_error("ita_mha_sync_check: null handle");
    return ITA_ERR_INVALID;
  }
  DeviceGuard guard(h->device);
  return check_abort(h, (cudaStream_t)cuda_stream, "ita_mha_sync_check");
}

int ita_mha_step(ita_mha_handle *h, int32_t step, const void *in0, const void *in1, void *out, int64_t batch, void *cuda_stream) {
  if (!h || !in0 || !out || batch < 0) {
    set_error("ita_mha_step: null handle/buffer or negative batch");
    return ITA_ERR_INVALID;
  }
  if ((step == ITA_STEP4_QK || step == ITA_STEP5_AV) && !in1) {
    set_error("ita_mha_step: step %d needs two inputs", step);
    return ITA_ERR_INVALID;
  }
  DeviceGuard guard(h->device);
  return staged_step(h, step, in0, in1, out, batch, (cudaStream_t)cuda_stream);
}

int ita_ffn_forward_path(ita_mha_handle *h, const int8_t *x_dev, int8_t *f1_dev, int8_t *f2_dev, int64_t batch, int32_t path,
                         void *cuda_stream) {
  int rc = check_fwd_args("ita_ffn_forward", h, x_dev, f2_dev, batch);
  if (rc) return rc;
  DeviceGuard guard(h->device);
  if (path == ITA_PATH_AUTO) path = h->ffn ? ITA_PATH_FUSED : ITA_PATH_STAGED;
  if (path == ITA_PATH_FUSED) return ffn_fused_forward(h, x_dev, f1_dev, f2_dev, batch, (cudaStream_t)cuda_stream);
  if (path == ITA_PATH_STAGED) return staged_ffn(h, x_dev, f1_dev, f2_dev, batch, (cudaStream_t)cuda_stream);
  set_error("unknown path %d", path);
  return ITA_ERR_INVALID;
}

int ita_ffn_forward(ita_mha_handle *h, const int8_t *x_dev, int8_t *f1_dev, int8_t *f2_dev, int64_t batch, void *cuda_stream) {
  return ita_ffn_forward_path(h, x_dev, f1_dev, f2_dev, batch, ITA_PATH_AUTO, cuda_stream);
}

int ita_ffn_fused_supported(const ita_mha_handle *h) {
  if (!h) return 0;
  if (h->ffn) return 1;
  set_error("%s", h->ffn_reason.c_str());
  return 0;
}

int ita_residual_layernorm(const float *x_dev, const int8_t *q_dev, float q_scale, const float *ln_w_dev, const float *ln_b_dev, float eps,
                           float *y_dev, int8_t *yq_dev, float yq_scale, int64_t rows, int32_t E, void *cuda_stream) {
  if (rows < 0 || (rows > 0 && (!x_dev || !q_dev || !ln_w_dev || !ln_b_dev || !y_dev))) {
    set_error("ita_residual_layernorm: bad arguments");
    return ITA_ERR_INVALID;
  }
  return launch_resid_ln(x_dev, q_dev, q_scale, ln_w_dev, ln_b_dev, eps, y_dev, yq_dev, yq_scale, rows, E, (cudaStream_t)cuda_stream);
}

int ita_encoder_layer_forward(ita_mha_handle *h, const ita_layer_params *lp, const float *x_dev, float *y_dev, int64_t batch,
                              void *cuda_stream) {
  if (!h || !lp || batch < 0 || (batch > 0 && (!x_dev || !y_dev)) || !lp->ln1_w || !lp->ln1_b || !lp->ln2_w || !lp->ln2_b) {
    set_error("ita_encoder_layer_forward: null handle / parameter / buffer or negative batch");
    return ITA_ERR_INVALID;
  }
  if (!(lp->s_attn_in > 0.f) || !(lp->s_ffn_in > 0.f)) {
    set_error("ita_encoder_layer_forward: quantisation scales must be positive");
    return ITA_ERR_INVALID;
  }
  DeviceGuard guard(h->device);
  return layer_forward(h, lp, x_dev, y_dev, batch, (cudaStream_t)cuda_stream);
}

int ita_tokenizer_forward(const void *img_dev, int32_t img_is_u8, int64_t batch, int32_t E, const float *conv_w_dev, const float *conv_b_dev,
                          const float *ln_w_dev, const float *ln_b_dev, float ln_eps, float *tokens_dev, int8_t *tokens_q_dev, float q_scale,
                          void *cuda_stream) {
  if (batch < 0 || (batch > 0 && (!img_dev || !conv_w_dev || !conv_b_dev || !ln_w_dev || !ln_b_dev || (!tokens_dev && !tokens_q_dev)))) {
    set_error("ita_tokenizer_forward: bad arguments");
    return ITA_ERR_INVALID;
  }
  return launch_tokenizer(img_dev, img_is_u8, batch, E, conv_w_dev, conv_b_dev, ln_w_dev, ln_b_dev, ln_eps, tokens_dev, tokens_q_dev, q_scale,
                          (cudaStream_t)cuda_stream);
}

int ita_fusion_conv_forward(const float *x_dev, int64_t batch, int32_t E, const float *conv_w_dev, const float *conv_b_dev, float *out_dev,
                            void *cuda_stream) {
  if (batch < 0 || (batch > 0 && (!x_dev || !conv_w_dev || !conv_b_dev || !out_dev))) {
    set_error("ita_fusion_conv_forward: bad arguments");
    return ITA_ERR_INVALID;
  }
  return launch_fuse_conv(x_dev, batch, E, conv_w_dev, conv_b_dev, out_dev, (cudaStream_t)cuda_stream);
}

int ita_lstm_cell_forward(const float *gates_dev, const float *c_prev_dev, float *h_dev, float *c_dev, int64_t batch, int32_t hidden,
                          void *cuda_stream) {
  if (batch < 0 || (batch > 0 && (!gates_dev || !c_prev_dev || !h_dev || !c_dev))) {
    set_error("ita_lstm_cell_forward: bad arguments");
    return ITA_ERR_INVALID;
  }
  return launch_lstm_cell(gates_dev, c_prev_dev, h_dev, c_dev, batch, hidden, (cudaStream_t)cuda_stream);
}

int ita_quantize_f32(const float *x_dev, int8_t *q_dev, int64_t n, float scale, void *cuda_stream) {
  if (n < 0 || (n > 0 && (!x_dev || !q_dev)) || !(scale > 0.f)) {
    set_error("ita_quantize_f32: bad arguments");
    return ITA_ERR_INVALID;
  }
  return launch_quantize(x_dev, q_dev, n, scale, (cudaStream_t)cuda_stream);
}
int ita_dequantize_s8(const int8_t *q_dev, float *y_dev, int64_t n, float scale, void *cuda_stream) {
  if (n < 0 || (n > 0 && (!q_dev || !y_dev))) {
    set_error("ita_dequantize_s8: bad arguments");
    return ITA_ERR_INVALID;
  }
  return launch_dequantize(q_dev, y_dev, n, scale, (cudaStream_t)cuda_stream);
}

int ita_selftest(int32_t device, char *report, int64_t report_len) {
  std::string rep;
  int rc = selftest_run(device, rep);
  if (report && report_len > 0) {
    strncpy(report, rep.c_str(), (size_t)report_len - 1);
    report[report_len - 1] = 0;
  }
  return rc;
}

int ita_microbench(int32_t device, int32_t kind, double *out, int32_t n_out) {
  if (!out || n_out < 1) {
    set_error("ita_microbench: bad arguments");
    return ITA_ERR_INVALID;
  }
  return microbench_run(device, kind, out, n_out);
}

}  // extern "C"
