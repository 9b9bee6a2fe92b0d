// layer_static.cu -- ita_encoder_layer_forward_static: one encoder layer whose intermediates live in memory that is part
// of the loaded library instead of memory obtained from the CUDA allocator.  No counterpart in the reference.
//
// Who needs this.  ita_encoder_layer_forward keeps five intermediates per frame (the quantised input, the attention
// output, the tokens after the first LayerNorm in f32 and int8, the FFN output) in a buffer it cudaMalloc's on the first
// call and grows when a larger batch arrives.  Some hosts of this operator cannot let a dispatch do that:
//   * the drone's inference loop (one frame per UDP packet, velocity command back within the control period) must not hit
//     cudaMalloc / cudaFree on its hot path, because both synchronise the whole device and take the driver's allocator lock;
//   * a runtime that owns all device memory itself (an IREE HAL allocator with a fixed pool, a process running under a
//     memory budget that is already handed out) has nothing left for a plugin to ask the driver for, and a failed
//     allocation in the middle of a flight is not an error it can handle;
//   * code that runs inside somebody else's stream-ordered region, where a device-wide synchronisation is not allowed.
// For them the intermediates sit in the file-scope array below.  It is a fixed ITA_LAYER_STATIC_SCRATCH_BYTES (16 MiB) of
// device memory on EVERY GPU that loads libita_b200.so, held until the process exits, whether or not this entry point is
// ever called; that is the price, and the reason the limit is modest.
//
// What the caller must know:
//   * batch * 8 * S * E must fit into the array (256 frames at S = 128, E = 64; 128 frames at E = 128); larger batches are
//     rejected with ITA_ERR_INVALID, never split and never spilled to an allocation.
//   * the handle must run the single-head tcgen05 kernels for both int8 blocks (ita_mha_fused_supported and
//     ita_ffn_fused_supported, H == 1): those launch with no per-call allocation.  The multi-head and the staged kernels
//     keep allocator-backed scratch of their own, so handles that need them are rejected here.
//   * the FIRST call on each device looks the array's device address up (cudaGetSymbolAddress, which may also make the
//     CUDA runtime load this library's device code on that GPU).  Make that first call during start-up, outside any stream
//     capture; later calls only launch kernels.
//   * there is ONE array per GPU for the whole process.  Every call, from every handle and every stream, writes its
//     intermediates to the same bytes: two calls whose kernels can overlap on the GPU (different streams, or different
//     host threads) corrupt each other.  Order them on one stream, or with events.
//   * nothing in the array has to start at zero: each of the five intermediates is fully written by the kernel that
//     produces it before the next kernel reads it, so the previous call's leftovers are never observed (and are not
//     cleared).  Do not treat the array as a place to read results from.
#include <mutex>

#include "common.cuh"

namespace ita {

// In host code `g_layer_static` is only a handle for the runtime's symbol table, not a device address: the address a kernel
// can use comes from cudaGetSymbolAddress (static_scratch_ptr below).
static __device__ __align__(256) unsigned char g_layer_static[ITA_LAYER_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
static std::mutex g_static_mu;
static int8_t *g_static_ptr[kMaxDevices] = {};   // per device ordinal, filled on the first call with that device current

// Device address of the array on `device`, which must be the current device.
static int static_scratch_ptr(int device, int8_t **out) {
  if (device < 0 || device >= kMaxDevices) {
    set_error("ita_encoder_layer_forward_static: device ordinal %d outside 0..%d", device, kMaxDevices - 1);
    return ITA_ERR_INVALID;
  }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[device]) {
    int cur = -1;
    ITA_CUDA_TRY(cudaGetDevice(&cur));
    if (cur != device) {
      set_error("ita_encoder_layer_forward_static: device %d is not current (current is %d)", device, cur);
      return ITA_ERR_CUDA;
    }
    void *p = nullptr;
    ITA_CUDA_TRY(cudaGetSymbolAddress(&p, g_layer_static));
    if (!p || (reinterpret_cast<uintptr_t>(p) & 255)) {
      set_error("ita_encoder_layer_forward_static: unusable address %p for the static scratch", p);
      return ITA_ERR_CUDA;
    }
    g_static_ptr[device] = static_cast<int8_t *>(p);
  }
  *out = g_static_ptr[device];
  return ITA_OK;
}

struct CurrentDevice {   // make `dev` current for the duration of the call, restore the caller's device afterwards
  int prev = -1;
  bool ok = true;
  explicit CurrentDevice(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; prev = -1; return; }
    if (dev != prev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~CurrentDevice() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace ita

using namespace ita;

extern "C" ITA_API int ita_encoder_layer_forward_static(ita_mha_handle *h, const ita_layer_params *lp, const float *x_dev, float *y_dev,
                                                        int64_t batch, void *cuda_stream) {
  const char *fn = "ita_encoder_layer_forward_static";
  if (!h || !lp || batch < 0 || (batch > 0 && (!x_dev || !y_dev)) || !lp->ln1_w || !lp->ln1_b || !lp->ln2_w || !lp->ln2_b) {
    set_error("%s: null handle / parameter / buffer or negative batch", fn);
    return ITA_ERR_INVALID;
  }
  if (!(lp->s_attn_in > 0.f) || !(lp->s_ffn_in > 0.f)) {
    set_error("%s: quantisation scales must be positive", fn);
    return ITA_ERR_INVALID;
  }
  if (!h->F) {
    set_error("%s: the handle was created without FFN parameters (F == 0)", fn);
    return ITA_ERR_INVALID;
  }
  if (h->mh || !h->fused || !h->ffn) {
    set_error("%s: needs the single-head tcgen05 attention and FFN kernels, which launch without allocating (attention: %s; ffn: %s)",
              fn, h->mh ? "multi-head kernels keep an allocated context buffer" : (h->fused ? "ok" : h->fused_reason.c_str()),
              h->ffn ? "ok" : h->ffn_reason.c_str());
    return ITA_ERR_UNSUPPORTED;
  }
  const size_t SE = (size_t)h->S * h->E;
  const size_t per_frame = 8 * SE;   // x1 (f32) + xq, yq, x1q, f2q (int8)
  const int64_t max_batch = (int64_t)(ITA_LAYER_STATIC_SCRATCH_BYTES / per_frame);
  if (batch > max_batch) {
    set_error("%s: batch %lld needs %zu bytes of intermediates, the static scratch holds %d (at most %lld frames of S=%d, E=%d); "
              "use ita_encoder_layer_forward for larger batches",
              fn, (long long)batch, per_frame * (size_t)batch, (int)ITA_LAYER_STATIC_SCRATCH_BYTES, (long long)max_batch, h->S, h->E);
    return ITA_ERR_INVALID;
  }
  if (batch == 0) return ITA_OK;
  CurrentDevice guard(h->device);
  if (!guard.ok) {
    set_error("%s: cannot make device %d current", fn, h->device);
    return ITA_ERR_CUDA;
  }
  int8_t *base = nullptr;
  int rc = static_scratch_ptr(h->device, &base);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)cuda_stream;
  const size_t n = (size_t)batch * SE;
  float *x1 = reinterpret_cast<float *>(base);
  int8_t *xq = base + 4 * n, *yq = xq + n, *x1q = yq + n, *f2q = x1q + n;
  // same five launches, same order and same arithmetic as layer_forward (layer.cu); only the home of the intermediates differs
  rc = launch_quantize(x_dev, xq, (int64_t)n, lp->s_attn_in, s);
  if (rc) return rc;
  rc = fused_forward(h, xq, yq, batch, nullptr, s);
  if (rc) return rc;
  rc = launch_resid_ln(x_dev, yq, lp->s_attn_out, lp->ln1_w, lp->ln1_b, lp->ln_eps, x1, x1q, lp->s_ffn_in, batch * h->S, h->E, s);
  if (rc) return rc;
  rc = ffn_fused_forward(h, x1q, nullptr, f2q, batch, s);
  if (rc) return rc;
  return launch_resid_ln(x1, f2q, lp->s_ffn_out, lp->ln2_w, lp->ln2_b, lp->ln_eps, y_dev, nullptr, 0.f, batch * h->S, h->E, s);
}
