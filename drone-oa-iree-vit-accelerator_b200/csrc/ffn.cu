// ffn.cu -- host side of the fused feed-forward kernel (ffn_kernel.cuh): eligibility, weight panels with the bias-digit
// K columns, per-stage requant constants, launch; plus the kernel instantiations.
#include <math.h>
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "ffn_kernel.cuh"

namespace ita {

ITA_FFN_DEFINE_SHAPE(64, 256)
ITA_FFN_DEFINE_SHAPE(128, 256)
static const FfnShape *const kFfnShapes[] = {&ffn_shape_64_256, &ffn_shape_128_256};

struct FfnPlan {
  int E, F, mode, smem_bytes, slots, threads;
  int8_t *wpanels = nullptr;
  FRq rq[2];
  FfnKernelFn fn = nullptr, fn_dump = nullptr;
};

template <class Fn>
static void to_panels_ffn(int8_t *dst, int R, int Ccols, Fn at) {
  for (int c = 0; c < Ccols; ++c)
    for (int r = 0; r < R; ++r) dst[(size_t)(c / 16) * R * 16 + (size_t)r * 16 + (c % 16)] = at(r, c);
}

int ffn_plan_create(ita_mha_handle *h, const ita_mha_desc *d) {
  h->ffn = nullptr;
  auto no = [&](const std::string &why) {
    h->ffn_reason = why;
    return ITA_OK;
  };
  if (!h->F) return no("handle has no FFN parameters");
  if (h->S != kS) return no("fused FFN kernel needs S == 128");
  const int E = h->E, F = h->F;
  const FfnShape *shape = nullptr;
  for (const FfnShape *s : kFfnShapes)
    if (s->E == E && s->F == F) shape = s;
  if (!shape) return no("fused FFN kernel is instantiated for (E,F) in {(64,256),(128,256)}");
  int max_smem = 0, cc_major = 0;
  ITA_CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
  ITA_CUDA_TRY(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, h->device));
  if (cc_major != 10 || shape->smem_bytes > max_smem) return no("fused FFN kernel needs an sm_100 device with 227 KB shared memory");

  const Rq &r1 = h->rq_host[ITA_ST_FC1 * h->H], &rr = h->rq_host[ITA_ST_RELU * h->H], &r2 = h->rq_host[ITA_ST_FC2 * h->H];
  // The second requant after the ReLU (golden model, W_B.py:278-301) must be the identity on [0, 127] -- it is for every
  // converted model (a quantised ReLU keeps its input's scale); anything else takes the staged kernels.
  if (rr.mult != 0) {
    for (int y = 0; y <= 127; ++y) {
      long long z;
      if (h->round_mode == ITA_ROUND_TORCH_RNE_F32) z = (long long)nearbyintf((float)y * rr.eff) + rr.add;
      else z = (((long long)y * rr.mult + rr.rnd) >> rr.shift) + rr.add;
      if (sat8(z) != y) return no("the ReLU requantisation is not the identity");
    }
  }
  // ---- accumulator bounds: 128 * sum_k |W[k][n]| + |b[n]|
  long long bound[2], bmax[2];
  {
    const int8_t *Wm[2] = {d->Wff, d->Wff2};
    const int32_t *Bs[2] = {d->Bff, d->Bff2};
    const int n[2] = {F, E}, Kd[2] = {E, F};
    for (int t = 0; t < 2; ++t) {
      long long worst = 0, bm = 0;
      for (int j = 0; j < n[t]; ++j) {
        long long a = 0;
        for (int k = 0; k < Kd[t]; ++k) {
          const int w = Wm[t][(size_t)k * n[t] + j];
          a += w < 0 ? -w : w;
        }
        const long long bb = Bs[t][j] < 0 ? -(long long)Bs[t][j] : Bs[t][j];
        worst = std::max(worst, a * 128 + bb);
        bm = std::max(bm, bb);
      }
      bound[t] = worst;
      bmax[t] = bm;
      if (worst >= (1LL << 30)) return no("bias magnitude too large for the s32 accumulate");
    }
  }
  const Rq *rs[2] = {&r1, &r2};
  const long long caps[2] = {digit_capacity(kBiasK), digit_capacity(kBiasKO)};
  FRq fp[2] = {};
  int mode;
  const char *env = getenv("ITA_FUSED_FP");
  const int envv = env ? atoi(env) : 2;
  if (h->round_mode == ITA_ROUND_TORCH_RNE_F32) {
    mode = RQM_TORCH;
    if (!(r1.eff > 0.f) || !(r2.eff > 0.f)) return no("fused FFN kernel needs positive effective scales");
    for (int form = envv >= 2 ? 2 : 1; form >= 1 && envv != 0 && mode == RQM_TORCH; --form)
      if (plan_torch_stage(r1, bound[0], bmax[0], caps[0], form, &fp[0]) && plan_torch_stage(r2, bound[1], bmax[1], caps[1], form, &fp[1]))
        mode = form == 2 ? RQM_TORCH_FP2 : RQM_TORCH_FP;
  } else {
    mode = RQM_INT_GEN;
    for (int t = 0; t < 2; ++t)
      if (rs[t]->shift < 0 || rs[t]->shift > 62 || rs[t]->mult < 0 || rs[t]->mult > 32767) return no("requant shift/mult outside the fused FFN kernel's exact range");
    if (envv != 0 && plan_fp_stage(r1, bound[0], bmax[0], caps[0], &fp[0]) && plan_fp_stage(r2, bound[1], bmax[1], caps[1], &fp[1])) mode = RQM_FP;
  }
  const bool fpm = rqm_offset(mode);

  FfnPlan *pl = new FfnPlan();
  pl->E = E; pl->F = F; pl->mode = mode; pl->smem_bytes = shape->smem_bytes; pl->slots = shape->slots; pl->threads = shape->threads;
  const int W1 = F * (E + kBiasK), W2 = E * (F + kBiasKO);
  const size_t wbytes = (size_t)W1 + W2 + 8 * 64;
  std::vector<int8_t> host(wbytes, 0);
  bool digits_ok = true;
  std::vector<int8_t> dig((size_t)std::max(E, F) * kBiasKO, 0);
  // W1^T [N = F rows, K = E + 32]: row n = column n of Wff (golden layout [E,F]) then the digits of its bias (+ offset)
  for (int n = 0; n < F; ++n) {
    int8_t *g = &dig[(size_t)n * kBiasK];
    digits_ok = (fpm ? offset_digits((long long)d->Bff[n] + fp[0].init, kBiasK, g) : bias_digits(d->Bff[n], g)) && digits_ok;
  }
  to_panels_ffn(host.data(), F, E + kBiasK, [&](int n, int k) { return k < E ? d->Wff[(size_t)k * F + n] : dig[(size_t)n * kBiasK + (k - E)]; });
  std::fill(dig.begin(), dig.end(), 0);
  // W2^T [N = E rows, K = F + 64]
  for (int n = 0; n < E; ++n) {
    int8_t *g = &dig[(size_t)n * kBiasKO];
    digits_ok = (fpm ? offset_digits((long long)d->Bff2[n] + fp[1].init, kBiasKO, g) : bias_digits(d->Bff2[n], g)) && digits_ok;
  }
  to_panels_ffn(host.data() + W1, E, F + kBiasKO, [&](int n, int k) { return k < F ? d->Wff2[(size_t)k * E + n] : dig[(size_t)n * kBiasKO + (k - F)]; });
  {
    int8_t row[64];
    for (int k = 0; k < 64; ++k) row[k] = (int8_t)(uint8_t)(k == 0 ? 1 : (fpm ? kBiasRadixU : kBiasRadix));
    const_tile(host.data() + W1 + W2, 64, row);
  }
  if (!digits_ok) {
    delete pl;
    return no("a bias exceeds the range the fused FFN kernel folds into the MMA");
  }
  for (int t = 0; t < 2; ++t) {
    FRq f{};
    if (mode == RQM_TORCH) {
      f.add = rs[t]->add; f.eff = rs[t]->eff;
    } else if (fpm) {
      f = fp[t];
    } else {
      f.multp = rs[t]->mult; f.shift = rs[t]->shift; f.add = rs[t]->add; f.c = rs[t]->rnd;
    }
    pl->rq[t] = f;
  }
  pl->fn = shape->pick(mode, false);
  pl->fn_dump = shape->pick(mode, true);
  if (!pl->fn || !pl->fn_dump) {
    delete pl;
    return no("no fused FFN kernel instance for this mode");
  }
  if (cudaMalloc(&pl->wpanels, wbytes) != cudaSuccess) {
    delete pl;
    set_error("cudaMalloc(%zu) for the FFN weight panels failed", wbytes);
    return ITA_ERR_NOMEM;
  }
  if (cudaMemcpy(pl->wpanels, host.data(), wbytes, cudaMemcpyHostToDevice) != cudaSuccess) {
    cudaFree(pl->wpanels);
    delete pl;
    set_error("cudaMemcpy of the FFN weight panels failed");
    return ITA_ERR_CUDA;
  }
  cudaError_t e1 = cudaFuncSetAttribute((const void *)pl->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, pl->smem_bytes);
  cudaError_t e2 = cudaFuncSetAttribute((const void *)pl->fn_dump, cudaFuncAttributeMaxDynamicSharedMemorySize, pl->smem_bytes);
  if (e1 != cudaSuccess || e2 != cudaSuccess) {
    cudaGetLastError();
    cudaFree(pl->wpanels);
    delete pl;
    return no(std::string("cudaFuncSetAttribute(max dynamic smem) failed: ") + cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
  }
  h->ffn = pl;
  return ITA_OK;
}

const char *rqm_name(int mode);   // fused.cu

std::string describe_ffn(const ita_mha_handle *h) {
  char b[256];
  if (!h->ffn) snprintf(b, sizeof(b), "ffn: staged CUDA-core kernels (%s)", h->ffn_reason.c_str());
  else
    snprintf(b, sizeof(b), "ffn: k_ita_ffn<E=%d,F=%d> %s, %d frames in flight per SM, %d threads, %d B smem", h->ffn->E, h->ffn->F,
             rqm_name(h->ffn->mode), h->ffn->slots, h->ffn->threads, h->ffn->smem_bytes);
  return b;
}

void ffn_plan_destroy(ita_mha_handle *h) {
  if (h->ffn) {
    cudaFree(h->ffn->wpanels);
    delete h->ffn;
    h->ffn = nullptr;
  }
}

int ffn_fused_forward(ita_mha_handle *h, const int8_t *x, int8_t *f1, int8_t *f2, int64_t B, cudaStream_t s) {
  FfnPlan *pl = h->ffn;
  if (!pl) {
    set_error("fused FFN path unavailable: %s", h->ffn_reason.c_str());
    return ITA_ERR_UNSUPPORTED;
  }
  if (B <= 0) return ITA_OK;
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(f2) & 15) || (reinterpret_cast<uintptr_t>(f1) & 15)) {
    set_error("fused FFN path needs 16-byte aligned buffers");
    return ITA_ERR_INVALID;
  }
  if (B * kS > 0x7FFFFFFFLL) {
    set_error("fused FFN path: batch too large for one launch (batch*S must fit int32)");
    return ITA_ERR_INVALID;
  }
  CUtensorMap map;
  int rc = make_x_tensor_map(&map, x, B * kS, pl->E);
  if (rc) return rc;
  FfnParams prm{};
  prm.wpanels = pl->wpanels;
  prm.y = f2;
  prm.f1 = f1;
  prm.batch = B;
  prm.abort_flag = h->abort_flag;
  prm.rq[0] = pl->rq[0];
  prm.rq[1] = pl->rq[1];
  long long ctas = (B + pl->slots - 1) / pl->slots;
  if (ctas > h->sm_count) ctas = h->sm_count;
  FfnKernelFn fn = f1 ? pl->fn_dump : pl->fn;
  fn<<<(unsigned)ctas, pl->threads, pl->smem_bytes, s>>>(map, prm);
  count_launch();
  ITA_CUDA_TRY(cudaGetLastError());
  return ITA_OK;
}

}  // namespace ita
