// mh.cu -- host side of the multi-head S = 256 kernels (mh_kernel.cuh): requant form selection over every head, per-head
// weight blocks / offset tiles / output-projection panels, launch of k_ita_mh followed by k_ita_mh_out.
#include <cuda.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "mh_kernel.cuh"

namespace ita {

const char *rqm_name(int mode);   // fused.cu

// instantiations live in mh_inst.cu
MhKernelFn mh_pick(int E, int P, int mode, int softmax, bool dump);
MoKernelFn mo_pick(int E, int P, int mode, bool per_head, bool dump);
int mh_smem_bytes(int E, int P);

struct MhPlan {
  int E, P, H, mode, softmax;
  bool per_head;
  int smem_a, smem_b, wo_bytes;
  int torch_fast = 1;
  int8_t *wblocks = nullptr, *consts = nullptr, *wo = nullptr;
  FRq *rq_dev = nullptr;      // [H][6]
  FRq *rq_o_dev = nullptr;    // [H] stage O
  Rq rq_sum{};
  int8_t *o_scratch = nullptr;
  size_t o_bytes = 0;
  MhKernelFn fa = nullptr, fa_dump = nullptr;
  MoKernelFn fb = nullptr, fb_dump = nullptr;
};

template <class F>
static void to_panels(int8_t *dst, int R, int Ccols, F at) {
  for (int c = 0; c < Ccols; ++c)
    for (int r = 0; r < R; ++r) dst[(size_t)(c / 16) * R * 16 + (size_t)r * 16 + (c % 16)] = at(r, c);
}

static void mh_free(MhPlan *pl) {
  if (!pl) return;
  cudaFree(pl->wblocks); cudaFree(pl->consts); cudaFree(pl->wo); cudaFree(pl->rq_dev); cudaFree(pl->rq_o_dev); cudaFree(pl->o_scratch);
  delete pl;
}

int mh_plan_create(ita_mha_handle *h, const ita_mha_desc *d) {
  h->mh = nullptr;
  auto no = [&](const std::string &why) {
    h->fused_reason = why;
    return ITA_OK;
  };
  const int E = h->E, P = h->P, H = h->H, S = h->S;
  if (S != kMhS) return no("multi-head tcgen05 kernel needs S == 256");
  if (!(P == 64 && E == 128)) return no("multi-head tcgen05 kernel is instantiated for E = 128, P = 64");
  // k_ita_mh itself could walk up to MhCfg::kMaxH heads, but the output kernel is budgeted (three context tiles of H * P
  // columns in shared memory; per head, H * E / 2 tensor-memory columns) and checked for at most MoCfg::kMaxH
  static_assert(MoCfg<128, 64>::kMaxH <= MhCfg<128, 64>::kMaxH, "the output kernel's head limit is the pipeline's");
  if (H < 1 || H > MoCfg<128, 64>::kMaxH)
    return no("multi-head tcgen05 kernels run 1 <= H <= 4 heads (k_ita_mh_out keeps three context tiles of H * P columns in shared memory and, "
              "in the per-head form, H * E / 2 tensor-memory columns per unit)");
  const bool per_head = (h->head_mode == ITA_HEADS_PYITA) && H > 1;
  if (per_head && H * E / 2 > 256) return no("per-head output projection needs H * E / 2 <= 256 tensor-memory columns per unit");
  int cc_major = 0, max_smem = 0;
  ITA_CUDA_TRY(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, h->device));
  ITA_CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
  if (cc_major != 10) return no("multi-head tcgen05 kernel needs an sm_100 device");
  const int smem_a = mh_smem_bytes(E, P);
  const int smem_b = MoCfg<128, 64>::smem_bytes(H, per_head);
  if (smem_a > max_smem || smem_b > max_smem) return no("multi-head tcgen05 kernel needs more shared memory than the device offers");

  // ---- worst-case accumulator magnitudes per head and stage (Q K V QK AV O), largest |bias| per stage
  const int stages[6] = {ITA_ST_Q, ITA_ST_K, ITA_ST_V, ITA_ST_QK, ITA_ST_AV, ITA_ST_O};
  std::vector<long long> bound((size_t)H * 6, 0), bmax((size_t)H * 6, 0);
  for (int hd = 0; hd < H; ++hd) {
    const int8_t *Wm[3] = {d->Wq + (size_t)hd * E * P, d->Wk + (size_t)hd * E * P, d->Wv + (size_t)hd * E * P};
    const int32_t *Bs[3] = {d->Bq + (size_t)hd * P, d->Bk + (size_t)hd * P, d->Bv + (size_t)hd * P};
    for (int t = 0; t < 3; ++t) {
      long long worst = 0, bm = 0;
      for (int j = 0; j < P; ++j) {
        long long a = 0;
        for (int k = 0; k < E; ++k) a += std::abs((int)Wm[t][(size_t)k * P + j]);
        const long long bb = std::llabs((long long)Bs[t][j]);
        worst = std::max(worst, a * 128 + bb);
        bm = std::max(bm, bb);
      }
      bound[hd * 6 + t] = worst;
      bmax[hd * 6 + t] = bm;
    }
    bound[hd * 6 + 3] = (long long)P * 128 * 128;
    bound[hd * 6 + 4] = (h->softmax_mode == ITA_SOFTMAX_REPO_GLOBAL) ? 255LL * 128 : (long long)kMhS * 255 * 128;
  }
  // output projection: per head, or (concatenated form) one accumulator over every head with head 0's bias and constants
  for (int hd = 0; hd < H; ++hd) {
    long long worst = 0, bm = 0;
    for (int n = 0; n < E; ++n) {
      long long a = 0;
      for (int h2 = (per_head ? hd : 0); h2 < (per_head ? hd + 1 : H); ++h2)
        for (int k = 0; k < P; ++k) a += std::abs((int)d->Wo[((size_t)h2 * P + k) * E + n]);
      const long long bb = std::llabs((long long)d->Bo[(size_t)(per_head ? hd : 0) * E + n]);
      worst = std::max(worst, a * 128 + bb);
      bm = std::max(bm, bb);
    }
    bound[hd * 6 + 5] = worst;
    bmax[hd * 6 + 5] = bm;
  }
  for (long long b : bound)
    if (b >= (1LL << 30)) return no("bias magnitude too large for the s32 accumulate");

  // ---- requant form: the first one every head and stage admits
  const long long cap32 = digit_capacity(kBiasK), cap64 = digit_capacity(kBiasKO), cap96 = digit_capacity(96);
  std::vector<FRq> fr((size_t)H * 6);
  auto try_offset = [&](int mode) {
    for (int hd = 0; hd < H; ++hd)
      for (int i = 0; i < 6; ++i) {
        const Rq &r = h->rq_host[stages[i] * H + ((i == 5 && !per_head) ? 0 : hd)];
        const long long cap = (i == 3) ? (mode == RQM_TORCH_FP2 ? cap96 : cap64) : (i == 5 ? cap64 : cap32);
        FRq f{};
        const bool ok = mode == RQM_FP ? plan_fp_stage(r, bound[hd * 6 + i], bmax[hd * 6 + i], cap, &f)
                                       : plan_torch_stage(r, bound[hd * 6 + i], bmax[hd * 6 + i], cap, mode == RQM_TORCH_FP2 ? 2 : 1, &f);
        if (!ok) return false;
        fr[hd * 6 + i] = f;
      }
    return true;
  };
  int mode = -1;
  int knob = 2;
  if (const char *e = getenv("ITA_FUSED_FP")) knob = atoi(e);
  if (h->round_mode == ITA_ROUND_TORCH_RNE_F32) {
    for (int hd = 0; hd < H; ++hd)
      for (int i = 0; i < 6; ++i)
        if (!(h->rq_host[stages[i] * H + hd].eff > 0.f)) return no("multi-head tcgen05 kernel needs positive effective scales");
    if (knob >= 2 && try_offset(RQM_TORCH_FP2)) mode = RQM_TORCH_FP2;
    else if (knob >= 1 && try_offset(RQM_TORCH_FP)) mode = RQM_TORCH_FP;
    else mode = RQM_TORCH;
  } else {
    if (knob >= 1 && try_offset(RQM_FP)) mode = RQM_FP;
    else {
      mode = RQM_INT_GEN;
      for (int hd = 0; hd < H; ++hd)
        for (int i = 0; i < 6; ++i) {
          const Rq &r = h->rq_host[stages[i] * H + hd];
          if (r.shift < 3 || r.shift > 62 || r.mult < 0 || r.mult > 32767) return no("requant shift/mult outside the tcgen05 kernels' exact range");
        }
    }
  }
  if (!rqm_offset(mode))
    for (int hd = 0; hd < H; ++hd)
      for (int i = 0; i < 6; ++i) {
        const Rq &r = h->rq_host[stages[i] * H + ((i == 5 && !per_head) ? 0 : hd)];
        FRq f{};
        if (mode == RQM_TORCH) {
          f.add = r.add; f.eff = r.eff;
        } else {
          f.multp = r.mult; f.shift = r.shift; f.add = r.add; f.c = r.rnd;
        }
        fr[hd * 6 + i] = f;
      }
  const bool fpm = rqm_offset(mode);

  MhPlan *pl = new MhPlan();
  pl->E = E; pl->P = P; pl->H = H; pl->mode = mode; pl->softmax = h->softmax_mode; pl->per_head = per_head;
  pl->smem_a = smem_a; pl->smem_b = smem_b;
  if (const char *e = getenv("ITA_FUSED_TORCH_FAST")) pl->torch_fast = atoi(e);
  pl->rq_sum = h->rq_host[ITA_ST_SUM * H];
  pl->fa = mh_pick(E, P, mode, h->softmax_mode, false);
  pl->fa_dump = mh_pick(E, P, mode, h->softmax_mode, true);
  pl->fb = mo_pick(E, P, mode, per_head, false);
  pl->fb_dump = mo_pick(E, P, mode, per_head, per_head);
  if (!pl->fa || !pl->fa_dump || !pl->fb || !pl->fb_dump) {
    delete pl;
    return no("no multi-head kernel instance for this mode");
  }

  // ---- per-head weight blocks: [Wq_h | Wk_h | Wv_h]^T as [N = 3P rows, K = E + 32 cols] panels (digits: bias (+ offset))
  using CA = MhCfg<128, 64>;
  const int WB = CA::WB_BYTES, PW = 3 * P;
  std::vector<int8_t> wb((size_t)H * WB, 0), consts((size_t)1024 + (size_t)H * CA::OT_BYTES, 0);
  bool digits_ok = true;
  std::vector<int8_t> dig((size_t)std::max(PW, E) * kBiasKO, 0);
  for (int hd = 0; hd < H; ++hd) {
    const int8_t *Ws[3] = {d->Wq + (size_t)hd * E * P, d->Wk + (size_t)hd * E * P, d->Wv + (size_t)hd * E * P};
    const int32_t *Bs[3] = {d->Bq + (size_t)hd * P, d->Bk + (size_t)hd * P, d->Bv + (size_t)hd * P};
    std::fill(dig.begin(), dig.end(), 0);
    for (int t = 0; t < 3; ++t)
      for (int n = 0; n < P; ++n) {
        int8_t *g = &dig[(size_t)(t * P + n) * kBiasK];
        digits_ok = (fpm ? offset_digits((long long)Bs[t][n] + fr[hd * 6 + t].init, kBiasK, g) : bias_digits(Bs[t][n], g)) && digits_ok;
      }
    to_panels(wb.data() + (size_t)hd * WB, PW, E + kBiasK,
              [&](int n, int k) { return k < E ? Ws[n / P][(size_t)k * P + (n % P)] : dig[(size_t)n * kBiasK + (k - E)]; });
    if (fpm) {
      int8_t row[96];
      for (int k = 0; k < 96; ++k) row[k] = 0;
      digits_ok = offset_digits(fr[hd * 6 + 3].init, 32 * rqm_qk_steps(mode), row) && digits_ok;
      const_tile(consts.data() + 1024 + (size_t)hd * CA::OT_BYTES, 96, row);
      for (int k = 0; k < 96; ++k) row[k] = 0;
      digits_ok = offset_digits(fr[hd * 6 + 4].init, 32, row) && digits_ok;
      const_tile(consts.data() + 1024 + (size_t)hd * CA::OT_BYTES + CA::OT_DAV, 32, row);
    }
  }
  {
    int8_t row[96];
    for (int k = 0; k < 96; ++k) row[k] = (int8_t)(uint8_t)(k == 0 ? 1 : (fpm ? kBiasRadixU : kBiasRadix));
    const_tile(consts.data(), 96, row);
  }
  // ---- output projection panels: Wo^T [E rows, K = H*P] | digit blocks [ND][E x 64] | constant A tile
  const int nd = per_head ? H : 1;
  pl->wo_bytes = MoCfg<128, 64>::w_bytes(H, per_head);
  std::vector<int8_t> wo((size_t)pl->wo_bytes, 0);
  to_panels(wo.data(), E, H * P, [&](int n, int k) { return d->Wo[((size_t)(k / P) * P + (k % P)) * E + n]; });
  for (int b = 0; b < nd; ++b) {
    std::fill(dig.begin(), dig.end(), 0);
    for (int n = 0; n < E; ++n) {
      int8_t *g = &dig[(size_t)n * kBiasKO];
      const int32_t bias = d->Bo[(size_t)b * E + n];
      digits_ok = (fpm ? offset_digits((long long)bias + fr[b * 6 + 5].init, kBiasKO, g) : bias_digits(bias, g)) && digits_ok;
    }
    to_panels(wo.data() + (size_t)E * H * P + (size_t)b * E * kBiasKO, E, kBiasKO, [&](int n, int k) { return dig[(size_t)n * kBiasKO + k]; });
  }
  memcpy(wo.data() + (size_t)E * H * P + (size_t)nd * E * kBiasKO, consts.data(), kConstABytes);
  if (!digits_ok) {
    delete pl;
    return no("a bias exceeds the range the tcgen05 kernels fold into the MMA");
  }
  std::vector<FRq> fo((size_t)H);
  for (int hd = 0; hd < H; ++hd) fo[hd] = fr[hd * 6 + 5];

  auto up = [&](void **dst, const void *src, size_t bytes) {
    if (cudaMalloc(dst, bytes) != cudaSuccess) return false;
    return cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice) == cudaSuccess;
  };
  if (!up((void **)&pl->wblocks, wb.data(), wb.size()) || !up((void **)&pl->consts, consts.data(), consts.size()) ||
      !up((void **)&pl->wo, wo.data(), wo.size()) || !up((void **)&pl->rq_dev, fr.data(), fr.size() * sizeof(FRq)) ||
      !up((void **)&pl->rq_o_dev, fo.data(), fo.size() * sizeof(FRq))) {
    cudaGetLastError();
    mh_free(pl);
    set_error("device allocation for the multi-head kernel parameters failed");
    return ITA_ERR_NOMEM;
  }
  cudaError_t e = cudaSuccess;
  for (const void *f : {(const void *)pl->fa, (const void *)pl->fa_dump}) {
    cudaError_t r = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_a);
    if (r != cudaSuccess) e = r;
  }
  for (const void *f : {(const void *)pl->fb, (const void *)pl->fb_dump}) {
    cudaError_t r = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_b);
    if (r != cudaSuccess) e = r;
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    mh_free(pl);
    return no(std::string("cudaFuncSetAttribute(max dynamic smem) failed: ") + cudaGetErrorString(e));
  }
  h->mh = pl;
  return ITA_OK;
}

void mh_plan_destroy(ita_mha_handle *h) {
  mh_free(h->mh);
  h->mh = nullptr;
}

std::string describe_mh(const ita_mha_handle *h) {
  char b[320];
  const MhPlan *pl = h->mh;
  snprintf(b, sizeof(b),
           "attention: k_ita_mh<E=%d,P=%d> (S=256, H=%d) %s softmax=%d, %d threads, %d B smem + k_ita_mh_out (%s output projection, %d B smem)",
           pl->E, pl->P, pl->H, rqm_name(pl->mode), pl->softmax, kMhThreads, pl->smem_a, pl->per_head ? "per-head + head sum" : "concatenated",
           pl->smem_b);
  return b;
}

int mh_forward(ita_mha_handle *h, const int8_t *x, int8_t *y, int64_t B, const ita_mha_stages *dump, cudaStream_t s) {
  MhPlan *pl = h->mh;
  if (B <= 0) return ITA_OK;
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(y) & 15)) {
    set_error("tcgen05 path needs 16-byte aligned x and y");
    return ITA_ERR_INVALID;
  }
  const int E = pl->E, P = pl->P, H = pl->H;
  if (B * H * kMhS > 0x7FFFFFFFLL) {
    set_error("tcgen05 path: batch too large for one launch (batch*H*S must fit int32)");
    return ITA_ERR_INVALID;
  }
  int8_t *o = dump && dump->O ? dump->O : nullptr;
  if (!o) {
    const size_t need = (size_t)B * H * kMhS * P;
    if (need > pl->o_bytes) {
      if (pl->o_scratch) cudaFree(pl->o_scratch);
      pl->o_scratch = nullptr;
      pl->o_bytes = 0;
      if (cudaMalloc(&pl->o_scratch, need) != cudaSuccess) {
        cudaGetLastError();
        set_error("cudaMalloc(%zu) for the context scratch failed", need);
        return ITA_ERR_NOMEM;
      }
      pl->o_bytes = need;
    }
    o = pl->o_scratch;
  }
  CUtensorMap xmap, omap;
  int rc = make_x_tensor_map(&xmap, x, B * kMhS, E);
  if (rc) return rc;
  rc = make_x_tensor_map(&omap, o, B * H * kMhS, P);
  if (rc) return rc;
  MhParams pa{};
  pa.wblocks = pl->wblocks; pa.consts = pl->consts; pa.rq = pl->rq_dev; pa.x = x; pa.o = o; pa.batch = B; pa.H = H;
  pa.abort_flag = h->abort_flag; pa.torch_fast = pl->torch_fast;
  if (dump) {
    pa.dQ = dump->Qp; pa.dK = dump->Kp; pa.dV = dump->Vp; pa.dA = dump->A; pa.dAsm = dump->Asm;
  }
  long long ctas = std::min<long long>(B, h->sm_count);
#ifdef ITA_TRACE
  static long long *trace_dev = nullptr;
  const size_t trace_n = (size_t)(640 / 32) * (kTraceIters + 1) * 32 * 4;
  const char *trace_file = getenv("ITA_TRACE_FILE");
  if (!trace_dev) cudaMalloc(&trace_dev, trace_n * sizeof(long long));
  if (trace_file) cudaMemsetAsync(trace_dev, 0, trace_n * sizeof(long long), s);
  pa.trace = trace_dev;
#endif
  (dump ? pl->fa_dump : pl->fa)<<<(unsigned)ctas, kMhThreads, pl->smem_a, s>>>(xmap, pa);
  count_launch();
  ITA_CUDA_TRY(cudaGetLastError());
#ifdef ITA_TRACE
  if (trace_file) {
    std::vector<long long> t(trace_n);
    cudaStreamSynchronize(s);
    cudaMemcpy(t.data(), trace_dev, trace_n * sizeof(long long), cudaMemcpyDeviceToHost);
    if (FILE *fp = fopen(trace_file, "wb")) {
      fwrite(t.data(), sizeof(long long), trace_n, fp);
      fclose(fp);
    }
  }
#endif
  MoParams pb{};
  pb.wo = pl->wo; pb.rq = pl->rq_o_dev; pb.rq_sum = pl->rq_sum; pb.y = y; pb.dOut = dump ? dump->Out : nullptr; pb.batch = B; pb.H = H;
  pb.abort_flag = h->abort_flag;
  ctas = std::min<long long>(B * kMhTiles, h->sm_count);
  (dump ? pl->fb_dump : pl->fb)<<<(unsigned)ctas, kMoThreads, pl->smem_b, s>>>(omap, pb);
  count_launch();
  ITA_CUDA_TRY(cudaGetLastError());
  if (dump && dump->Out && !pl->per_head)      // concatenated form: Out [B,1,S,E] is the layer output
    ITA_CUDA_TRY(cudaMemcpyAsync(dump->Out, y, (size_t)B * kMhS * E, cudaMemcpyDeviceToDevice, s));
  return ITA_OK;
}

}  // namespace ita
