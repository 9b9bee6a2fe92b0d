// fused.cu -- host side of the fused attention kernel (fused_kernel.cuh): parameter packing (weight panels with the
// bias-digit K columns, per-stage uniform requant constants), shape/parameter eligibility, launch.
#include <cuda.h>
#include <math.h>
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "fused_kernel.cuh"

namespace ita {

#ifdef ITA_FUSED_FAST_BUILD   // developer builds: only the flagship shape
extern const FusedShape fused_shape_64_192;
static const FusedShape *const kShapes[] = {&fused_shape_64_192};
#else
extern const FusedShape fused_shape_64_192, fused_shape_64_64, fused_shape_64_128, fused_shape_128_64, fused_shape_128_192, fused_shape_128_128;
static const FusedShape *const kShapes[] = {&fused_shape_64_192,  &fused_shape_64_64,   &fused_shape_64_128,
                                            &fused_shape_128_64, &fused_shape_128_192, &fused_shape_128_128};
#endif

typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  }
  return fn;
}

int make_x_tensor_map(CUtensorMap *map, const void *x, long long rows, int row_bytes) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled is unavailable");
    return ITA_ERR_CUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)row_bytes, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)row_bytes};
  cuuint32_t box[2] = {16, (cuuint32_t)kS};
  cuuint32_t estr[2] = {1, 1};
  CUresult cr = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(x), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (cr != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)cr);
    return ITA_ERR_CUDA;
  }
  return ITA_OK;
}

struct FusedPlan {
  int E, P, mode, softmax, alt, slots, threads;
  int torch_fast = 1;          // RQM_TORCH_FP: fast ITAmax rows (tuning knob ITA_FUSED_TORCH_FAST, read at handle creation)
  int smem_bytes;
  int8_t *wpanels = nullptr;   // device
  FRq rq[6];
  KernelFn fn = nullptr, fn_dump = nullptr;
};

// panel layout of a [R x C] matrix given element accessor
template <class F>
static void to_panels(int8_t *dst, int R, int Ccols, F at) {
  for (int c = 0; c < Ccols; ++c)
    for (int r = 0; r < R; ++r) dst[(size_t)(c / 16) * R * 16 + (size_t)r * 16 + (c % 16)] = at(r, c);
}

// int32 bias -> kBiasK signed digits g with  b = g[0] + kBiasRadix * (g[1] + ... + g[kBiasK-1]),  every g in [-127, 127]
bool bias_digits(int32_t b, int8_t *g) {
  long long q = (b >= 0 ? (b + kBiasRadix / 2) : -((-(long long)b + kBiasRadix / 2))) / kBiasRadix;   // round to nearest
  long long rem = (long long)b - q * kBiasRadix;                                                      // |rem| <= 63
  g[0] = (int8_t)rem;
  for (int j = 1; j < kBiasK; ++j) {
    long long t = q > 127 ? 127 : (q < -127 ? -127 : q);
    g[j] = (int8_t)t;
    q -= t;
  }
  return q == 0;
}

// non-negative value -> n unsigned digits g with  v = g[0] + kBiasRadixU * (g[1] + ... + g[n-1]),  every g in [0, 255]
bool offset_digits(long long v, int n, int8_t *g) {
  if (v < 0) return false;
  long long q = v / kBiasRadixU;
  g[0] = (int8_t)(uint8_t)(v % kBiasRadixU);
  for (int j = 1; j < n; ++j) {
    long long t = q > 255 ? 255 : q;
    g[j] = (int8_t)(uint8_t)t;
    q -= t;
  }
  return q == 0;
}
// 8-row constant tile [8 x ncols] whose rows are all `row`, in the panel layout ([ncols/16][8][16])
void const_tile(int8_t *dst, int ncols, const int8_t *row) {
  for (int c = 0; c < ncols; ++c)
    for (int r = 0; r < 8; ++r) dst[(size_t)(c / 16) * 128 + r * 16 + (c % 16)] = row[c];
}


long long digit_capacity(int ncols) { return (long long)kBiasRadixU - 1 + (long long)kBiasRadixU * 255 * (ncols - 1); }

// RQM_FP constants of one stage (requantisation on the FMA pipe, fused_kernel.cuh): an accumulator offset INIT with
// INIT*mult == rnd (mod 2^shift), INIT > |acc|max (bnd), INIT + |acc|max < 2^24, and C = (rnd - INIT*mult)/2^shift + add + 128
// with C < 0 and 257 - C < 2^24 (the ITAmax distance constant m + 129 - C must stay in the linear f32 range); the offset plus
// the largest bias (bmax) must fit the unsigned digits of the stage's digit K steps (cap).
bool plan_fp_stage(const Rq &r, long long bnd, long long bmax, long long cap, FRq *out) {
  if (r.mult <= 0 || r.shift < 1 || r.shift > 30) return false;
  int t = 0;
  while (((r.mult >> t) & 1) == 0) ++t;
  const int mb = r.shift - t;                                  // INIT is determined modulo 2^mb
  const long long rnd = r.rnd;
  if (mb < 0 || (rnd & ((1LL << t) - 1))) return false;
  const long long lo = (bnd + 1 > (1LL << 20)) ? bnd + 1 : (1LL << 20);
  long long init = lo;
  if (mb > 0) {
    const long long mo = r.mult >> t, mask = (1LL << mb) - 1;
    unsigned long long inv = 1;
    for (int k = 0; k < 6; ++k) inv = inv * (2ULL - (unsigned long long)mo * inv);   // Newton: inverse of the odd part modulo 2^64
    const long long k0 = (long long)((((unsigned long long)(rnd >> t) & mask) * (inv & mask)) & mask);
    init = lo + ((k0 - lo) & mask);                            // smallest value >= lo congruent to k0
  }
  if ((init * r.mult - rnd) & ((1LL << r.shift) - 1)) return false;
  const long long Cc = (rnd - init * r.mult) / (1LL << r.shift) + r.add + 128;
  if (init + bnd >= (1LL << 24) || init + bmax > cap || Cc >= 0 || 257 - Cc >= (1LL << 24)) return false;
  out->init = (int32_t)init;
  out->negc = (int32_t)-Cc;
  out->fb = ldexpf((float)r.mult, -r.shift);
  out->fc = ldexpf((float)Cc, -149);
  return (long long)ldexpf(out->fc, 149) == Cc;
}

// The fp32 requant form on offset accumulators.  form 2 (RQM_TORCH_FP2, two FFMAs per element): the offset is a power of
// two; form 1 (RQM_TORCH_FP, three): any INIT > |acc|max that fits.
bool plan_torch_stage(const Rq &r, long long bnd, long long bmax, long long cap, int form, FRq *out) {
  long long init = (bnd + 1 > (1LL << 20)) ? bnd + 1 : (1LL << 20);
  if (form == 2)
    while (init & (init - 1)) init += init & -init;          // round up to a power of two
  const long long Cc = (long long)r.add + 128;
  // (zp + 128 must be even: the final FFMA rounds t + zp + 128 to nearest-even, which equals rne(t) + zp + 128 on ties only
  //  then; the reference's zero-points are all 0 -- symmetric quantisation, layers.py:10-21)
  if (init + bnd >= (1LL << 24) || init + bmax > cap || Cc < 0 || (Cc & 1) || Cc >= (1LL << 23) || !(r.eff > 0.f)) return false;
  out->init = (int32_t)init;
  out->negc = (int32_t)-Cc;
  out->fc = ldexpf((float)Cc, -149);
  if (form == 2) {
    // fb = eff * 2^85, eff = -INIT * eff * 2^-64 (exact: INIT is a power of two); both must be normal and exact scalings
    const float b1 = ldexpf(r.eff, 85), c1 = -ldexpf(r.eff * (float)init, -64);
    if (!isfinite(b1) || ldexpf(b1, -85) != r.eff || !(fabsf(c1) >= 1e-30f) || -ldexpf(c1, 64) != r.eff * (float)init || r.eff < 1e-9f ||
        r.eff > 256.f)
      return false;
    out->fb = b1;
    out->eff = c1;
  } else {
    const float effs = ldexpf(r.eff, 23);
    if (!isfinite(effs) || ldexpf(effs, -23) != r.eff) return false;
    out->fb = -ldexpf((float)init, -23);
    out->eff = effs;
  }
  return true;
}

int fused_plan_create(ita_mha_handle *h, const ita_mha_desc *d) {
  h->fused = nullptr;
  auto no = [&](const std::string &why) {
    h->fused_reason = why;
    return ITA_OK;
  };
  if (h->S == 256) return mh_plan_create(h, d);      // multi-head kernels (mh.cu)
  if (h->S != kS) return no("fused kernel needs S == 128 (or S == 256 for the multi-head kernels)");
  if (h->H != 1) return no("fused kernel needs H == 1");
  const int E = h->E, P = h->P;
  const FusedShape *shape = nullptr;
  for (const FusedShape *s : kShapes)
    if (s->E == E && s->P == P) shape = s;
  if (!shape) return no("fused kernel is instantiated for (E,P) in {(64,192),(64,64),(64,128),(128,64),(128,192),(128,128)}");
  int max_smem = 0;
  ITA_CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
  if (shape->smem_bytes > max_smem) return no("fused kernel needs more shared memory than the device offers");
  int cc_major = 0;
  ITA_CUDA_TRY(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, h->device));
  if (cc_major != 10) return no("fused kernel needs an sm_100 device");
  if (!get_encode()) return no("cuTensorMapEncodeTiled is unavailable");

  // ---- requant mode: every stage must be representable
  int mode;
  const int stages[6] = {ITA_ST_Q, ITA_ST_K, ITA_ST_V, ITA_ST_QK, ITA_ST_AV, ITA_ST_O};
  if (h->round_mode == ITA_ROUND_TORCH_RNE_F32) {
    mode = RQM_TORCH;
    for (int i = 0; i < 6; ++i)
      if (!(h->rq_host[stages[i] * h->H].eff > 0.f)) return no("fused kernel needs positive effective scales");
  } else {
    mode = RQM_INT_HI;
    for (int i = 0; i < 6; ++i) {
      const Rq &r = h->rq_host[stages[i] * h->H];
      if (r.shift < 9 || r.shift > 32 || r.mult < 0 || r.mult > 255) mode = RQM_INT_GEN;
    }
    if (mode == RQM_INT_GEN)
      for (int i = 0; i < 6; ++i) {
        const Rq &r = h->rq_host[stages[i] * h->H];
        if (r.shift < 3 || r.shift > 62 || r.mult < 0 || r.mult > 32767) return no("requant shift/mult outside the fused kernel's exact range");
      }
  }
  // ---- worst-case accumulator magnitude per stage (bias included); must stay well inside s32
  long long bound[6], bmaxs[6] = {0, 0, 0, 0, 0, 0};      // stage order Q K V QK AV O; bmaxs = largest |bias| of the stage
  {
    // |x| <= 128, so a column's accumulator is bounded by 128 * sum_k |W[k][n]| + |b[n]| (far below the generic K * 2^14
    // for real weights: the offsets of RQM_FP / RQM_TORCH_FP2 then fit one digit K step)
    const int8_t *Wm[4] = {d->Wq, d->Wk, d->Wv, d->Wo};
    const int32_t *Bs[4] = {d->Bq, d->Bk, d->Bv, d->Bo};
    const int n[4] = {P, P, P, E}, Kd[4] = {E, E, E, P}, stage[4] = {0, 1, 2, 5};
    for (int t = 0; t < 4; ++t) {
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
      bound[stage[t]] = worst;
      bmaxs[stage[t]] = bm;
    }
    bound[3] = (long long)P * 128 * 128;
    // A.V: with the repo ITAmax the probabilities of a row sum to <= 255 (sum_k (n_k*inv >> 16) <= 255*sum n/sum n)
    bound[4] = (h->softmax_mode == ITA_SOFTMAX_REPO_GLOBAL) ? 255LL * 128 : (long long)kS * 255 * 128;
    for (int i = 0; i < 6; ++i)
      if (bound[i] >= (1LL << 30)) return no("bias magnitude too large for the fused kernel's s32 accumulate");
  }

  // ---- RQM_FP (requantisation on the FMA pipe, fused_kernel.cuh): per stage an accumulator offset INIT with
  //      INIT*mult == rnd (mod 2^shift), INIT > |acc|max, INIT + |acc|max < 2^24, and C = (rnd - INIT*mult)/2^shift + add + 128
  //      with C < 0 and 257 - C < 2^24 (the ITAmax distance constant m + 129 - C must stay in the linear f32 range); the offset
  //      (plus the bias) must fit the unsigned digits of the stage's digit K steps.
  FRq fp[6] = {};
  bool fp_ok = (mode == RQM_INT_HI);
  if (const char *e = getenv("ITA_FUSED_FP")) fp_ok = fp_ok && atoi(e) != 0;    // tuning knob: 0 = keep the integer forms
  const long long cap32 = digit_capacity(kBiasK), cap64 = digit_capacity(kBiasKO), cap96 = digit_capacity(96);
  for (int i = 0; i < 6 && fp_ok; ++i) {
    // the stage's digit capacity: bias + offset (projections: one step; output projection, logits: two; context: one)
    fp_ok = plan_fp_stage(h->rq_host[stages[i] * h->H], bound[i], bmaxs[i], (i == 3 || i == 5) ? cap64 : cap32, &fp[i]);
  }
  if (fp_ok) mode = RQM_FP;
  // ---- the fp32 requant form on offset accumulators.  RQM_TORCH_FP2 (two FFMAs per element) needs every stage's offset to
  //      be a power of two that fits the stage's digit steps (one for the projections and A.V, two for the output projection,
  //      three for Q.K^T); RQM_TORCH_FP (three) takes any INIT > |acc|max that fits (two steps for Q.K^T).
  if (mode == RQM_TORCH) {
    bool allow = true;
    if (const char *e = getenv("ITA_FUSED_FP")) allow = atoi(e) != 0;   // 0: conversion form; 1: three-FFMA form; default / 2: two-FFMA form when eligible
    const bool want2 = !getenv("ITA_FUSED_FP") || atoi(getenv("ITA_FUSED_FP")) >= 2;
    for (int form = want2 ? 2 : 1; form >= 1 && allow && mode == RQM_TORCH; --form) {
      bool ok = true;
      for (int i = 0; i < 6 && ok; ++i) {
        const long long cap = (i == 3) ? (form == 2 ? cap96 : cap64) : (i == 5 ? cap64 : cap32);
        ok = plan_torch_stage(h->rq_host[stages[i] * h->H], bound[i], bmaxs[i], cap, form, &fp[i]);
      }
      if (ok) mode = form == 2 ? RQM_TORCH_FP2 : RQM_TORCH_FP;
    }
  }

  FusedPlan *pl = new FusedPlan();
  pl->E = E; pl->P = P; pl->mode = mode; pl->softmax = h->softmax_mode; pl->smem_bytes = shape->smem_bytes;
  pl->slots = shape->slots; pl->threads = shape->threads;
  if (const char *e = getenv("ITA_FUSED_TORCH_FAST")) pl->torch_fast = atoi(e);
  const int WA = P * (E + kBiasK), WO = E * (P + kBiasKO);
  const size_t wbytes = (size_t)3 * WA + WO + kConstBytes;
  std::vector<int8_t> host(wbytes, 0);
  const bool fpm = rqm_offset(mode);
  // stage index (Q K V QK AV O) of the three concatenated projections Q | K | V
  const int cat_stage[3] = {0, 1, 2};
  // [Wq | Wk | Wv]^T as ONE matrix [N = 3P rows, K = E + 32 cols] (the kernel's projection chunks are 128 consecutive rows of
  // it): row n = t*P + j holds column j of W_t (golden layout W [E,P]), then the digits of that column's bias (RQM_FP:
  // bias + the stage's offset, unsigned)
  const int8_t *Ws[3] = {d->Wq, d->Wk, d->Wv};
  const int32_t *Bs[3] = {d->Bq, d->Bk, d->Bv};
  bool digits_ok = true;
  std::vector<int8_t> dig((size_t)(3 * P > E ? 3 * P : E) * kBiasKO, 0);
  for (int t = 0; t < 3; ++t)
    for (int n = 0; n < P; ++n) {
      int8_t *g = &dig[(size_t)(t * P + n) * kBiasK];
      digits_ok = (fpm ? offset_digits((long long)Bs[t][n] + fp[cat_stage[t]].init, kBiasK, g) : bias_digits(Bs[t][n], g)) && digits_ok;
    }
  to_panels(host.data(), 3 * P, E + kBiasK, [&](int n, int k) {
    return k < E ? Ws[n / P][(size_t)k * P + (n % P)] : dig[(size_t)n * kBiasK + (k - E)];
  });
  // Wo^T as [N = E rows, K = P + 64 cols]: element (n, k) = Wo[k][n]  (golden layout Wo [P,E]); 64 digit columns (the integer
  // forms use the first 32)
  std::fill(dig.begin(), dig.end(), 0);
  for (int n = 0; n < E; ++n) {
    int8_t *g = &dig[(size_t)n * kBiasKO];
    digits_ok = (fpm ? offset_digits((long long)d->Bo[n] + fp[5].init, kBiasKO, g) : bias_digits(d->Bo[n], g)) && digits_ok;
  }
  to_panels(host.data() + (size_t)3 * WA, E, P + kBiasKO,
            [&](int n, int k) { return k < P ? d->Wo[(size_t)k * E + n] : dig[(size_t)n * kBiasKO + (k - P)]; });
  // constant tiles: A operand of the digit steps [1, R, ..., R]; RQM_FP: the offset digits of Q.K^T and A.V
  {
    int8_t row[96];
    for (int k = 0; k < 96; ++k) row[k] = (int8_t)(uint8_t)(k == 0 ? 1 : (fpm ? kBiasRadixU : kBiasRadix));
    int8_t *ca = host.data() + (size_t)3 * WA + WO;
    const_tile(ca, 96, row);
    if (fpm) {
      for (int k = 0; k < 96; ++k) row[k] = 0;
      digits_ok = offset_digits(fp[3].init, 32 * rqm_qk_steps(mode), row) && digits_ok;
      const_tile(ca + kConstABytes, 96, row);
      digits_ok = offset_digits(fp[4].init, 32, row) && digits_ok;
      const_tile(ca + kConstABytes + kConstQKBytes, 32, row);
    }
  }
  if (!digits_ok) {
    delete pl;
    return no("a bias exceeds the range the fused kernel folds into the MMA (|b| <= 127*31*127)");
  }

  // ---- per-stage uniform requant constants; the 32-bit INT_HI form needs (|acc|max * mult + c32) and the softmax
  //      distance constant ((128 + 1) << shift) to stay inside s32
  bool alt_ok = (mode == RQM_INT_HI);
  for (int i = 0; i < 6; ++i) {
    const Rq &r = h->rq_host[stages[i] * h->H];
    FRq f{};
    if (mode == RQM_TORCH) {
      f.add = r.add; f.eff = r.eff;
    } else if (rqm_offset(mode)) {
      f = fp[i];
    } else if (mode == RQM_INT_HI) {
      f.multp = (int32_t)((long long)r.mult << (32 - r.shift));
      f.shift = r.shift; f.add = r.add;
      f.c = (r.rnd << (32 - r.shift)) + ((long long)r.add << 32);
      const long long c32 = r.rnd + ((long long)r.add << r.shift);
      const long long worst = bound[i] * r.mult + (c32 < 0 ? -c32 : c32) + (258LL << r.shift);
      if (r.shift > 30 || worst >= (1LL << 31)) {
        alt_ok = false;
      } else {
        f.mult32 = r.mult;
        f.c32 = (int32_t)c32;
      }
    } else {
      f.multp = r.mult; f.shift = r.shift; f.add = r.add; f.c = r.rnd;
    }
    pl->rq[i] = f;
  }
  int alt = alt_ok ? 2 : 0;
  if (const char *e = getenv("ITA_FUSED_ALT")) {   // tuning knob: force the INT_HI element form (0, 1 or 2)
    const int want = atoi(e);
    if (want == 0 || (alt_ok && (want == 1 || want == 2))) alt = want;
  }
  pl->fn = shape->pick(mode, h->softmax_mode, alt, false);
  if (!pl->fn) {
    alt = alt_ok ? 2 : 0;
    pl->fn = shape->pick(mode, h->softmax_mode, alt, false);
  }
  pl->alt = alt;
  pl->fn_dump = shape->pick(mode, h->softmax_mode, 0, true);
  if (!pl->fn || !pl->fn_dump) {
    delete pl;
    return no("no fused kernel instance for this mode");
  }

  if (cudaMalloc(&pl->wpanels, wbytes) != cudaSuccess) {
    delete pl;
    set_error("cudaMalloc(%zu) for the fused weight panels failed", wbytes);
    return ITA_ERR_NOMEM;
  }
  if (cudaMemcpy(pl->wpanels, host.data(), wbytes, cudaMemcpyHostToDevice) != cudaSuccess) {
    cudaFree(pl->wpanels);
    delete pl;
    set_error("cudaMemcpy of the fused weight panels failed");
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
  h->fused = pl;
  return ITA_OK;
}

static const char *const kModeNames[] = {"RQM_INT_HI", "RQM_INT_GEN", "RQM_TORCH", "RQM_FP", "RQM_TORCH_FP", "RQM_TORCH_FP2"};
const char *rqm_name(int mode) { return mode >= 0 && mode <= 5 ? kModeNames[mode] : "?"; }

std::string describe_fused(const ita_mha_handle *h) {
  char b[256];
  if (h->mh) return describe_mh(h);
  if (!h->fused) {
    snprintf(b, sizeof(b), "attention: staged CUDA-core kernels (%s)", h->fused_reason.c_str());
  } else {
    const FusedPlan *pl = h->fused;
    snprintf(b, sizeof(b), "attention: k_ita_fused<E=%d,P=%d> %s softmax=%d alt=%d, %d frame(s) in flight per SM, %d threads, %d B smem",
             pl->E, pl->P, rqm_name(pl->mode), pl->softmax, pl->alt, pl->slots, pl->threads, pl->smem_bytes);
  }
  return b;
}

void fused_plan_destroy(ita_mha_handle *h) {
  mh_plan_destroy(h);
  if (h->fused) {
    cudaFree(h->fused->wpanels);
    delete h->fused;
    h->fused = nullptr;
  }
}

int fused_forward(ita_mha_handle *h, const int8_t *x, int8_t *y, int64_t B, const ita_mha_stages *dump, cudaStream_t s) {
  if (h->mh) return mh_forward(h, x, y, B, dump, s);
  FusedPlan *pl = h->fused;
  if (!pl) {
    set_error("fused path unavailable: %s", h->fused_reason.c_str());
    return ITA_ERR_UNSUPPORTED;
  }
  if (B <= 0) return ITA_OK;
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(y) & 15)) {
    set_error("fused path needs 16-byte aligned x and y");
    return ITA_ERR_INVALID;
  }
  if (B * kS > 0x7FFFFFFFLL) {
    set_error("fused path: batch too large for one launch (batch*S must fit int32)");
    return ITA_ERR_INVALID;
  }
  const int E = pl->E;
  // X [B*S rows, E bytes] as a rank-2 tensor; a (16 B x 128 rows) box is one panel of a frame's tile
  CUtensorMap map;
  int rc = make_x_tensor_map(&map, x, B * kS, E);
  if (rc) return rc;
  FusedParams prm{};
  prm.wpanels = pl->wpanels;
  prm.x = x;
  prm.y = y;
  prm.batch = B;
  prm.abort_flag = h->abort_flag;
  prm.torch_fast = pl->torch_fast;
  for (int i = 0; i < 6; ++i) prm.rq[i] = pl->rq[i];
  if (dump) {
    prm.dQ = dump->Qp; prm.dK = dump->Kp; prm.dV = dump->Vp; prm.dA = dump->A; prm.dAsm = dump->Asm; prm.dO = dump->O;
  }
  long long ctas = (B + pl->slots - 1) / pl->slots;
  if (ctas > h->sm_count) ctas = h->sm_count;
  KernelFn fn = dump ? pl->fn_dump : pl->fn;
#ifdef ITA_TRACE
  static long long *trace_dev = nullptr;
  const size_t trace_n = (size_t)(640 / 32) * (kTraceIters + 1) * 32 * 4;
  const char *trace_file = getenv("ITA_TRACE_FILE");
  if (!trace_dev) cudaMalloc(&trace_dev, trace_n * sizeof(long long));
  if (trace_file) cudaMemsetAsync(trace_dev, 0, trace_n * sizeof(long long), s);
  prm.trace = trace_dev;
#endif
  fn<<<(unsigned)ctas, pl->threads, pl->smem_bytes, s>>>(map, prm);
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
  return ITA_OK;
}

}  // namespace ita
