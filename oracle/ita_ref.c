/* ita_ref (C) -- CPU restatement of the ITA int8 multi-head-attention integer specification.
 *
 * TEST INFRASTRUCTURE ONLY: the checker for tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs.  The product library (libita_b200.so) never links, loads or calls this file.
 *
 * It is the C twin of oracle/ita_ref.py (same arithmetic, same enumerations) with OpenMP over frames so that it can
 * serve as the timed host-CPU baseline.  Parity status is stated in the header of oracle/ita_ref.py:
 *   - ITAmax "repo_global", torch-rounding requant of the linear / A.V stages, hw-param derivation: PINNED against
 *     the importable reference (fixtures under tests/golden/).
 *   - round-half-up requant, streaming ITAmax, head-sum: restated from the absent golden model
 *     (OpenHardware-Initiative/ITA-FPGA @ 9529822b, PyITA/ITA.py) -- parity unpinned.
 *
 * Reference lines followed:
 *   stage order / layouts      models/ITA/QAT/layers.py:101-127, tests/export_and_validation_W_B.py:479-491,518-542
 *   ITAmax                     models/ITA/QAT/ITA_softmax.py:28-61
 *   u8 x s8 A.V + requant      tests/export_and_validation_W_B.py:120-151
 *   mult/shift requant form    scripts/export_model.py:132-135 (floor), SURVEY.md App. A.2 (half-up, torch)
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif

enum { ROUND_HALF_UP = 0, ROUND_FLOOR = 1, ROUND_TORCH_RNE_F32 = 2 };
enum { SOFTMAX_REPO_GLOBAL = 0, SOFTMAX_STREAM16 = 1 };
enum { HEADS_PYITA = 0, HEADS_TORCH = 1 };
enum { ST_Q = 0, ST_K, ST_V, ST_QK, ST_AV, ST_O, ST_SUM, ST_FC1, ST_RELU, ST_FC2, NUM_STAGES };

typedef struct ita_ref_params {
  int32_t S, E, P, H, F;
  int32_t round_mode, softmax_mode, head_mode;
  const int8_t *Wq, *Wk, *Wv;      /* [H,E,P] */
  const int8_t *Wo;                /* [H,P,E] */
  const int32_t *Bq, *Bk, *Bv;     /* [H,P]   */
  const int32_t *Bo;               /* [H,E]   */
  const int8_t *Wff, *Wff2;        /* [1,E,F], [1,F,E] (may be NULL) */
  const int32_t *Bff, *Bff2;       /* [1,F], [1,E]                   */
  const int32_t *mult, *shift, *add; /* [NUM_STAGES,H] */
  const float *eff;                /* [NUM_STAGES,H] (torch rounding only; may be NULL otherwise) */
} ita_ref_params;

typedef struct { int32_t mult, shift, add, mode; float eff; } rq_t;

static inline rq_t rq_of(const ita_ref_params *p, int st, int h) {
  rq_t r;
  r.mult = p->mult[st * p->H + h];
  r.shift = p->shift[st * p->H + h];
  r.add = p->add[st * p->H + h];
  r.mode = p->round_mode;
  r.eff = p->eff ? p->eff[st * p->H + h] : 0.f;
  return r;
}

static inline int8_t sat8(int64_t v) { return (int8_t)(v < -128 ? -128 : (v > 127 ? 127 : v)); }

/* SURVEY.md App. A.2 */
static inline int8_t requant(int64_t acc, rq_t r) {
  int64_t y;
  if (r.mode == ROUND_TORCH_RNE_F32) {
    float f = (float)acc * r.eff;           /* f32(acc) * f32(eff), one IEEE multiply           */
    y = (int64_t)nearbyintf(f) + r.add;     /* default rounding mode = round-half-even (rne)    */
  } else {
    int64_t prod = acc * (int64_t)r.mult;
    if (r.mode == ROUND_HALF_UP && r.shift > 0) prod += (int64_t)1 << (r.shift - 1);
    y = (prod >> r.shift) + r.add;          /* arithmetic shift == floor division by 2^shift    */
  }
  return sat8(y);
}

/* C[M,N] (s32) = A[M,K] (s8, or u8 when a_unsigned) . B[K,N] (s8) */
static void gemm_i8(const void *A_, int a_unsigned, const int8_t *B, int32_t *C, int M, int K, int N) {
  const int8_t *As = (const int8_t *)A_;
  const uint8_t *Au = (const uint8_t *)A_;
  for (int i = 0; i < M; ++i) {
    int32_t *c = C + (size_t)i * N;
    for (int j = 0; j < N; ++j) c[j] = 0;
    for (int k = 0; k < K; ++k) {
      const int32_t a = a_unsigned ? (int32_t)Au[(size_t)i * K + k] : (int32_t)As[(size_t)i * K + k];
      const int8_t *b = B + (size_t)k * N;
      for (int j = 0; j < N; ++j) c[j] += a * (int32_t)b[j];
    }
  }
}

/* models/ITA/QAT/ITA_softmax.py:51-61 -- one row */
static void itamax_repo_row(const int8_t *x, uint8_t *out, int n) {
  int m = -128;
  for (int j = 0; j < n; ++j) m = x[j] > m ? x[j] : m;
  int64_t tot = 0;
  for (int j = 0; j < n; ++j) { int d = m - x[j]; tot += d <= 8 ? (256 >> d) : 0; }
  if (tot < 1) tot = 1;
  const int64_t inv = (255LL * 65536LL) / tot;
  for (int j = 0; j < n; ++j) {
    int d = m - x[j];
    int64_t num = d <= 8 ? (256 >> d) : 0;
    out[j] = (uint8_t)((num * inv) >> 16);
  }
}

/* recalled upstream PyITA streamingPartialSoftmax (integerised, 16-column groups) -- one row; parity unpinned */
static void itamax_stream_row(const int8_t *x, uint8_t *out, int n) {
  int m = -128;
  int64_t tot = 0;
  for (int g = 0; g < n; g += 16) {
    int cm = -128;
    for (int j = g; j < g + 16; ++j) cm = x[j] > cm ? x[j] : cm;
    if (cm > m) { tot >>= ((cm - m + 16) >> 5); m = cm; }
    for (int j = g; j < g + 16; ++j) { int sh = (m - x[j] + 16) >> 5; tot += 256 >> sh; }
  }
  if (tot < 1) tot = 1;
  const int64_t inv = (255LL * 256LL) / tot;
  for (int j = 0; j < n; ++j) { int sh = (m - x[j] + 16) >> 5; out[j] = (uint8_t)(inv >> sh); }
}

static void frame_mha(const ita_ref_params *p, const int8_t *X, int8_t *Y,
                      int8_t *Qp, int8_t *Kp, int8_t *Vp, int8_t *A, uint8_t *Asm, int8_t *O, int8_t *Out,
                      int32_t *acc, int32_t *accsum, int8_t *Kt) {
  const int S = p->S, E = p->E, P = p->P, H = p->H;
  for (int h = 0; h < H; ++h) {
    int8_t *q = Qp + (size_t)h * S * P, *k = Kp + (size_t)h * S * P, *v = Vp + (size_t)h * S * P;
    int8_t *a = A + (size_t)h * S * S, *o = O + (size_t)h * S * P;
    uint8_t *asm_ = Asm + (size_t)h * S * S;
    const int8_t *W[3] = {p->Wq + (size_t)h * E * P, p->Wk + (size_t)h * E * P, p->Wv + (size_t)h * E * P};
    const int32_t *Bi[3] = {p->Bq + (size_t)h * P, p->Bk + (size_t)h * P, p->Bv + (size_t)h * P};
    int8_t *dst[3] = {q, k, v};
    for (int t = 0; t < 3; ++t) {                       /* step1_Qp / step2_Kp / step3_Vp */
      rq_t r = rq_of(p, ST_Q + t, h);
      gemm_i8(X, 0, W[t], acc, S, E, P);
      for (int i = 0; i < S; ++i)
        for (int j = 0; j < P; ++j)
          dst[t][(size_t)i * P + j] = requant((int64_t)acc[(size_t)i * P + j] + Bi[t][j], r);
    }
    for (int i = 0; i < S; ++i)                         /* K^T for the logits */
      for (int j = 0; j < P; ++j) Kt[(size_t)j * S + i] = k[(size_t)i * P + j];
    {                                                   /* step4_QK: requant(Q.K^T), no 1/sqrt(d) */
      rq_t r = rq_of(p, ST_QK, h);
      gemm_i8(q, 0, Kt, acc, S, P, S);
      for (size_t i = 0; i < (size_t)S * S; ++i) a[i] = requant(acc[i], r);
    }
    for (int i = 0; i < S; ++i) {                       /* ITAmax */
      if (p->softmax_mode == SOFTMAX_REPO_GLOBAL) itamax_repo_row(a + (size_t)i * S, asm_ + (size_t)i * S, S);
      else itamax_stream_row(a + (size_t)i * S, asm_ + (size_t)i * S, S);
    }
    {                                                   /* step5_AV: u8 x s8 */
      rq_t r = rq_of(p, ST_AV, h);
      gemm_i8(asm_, 1, v, acc, S, S, P);
      for (size_t i = 0; i < (size_t)S * P; ++i) o[i] = requant(acc[i], r);
    }
    gemm_i8(o, 0, p->Wo + (size_t)h * P * E, acc, S, P, E);   /* step6_O */
    if (p->head_mode == HEADS_PYITA) {
      rq_t r = rq_of(p, ST_O, h);
      int8_t *oo = Out + (size_t)h * S * E;
      for (int i = 0; i < S; ++i)
        for (int j = 0; j < E; ++j) {
          int8_t yv = requant((int64_t)acc[(size_t)i * E + j] + p->Bo[(size_t)h * E + j], r);
          oo[(size_t)i * E + j] = yv;
          accsum[(size_t)i * E + j] = (h == 0 ? 0 : accsum[(size_t)i * E + j]) + yv;
        }
    } else {
      for (size_t i = 0; i < (size_t)S * E; ++i) accsum[i] = (h == 0 ? 0 : accsum[i]) + acc[i];
    }
  }
  if (p->head_mode == HEADS_PYITA) {
    if (H > 1) {                                        /* step7_Osum */
      rq_t r = rq_of(p, ST_SUM, 0);
      for (size_t i = 0; i < (size_t)S * E; ++i) Y[i] = requant(accsum[i], r);
    } else {
      memcpy(Y, Out, (size_t)S * E);
    }
  } else {
    rq_t r = rq_of(p, ST_O, 0);
    for (int i = 0; i < S; ++i)
      for (int j = 0; j < E; ++j) Y[(size_t)i * E + j] = requant((int64_t)accsum[(size_t)i * E + j] + p->Bo[j], r);
    memcpy(Out, Y, (size_t)S * E);
  }
}

/* X [B,S,E] -> Y [B,S,E]; intermediates (any may be NULL): Qp/Kp/Vp/O [B,H,S,P], A/Asm [B,H,S,S],
 * Out [B,H,S,E] (HEADS_TORCH: [B,1,S,E]).  nthreads <= 0: all available.  Returns 0, or -1 on allocation failure. */
int ita_ref_mha(const ita_ref_params *p, const int8_t *X, int64_t B, int8_t *Y,
                int8_t *Qp, int8_t *Kp, int8_t *Vp, int8_t *A, uint8_t *Asm, int8_t *O, int8_t *Out, int nthreads) {
  const size_t S = p->S, E = p->E, P = p->P, H = p->H;
  const size_t nOut = p->head_mode == HEADS_PYITA ? H : 1;
  const size_t mx = S > P ? S : P;
  int fail = 0;
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel num_threads(nthreads)
#endif
  {
    /* per-thread scratch: intermediates that the caller did not ask for, plus accumulators */
    size_t szHP = H * S * P, szHS = H * S * S, szHE = H * S * E;
    int8_t *scr = (int8_t *)malloc(4 * szHP + 2 * szHS + szHE + P * S);
    int32_t *acc = (int32_t *)malloc(sizeof(int32_t) * (S * mx + S * E));
    if (!scr || !acc) {
#ifdef _OPENMP
#pragma omp atomic write
#endif
      fail = 1;
    } else {
      int8_t *sq = scr, *sk = sq + szHP, *sv = sk + szHP, *so = sv + szHP, *sa = so + szHP;
      uint8_t *ssm = (uint8_t *)(sa + szHS);
      int8_t *sout = (int8_t *)ssm + szHS, *kt = sout + szHE;
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
      for (int64_t b = 0; b < B; ++b) {
        frame_mha(p, X + b * S * E, Y + b * S * E,
                  Qp ? Qp + b * szHP : sq, Kp ? Kp + b * szHP : sk, Vp ? Vp + b * szHP : sv,
                  A ? A + b * szHS : sa, Asm ? Asm + b * szHS : ssm, O ? O + b * szHP : so,
                  Out ? Out + b * nOut * S * E : sout, acc, acc + S * mx, kt);
      }
    }
    free(scr);
    free(acc);
  }
  return fail ? -1 : 0;
}

/* FFN: F1 [B,S,F] = relu_requant(max(0, requant_fc1(X.Wff + Bff))), F2 [B,S,E] = requant_fc2(F1.Wff2 + Bff2).
 * mult[ST_RELU] == 0 means plain ReLU.  (reference call site tests/export_and_validation_W_B.py:544-553) */
int ita_ref_ffn(const ita_ref_params *p, const int8_t *X, int64_t B, int8_t *F1, int8_t *F2, int nthreads) {
  const size_t S = p->S, E = p->E, F = p->F;
  if (!F || !p->Wff || !p->Wff2) return -2;
  int fail = 0;
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel num_threads(nthreads)
#endif
  {
    int32_t *acc = (int32_t *)malloc(sizeof(int32_t) * S * (F > E ? F : E));
    int8_t *f1s = (int8_t *)malloc(S * F);
    if (!acc || !f1s) {
#ifdef _OPENMP
#pragma omp atomic write
#endif
      fail = 1;
    } else {
      rq_t r1 = rq_of(p, ST_FC1, 0), rr = rq_of(p, ST_RELU, 0), r2 = rq_of(p, ST_FC2, 0);
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
      for (int64_t b = 0; b < B; ++b) {
        int8_t *f1 = F1 ? F1 + b * S * F : f1s;
        gemm_i8(X + b * S * E, 0, p->Wff, acc, (int)S, (int)E, (int)F);
        for (size_t i = 0; i < S; ++i)
          for (size_t j = 0; j < F; ++j) {
            int64_t v = requant((int64_t)acc[i * F + j] + p->Bff[j], r1);
            v = v > 0 ? v : 0;
            f1[i * F + j] = rr.mult ? requant(v, rr) : (int8_t)v;
          }
        gemm_i8(f1, 0, p->Wff2, acc, (int)S, (int)F, (int)E);
        for (size_t i = 0; i < S; ++i)
          for (size_t j = 0; j < E; ++j)
            F2[b * S * E + i * E + j] = requant((int64_t)acc[i * E + j] + p->Bff2[j], r2);
      }
    }
    free(acc);
    free(f1s);
  }
  return fail ? -1 : 0;
}

/* standalone ITAmax over rows: x [rows,n] s8 -> out [rows,n] u8 */
void ita_ref_itamax(const int8_t *x, uint8_t *out, int64_t rows, int n, int mode) {
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
  for (int64_t r = 0; r < rows; ++r) {
    if (mode == SOFTMAX_REPO_GLOBAL) itamax_repo_row(x + r * n, out + r * n, n);
    else itamax_stream_row(x + r * n, out + r * n, n);
  }
}

/* standalone requant over a flat s32 array (s64 math inside) */
void ita_ref_requant(const int32_t *acc, int8_t *out, int64_t n, int mult, int shift, int add, int mode, float eff) {
  rq_t r = {mult, shift, add, mode, eff};
  for (int64_t i = 0; i < n; ++i) out[i] = requant(acc[i], r);
}

int ita_ref_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
