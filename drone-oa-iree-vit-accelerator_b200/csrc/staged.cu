// staged.cu -- one kernel per golden-model step, every intermediate in HBM.
//
// This is the "any shape, every intermediate" path behind ita_mha_forward_stages / ita_mha_step (the per-stage attributes
// the reference harness reads from the golden model, tests/export_and_validation_W_B.py:520-542) and the route for shapes
// or parameters the fused tcgen05 kernel (fused.cu) does not cover.  Plain CUDA-core int8 (dp4a) GEMM tiles with the
// requantisation epilogue; exact for every parameter value.  The throughput path is fused.cu.
#include "common.cuh"

namespace ita {

// ---------------------------------------------------------------------------------------------------------------- GEMM
// C[z][m][n] = requant( sum_k A[z][m][k] * B[z][k][n] + bias[h][n] ),  z = b*H + h.
//   A: s8 or u8, row-major [M,K] at A + b*sAb + h*sAh
//   B: s8; B_NT == false: row-major [K,N] ("NN": weights [E,P], V [S,P]);  B_NT == true: row-major [N,K] ("NT": K for Q.K^T)
//   C: s8 row-major [M,N] at C + b*sCb + h*sCh
struct GemmArgs {
  const int8_t *A, *B;
  int8_t *C;
  const int32_t *bias;   // [H,N] or nullptr
  const Rq *rq;          // [H] for this stage
  long long sAb, sAh, sBb, sBh, sCb, sCh;
  int M, N, K, H;
  int a_unsigned, b_nt, relu;
  const Rq *rq_relu;     // optional second requant after ReLU (FFN), may be nullptr
};

constexpr int TM = 64, TN = 64, TK = 32;

// 4-way dot product, a = 4 x u8, b = 4 x s8 (the A.V stage: softmax output is unsigned)
__device__ __forceinline__ int dp4a_us(int a, int b, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

__global__ void __launch_bounds__(256) k_gemm_rq(GemmArgs g) {
  __shared__ __align__(16) int8_t As[TM][TK + 4];
  __shared__ __align__(16) int8_t Bs[TN][TK + 4];   // stored [n][k]
  const int z = blockIdx.z, b = z / g.H, h = z % g.H;
  const int8_t *A = g.A + b * g.sAb + h * g.sAh;
  const int8_t *Bm = g.B + b * g.sBb + h * g.sBh;
  int8_t *C = g.C + b * g.sCb + h * g.sCh;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  const int tid = threadIdx.x, tx = tid % 16, ty = tid / 16;   // thread -> rows ty*4..+3, cols tx*4..+3
  int acc[4][4] = {};
  for (int k0 = 0; k0 < g.K; k0 += TK) {
    for (int i = tid; i < TM * TK; i += 256) {
      int r = i / TK, c = i % TK;
      int gm = m0 + r, gk = k0 + c;
      As[r][c] = (gm < g.M && gk < g.K) ? A[(long long)gm * g.K + gk] : (int8_t)0;
    }
    if (g.b_nt) {
      for (int i = tid; i < TN * TK; i += 256) {
        int r = i / TK, c = i % TK;
        int gn = n0 + r, gk = k0 + c;
        Bs[r][c] = (gn < g.N && gk < g.K) ? Bm[(long long)gn * g.K + gk] : (int8_t)0;
      }
    } else {
      for (int i = tid; i < TN * TK; i += 256) {
        int c = i / TN, r = i % TN;   // consecutive threads walk n (contiguous in global)
        int gn = n0 + r, gk = k0 + c;
        Bs[r][c] = (gn < g.N && gk < g.K) ? Bm[(long long)gk * g.N + gn] : (int8_t)0;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TK; kk += 4) {
      int a4[4], b4[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a4[i] = *reinterpret_cast<const int *>(&As[ty * 4 + i][kk]);
#pragma unroll
      for (int j = 0; j < 4; ++j) b4[j] = *reinterpret_cast<const int *>(&Bs[tx * 4 + j][kk]);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          acc[i][j] = g.a_unsigned ? dp4a_us(a4[i], b4[j], acc[i][j]) : __dp4a(a4[i], b4[j], acc[i][j]);
    }
    __syncthreads();
  }
  const Rq r = g.rq[h];
  Rq rr;
  if (g.rq_relu) rr = g.rq_relu[0];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int gm = m0 + ty * 4 + i;
    if (gm >= g.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int gn = n0 + tx * 4 + j;
      if (gn >= g.N) continue;
      int a = acc[i][j] + (g.bias ? g.bias[h * g.N + gn] : 0);
      int y = requant_general(a, r);
      if (g.relu) {
        y = y > 0 ? y : 0;
        if (g.rq_relu && rr.mult != 0) y = requant_general(y, rr);
      }
      C[(long long)gm * g.N + gn] = (int8_t)y;
    }
  }
}

static int launch_gemm(const GemmArgs &g, int64_t B, cudaStream_t s) {
  if (B <= 0) return ITA_OK;
  const int64_t zmax = 65535;
  // gridDim.z is limited to 65535: walk the batch in chunks
  for (int64_t b0 = 0; b0 < B; b0 += zmax / g.H) {
    int64_t nb = B - b0 < zmax / g.H ? B - b0 : zmax / g.H;
    GemmArgs a = g;
    a.A += b0 * g.sAb;
    a.B += b0 * g.sBb;
    a.C += b0 * g.sCb;
    dim3 grid((g.N + TN - 1) / TN, (g.M + TM - 1) / TM, (unsigned)(nb * g.H));
    k_gemm_rq<<<grid, 256, 0, s>>>(a);
    count_launch();
  }
  ITA_CUDA_TRY(cudaGetLastError());
  return ITA_OK;
}

// -------------------------------------------------------------------------------------------------------------- ITAmax
// one thread per row (rows = B*H*S); n = row length.  Both variants of SURVEY.md App. A.3(3).
__global__ void k_itamax(const int8_t *__restrict__ x, uint8_t *__restrict__ out, long long rows, int n, int mode) {
  long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const int8_t *xr = x + r * n;
  uint8_t *o = out + r * n;
  if (mode == ITA_SOFTMAX_REPO_GLOBAL) {
    int m = -128;
    for (int j = 0; j < n; ++j) m = max(m, (int)xr[j]);
    unsigned tot = 0;
    for (int j = 0; j < n; ++j) {
      int d = m - xr[j];
      tot += d <= 8 ? (256u >> d) : 0u;
    }
    tot = tot < 1 ? 1 : tot;
    const unsigned inv = (255u * 65536u) / tot;
    for (int j = 0; j < n; ++j) {
      int d = m - xr[j];
      unsigned num = d <= 8 ? (256u >> d) : 0u;
      o[j] = (uint8_t)((num * inv) >> 16);
    }
  } else {
    int m = -128;
    unsigned tot = 0;
    for (int g = 0; g < n; g += 16) {
      int cm = -128;
      for (int j = g; j < g + 16; ++j) cm = max(cm, (int)xr[j]);
      if (cm > m) {
        tot >>= ((cm - m + 16) >> 5);
        m = cm;
      }
      for (int j = g; j < g + 16; ++j) tot += 256u >> ((m - xr[j] + 16) >> 5);
    }
    tot = tot < 1 ? 1 : tot;
    const unsigned inv = (255u * 256u) / tot;
    for (int j = 0; j < n; ++j) o[j] = (uint8_t)(inv >> ((m - xr[j] + 16) >> 5));
  }
}

// ------------------------------------------------------------------------------------------------------------ head sum
__global__ void k_headsum(const int8_t *__restrict__ out_h, int8_t *__restrict__ y, long long B, int H, long long SE, const Rq *rq) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * SE) return;
  long long b = i / SE, e = i % SE;
  int acc = 0;
  for (int h = 0; h < H; ++h) acc += out_h[(b * H + h) * SE + e];
  y[i] = (int8_t)requant_general(acc, rq[0]);
}

// HEADS_TORCH output projection: Y = requant( sum_h O[h].Wo[h] + Bo[0] ) == one GEMM over the concatenated context.
// O is stored [B,H,S,P]; gather K = H*P through the (h,p) index.
__global__ void __launch_bounds__(256) k_outproj_torch(const int8_t *__restrict__ O, const int8_t *__restrict__ Wo,
                                                       const int32_t *__restrict__ Bo, int8_t *__restrict__ Y, int S, int P,
                                                       int E, int H, const Rq *rq) {
  const long long b = blockIdx.y;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= S * E) return;
  const int s = idx / E, e = idx % E;
  int acc = Bo[e];
  for (int h = 0; h < H; ++h) {
    const int8_t *o = O + ((b * H + h) * S + s) * P;
    const int8_t *w = Wo + (long long)h * P * E + e;
    for (int p = 0; p < P; ++p) acc += (int)o[p] * (int)w[(long long)p * E];
  }
  Y[(b * S + s) * E + e] = (int8_t)requant_general(acc, rq[0]);
}

// ---------------------------------------------------------------------------------------------------- float boundary
__global__ void k_quantize(const float *__restrict__ x, int8_t *__restrict__ q, long long n, float scale) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  // torch.quantize_per_tensor: clamp(nearbyint(x * (1/scale)) + zp); inv_scale is formed in fp32 like ATen does
  float inv = 1.0f / scale;
  float v = nearbyintf(x[i] * inv);
  v = fminf(fmaxf(v, -128.f), 127.f);
  q[i] = (int8_t)(int)v;
}
__global__ void k_dequantize(const int8_t *__restrict__ q, float *__restrict__ y, long long n, float scale) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  y[i] = (float)q[i] * scale;
}

// 16 elements per thread: four 16-byte loads in flight, one 16-byte store (the stubs are pure HBM streams)
__global__ void __launch_bounds__(256) k_quantize16(const float4 *__restrict__ x, uint4 *__restrict__ q, long long n16, float scale) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n16) return;
  const float inv = 1.0f / scale;
  float4 a[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) a[k] = x[i * 4 + k];
  uint32_t w[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int v0 = (int)fminf(fmaxf(nearbyintf(a[k].x * inv), -128.f), 127.f), v1 = (int)fminf(fmaxf(nearbyintf(a[k].y * inv), -128.f), 127.f);
    const int v2 = (int)fminf(fmaxf(nearbyintf(a[k].z * inv), -128.f), 127.f), v3 = (int)fminf(fmaxf(nearbyintf(a[k].w * inv), -128.f), 127.f);
    w[k] = (uint32_t)(v0 & 0xFF) | ((uint32_t)(v1 & 0xFF) << 8) | ((uint32_t)(v2 & 0xFF) << 16) | ((uint32_t)(v3 & 0xFF) << 24);
  }
  q[i] = make_uint4(w[0], w[1], w[2], w[3]);
}
__global__ void __launch_bounds__(256) k_dequantize16(const uint4 *__restrict__ q, float4 *__restrict__ y, long long n16, float scale) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n16) return;
  const uint4 v = q[i];
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int k = 0; k < 4; ++k)
    y[i * 4 + k] = make_float4((float)(int8_t)(w[k] & 0xFF) * scale, (float)(int8_t)((w[k] >> 8) & 0xFF) * scale,
                               (float)(int8_t)((w[k] >> 16) & 0xFF) * scale, (float)(int8_t)(w[k] >> 24) * scale);
}

int launch_quantize(const float *x, int8_t *q, int64_t n, float scale, cudaStream_t s) {
  if (n <= 0) return ITA_OK;
  if (n % 16 == 0 && !(reinterpret_cast<uintptr_t>(x) & 15) && !(reinterpret_cast<uintptr_t>(q) & 15)) {
    k_quantize16<<<(unsigned)((n / 16 + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4 *>(x), reinterpret_cast<uint4 *>(q), n / 16, scale);
    count_launch();
    ITA_CUDA_TRY(cudaGetLastError());
    return ITA_OK;
  }
  k_quantize<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(x, q, n, scale);
  count_launch();
  ITA_CUDA_TRY(cudaGetLastError());
  return ITA_OK;
}
int launch_dequantize(const int8_t *q, float *y, int64_t n, float scale, cudaStream_t s) {
  if (n <= 0) return ITA_OK;
  if (n % 16 == 0 && !(reinterpret_cast<uintptr_t>(y) & 15) && !(reinterpret_cast<uintptr_t>(q) & 15)) {
    k_dequantize16<<<(unsigned)((n / 16 + 255) / 256), 256, 0, s>>>(reinterpret_cast<const uint4 *>(q), reinterpret_cast<float4 *>(y), n / 16, scale);
    count_launch();
    ITA_CUDA_TRY(cudaGetLastError());
    return ITA_OK;
  }
  k_dequantize<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(q, y, n, scale);
  count_launch();
  ITA_CUDA_TRY(cudaGetLastError());
  return ITA_OK;
}

// --------------------------------------------------------------------------------------------------------------- steps
int staged_step(ita_mha_handle *h, int step, const void *in0, const void *in1, void *out, int64_t B, cudaStream_t s) {
  const long long S = h->S, E = h->E, P = h->P, H = h->H;
  if (B <= 0) return ITA_OK;
  GemmArgs g{};
  g.H = (int)H;
  switch (step) {
    case ITA_STEP1_QP:
    case ITA_STEP2_KP:
    case ITA_STEP3_VP: {
      const int t = step - ITA_STEP1_QP;
      const int8_t *W[3] = {h->Wq, h->Wk, h->Wv};
      const int32_t *Bi[3] = {h->Bq, h->Bk, h->Bv};
      g.A = (const int8_t *)in0; g.sAb = S * E; g.sAh = 0;
      g.B = W[t]; g.sBb = 0; g.sBh = E * P;
      g.C = (int8_t *)out; g.sCb = H * S * P; g.sCh = S * P;
      g.bias = Bi[t]; g.rq = h->rq_dev + (ITA_ST_Q + t) * H;
      g.M = (int)S; g.N = (int)P; g.K = (int)E;
      return launch_gemm(g, B, s);
    }
    case ITA_STEP4_QK:
      g.A = (const int8_t *)in0; g.sAb = H * S * P; g.sAh = S * P;
      g.B = (const int8_t *)in1; g.sBb = H * S * P; g.sBh = S * P; g.b_nt = 1;
      g.C = (int8_t *)out; g.sCb = H * S * S; g.sCh = S * S;
      g.rq = h->rq_dev + ITA_ST_QK * H;
      g.M = (int)S; g.N = (int)S; g.K = (int)P;
      return launch_gemm(g, B, s);
    case ITA_STEP4_SOFTMAX: {
      long long rows = B * H * S;
      k_itamax<<<(unsigned)((rows + 127) / 128), 128, 0, s>>>((const int8_t *)in0, (uint8_t *)out, rows, (int)S, h->softmax_mode);
      count_launch();
      ITA_CUDA_TRY(cudaGetLastError());
      return ITA_OK;
    }
    case ITA_STEP5_AV:
      g.A = (const int8_t *)in0; g.sAb = H * S * S; g.sAh = S * S; g.a_unsigned = 1;
      g.B = (const int8_t *)in1; g.sBb = H * S * P; g.sBh = S * P;
      g.C = (int8_t *)out; g.sCb = H * S * P; g.sCh = S * P;
      g.rq = h->rq_dev + ITA_ST_AV * H;
      g.M = (int)S; g.N = (int)P; g.K = (int)S;
      return launch_gemm(g, B, s);
    case ITA_STEP6_O:
      if (h->head_mode == ITA_HEADS_PYITA) {
        g.A = (const int8_t *)in0; g.sAb = H * S * P; g.sAh = S * P;
        g.B = h->Wo; g.sBb = 0; g.sBh = P * E;
        g.C = (int8_t *)out; g.sCb = H * S * E; g.sCh = S * E;
        g.bias = h->Bo; g.rq = h->rq_dev + ITA_ST_O * H;
        g.M = (int)S; g.N = (int)E; g.K = (int)P;
        return launch_gemm(g, B, s);
      } else {
        dim3 grid((unsigned)((S * E + 255) / 256), 1, 1);
        for (int64_t b0 = 0; b0 < B; b0 += 65535) {
          int64_t nb = B - b0 < 65535 ? B - b0 : 65535;
          grid.y = (unsigned)nb;
          k_outproj_torch<<<grid, 256, 0, s>>>((const int8_t *)in0 + b0 * H * S * P, h->Wo, h->Bo, (int8_t *)out + b0 * S * E,
                                               (int)S, (int)P, (int)E, (int)H, h->rq_dev + ITA_ST_O * H);
          count_launch();
        }
        ITA_CUDA_TRY(cudaGetLastError());
        return ITA_OK;
      }
    case ITA_STEP7_OSUM: {
      long long n = B * S * E;
      k_headsum<<<(unsigned)((n + 255) / 256), 256, 0, s>>>((const int8_t *)in0, (int8_t *)out, B, (int)H, S * E,
                                                           h->rq_dev + ITA_ST_SUM * H);
      count_launch();
      ITA_CUDA_TRY(cudaGetLastError());
      return ITA_OK;
    }
    default:
      set_error("ita_mha_step: unknown step %d", step);
      return ITA_ERR_INVALID;
  }
}

int staged_forward(ita_mha_handle *h, const int8_t *x, int8_t *y, int64_t B, const ita_mha_stages *st, cudaStream_t s) {
  const size_t S = h->S, E = h->E, P = h->P, H = h->H;
  if (B <= 0) return ITA_OK;
  ita_mha_stages w{};
  if (st) w = *st;
  // Intermediates the caller did not ask for live in the handle's scratch; the batch is walked in chunks so that the
  // scratch stays bounded (<= ~1 GiB) whatever the batch.
  const size_t per_frame = 4 * H * S * P + 2 * H * S * S + H * S * E;
  int64_t chunk = (int64_t)((size_t(1) << 30) / per_frame);
  if (chunk < 1) chunk = 1;
  if (chunk > B) chunk = B;
  int rc = ensure_scratch(h, per_frame * (size_t)chunk);
  if (rc) return rc;
  for (int64_t b0 = 0; b0 < B; b0 += chunk) {
    const int64_t nb = B - b0 < chunk ? B - b0 : chunk;
    int8_t *p = h->scratch;
    auto take = [&](size_t per, void *user, size_t user_per) -> void * {
      void *r = user ? (void *)((int8_t *)user + (size_t)b0 * user_per) : (void *)p;
      p += per * (size_t)chunk;
      return r;
    };
    int8_t *Qp = (int8_t *)take(H * S * P, w.Qp, H * S * P);
    int8_t *Kp = (int8_t *)take(H * S * P, w.Kp, H * S * P);
    int8_t *Vp = (int8_t *)take(H * S * P, w.Vp, H * S * P);
    int8_t *A = (int8_t *)take(H * S * S, w.A, H * S * S);
    uint8_t *Asm = (uint8_t *)take(H * S * S, w.Asm, H * S * S);
    int8_t *O = (int8_t *)take(H * S * P, w.O, H * S * P);
    const bool pyita = h->head_mode == ITA_HEADS_PYITA;
    const size_t out_per = (pyita ? H : 1) * S * E;
    int8_t *Out = (int8_t *)take(H * S * E, w.Out, out_per);
    const int8_t *xb = x + (size_t)b0 * S * E;
    int8_t *yb = y ? y + (size_t)b0 * S * E : nullptr;
    if ((rc = staged_step(h, ITA_STEP1_QP, xb, nullptr, Qp, nb, s))) return rc;
    if ((rc = staged_step(h, ITA_STEP2_KP, xb, nullptr, Kp, nb, s))) return rc;
    if ((rc = staged_step(h, ITA_STEP3_VP, xb, nullptr, Vp, nb, s))) return rc;
    if ((rc = staged_step(h, ITA_STEP4_QK, Qp, Kp, A, nb, s))) return rc;
    if ((rc = staged_step(h, ITA_STEP4_SOFTMAX, A, nullptr, Asm, nb, s))) return rc;
    if ((rc = staged_step(h, ITA_STEP5_AV, Asm, Vp, O, nb, s))) return rc;
    if (pyita && H > 1) {
      if ((rc = staged_step(h, ITA_STEP6_O, O, nullptr, Out, nb, s))) return rc;
      if (yb && (rc = staged_step(h, ITA_STEP7_OSUM, Out, nullptr, yb, nb, s))) return rc;
    } else {
      // single head (or torch convention): step 6 produces Y directly; mirror it into Out when the caller wants both
      int8_t *dst = yb ? yb : Out;
      if ((rc = staged_step(h, ITA_STEP6_O, O, nullptr, dst, nb, s))) return rc;
      if (yb && w.Out) ITA_CUDA_TRY(cudaMemcpyAsync(Out, yb, (size_t)nb * S * E, cudaMemcpyDeviceToDevice, s));
    }
  }
  return ITA_OK;
}

int staged_ffn(ita_mha_handle *h, const int8_t *x, int8_t *f1, int8_t *f2, int64_t B, cudaStream_t s) {
  const long long S = h->S, E = h->E, F = h->F;
  if (!F || !h->Wff || !h->Wff2) {
    set_error("ita_ffn_forward: handle was created without FFN parameters (F == 0)");
    return ITA_ERR_INVALID;
  }
  if (B <= 0) return ITA_OK;
  int8_t *f1buf = f1;
  int64_t chunk = B;
  if (!f1buf) {
    chunk = (int64_t)((size_t(1) << 29) / (size_t)(S * F));
    if (chunk < 1) chunk = 1;
    if (chunk > B) chunk = B;
    int rc = ensure_scratch(h, (size_t)chunk * S * F);
    if (rc) return rc;
  }
  for (int64_t b0 = 0; b0 < B; b0 += chunk) {
    const int64_t nb = B - b0 < chunk ? B - b0 : chunk;
    int8_t *f1b = f1 ? f1 + b0 * S * F : h->scratch;
    GemmArgs g{};
    g.H = 1;
    g.A = x + b0 * S * E; g.sAb = S * E;
    g.B = h->Wff;
    g.C = f1b; g.sCb = S * F;
    g.bias = h->Bff; g.rq = h->rq_dev + ITA_ST_FC1 * h->H;
    g.relu = 1; g.rq_relu = h->rq_dev + ITA_ST_RELU * h->H;
    g.M = (int)S; g.N = (int)F; g.K = (int)E;
    int rc = launch_gemm(g, nb, s);
    if (rc) return rc;
    GemmArgs g2{};
    g2.H = 1;
    g2.A = f1b; g2.sAb = S * F;
    g2.B = h->Wff2;
    g2.C = f2 + b0 * S * E; g2.sCb = S * E;
    g2.bias = h->Bff2; g2.rq = h->rq_dev + ITA_ST_FC2 * h->H;
    g2.M = (int)S; g2.N = (int)E; g2.K = (int)F;
    if ((rc = launch_gemm(g2, nb, s))) return rc;
  }
  return ITA_OK;
}

}  // namespace ita
