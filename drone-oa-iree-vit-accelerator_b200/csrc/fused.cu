// fused.cu -- the hot path: one persistent sm_100a kernel that runs a whole ITA attention layer per depth frame
// (Q/K/V projections, Q.K^T, ITAmax, A.V, output projection, all requantisations) without any intermediate leaving
// the SM: int8 operands in shared memory, s32 accumulators in tensor memory, tcgen05.mma kind::i8.
//
// Data layout in shared memory ("panel" layout, the UMMA no-swizzle canonical layout):
//   a [R x C] int8 matrix is stored as C/16 panels; panel c holds the 16-byte column chunk c of every row, rows
//   consecutive:   addr(r, col) = (col/16) * R*16 + r*16 + col%16.
//   * read as a K-major operand (rows = M or N index, columns = K):  SBO = 128 B (next 8 rows), LBO = R*16 (next 16 k)
//   * read as an MN-major operand (rows = K index, columns = N):     SBO = R*16 (next 16 n),  LBO = 128 B (next 8 k)
//   An epilogue thread owns one row (= one TMEM lane) and writes 16 B per 16 columns; a warp's 32 rows are 512
//   contiguous bytes, so the writes are bank-conflict free, and the same buffer serves as the next MMA's operand:
//     Q [S,P] -> A operand of Q.K^T (K-major)        K [S,P] -> B operand of Q.K^T (K-major, N = S)
//     V [S,P] -> B operand of A.V (MN-major: rows are K) Asm [S,S] u8 -> A operand of A.V     O [S,P] -> A operand of O.Wo
//   Weights are pre-arranged on the host into the same panel layout (W^T as [N rows, K cols]) and bulk-copied once per CTA.
//
// Pipeline.  A frame is cut into NCH = 4*P/64 + 4 "chunks", each one accumulator tile of <= 64 columns:
//     Q[0..NQ) K[0..NQ) V[0..NQ)  S0 S1  AV[0..NQ)  O0 O1            (NQ = P/64; S chunks are 64 key columns, O chunks E/2)
//   Two frames ("slots") are in flight per SM.  Per slot, TMEM holds a ring of 4 regions x 64 columns; chunk c of every
//   frame uses region c%4 and is drained by epilogue warpgroup c%2 of the slot (4 warps = the 4 TMEM lane quadrants), so
//   every warpgroup owns two regions and the MMA of its next chunk runs while it requantises the current one.  One MMA
//   issuing thread per slot walks the chunks in order, waiting only for (a) the region to be drained and (b) the operand
//   buffers of the chunk to be written (mbarriers armed by the producing epilogues); a loader thread prefetches the next
//   frame's X tile by TMA as soon as the V projections have consumed the current one.  tcgen05 MMAs of one thread
//   complete in issue order, which is what makes the reuse of the operand buffers (Asm over K, O over Q) safe.
//
// Thread roles (608 threads): warps 0-7 = epilogue of slot 0 (warp w: TMEM lanes 32*(w%4).., warpgroup (w/4)%2),
// warps 8-15 = slot 1, warps 16/17 = MMA issue for slot 0/1 (one elected thread), warp 18 = TMA loader + TMEM allocator.
//
// Reference semantics: models/ITA/QAT/layers.py:103-126 and tests/export_and_validation_W_B.py:518-542 (stage order,
// layouts), models/ITA/QAT/ITA_softmax.py:51-61 (ITAmax), SURVEY.md App. A (integer spec).
#include <cuda.h>

#include <type_traits>
#include <vector>

#include "common.cuh"
#include "ptx.cuh"

namespace ita {

constexpr int kS = 128;                 // rows per tile = tcgen05 M
constexpr int kSlots = 2;
constexpr int kEpiWarpsPerSlot = 8;
constexpr int kWarpMma = kSlots * kEpiWarpsPerSlot;     // + slot
constexpr int kWarpLoad = kWarpMma + kSlots;
constexpr int kThreads = (kWarpLoad + 1) * 32;
constexpr int kRegions = 4;             // accumulator ring per slot
constexpr int kRegionCols = 64;
constexpr int kTmemColsPerSlot = kRegions * kRegionCols;
constexpr int kBarsPerSlot = 18;

enum { RQM_INT_HI = 0, RQM_INT_GEN = 1, RQM_TORCH = 2 };
enum { STG_Q = 0, STG_K, STG_V, STG_S, STG_AV, STG_O };
// per-slot mbarriers
enum { BAR_X = 0, BAR_XFREE = 1, BAR_FULL = 2, BAR_EMPTY = 6, BAR_DEP = 10 /* +STG_Q, STG_K, STG_V, STG_S, STG_AV */,
       BAR_SM = 15 /* two softmax exchange points between the warpgroups of a slot */ };

// Per-stage uniform requant parameters as the fused kernel consumes them.
struct FRq {
  int32_t multp;      // INT_HI: mult << (32 - shift);  INT_GEN: mult
  int32_t shift;      // INT_GEN only
  int32_t add;        // INT_GEN / TORCH
  float eff;          // TORCH
  long long c;        // bias-free stages: INT_HI: (rnd << (32-shift)) + (add << 32);  INT_GEN: rnd
};

struct FusedParams {
  const int8_t *wpanels;     // device: [Wq^T | Wk^T | Wv^T | Wo^T] in panel layout, then per-column constants
  int8_t *y;
  long long batch;
  int *abort_flag;
  FRq rq[6];                 // Q K V QK AV O
  // debug dump (natural layouts), any may be null
  int8_t *dQ, *dK, *dV, *dA;
  uint8_t *dAsm;
  int8_t *dO;
};

template <int E, int P>
struct Cfg {
  static constexpr int S = kS;
  static constexpr int NQ = P / 64;                                      // chunks per projection / per A.V
  static constexpr int NCH = 4 * NQ + 4;                                 // chunks per frame
  static constexpr int OC = E / 2;                                       // columns per output-projection chunk
  static constexpr int W_BYTES = P * E;                                  // each of Wq^T, Wk^T, Wv^T, Wo^T
  static constexpr int CONST_ENTRIES = 3 * P + E;                        // per-column 8-byte requant constants
  static constexpr int WPANEL_BYTES = 4 * W_BYTES + CONST_ENTRIES * 8;   // what is bulk-copied at kernel start
  static constexpr int X_BYTES = S * E;
  static constexpr int SP_BYTES = S * P;
  static constexpr int KA_BYTES = (S * P > S * S) ? S * P : S * S;       // K, later the softmax output
  static constexpr int SLOT_BYTES = X_BYTES + SP_BYTES + KA_BYTES + SP_BYTES;
  static constexpr int OFF_W = 0;
  static constexpr int OFF_CONST = 4 * W_BYTES;
  static constexpr int OFF_SLOT = ((WPANEL_BYTES + 127) / 128) * 128;
  static constexpr int OFF_XCHG = OFF_SLOT + kSlots * SLOT_BYTES;        // softmax exchange: [slot][2][S] u32 x2
  static constexpr int XCHG_BYTES = kSlots * 2 * S * 4 * 2;
  static constexpr int OFF_BAR = OFF_XCHG + XCHG_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 512;
  static_assert(E % 32 == 0 && P % 64 == 0, "E must be a multiple of 32 (tcgen05 kind::i8 K) and P of 64 (chunk width)");
  static_assert(OC % 16 == 0 && OC <= kRegionCols, "output-projection chunks must be 16..64 columns");
  static_assert(NCH % kRegions == 0, "chunks per frame must tile the accumulator ring");
  static_assert((1 + kSlots * kBarsPerSlot) * 8 + 8 <= 512, "barrier area");
  // chunk c of a frame -> stage, index within the stage, first column, width
  static constexpr int stage_of(int c) {
    return c < NQ ? STG_Q : c < 2 * NQ ? STG_K : c < 3 * NQ ? STG_V : c < 3 * NQ + 2 ? STG_S : c < 4 * NQ + 2 ? STG_AV : STG_O;
  }
  static constexpr int idx_of(int c) {
    return c < 3 * NQ ? c % NQ : c < 3 * NQ + 2 ? c - 3 * NQ : c < 4 * NQ + 2 ? c - 3 * NQ - 2 : c - 4 * NQ - 2;
  }
  static constexpr int ncols_of(int c) { return stage_of(c) == STG_O ? OC : 64; }
  static constexpr int col0_of(int c) { return idx_of(c) * ncols_of(c); }
  // arrivals (warps) that complete a stage's "operand written" barrier
  static constexpr int dep_count(int stage) { return (stage == STG_S ? 2 : NQ) * 4; }
};

// ---------------------------------------------------------------------------------------------------- requant forms
template <int MODE>
__device__ __forceinline__ int rq_apply(int acc, const FRq &r, long long c) {
  if constexpr (MODE == RQM_INT_HI) {
    // ((acc + bias)*mult + rnd) >> shift + add  ==  hi32( acc * (mult << (32-shift)) + c ),  c folded on the host
    long long p = (long long)acc * (long long)r.multp + c;
    return (int)(p >> 32);
  } else if constexpr (MODE == RQM_INT_GEN) {
    long long p = (long long)acc * (long long)r.multp + c;
    return (int)(p >> r.shift) + r.add;
  } else {
    // c carries the int32 bias in its low word
    float f = __fmul_rn(__int2float_rn(acc + (int)c), r.eff);
    return __float2int_rn(f) + r.add;
  }
}

__device__ __forceinline__ uint32_t pack_s8x4(int y0, int y1, int y2, int y3) {
  uint32_t t, d;
  asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(y3), "r"(y2), "r"(0));
  asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(y1), "r"(y0), "r"(t));
  return d;
}
__device__ __forceinline__ uint32_t pack_u8x4(int y0, int y1, int y2, int y3) {
  uint32_t t, d;
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(y3), "r"(y2), "r"(0));
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(y1), "r"(y0), "r"(t));
  return d;
}
// X tile [kS rows x E bytes] -> panel layout: one 2-D TMA box (16 B x kS rows) per 16-byte column chunk
template <int E>
__device__ __forceinline__ void load_x_tile(void *dst, const CUtensorMap *map, long long frame, uint64_t *bar) {
  mbar_arrive_expect_tx(bar, kS * E);
#pragma unroll
  for (int c = 0; c < E / 16; ++c) tma_load_2d(reinterpret_cast<uint8_t *>(dst) + c * kS * 16, map, c * 16, (int)(frame * kS), bar);
}

__device__ __forceinline__ uint32_t shr_clamp(uint32_t v, uint32_t s) {   // PTX shr: shift amounts > 31 give 0
  uint32_t r;
  asm("shr.u32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(s));
  return r;
}

// compile-time loop
template <int I, int N, class F>
__device__ __forceinline__ void static_for(F &&f) {
  if constexpr (I < N) {
    f(std::integral_constant<int, I>{});
    static_for<I + 1, N>(f);
  }
}

// Load NC (32 or 64) accumulator columns of this thread's row, then hand the TMEM region back to the MMA thread.
template <int NC>
__device__ __forceinline__ void load_acc(uint32_t taddr, uint32_t (&v)[NC], uint64_t *bar_empty, int lane) {
  static_assert(NC == 32 || NC == 64, "");
  tmem_ld32(taddr, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
  if constexpr (NC == 64) tmem_ld32(taddr + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
  tmem_ld_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(bar_empty);
}

// Requantise NC accumulators of one row -> int8 -> panel buffer (and/or global).
//   consts     per-column constants for these columns (nullptr: bias-free stage, uniform constant r.c)
//   panel      shared-memory panel buffer of a [kS x *] matrix (nullptr: do not store to smem)
//   grow       global row pointer (natural layout, first of these columns) or nullptr
template <int MODE, int NC>
__device__ __forceinline__ void requant_store(const uint32_t (&v)[NC], const FRq &r, const long long *consts, int8_t *panel,
                                              int row, int col0, int8_t *grow) {
#pragma unroll
  for (int j = 0; j < NC; j += 16) {
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      int y[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) y[i] = rq_apply<MODE>((int)v[j + q * 4 + i], r, consts ? consts[j + q * 4 + i] : r.c);
      w[q] = pack_s8x4(y[0], y[1], y[2], y[3]);
    }
    const uint4 o = make_uint4(w[0], w[1], w[2], w[3]);
    if (panel) *reinterpret_cast<uint4 *>(panel + ((col0 + j) >> 4) * (kS * 16) + row * 16) = o;
    if (grow) *reinterpret_cast<uint4 *>(grow + j) = o;
  }
}

// ---------------------------------------------------------------------------------------------------- the kernel
struct SlotCtx {
  uint64_t *bars;       // this slot's kBarsPerSlot mbarriers
  uint8_t *smem;        // this slot's operand buffers: X | Q(O) | K(Asm) | V
  uint32_t tmem;        // first TMEM column of this slot's accumulator ring
};

// ---- epilogue of one warpgroup (W2 = 0/1: the chunks c with c % 2 == W2) of one slot
template <int E, int P, int MODE, int SOFTMAX, bool DUMP, int W2>
__device__ __forceinline__ void epilogue_loop(const FusedParams &prm, const SlotCtx &sc, const long long *s_const, uint32_t *x_max,
                                              uint32_t *x_sum, int slot, int quad, int lane, long long iters, long long frame0,
                                              long long frame_stride) {
  using C = Cfg<E, P>;
  const int row = quad * 32 + lane;
  const uint32_t t_lane = sc.tmem + ((uint32_t)(quad * 32) << 16);
  int8_t *q_buf = reinterpret_cast<int8_t *>(sc.smem) + C::X_BYTES, *k_buf = q_buf + C::SP_BYTES, *v_buf = k_buf + C::KA_BYTES;
  bool ok = true;
  for (long long it = 0; it < iters && ok; ++it) {
    const long long f = frame0 + it * frame_stride;
    static_for<0, C::NCH>([&](auto cc) {
      constexpr int c = decltype(cc)::value;
      if constexpr (c % 2 == W2) {
        constexpr int stage = C::stage_of(c), idx = C::idx_of(c), NC = C::ncols_of(c), col0 = C::col0_of(c), r = c % kRegions;
        if (!ok) return;
        const uint32_t use = (uint32_t)(it * (C::NCH / kRegions)) + c / kRegions;
        ok = mbar_wait(sc.bars + BAR_FULL + r, use & 1, prm.abort_flag);
        if (!ok) return;
        tc_fence_after();
        const uint32_t taddr = t_lane + r * kRegionCols;
        if constexpr (stage != STG_S) {
          uint32_t v[NC];
          load_acc<NC>(taddr, v, sc.bars + BAR_EMPTY + r, lane);
          if constexpr (stage <= STG_V) {
            int8_t *dst = stage == STG_Q ? q_buf : (stage == STG_K ? k_buf : v_buf);
            int8_t *g = nullptr;
            if constexpr (DUMP) {
              int8_t *base = stage == STG_Q ? prm.dQ : (stage == STG_K ? prm.dK : prm.dV);
              if (base) g = base + (f * kS + row) * P + col0;
            }
            requant_store<MODE, NC>(v, prm.rq[stage], s_const + stage * P + col0, dst, row, col0, g);
          } else if constexpr (stage == STG_AV) {
            int8_t *g = nullptr;
            if constexpr (DUMP) {
              if (prm.dO) g = prm.dO + (f * kS + row) * P + col0;
            }
            requant_store<MODE, NC>(v, prm.rq[4], nullptr, q_buf, row, col0, g);
          } else {
            requant_store<MODE, NC>(v, prm.rq[5], s_const + 3 * P + col0, nullptr, row, col0, prm.y + (f * kS + row) * E + col0);
          }
          if constexpr (stage != STG_O) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(sc.bars + BAR_DEP + stage);
          }
        } else {
          // ---- logits chunk (64 key columns of this row) -> ITAmax -> Asm (u8) into the K buffer
          constexpr int SH = 64, half = idx;
          uint32_t pk[SH / 4];   // biased logits a + 128 in [0,255], 4 per word; later the distances d = max - a
          int mx = 0;
          {
            uint32_t v[SH];
            load_acc<SH>(taddr, v, sc.bars + BAR_EMPTY + r, lane);
            const FRq rqk = prm.rq[3];
#pragma unroll
            for (int j = 0; j < SH; j += 4) {
              int y[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                y[i] = rq_apply<MODE>((int)v[j + i], rqk, rqk.c) + 128;
                mx = max(mx, y[i]);
              }
              pk[j >> 2] = pack_u8x4(y[0], y[1], y[2], y[3]);
            }
          }
          mx = min(mx, 255);
          if constexpr (DUMP) {
            if (prm.dA) {
              uint32_t *g = reinterpret_cast<uint32_t *>(prm.dA + (f * kS + row) * kS + half * SH);
#pragma unroll
              for (int i = 0; i < SH / 4; ++i) g[i] = pk[i] ^ 0x80808080u;
            }
          }
          uint32_t out[SH / 4];
          // exchange point between the two warpgroups of the slot (all 256 threads arrive; bounded wait)
          auto xsync = [&](int which) {
            mbar_arrive(sc.bars + BAR_SM + which);
            ok = mbar_wait(sc.bars + BAR_SM + which, (uint32_t)(it & 1), prm.abort_flag) && ok;
          };
          if constexpr (SOFTMAX == ITA_SOFTMAX_REPO_GLOBAL) {
            // models/ITA/QAT/ITA_softmax.py:51-61 -- global row max, numerator 256 >> (max - a), inv = floor(255*2^16 / sum)
            x_max[half * kS + row] = (uint32_t)mx;
            xsync(0);                                          // both S chunks' MMAs are complete past this point
            const uint32_t m = max((uint32_t)mx, x_max[(half ^ 1) * kS + row]);
            const uint32_t m4 = m * 0x01010101u;
            uint32_t sum = 0;
#pragma unroll
            for (int i = 0; i < SH / 4; ++i) {
              const uint32_t d4 = m4 - pk[i];   // per byte d >= 0: no borrows
              pk[i] = d4;
              sum += shr_clamp(256u, d4 & 0xFF) + shr_clamp(256u, (d4 >> 8) & 0xFF) + shr_clamp(256u, (d4 >> 16) & 0xFF) +
                     shr_clamp(256u, d4 >> 24);
            }
            x_sum[half * kS + row] = sum;
            xsync(1);
            const uint32_t tot = sum + x_sum[(half ^ 1) * kS + row];
            const uint32_t inv8 = ((255u * 65536u) / tot) >> 8;   // (n*inv) >> 16 == inv >> (8 + d) for n = 256 >> d
#pragma unroll
            for (int i = 0; i < SH / 4; ++i) {
              const uint32_t d4 = pk[i];
              out[i] = shr_clamp(inv8, d4 & 0xFF) | (shr_clamp(inv8, (d4 >> 8) & 0xFF) << 8) |
                       (shr_clamp(inv8, (d4 >> 16) & 0xFF) << 16) | (shr_clamp(inv8, d4 >> 24) << 24);
            }
          } else {
            // recalled golden-model streaming softmax: 16-column groups, running max, eps = 1/32.
            // The row is walked in column order; the thread owning the upper half continues from the lower half's state.
            uint32_t m = 0;      // biased running max (== -128)
            uint32_t tot = 0;
            if (half == 1) {
              xsync(0);
              m = x_max[row];
              tot = x_sum[row];
            }
#pragma unroll
            for (int g = 0; g < SH / 16; ++g) {
              uint32_t cm = 0;
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint32_t w = pk[g * 4 + i];
                cm = max(max(cm, w & 0xFF), max((w >> 8) & 0xFF, max((w >> 16) & 0xFF, w >> 24)));
              }
              if (cm > m) {
                tot = shr_clamp(tot, (cm - m + 16) >> 5);
                m = cm;
              }
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint32_t w = pk[g * 4 + i];
                tot += (256u >> ((m - (w & 0xFF) + 16) >> 5)) + (256u >> ((m - ((w >> 8) & 0xFF) + 16) >> 5)) +
                       (256u >> ((m - ((w >> 16) & 0xFF) + 16) >> 5)) + (256u >> ((m - (w >> 24) + 16) >> 5));
              }
            }
            if (half == 0) {
              x_max[row] = m;
              x_sum[row] = tot;
              mbar_arrive(sc.bars + BAR_SM);
            } else {
              x_max[kS + row] = m;
              x_sum[kS + row] = tot;
            }
            xsync(1);                                          // both S chunks' MMAs are complete past this point
            m = x_max[kS + row];
            tot = x_sum[kS + row];
            const uint32_t inv = (255u * 256u) / (tot < 1 ? 1 : tot);
#pragma unroll
            for (int i = 0; i < SH / 4; ++i) {
              const uint32_t w = pk[i];
              out[i] = (inv >> ((m - (w & 0xFF) + 16) >> 5)) | ((inv >> ((m - ((w >> 8) & 0xFF) + 16) >> 5)) << 8) |
                       ((inv >> ((m - ((w >> 16) & 0xFF) + 16) >> 5)) << 16) | ((inv >> ((m - (w >> 24) + 16) >> 5)) << 24);
            }
          }
#pragma unroll
          for (int i = 0; i < SH / 16; ++i) {
            const uint4 o = make_uint4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]);
            *reinterpret_cast<uint4 *>(k_buf + ((half * SH + i * 16) >> 4) * (kS * 16) + row * 16) = o;
            if constexpr (DUMP) {
              if (prm.dAsm) *reinterpret_cast<uint4 *>(prm.dAsm + (f * kS + row) * kS + half * SH + i * 16) = o;
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(sc.bars + BAR_DEP + STG_S);
        }
      }
    });
  }
}

template <int E, int P, int MODE, int SOFTMAX, bool DUMP>
__global__ void __launch_bounds__(kThreads, 1) k_ita_fused(const __grid_constant__ CUtensorMap xmap, const FusedParams prm) {
  using C = Cfg<E, P>;
  extern __shared__ __align__(1024) uint8_t smem[];
  int8_t *s_w = reinterpret_cast<int8_t *>(smem + C::OFF_W);
  const long long *s_const = reinterpret_cast<const long long *>(smem + C::OFF_CONST);
  uint32_t *s_xchg = reinterpret_cast<uint32_t *>(smem + C::OFF_XCHG);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + C::OFF_BAR);
  uint64_t *bar_w = bars;                          // weights + constants landed
  uint64_t *slot_bars = bars + 1;                  // [slot][kBarsPerSlot]
  uint32_t *s_tmem = reinterpret_cast<uint32_t *>(bars + 1 + kSlots * kBarsPerSlot);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long nframes = prm.batch;
  // frame f of slot s in iteration it:  f = (it * gridDim.x + blockIdx.x) * kSlots + s
  const long long frame_stride = (long long)gridDim.x * kSlots;
  auto iters_of = [&](int slot) -> long long {
    long long first = (long long)blockIdx.x * kSlots + slot;
    if (first >= nframes) return 0;
    return (nframes - first + frame_stride - 1) / frame_stride;
  };

  if (threadIdx.x == 0) {
    mbar_init(bar_w, 1);
    for (int s = 0; s < kSlots; ++s) {
      uint64_t *b = slot_bars + s * kBarsPerSlot;
      mbar_init(b + BAR_X, 1);
      mbar_init(b + BAR_XFREE, 1);
      for (int r = 0; r < kRegions; ++r) {
        mbar_init(b + BAR_FULL + r, 1);
        mbar_init(b + BAR_EMPTY + r, 4);
      }
      for (int st = STG_Q; st <= STG_AV; ++st) mbar_init(b + BAR_DEP + st, C::dep_count(st));
      for (int i = 0; i < 2; ++i) mbar_init(b + BAR_SM + i, kEpiWarpsPerSlot * 32);
    }
    fence_mbar_init();
  }
  if (warp == kWarpLoad) tmem_alloc<512>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == kWarpLoad) {
    // ================================================================= loader: weights once, then the X tiles
    if (lane == 0) {
      tma_prefetch_desc(&xmap);
      mbar_arrive_expect_tx(bar_w, C::WPANEL_BYTES);
      bulk_g2s(smem + C::OFF_W, prm.wpanels, C::WPANEL_BYTES, bar_w);
      long long iters[kSlots];
      long long mx = 0;
      for (int s = 0; s < kSlots; ++s) {
        iters[s] = iters_of(s);
        mx = iters[s] > mx ? iters[s] : mx;
        if (iters[s] > 0)
          load_x_tile<E>(smem + C::OFF_SLOT + s * C::SLOT_BYTES, &xmap, (long long)blockIdx.x * kSlots + s,
                         slot_bars + s * kBarsPerSlot + BAR_X);
      }
      bool ok = true;
      for (long long it = 0; it + 1 < mx && ok; ++it)
        for (int s = 0; s < kSlots && ok; ++s)
          if (it + 1 < iters[s]) {
            // the V projections of frame `it` have consumed this slot's X tile
            ok = mbar_wait(slot_bars + s * kBarsPerSlot + BAR_XFREE, (uint32_t)(it & 1), prm.abort_flag);
            if (ok)
              load_x_tile<E>(smem + C::OFF_SLOT + s * C::SLOT_BYTES, &xmap, (long long)blockIdx.x * kSlots + s + (it + 1) * frame_stride,
                             slot_bars + s * kBarsPerSlot + BAR_X);
          }
    }
    __syncwarp();
  } else if (warp >= kWarpMma) {
    // ================================================================= MMA issue for one slot (one elected thread)
    const int slot = warp - kWarpMma;
    const long long iters = iters_of(slot);
    if (lane == 0 && iters > 0) {
      uint64_t *b = slot_bars + slot * kBarsPerSlot;
      bool ok = mbar_wait(bar_w, 0, prm.abort_flag);
      // instruction descriptors (M = 128)
      constexpr uint32_t ID_PROJ = make_idesc_i8(kS, 64, 1, 1, 0, 0);      // X . W^T           64 of the P columns
      constexpr uint32_t ID_QK = make_idesc_i8(kS, 64, 1, 1, 0, 0);        // Q . K^T           64 key columns
      constexpr uint32_t ID_AV = make_idesc_i8(kS, 64, 0, 1, 0, 1);        // Asm(u8) . V       64 of the P columns, B MN-major
      constexpr uint32_t ID_O = make_idesc_i8(kS, C::OC, 1, 1, 0, 0);      // O . Wo            E/2 columns
      const uint32_t w_addr = smem_u32(s_w);
      const uint32_t slot_addr = smem_u32(smem + C::OFF_SLOT + slot * C::SLOT_BYTES);
      const uint32_t x_addr = slot_addr, q_addr = slot_addr + C::X_BYTES, k_addr = q_addr + C::SP_BYTES, v_addr = k_addr + C::KA_BYTES;
      const uint32_t d_base = tmem_base + slot * kTmemColsPerSlot;
      for (long long it = 0; it < iters && ok; ++it) {
        const uint32_t ph = (uint32_t)(it & 1);
        static_for<0, C::NCH>([&](auto cc) {
          constexpr int c = decltype(cc)::value;
          constexpr int stage = C::stage_of(c), idx = C::idx_of(c), r = c % kRegions;
          if (!ok) return;
          const uint32_t use = (uint32_t)(it * (C::NCH / kRegions)) + c / kRegions;
          if (use > 0) ok = mbar_wait(b + BAR_EMPTY + r, (use - 1) & 1, prm.abort_flag);
          if constexpr (idx == 0) {
            if constexpr (stage == STG_Q) ok = ok && mbar_wait(b + BAR_X, ph, prm.abort_flag);
            if constexpr (stage == STG_S) ok = ok && mbar_wait(b + BAR_DEP + STG_Q, ph, prm.abort_flag) && mbar_wait(b + BAR_DEP + STG_K, ph, prm.abort_flag);
            if constexpr (stage == STG_AV) ok = ok && mbar_wait(b + BAR_DEP + STG_V, ph, prm.abort_flag) && mbar_wait(b + BAR_DEP + STG_S, ph, prm.abort_flag);
            if constexpr (stage == STG_O) ok = ok && mbar_wait(b + BAR_DEP + STG_AV, ph, prm.abort_flag);
          }
          if (!ok) return;
          tc_fence_after();
          const uint32_t d_tmem = d_base + r * kRegionCols;
          if constexpr (stage <= STG_V) {
            // projections: A = X [S rows, K = E] K-major; B = rows [64 idx, +64) of W_t^T [P rows, K = E] K-major
            const uint32_t wt = w_addr + stage * C::W_BYTES + idx * 64 * 16;
#pragma unroll
            for (int k = 0; k < E / 32; ++k)
              umma_i8(d_tmem, make_smem_desc(x_addr + k * 2 * kS * 16, kS * 16, 128), make_smem_desc(wt + k * 2 * P * 16, P * 16, 128),
                      ID_PROJ, k > 0);
          } else if constexpr (stage == STG_S) {
            // logits: A = Q [S, K = P] K-major; B = rows (keys) [64 idx, +64) of K [S rows, K = P] K-major
#pragma unroll
            for (int k = 0; k < P / 32; ++k)
              umma_i8(d_tmem, make_smem_desc(q_addr + k * 2 * kS * 16, kS * 16, 128),
                      make_smem_desc(k_addr + idx * 64 * 16 + k * 2 * kS * 16, kS * 16, 128), ID_QK, k > 0);
          } else if constexpr (stage == STG_AV) {
            // context: A = Asm [S, K = S] u8 K-major (in the K buffer); B = columns [64 idx, +64) of V [K = S rows, N = P] MN-major
#pragma unroll
            for (int k = 0; k < kS / 32; ++k)
              umma_i8(d_tmem, make_smem_desc(k_addr + k * 2 * kS * 16, kS * 16, 128),
                      make_smem_desc(v_addr + idx * 4 * kS * 16 + k * 32 * 16, 128, kS * 16), ID_AV, k > 0);
          } else {
            // output projection: A = O [S, K = P] K-major (in the Q buffer); B = rows [OC idx, +OC) of Wo^T [E rows, K = P] K-major
            const uint32_t wo = w_addr + 3 * C::W_BYTES + idx * C::OC * 16;
#pragma unroll
            for (int k = 0; k < P / 32; ++k)
              umma_i8(d_tmem, make_smem_desc(q_addr + k * 2 * kS * 16, kS * 16, 128), make_smem_desc(wo + k * 2 * E * 16, E * 16, 128),
                      ID_O, k > 0);
          }
          umma_commit(b + BAR_FULL + r);
          if constexpr (stage == STG_V && idx == C::NQ - 1) umma_commit(b + BAR_XFREE);   // X tile may be overwritten
        });
      }
    }
    __syncwarp();
  } else {
    // ================================================================= epilogue warps
    const int slot = warp / kEpiWarpsPerSlot;
    const int quad = warp & 3;
    const int w2 = (warp >> 2) & 1;
    SlotCtx sc;
    sc.bars = slot_bars + slot * kBarsPerSlot;
    sc.smem = smem + C::OFF_SLOT + slot * C::SLOT_BYTES;
    sc.tmem = tmem_base + slot * kTmemColsPerSlot;
    uint32_t *x_max = s_xchg + (slot * 2) * kS;            // [half][row]
    uint32_t *x_sum = s_xchg + (kSlots * 2 + slot * 2) * kS;
    const long long iters = iters_of(slot);
    if (iters > 0 && mbar_wait(bar_w, 0, prm.abort_flag)) {
      const long long frame0 = (long long)blockIdx.x * kSlots + slot;
      if (w2 == 0)
        epilogue_loop<E, P, MODE, SOFTMAX, DUMP, 0>(prm, sc, s_const, x_max, x_sum, slot, quad, lane, iters, frame0, frame_stride);
      else
        epilogue_loop<E, P, MODE, SOFTMAX, DUMP, 1>(prm, sc, s_const, x_max, x_sum, slot, quad, lane, iters, frame0, frame_stride);
    }
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == kWarpLoad) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------------ host side
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

typedef void (*KernelFn)(const CUtensorMap, const FusedParams);

struct FusedPlan {
  int E, P, mode, softmax;
  int smem_bytes;
  int8_t *wpanels = nullptr;   // device
  FRq rq[6];
  KernelFn fn = nullptr, fn_dump = nullptr;
};

template <int E, int P>
static void pick_kernels(int mode, int softmax, KernelFn &fn, KernelFn &fnd) {
#define ITA_PICK(M, SM)                                     \
  if (mode == M && softmax == SM) {                         \
    fn = k_ita_fused<E, P, M, SM, false>;                   \
    fnd = k_ita_fused<E, P, M, SM, true>;                   \
  }
  ITA_PICK(RQM_INT_HI, 0) ITA_PICK(RQM_INT_HI, 1) ITA_PICK(RQM_INT_GEN, 0) ITA_PICK(RQM_INT_GEN, 1) ITA_PICK(RQM_TORCH, 0)
  ITA_PICK(RQM_TORCH, 1)
#undef ITA_PICK
}

// panel layout of a [R x C] matrix given element accessor
template <class F>
static void to_panels(int8_t *dst, int R, int Ccols, F at) {
  for (int c = 0; c < Ccols; ++c)
    for (int r = 0; r < R; ++r) dst[(size_t)(c / 16) * R * 16 + (size_t)r * 16 + (c % 16)] = at(r, c);
}

int fused_plan_create(ita_mha_handle *h, const ita_mha_desc *d) {
  h->fused = nullptr;
  auto no = [&](const std::string &why) {
    h->fused_reason = why;
    return ITA_OK;
  };
  if (h->S != kS) return no("fused kernel needs S == 128");
  if (h->H != 1) return no("fused kernel needs H == 1");
  const int E = h->E, P = h->P;
  int smem_bytes = 0;
  bool shape_ok = false;
#define ITA_SHAPE(EE, PP)                        \
  if (E == EE && P == PP) {                      \
    shape_ok = true;                             \
    smem_bytes = Cfg<EE, PP>::SMEM_BYTES;        \
  }
  ITA_SHAPE(64, 192) ITA_SHAPE(64, 64) ITA_SHAPE(64, 128) ITA_SHAPE(128, 64)
#undef ITA_SHAPE
  if (!shape_ok) return no("fused kernel is instantiated for (E,P) in {(64,192),(64,64),(64,128),(128,64)}");
  int max_smem = 0;
  ITA_CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
  if (smem_bytes > max_smem) return no("fused kernel needs more shared memory than the device offers");
  int cc_major = 0;
  ITA_CUDA_TRY(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, h->device));
  if (cc_major != 10) return no("fused kernel needs an sm_100 device");
  if (!get_encode()) return no("cuTensorMapEncodeTiled is unavailable");

  // ---- requant mode: every stage must be representable
  int mode;
  const int stages[6] = {ITA_ST_Q, ITA_ST_K, ITA_ST_V, ITA_ST_QK, ITA_ST_AV, ITA_ST_O};
  if (h->round_mode == ITA_ROUND_TORCH_RNE_F32) {
    mode = RQM_TORCH;
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
  // accumulators + bias must stay inside s32 with margin (they do for every int8 model; checked, not assumed)
  {
    long long worst = 0;
    const int32_t *Bs[4] = {d->Bq, d->Bk, d->Bv, d->Bo};
    const int n[4] = {P, P, P, E};
    const long long kdim[4] = {E, E, E, P};
    for (int t = 0; t < 4; ++t)
      for (int j = 0; j < n[t]; ++j) {
        long long b = Bs[t][j] < 0 ? -(long long)Bs[t][j] : Bs[t][j];
        long long w = b + kdim[t] * 128 * 128;
        worst = w > worst ? w : worst;
      }
    if (worst >= (1LL << 30)) return no("bias magnitude too large for the fused kernel's s32 accumulate");
  }

  FusedPlan *pl = new FusedPlan();
  pl->E = E; pl->P = P; pl->mode = mode; pl->softmax = h->softmax_mode; pl->smem_bytes = smem_bytes;
  const size_t wbytes = (size_t)4 * P * E + (size_t)(3 * P + E) * 8;
  std::vector<int8_t> host(wbytes);
  // W_t^T as [N = P rows, K = E cols]: element (n, k) = W[k][n]  (golden layout W [E,P])
  const int8_t *Ws[3] = {d->Wq, d->Wk, d->Wv};
  for (int t = 0; t < 3; ++t) to_panels(host.data() + (size_t)t * P * E, P, E, [&](int n, int k) { return Ws[t][(size_t)k * P + n]; });
  // Wo^T as [N = E rows, K = P cols]: element (n, k) = Wo[k][n]  (golden layout Wo [P,E])
  to_panels(host.data() + (size_t)3 * P * E, E, P, [&](int n, int k) { return d->Wo[(size_t)k * E + n]; });
  long long *consts = reinterpret_cast<long long *>(host.data() + (size_t)4 * P * E);
  const int32_t *Bs[3] = {d->Bq, d->Bk, d->Bv};
  for (int i = 0; i < 6; ++i) {
    const Rq &r = h->rq_host[stages[i] * h->H];
    FRq f{};
    if (mode == RQM_TORCH) {
      f.multp = 0; f.shift = 0; f.add = r.add; f.eff = r.eff; f.c = 0;
    } else if (mode == RQM_INT_HI) {
      f.multp = (int32_t)((long long)r.mult << (32 - r.shift));
      f.shift = r.shift; f.add = r.add; f.eff = 0.f;
      f.c = (r.rnd << (32 - r.shift)) + ((long long)r.add << 32);
    } else {
      f.multp = r.mult; f.shift = r.shift; f.add = r.add; f.eff = 0.f; f.c = r.rnd;
    }
    pl->rq[i] = f;
  }
  auto col_const = [&](int stage_i, int32_t bias) -> long long {
    const Rq &r = h->rq_host[stages[stage_i] * h->H];
    if (mode == RQM_TORCH) return (long long)(uint32_t)bias;
    if (mode == RQM_INT_HI) return (((long long)bias * r.mult + r.rnd) << (32 - r.shift)) + ((long long)r.add << 32);
    return (long long)bias * r.mult + r.rnd;
  };
  for (int t = 0; t < 3; ++t)
    for (int j = 0; j < P; ++j) consts[t * P + j] = col_const(t, Bs[t][j]);
  for (int j = 0; j < E; ++j) consts[3 * P + j] = col_const(5, d->Bo[j]);

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
#define ITA_SHAPE(EE, PP) \
  if (E == EE && P == PP) pick_kernels<EE, PP>(mode, h->softmax_mode, pl->fn, pl->fn_dump);
  ITA_SHAPE(64, 192) ITA_SHAPE(64, 64) ITA_SHAPE(64, 128) ITA_SHAPE(128, 64)
#undef ITA_SHAPE
  cudaError_t e1 = cudaFuncSetAttribute((const void *)pl->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  cudaError_t e2 = cudaFuncSetAttribute((const void *)pl->fn_dump, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  if (e1 != cudaSuccess || e2 != cudaSuccess) {
    cudaGetLastError();
    cudaFree(pl->wpanels);
    delete pl;
    return no(std::string("cudaFuncSetAttribute(max dynamic smem) failed: ") + cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
  }
  h->fused = pl;
  return ITA_OK;
}

void fused_plan_destroy(ita_mha_handle *h) {
  if (h->fused) {
    cudaFree(h->fused->wpanels);
    delete h->fused;
    h->fused = nullptr;
  }
}

int fused_forward(ita_mha_handle *h, const int8_t *x, int8_t *y, int64_t B, const ita_mha_stages *dump, cudaStream_t s) {
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
  prm.y = y;
  prm.batch = B;
  prm.abort_flag = h->abort_flag;
  for (int i = 0; i < 6; ++i) prm.rq[i] = pl->rq[i];
  if (dump) {
    prm.dQ = dump->Qp; prm.dK = dump->Kp; prm.dV = dump->Vp; prm.dA = dump->A; prm.dAsm = dump->Asm; prm.dO = dump->O;
  }
  long long ctas = (B + kSlots - 1) / kSlots;
  if (ctas > h->sm_count) ctas = h->sm_count;
  KernelFn fn = dump ? pl->fn_dump : pl->fn;
  fn<<<(unsigned)ctas, kThreads, pl->smem_bytes, s>>>(map, prm);
  count_launch();
  ITA_CUDA_TRY(cudaGetLastError());
  return ITA_OK;
}

}  // namespace ita
