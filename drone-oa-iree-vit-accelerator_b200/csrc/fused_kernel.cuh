// fused_kernel.cuh -- the hot path (device code; instantiated per shape by fused_inst_*.cu, host side in fused.cu): one persistent sm_100a kernel that runs a whole ITA attention layer per depth frame
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
// Biases ride on the tensor pipe: every weight matrix carries extra K columns holding the int32 bias of its output
// column as base-R digits (b = g0 + R * (g1 + ... + g31)), and the matching A operand of that extra K step is a constant
// tile whose rows are [1, R, ..., R] (8 rows in shared memory, SBO = 0 so that every row group of the operand reads them).
// The accumulator that reaches the epilogue already contains the bias, so requantisation needs only per-stage *uniform*
// constants (no per-column loads).  R = 127 with signed digits for the integer requant forms, 255 with unsigned digits
// for RQM_FP, where the digits also carry the stage's accumulator offset (see "requantisation on the FMA pipe").
//
// Pipeline.  The three projections are ONE GEMM against the concatenated weights [Wq | Wv | Wk] (3P output columns).  A
// frame is cut into NCH "chunks", each one accumulator tile of 128 (or a trailing 64) columns, in the order
//     P[0..NPC)  S  AV[0..NAVC)  O           (NPC = ceil(3P/128), S = all 128 keys, NAVC = ceil(P/128), O = E columns)
//   Two frames ("slots") are in flight per SM.  Per slot, TMEM holds a ring of 2 regions x 128 columns; the n-th chunk
//   of the slot uses region n % 2.  Each slot has two epilogue warpgroups (4 warps = the 4 TMEM lane quadrants); BOTH drain
//   every chunk, warpgroup w taking the sub-block [w*SW, (w+1)*SW) of its columns (SW = 64 or 32).  For the logits that
//   means a thread of warpgroup A holds keys 0..63 of its row and its twin in warpgroup B keys 64..127; they exchange the
//   row max / row sum through shared memory.  One MMA-issuing thread per slot walks the chunks in order, waiting only for
//   (a) the region to be drained and (b) the operand buffers of the chunk to be written (mbarriers armed by the
//   producing epilogues); a loader thread prefetches the next frame's X tile by TMA.  tcgen05 MMAs of one thread complete
//   in issue order, which is what makes the reuse of the operand buffers (Asm over K, O over Q, the next X tile in the
//   unused tail of the K buffer) safe.  The two slots are kept out of phase (BAR_TURN).
//
// Thread roles (608 threads): warps 0-7 = epilogue of slot 0 (warp w: TMEM lanes 32*(w%4).., warpgroup (w/4)%2),
// warps 8-15 = slot 1, warps 16/17 = MMA issue for slot 0/1 (one elected thread), warp 18 = TMA loader + TMEM allocator.
//
// Reference semantics: models/ITA/QAT/layers.py:103-126 and tests/export_and_validation_W_B.py:518-542 (stage order,
// layouts), models/ITA/QAT/ITA_softmax.py:51-61 (ITAmax), SURVEY.md App. A (integer spec).
#pragma once
#include <cuda.h>
#include <stdlib.h>

#include <type_traits>
#include <vector>

#include "common.cuh"
#include "ptx.cuh"

namespace ita {

constexpr int kS = 128;                 // rows per tile = tcgen05 M
constexpr int kMaxSlots = 2;             // frames in flight per SM (Cfg<E,P>::SLOTS: 2 when the operand buffers fit twice, else 1)
constexpr int kEpiWarpsPerSlot = 8;
constexpr int kSmemLimit = 232448;       // 227 KB: the per-block opt-in maximum of sm_100
constexpr int kRegions = 2;             // accumulator ring per slot
constexpr int kRegionCols = 128;
constexpr int kTmemColsPerSlot = kRegions * kRegionCols;
constexpr int kBarsPerSlot = 16;
constexpr int kBiasK = 32;              // extra K columns carrying the bias digits (one K step)
constexpr int kBiasKO = 64;             // the output projection reserves two such steps (RQM_FP needs the range)
constexpr int kBiasRadix = 127;         // signed digits (integer requant forms)
constexpr int kBiasRadixU = 255;        // unsigned digits (RQM_FP)
// Constant operand tiles, 8 rows each, read with SBO = 0 so that every 8-row group of the MMA operand aliases them:
//   CA  [8 x 96]  A operand of the digit K steps: rows [1, R, R, ..., R]  (R = 127 s8 / 255 u8)
//   DQK [8 x 96]  B operand of the offset steps of Q.K^T (two for RQM_FP / RQM_TORCH_FP, three for RQM_TORCH_FP2, whose
//                 offset is a power of two)                               DAV [8 x 32]  B operand of the offset step of A.V
constexpr int kConstABytes = 8 * 96, kConstQKBytes = 8 * 96, kConstAVBytes = 8 * 32;
constexpr int kConstBytes = kConstABytes + kConstQKBytes + kConstAVBytes;

enum { RQM_INT_HI = 0, RQM_INT_GEN = 1, RQM_TORCH = 2, RQM_FP = 3, RQM_TORCH_FP = 4, RQM_TORCH_FP2 = 5 };
__host__ __device__ constexpr bool rqm_offset(int mode) { return mode == RQM_FP || mode == RQM_TORCH_FP || mode == RQM_TORCH_FP2; }   // offset accumulators
__host__ __device__ constexpr bool rqm_torch_fp(int mode) { return mode == RQM_TORCH_FP || mode == RQM_TORCH_FP2; }
__host__ __device__ constexpr int rqm_qk_steps(int mode) { return mode == RQM_TORCH_FP2 ? 3 : 2; }   // offset digit K steps of Q.K^T
enum { STG_Q = 0, STG_K, STG_V, STG_S, STG_AV, STG_O };
// per-slot mbarriers
enum { BAR_X = 0, BAR_XFREE = 1, BAR_FULL = 2 /* + region */, BAR_EMPTY = 4 /* + region */, BAR_CHUNK = 6 /* + position: its epilogue has written its operand buffer */, BAR_TURN = 15 };
enum { CK_PROJ = 0, CK_S, CK_AV, CK_O };   // chunk kinds

// Per-stage uniform requant parameters as the fused kernel consumes them.
struct FRq {
  int32_t multp;      // INT_HI: mult << (32 - shift);  INT_GEN: mult
  int32_t shift;      // INT_GEN, and the 32-bit form of INT_HI
  int32_t add;        // INT_GEN / TORCH
  float eff;          // TORCH
  long long c;        // INT_HI: (rnd << (32-shift)) + (add << 32);  INT_GEN: rnd
  int32_t mult32;     // 32-bit form of INT_HI:  (acc * mult32 + c32) >> shift   (host-checked not to overflow)
  int32_t c32;        //   rnd + (add << shift)
  // FP form (see "requantisation on the FMA pipe" below)
  int32_t init;       // offset the digit K steps add to every accumulator of the stage
  int32_t negc;       // -C
  float fb;           // mult * 2^-shift
  float fc;           // C * 2^-149
};

struct FusedParams {
  const int8_t *wpanels;     // device: [Wq^T | Wk^T | Wv^T | Wo^T] (each with its bias-digit K columns) | constant A tile
  const int8_t *x;           // the input frames (the tiles are loaded through the tensor map; this is for the L2 prefetch)
  int8_t *y;
  long long batch;
  int *abort_flag;
  FRq rq[6];                 // Q K V QK AV O
  // debug dump (natural layouts), any may be null
  int8_t *dQ, *dK, *dV, *dA;
  uint8_t *dAsm;
  int8_t *dO;
  int torch_fast;            // RQM_TORCH_FP: take the one-FFMA-chain ITAmax distance on unsaturated rows (host knob)
  long long *trace;          // developer builds (-DITA_TRACE): [warp][iteration < 8][chunk][4] SM clock stamps of CTA 0
};

template <int E, int P>
struct Cfg {
  static constexpr int S = kS;
  static constexpr int PW = 3 * P;                                       // concatenated projection output: Q | K | V
  static constexpr int NQK = 2 * P / 128;                                // projection chunks with Q and K columns (P % 64 == 0)
  static constexpr int NVC = (P + 127) / 128;                            // projection chunks with V columns (the last may be 64 wide)
  static constexpr int NPC = NQK + NVC;                                  // projection chunks
  static constexpr int NAVC = NVC;                                       // A.V chunks
  static constexpr int NCH = NPC + NAVC + 2;                             // + logits + output projection
  // Schedule ("position" = order of MMA issue = order of the epilogue = order in the accumulator ring):
  //     QK[0..NQK)  V[0]  S  V[1..NVC)  AV[0..NAVC)  O
  // The first V chunk is issued BEFORE the logits and drained while the logits MMA runs; the remaining V chunk (P = 192)
  // runs on the tensor pipe during ITAmax and is drained while A.V[0] runs (which needs V[0] only).
  static constexpr int C_S = NQK + 1, C_AV0 = C_S + NVC, C_O = C_AV0 + NAVC;
  static constexpr int C_LASTPROJ = NVC > 1 ? C_AV0 - 1 : NQK;           // the last reader of the X tile
#ifdef ITA_C_TURN
  static constexpr int C_TURN = ITA_C_TURN;                              // tuning builds
#else
  static constexpr int C_TURN = NQK < 2 ? NQK : 2;                       // chunk whose issue lets the other slot start its frame
#endif
  static constexpr int WP_BYTES = PW * (E + kBiasK);                     // [Wq | Wk | Wv]^T with bias digits
  static constexpr int WO_BYTES = E * (P + kBiasKO);                     // Wo^T with bias digits
  static constexpr int OFF_W = 0;
  static constexpr int OFF_WO = WP_BYTES;
  static constexpr int OFF_CA = WP_BYTES + WO_BYTES, OFF_DQK = OFF_CA + kConstABytes, OFF_DAV = OFF_DQK + kConstQKBytes;
  static constexpr int WPANEL_BYTES = OFF_CA + kConstBytes;              // what is bulk-copied at kernel start
  static constexpr int X_BYTES = S * E;
  static constexpr int SP_BYTES = S * P;
  static constexpr int KA_BYTES = (S * P > S * S) ? S * P : S * S;       // K, later the softmax output
  // The X tile lives in the columns of the V buffer that the LAST V chunk writes (its epilogue runs after its own MMA, the
  // last reader of X), when it fits there; the next frame's tile is loaded once the last A.V MMA has read V.
  static constexpr int VLAST_COL0 = 128 * (NVC - 1), VLAST_W = P - VLAST_COL0;
  static constexpr bool X_IN_V = X_BYTES <= S * VLAST_W;
  static constexpr int OFF_Q = 0, OFF_K = SP_BYTES, OFF_V = OFF_K + KA_BYTES;
  static constexpr int OFF_X = X_IN_V ? OFF_V + S * VLAST_COL0 : OFF_V + SP_BYTES;
  static constexpr int C_XFREE = X_IN_V ? C_O - 1 : C_LASTPROJ;          // chunk whose completion frees the X buffer
  static constexpr int SLOT_BYTES = OFF_V + SP_BYTES + (X_IN_V ? 0 : X_BYTES);
  static constexpr int OFF_SLOT = ((WPANEL_BYTES + 127) / 128) * 128;
  // frames in flight per SM: two when weights + 2 x operand buffers fit the 227 KB, else one (E = 128 with P = 192)
  static constexpr int smem_for(int slots) { return OFF_SLOT + slots * SLOT_BYTES + slots * 2 * S * 8 + 512; }
  static constexpr int SLOTS = smem_for(kMaxSlots) <= kSmemLimit ? kMaxSlots : 1;
  static constexpr int WARP_MMA = SLOTS * kEpiWarpsPerSlot;              // + slot: the MMA-issuing warp of the slot
  static constexpr int WARP_LOAD = WARP_MMA + SLOTS;                     // TMA loader + TMEM allocator
  static constexpr int THREADS = (WARP_LOAD + 1) * 32;
  static constexpr int OFF_XCHG = OFF_SLOT + SLOTS * SLOT_BYTES;         // softmax pair exchange: [slot][2 halves][S] int2
  static constexpr int XCHG_BYTES = SLOTS * 2 * S * 8;
  static constexpr int OFF_BAR = OFF_XCHG + XCHG_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 512;
  static_assert(SMEM_BYTES <= kSmemLimit, "shape does not fit shared memory even with one frame in flight");
  static_assert(E % 64 == 0 && E <= 128 && P % 64 == 0, "E in {64, 128} (one output chunk), P a multiple of 64 (sub-block width)");
  static_assert((1 + kMaxSlots * kBarsPerSlot) * 8 + 8 <= 512, "barrier area");
  static_assert(BAR_CHUNK + NCH - 1 <= BAR_TURN, "one 'chunk written' barrier per position but the last");
  static_assert(WPANEL_BYTES % 16 == 0 && SLOT_BYTES % 128 == 0, "alignment");
  // position c of a frame -> kind, first column (in the concatenated Q|K|V space for projections), width (128 or 64)
  static constexpr int kind_of(int c) { return c == C_S ? CK_S : c < C_AV0 ? CK_PROJ : c < C_O ? CK_AV : CK_O; }
  static constexpr int pidx(int c) { return c < C_S ? c : c - 1; }       // projection chunk number of a CK_PROJ position
  static constexpr int pos_of_proj(int i) { return i <= NQK ? i : i + 1; }
  static constexpr int col0_of(int c) { return kind_of(c) == CK_PROJ ? pidx(c) * 128 : kind_of(c) == CK_AV ? (c - C_AV0) * 128 : 0; }
  static constexpr int ncols_of(int c) {
    const int k = kind_of(c);
    const int left = k == CK_PROJ ? PW - pidx(c) * 128 : k == CK_AV ? P - (c - C_AV0) * 128 : k == CK_S ? S : E;
    return left < 128 ? left : 128;
  }
  // Both epilogue warpgroups of the slot drain every chunk: warpgroup w takes the sub-block [w * SW, (w + 1) * SW) of the
  // chunk's columns, SW = ncols / 2 (64 or 32).  A sub-block never straddles two stages (P is a multiple of 64).
  static constexpr int sub_w(int c) { return ncols_of(c) / 2; }
  static constexpr int stage_at(int c, int col) {                        // stage that owns column `col` of position c
    const int k = kind_of(c);
    if (k != CK_PROJ) return k == CK_S ? STG_S : k == CK_AV ? STG_AV : STG_O;
    const int gc = pidx(c) * 128 + (col < ncols_of(c) ? col : 0);
    return gc < P ? STG_Q : gc < 2 * P ? STG_K : STG_V;
  }
  static constexpr int stage_of(int c, int w) { return stage_at(c, w * sub_w(c)); }
  static constexpr int stage_col0(int c, int w) {                        // first column of the sub-block inside its stage's matrix
    const int gc = col0_of(c) + w * sub_w(c);
    return kind_of(c) == CK_PROJ ? gc % P : gc;
  }
  // Operand dependencies of the MMA K steps, as the LAST position whose epilogue must have written its buffer (epilogues run
  // in position order, so one "chunk written" barrier per position orders everything before it as well):
  //   logits K step k (32 of the P columns of Q and K): the projection chunk holding K[:, 32k..]
  //   A.V chunk j: the logits (ITAmax output) and the V chunk with the same columns
  //   output projection K step k (32 context columns): the A.V chunk holding them
  static constexpr int dep_S(int k) { return pos_of_proj((P + 32 * k) / 128); }
  static constexpr int dep_AV(int j) { return j == 0 ? C_S : C_S + j; }
  static constexpr int dep_O(int k) { return C_AV0 + (32 * k) / 128; }
};

// ------------------------------------------------------------------------------- requantisation on the FMA pipe (RQM_FP)
// The integer forms below cost the two scarce pipes (IMAD.HI: 4 fmaheavy cycles per warp; IMAD + SHF: fmaheavy + alu).
// RQM_FP moves the multiply, the rounding and the shift of  y = ((acc*mult + rnd) >> shift) + add  into ONE FFMA.RM:
//   * the digit K steps add a per-stage offset INIT > |acc|max on the tensor pipe (unsigned radix-255 digits against the
//     constant tile: folded into the bias digits of the projections, two extra steps for Q.K^T, one for A.V), so TMEM holds
//     bits = INIT + acc with 0 < bits < 2^24.  Read as an f32 such a pattern (exponent field 0 or 1) is exactly bits * 2^-149;
//   * INIT is chosen per stage with INIT*mult == rnd (mod 2^shift), so C = (rnd - INIT*mult)/2^shift + add + 128 is an
//     integer and  fma.rm(bits_f32, mult*2^-shift, C*2^-149)  has the exact value ((acc*mult + rnd)/2^shift + add + 128) * 2^-149;
//     rounding toward -inf at the denormal ulp is the floor, so the result's BIT PATTERN is y + 128 when that is >= 0 and
//     has the sign bit set when it is negative;
//   * cvt.pack.sat.u8.s32 of the bit patterns clamps to [0, 255] = sat_s8(y) + 128, and one XOR 0x80808080 per four
//     elements turns that into s8.  (Checked exhaustively over +-4.3M accumulators x 59 parameter sets on B200:
//     tools/fp_requant_check.cu.)
// The ITAmax distance m - a + 1 = ceil((m + 129 - C) - bits*mult*2^-shift) is one FFMA.RP the same way.
// RQM_TORCH_FP gives the reference's fp32 form  clamp(rne(fp32(acc) * eff) + zp)  the same treatment without the two
// conversions (I2F, F2I run on the quarter-rate XU pipe): u = fma.rn(bits_f32, 2^126, -INIT*2^-23) = acc*2^-23 exactly,
// t = u * (eff*2^23) rounds exactly like fp32(acc)*eff (power-of-two scalings are exact), and fma.rn(t, 2^-149, (zp+128)*2^-149)
// rounds t to the nearest-even integer in the denormal ulp: its bit pattern is rne(t) + zp + 128 (for even zp, which keeps
// the parity of ties; the reference's zero-points are 0).
// (A first version preset the accumulators with tcgen05.st instead: a store needs distinct source registers, i.e. one
//  register copy per column, which cost more than the FFMA saved.)
// ---------------------------------------------------------------------------------------------------- requant forms
// ALT selects how INT_HI elements are computed: 0 = one IMAD.HI each; 1 = 32-bit IMAD + SHF each; 2 = alternate (even
// elements IMAD.HI, odd elements 32-bit) to balance the fma pipe (IMAD.HI occupies it twice as long) against issue slots.
template <int MODE, int ALT>
__device__ __forceinline__ int rq_apply(int acc, const FRq &r, int elem) {
  if constexpr (MODE == RQM_INT_HI) {
    if (ALT == 1 || (ALT == 2 && (elem & 1))) {
      return (acc * r.mult32 + r.c32) >> r.shift;
    } else {
      // (acc*mult + rnd) >> shift + add  ==  hi32( acc * (mult << (32-shift)) + c ),  c folded on the host
      long long p = (long long)acc * (long long)r.multp + r.c;
      return (int)(p >> 32);
    }
  } else if constexpr (MODE == RQM_INT_GEN) {
    long long p = (long long)acc * (long long)r.multp + r.c;
    return (int)(p >> r.shift) + r.add;
  } else {
    float f = __fmul_rn(__int2float_rn(acc), r.eff);
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
// RQM_FP: bit pattern of  y + 128  (sign bit set when negative) from the accumulator bit pattern
__device__ __forceinline__ int rq_fp_bits(uint32_t v, const FRq &r) { return __float_as_int(__fmaf_rd(__uint_as_float(v), r.fb, r.fc)); }
// RQM_TORCH_FP: the same for the fp32 requant form; fb = -INIT*2^-23, eff = eff*2^23, fc = (zp + 128)*2^-149
__device__ __forceinline__ int rq_tfp_bits(uint32_t v, const FRq &r) {
  const float u = __fmaf_rn(__uint_as_float(v), __uint_as_float(0x7E800000u) /* 2^126 */, r.fb);
  const float t = __fmul_rn(u, r.eff);
  return __float_as_int(__fmaf_rn(t, __uint_as_float(1u) /* 2^-149 */, r.fc));
}
// RQM_TORCH_FP2: the same in two FFMAs when the stage's offset INIT is a power of two (INIT*eff is then exactly
// representable):  t' = fma.rn(bits_f32, eff*2^85, -INIT*eff*2^-64) = rn(acc*eff) * 2^-64  (one rounding of the exact product,
// the same 24-bit rounding as fp32(acc)*eff -- everything stays in the normal range), then fma.rn(t', 2^-85, (zp+128)*2^-149)
// rounds t + zp + 128 to nearest-even in the denormal ulp.  fb = eff*2^85, eff = -INIT*eff*2^-64, fc = (zp + 128)*2^-149.
constexpr uint32_t kTwoM85 = 0x15000000u;        // 2^-85
__device__ __forceinline__ int rq_tfp2_bits(uint32_t v, const FRq &r) {
  const float t = __fmaf_rn(__uint_as_float(v), r.fb, r.eff);
  return __float_as_int(__fmaf_rn(t, __uint_as_float(kTwoM85), r.fc));
}
template <int MODE>
__device__ __forceinline__ int rq_off_bits(uint32_t v, const FRq &r) {
  if constexpr (MODE == RQM_TORCH_FP) return rq_tfp_bits(v, r);
  else if constexpr (MODE == RQM_TORCH_FP2) return rq_tfp2_bits(v, r);
  else return rq_fp_bits(v, r);
}
// ---- packed f32x2 forms (FFMA2 / FMUL2, sm_100): two elements per issued instruction, same IEEE results (no .ftz: the
//      denormal-range patterns are the point).  -DITA_NO_F32X2 keeps the scalar forms (A/B timing).
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk2(uint32_t lo, uint32_t hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi)); return r; }
__device__ __forceinline__ u64 pk2f(float v) { return pk2(__float_as_uint(v), __float_as_uint(v)); }
__device__ __forceinline__ void upk2(u64 v, int &lo, int &hi) { asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v)); }
__device__ __forceinline__ u64 ffma2_rn(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 ffma2_rm(u64 a, u64 b, u64 c) { u64 d; asm("fma.rm.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 ffma2_rp(u64 a, u64 b, u64 c) { u64 d; asm("fma.rp.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 fmul2_rn(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
template <int MODE>
__device__ __forceinline__ void rq_off_bits2(uint32_t v0, uint32_t v1, const FRq &r, int &y0, int &y1) {
#ifdef ITA_NO_F32X2
  y0 = rq_off_bits<MODE>(v0, r);
  y1 = rq_off_bits<MODE>(v1, r);
#else
  const u64 a = pk2(v0, v1);
  u64 d;
  if constexpr (MODE == RQM_TORCH_FP) {
    const u64 u = ffma2_rn(a, pk2(0x7E800000u, 0x7E800000u), pk2f(r.fb));
    d = ffma2_rn(fmul2_rn(u, pk2f(r.eff)), pk2(1u, 1u), pk2f(r.fc));
  } else if constexpr (MODE == RQM_TORCH_FP2) {
    d = ffma2_rn(ffma2_rn(a, pk2f(r.fb), pk2f(r.eff)), pk2(kTwoM85, kTwoM85), pk2f(r.fc));
  } else {
    d = ffma2_rm(a, pk2f(r.fb), pk2f(r.fc));
  }
  upk2(d, y0, y1);
#endif
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

#ifdef ITA_TRACE
// Developer builds: [warp][iteration < 8][chunk][4] SM clock stamps of CTA 0 (every lane stores to the same word; the
// condition is warp-uniform; launches that are not traced still get a buffer).
constexpr int kTraceIters = 8;
#define ITA_TR(prm, warp, it, c, k)                                                                                          \
  do {                                                                                                                       \
    if (blockIdx.x == 0 && (it) < kTraceIters)   /* warp-uniform: a predicated store, no divergence */                      \
      (prm).trace[((((long long)(warp)) * (kTraceIters + 1) + (it)) * 32 + (c)) * 4 + (k)] = clock64();                      \
  } while (0)
#else
#define ITA_TR(prm, warp, it, c, k) do { } while (0)
#endif

// compile-time loop
template <int I, int N, class F>
__device__ __forceinline__ void static_for(F &&f) {
  if constexpr (I < N) {
    f(std::integral_constant<int, I>{});
    static_for<I + 1, N>(f);
  }
}

// tcgen05.wait::ld that also "redefines" the 32 destination registers of the load it waits for, so that the compiler
// cannot move or copy them between the asynchronous load and the wait.
__device__ __forceinline__ void tmem_wait32(uint32_t (&v)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;\n"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                 "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]),
                 "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]),
                 "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
               :
               : "memory");
}

// Empty asm that "redefines" 32 registers: an ordering point for the compiler, no instruction.
__device__ __forceinline__ void tmem_pin32(uint32_t (&v)[32]) {
  asm volatile(""
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                 "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]),
                 "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]),
                 "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
               :
               : "memory");
}

__device__ __forceinline__ void tmem_pin16(uint32_t (&v)[16]) {
  asm volatile(""
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                 "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])
               :
               : "memory");
}

// Walk NG groups of 32 accumulator columns of this thread's row, software-pipelined: the TMEM load of group g+1 is in
// flight while f(g, v) works on group g.  addr(g) gives the TMEM address of group g.  After the last load has landed,
// `done.issue()` runs once, and `done.finish()` after the last group has been processed (together they hand the TMEM
// region back to the MMA thread).
struct NoRelease {
  __device__ __forceinline__ void issue() const {}
  __device__ __forceinline__ void finish() const {}
};
template <class I, class Fn>
struct Release {
  I issue_fn;
  Fn finish_fn;
  __device__ __forceinline__ void issue() { issue_fn(); }
  __device__ __forceinline__ void finish() { finish_fn(); }
};
template <class I, class Fn>
__device__ __forceinline__ Release<I, Fn> make_release(I i, Fn f) { return Release<I, Fn>{i, f}; }

template <int NG, class A, class F, class D>
__device__ __forceinline__ void tmem_walk(A &&addr, F &&f, D &&done) {
  uint32_t va[32], vb[32];
#if defined(ITA_EXP_NOLD)      // timing experiment (wrong results): no TMEM reads
#pragma unroll
  for (int i = 0; i < 32; ++i) va[i] = vb[i] = addr(0) + i;
  tmem_pin32(va);
  tmem_pin32(vb);
  static_for<0, NG>([&](auto gg) {
    constexpr int g = decltype(gg)::value;
    if constexpr (g + 1 == NG) done.issue();
    f(gg, (g & 1) ? vb : va);
    if constexpr (g + 1 == NG) done.finish();
  });
#else
  tmem_ld32(addr(0), va);
  tmem_wait32(va);
  static_for<0, NG>([&](auto gg) {
    constexpr int g = decltype(gg)::value;
    uint32_t(&cur)[32] = (g & 1) ? vb : va;
    uint32_t(&nxt)[32] = (g & 1) ? va : vb;
    if constexpr (g + 1 < NG) {
      tmem_ld32(addr(g + 1), nxt);
      tmem_pin32(cur);              // keeps the compiler from hoisting the work on `cur` above the load just issued
    }
    if constexpr (g + 1 == NG) done.issue();
#if defined(ITA_EXP_NOMATH)    // timing experiment (wrong results): TMEM reads and synchronisation only
    if (cur[0] == 0x12345678u && cur[31] == 0x9abcdef0u) f(gg, cur);
#else
    f(gg, cur);
#endif
    if constexpr (g + 1 < NG) tmem_wait32(nxt);
    if constexpr (g + 1 == NG) done.finish();
  });
#endif
}

// The same walk when the caller has already issued the load of group 0 into `va` (so that it is in flight across a barrier).
template <int NG, class A, class F, class D>
__device__ __forceinline__ void tmem_walk_pre(uint32_t (&va)[32], A &&addr, F &&f, D &&done) {
  uint32_t vb[32];
  tmem_wait32(va);
  static_for<0, NG>([&](auto gg) {
    constexpr int g = decltype(gg)::value;
    uint32_t(&cur)[32] = (g & 1) ? vb : va;
    uint32_t(&nxt)[32] = (g & 1) ? va : vb;
    if constexpr (g + 1 < NG) {
      tmem_ld32(addr(g + 1), nxt);
      tmem_pin32(cur);
    }
    if constexpr (g + 1 == NG) done.issue();
    f(gg, cur);
    if constexpr (g + 1 < NG) tmem_wait32(nxt);
    if constexpr (g + 1 == NG) done.finish();
  });
}

// Requantise 32 accumulators of one row -> int8 -> panel buffer (and/or global).
//   panel      shared-memory panel buffer of a [kS x *] matrix (nullptr: do not store to smem)
//   col        first of the 32 columns inside that matrix
//   grow       global pointer to the first of these 32 columns (natural layout) or nullptr
//   RELU       max(0, .) on the requantised value (FFN fc1): the sign bytes are replicated into a mask by one PRMT
//   N          accumulators per call (32 or 16);  PR = rows of the panel buffer (kS, or 256 for the multi-head kernel's Q/K/V)
template <int MODE, int ALT, bool RELU = false, int N = 32, int PR = kS>
__device__ __forceinline__ void requant_store32(const uint32_t (&v)[N], const FRq &r, int8_t *panel, int row, int col, int8_t *grow) {
#pragma unroll
  for (int j = 0; j < N; j += 16) {
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      int y[4];
      if constexpr (rqm_offset(MODE)) {
        rq_off_bits2<MODE>(v[j + q * 4], v[j + q * 4 + 1], r, y[0], y[1]);
        rq_off_bits2<MODE>(v[j + q * 4 + 2], v[j + q * 4 + 3], r, y[2], y[3]);
        const uint32_t u = pack_u8x4(y[0], y[1], y[2], y[3]);              // y + 128 per byte
        if constexpr (RELU) {
          uint32_t m;                                                       // 0xFF where y >= 0 (msb of the biased byte)
          asm("prmt.b32 %0, %1, %1, 0xba98;" : "=r"(m) : "r"(u));
          w[q] = (u ^ 0x80808080u) & m;
        } else {
          w[q] = u ^ 0x80808080u;
        }
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) y[i] = rq_apply<MODE, ALT>((int)v[j + q * 4 + i], r, i);
        w[q] = pack_s8x4(y[0], y[1], y[2], y[3]);
        if constexpr (RELU) {
          uint32_t m;                                                       // 0xFF where y < 0
          asm("prmt.b32 %0, %1, %1, 0xba98;" : "=r"(m) : "r"(w[q]));
          w[q] &= ~m;
        }
      }
    }
    const uint4 o = make_uint4(w[0], w[1], w[2], w[3]);
    if (panel) *reinterpret_cast<uint4 *>(panel + ((col + j) >> 4) * (PR * 16) + row * 16) = o;
    if (grow) *reinterpret_cast<uint4 *>(grow + j) = o;
  }
}

struct SlotCtx {
  uint64_t *bars;       // this slot's kBarsPerSlot mbarriers
  uint8_t *smem;        // this slot's operand buffers: Q(O) | K(Asm, next X) | V | [X]
  uint32_t tmem;        // first TMEM column of this slot's accumulator ring
};

// ---------------------------------------------------------------------------------------------------- ITAmax on a row
// A row of the logits accumulator is 128 columns in two TMEM regions; the thread of warpgroup A that owns the row's
// TMEM lane works on keys 0..63 (HALF 0), its twin in warpgroup B (same lane quadrant, same lane) on keys 64..127.
// The two warps of such a pair exchange their partial row max / row sum through shared memory (`xch`: [2 halves][128
// rows] of {max, sum}) and meet at a 64-thread named barrier.  The half row is read from TMEM up to three times
// instead of being kept in registers; `release()` runs when the last read has landed.  Writes the u8 probabilities
// into the K buffer (A operand of A.V) and, for DUMP, to global memory.
struct PairCtx {
  uint32_t bar_id;     // named barrier shared by the two warps of the pair (64 threads)
  int2 *xch;           // this slot's exchange area [2][kS]
};
template <int NPART = 2>
__device__ __forceinline__ void pair_sync(const PairCtx &pc) { named_bar_sync(pc.bar_id, 32 * NPART); }
// The NPART threads that share a row (NPART = 2: the pair above; 4: the multi-head kernel, 64 of 256 keys each) publish a
// partial row max / row sum and combine all of them (own value included) after the barrier.
template <int NPART>
__device__ __forceinline__ int xchg_max(const PairCtx &pc, int part, int row, int v) {
  pc.xch[part * kS + row].x = v;
  pair_sync<NPART>(pc);
#pragma unroll
  for (int p = 0; p < NPART; ++p) v = max(v, pc.xch[p * kS + row].x);
  return v;
}
template <int NPART>
__device__ __forceinline__ uint32_t xchg_sum(const PairCtx &pc, int part, int row, uint32_t v) {
  pc.xch[part * kS + row].y = (int)v;
  pair_sync<NPART>(pc);
  uint32_t t = 0;
#pragma unroll
  for (int p = 0; p < NPART; ++p) t += (uint32_t)pc.xch[p * kS + row].y;
  return t;
}

// repo_global (models/ITA/QAT/ITA_softmax.py:51-61): m = row max, d = m - a, n = 256 >> d, inv = floor(255*2^16 / sum n),
// out = (n * inv) >> 16 == (inv >> 8) >> d.  Requantisation is monotone, so the row max is taken on the raw
// accumulators (pass 1) and requantised once.  Passes 2 and 3 need d only; for INT_HI it comes out of a single
// multiply-add,  d = hi32(acc * -multp + c2),  c2 = (m << 32) - c + 2^32 - 1   (-floor(z / 2^32) == floor((-z + 2^32 - 1) / 2^32)),
// which ignores the s8 saturation of a.  That is exact whenever no logit of the row saturates high (unsaturated max
// <= 127) and low saturation cannot matter (m >= -119: a saturated element has d > 8 either way); other rows (and
// the other requant forms) take the general form with the explicit clamp.  The choice is made per warp; both warps of
// a pair hold the same rows and therefore choose alike.
template <int MODE, int ALT, bool DUMP, int NPART = 2, class R>
__device__ __forceinline__ void itamax_repo_half(int HALF, uint32_t t, const FRq &rq, int8_t *k_buf, int row, int8_t *gA, uint8_t *gAsm,
                                                 const PairCtx &pc, R &&release) {
  auto addr = [&](int g) { return t + (uint32_t)g * 32u; };
  // ---- pass 1: max of the raw accumulators
  int amax = (int)0x80000000;
  tmem_walk<2>(addr,
               [&](auto, const uint32_t(&v)[32]) {
#pragma unroll
                 for (int i = 0; i < 32; ++i) amax = max(amax, (int)v[i]);
               },
               NoRelease{});
  amax = xchg_max<NPART>(pc, HALF, row, amax);
  const int m_un = rq_apply<MODE, 0>(amax, rq, 0);
  const int m = min(max(m_un, -128), 127);
  bool fast = false;
  if constexpr (MODE == RQM_INT_HI && !DUMP) fast = __all_sync(0xffffffffu, m_un <= 127 && m >= -119);
  uint32_t sum = 0;
  if (fast) {
    const long long c2 = ((long long)m << 32) - rq.c + 0xFFFFFFFFLL;
    const int nmultp = -rq.multp;
    const int nmult32 = -rq.mult32;
    const int c32n = m * (int)(1u << rq.shift) - rq.c32 + (int)((1u << rq.shift) - 1u);
    auto dist = [&](int acc, int elem) -> uint32_t {
      if (ALT == 1 || (ALT == 2 && (elem & 1))) return (uint32_t)((acc * nmult32 + c32n) >> rq.shift);
      return (uint32_t)(((long long)acc * (long long)nmultp + c2) >> 32);
    };
    tmem_walk<2>(addr,
                 [&](auto, const uint32_t(&v)[32]) {
#pragma unroll
                   for (int i = 0; i < 32; ++i) sum += shr_clamp(256u, dist((int)v[i], i));
                 },
                 NoRelease{});
    sum = xchg_sum<NPART>(pc, HALF, row, sum);
    const uint32_t inv8 = ((255u * 65536u) / sum) >> 8;
    tmem_walk<2>(addr,
                 [&](auto gg, const uint32_t(&v)[32]) {
                   constexpr int g = decltype(gg)::value;
#pragma unroll
                   for (int j = 0; j < 32; j += 16) {
                     uint32_t w[4];
#pragma unroll
                     for (int q = 0; q < 4; ++q) {
                       int o[4];
#pragma unroll
                       for (int i = 0; i < 4; ++i) o[i] = (int)shr_clamp(inv8, dist((int)v[j + q * 4 + i], i));
                       w[q] = pack_u8x4(o[0], o[1], o[2], o[3]);
                     }
                     *reinterpret_cast<uint4 *>(k_buf + ((HALF * 64 + g * 32 + j) >> 4) * (kS * 16) + row * 16) =
                         make_uint4(w[0], w[1], w[2], w[3]);
                   }
                 },
                 release);
  } else {
    auto dist = [&](int acc) -> uint32_t {
      const int a = min(max(rq_apply<MODE, 0>(acc, rq, 0), -128), 127);
      return (uint32_t)(m - a);
    };
    tmem_walk<2>(addr,
                 [&](auto, const uint32_t(&v)[32]) {
#pragma unroll
                   for (int i = 0; i < 32; ++i) sum += shr_clamp(256u, dist((int)v[i]));
                 },
                 NoRelease{});
    sum = xchg_sum<NPART>(pc, HALF, row, sum);
    const uint32_t inv8 = ((255u * 65536u) / sum) >> 8;
    tmem_walk<2>(addr,
                 [&](auto gg, const uint32_t(&v)[32]) {
                   constexpr int g = decltype(gg)::value;
#pragma unroll
                   for (int j = 0; j < 32; j += 16) {
                     uint32_t w[4], wa[4];
#pragma unroll
                     for (int q = 0; q < 4; ++q) {
                       int o[4], a[4];
#pragma unroll
                       for (int i = 0; i < 4; ++i) {
                         const uint32_t d = dist((int)v[j + q * 4 + i]);
                         o[i] = (int)shr_clamp(inv8, d);
                         a[i] = m - (int)d;
                       }
                       w[q] = pack_u8x4(o[0], o[1], o[2], o[3]);
                       if constexpr (DUMP) wa[q] = pack_s8x4(a[0], a[1], a[2], a[3]);
                     }
                     const uint4 ov = make_uint4(w[0], w[1], w[2], w[3]);
                     *reinterpret_cast<uint4 *>(k_buf + ((HALF * 64 + g * 32 + j) >> 4) * (kS * 16) + row * 16) = ov;
                     if constexpr (DUMP) {
                       if (gAsm) *reinterpret_cast<uint4 *>(gAsm + HALF * 64 + g * 32 + j) = ov;
                       if (gA) *reinterpret_cast<uint4 *>(gA + HALF * 64 + g * 32 + j) = make_uint4(wa[0], wa[1], wa[2], wa[3]);
                     }
                   }
                 },
                 release);
  }
  // (the exchange area needs no closing barrier: a value is overwritten only after a later pair_sync that the partner
  //  reaches after its read of the old value)
}

// The same ITAmax with the requantisation on the FMA pipe (RQM_FP).  TMEM holds bits = INIT + acc (monotone in acc), so
// pass 1 takes the max of the raw patterns; z = rq_fp_bits(.) is a + 128 before saturation.  Fast rows (no logit
// saturates high, m >= -119) get the distance d = m - a from one FFMA.RP per element; other rows clamp explicitly.
template <int MODE, bool DUMP, int NPART = 2, class R>
__device__ __forceinline__ void itamax_repo_half_fp(bool allow_fast, int HALF, uint32_t t, const FRq &rq, int8_t *k_buf, int row, int8_t *gA, uint8_t *gAsm,
                                                    const PairCtx &pc, R &&release) {
  auto addr = [&](int g) { return t + (uint32_t)g * 32u; };
  int amax = 0;                                   // patterns are positive
  tmem_walk<2>(addr,
               [&](auto, const uint32_t(&v)[32]) {
#pragma unroll
                 for (int i = 0; i < 32; ++i) amax = max(amax, (int)v[i]);
               },
               NoRelease{});
#ifndef ITA_NO_PRELOAD
  uint32_t pre[32];                               // group 0 of the next pass, in flight while the row partners exchange
  tmem_ld32(addr(0), pre);
#endif
  amax = xchg_max<NPART>(pc, HALF, row, amax);
  const int zmax = rq_off_bits<MODE>((uint32_t)amax, rq);
  const int mp = min(max(zmax, 0), 255);          // m + 128
  bool fast = false;
#ifndef ITA_FP_NOFAST
  if constexpr (!DUMP) fast = allow_fast && __all_sync(0xffffffffu, zmax >= 9);
#endif
  // rows whose maximum saturates high (common with real scales: the observers put the top of the logit range at 127) keep the
  // one-FFMA distance with the CLAMPED maximum: an element that saturates as well then gets d + 1 <= 0, which one signed
  // max(., 1) repairs (any pattern with the sign bit set, -0.0 included, is a negative int).  Warp-uniform choice.
  const bool sat = !__all_sync(0xffffffffu, zmax <= 255);
  uint32_t sum = 0;
  auto store16 = [&](int g, int j, const uint4 &ov) {
    *reinterpret_cast<uint4 *>(k_buf + ((HALF * 64 + g * 32 + j) >> 4) * (kS * 16) + row * 16) = ov;
  };
  if (fast) {
    // d + 1 = ceil((m + 129 - C) - bits*b) is >= 1 for every element of the row (an exact value in (-1, 0) would round to
    // -0.0, whose pattern is not 0), so the shifts below work on 2*256 and 2*inv8
    // (RQM_TORCH_FP: d + 1 = rne(-t) + (m + 129 - zp - 128), the last FFMA of the requant with -t and that constant, which
    //  may be negative: sign-magnitude pattern -- and must be EVEN so that round-to-nearest-even keeps the parity of ties; an
    //  odd constant is raised by one, the row then works with d + 2, still >= 1: never -0.0)
    const int k2raw = mp + 1 + rq.negc;
    const uint32_t odd = rqm_torch_fp(MODE) ? (uint32_t)(k2raw & 1) : 0u;
    const int k2 = k2raw + (int)odd;
    const float c2 = __int_as_float(k2 >= 0 ? k2 : (int)(0x80000000u | (uint32_t)(-k2))), nfb = -rq.fb;
    const uint32_t one = 512u << odd;
    const int dmin = 1 + (int)odd;
    auto passes = [&](auto clamp_tag) {
    constexpr bool CLAMP = decltype(clamp_tag)::value;
    // two distances per FFMA2
    auto dist2 = [&](uint32_t v0, uint32_t v1, uint32_t &d0, uint32_t &d1) {
#ifdef ITA_NO_F32X2
      auto dist = [&](uint32_t v) -> uint32_t {
        if constexpr (MODE == RQM_TORCH_FP) {
          const float u = __fmaf_rn(__uint_as_float(v), __uint_as_float(0x7E800000u), rq.fb);
          const float t = __fmul_rn(u, rq.eff);
          return (uint32_t)__float_as_int(__fmaf_rn(-t, __uint_as_float(1u), c2));
        } else if constexpr (MODE == RQM_TORCH_FP2) {
          const float t = __fmaf_rn(__uint_as_float(v), rq.fb, rq.eff);
          return (uint32_t)__float_as_int(__fmaf_rn(-t, __uint_as_float(kTwoM85), c2));
        } else {
          return (uint32_t)__float_as_int(__fmaf_ru(__uint_as_float(v), nfb, c2));
        }
      };
      d0 = dist(v0);
      d1 = dist(v1);
#else
      const u64 a = pk2(v0, v1);
      u64 d;
      if constexpr (MODE == RQM_TORCH_FP) {
        const u64 u = ffma2_rn(a, pk2(0x7E800000u, 0x7E800000u), pk2f(rq.fb));
        d = ffma2_rn(fmul2_rn(u, pk2f(rq.eff)), pk2(0x80000001u, 0x80000001u) /* -2^-149 */, pk2f(c2));
      } else if constexpr (MODE == RQM_TORCH_FP2) {
        d = ffma2_rn(ffma2_rn(a, pk2f(rq.fb), pk2f(rq.eff)), pk2(0x80000000u | kTwoM85, 0x80000000u | kTwoM85) /* -2^-85 */, pk2f(c2));
      } else {
        d = ffma2_rp(a, pk2f(nfb), pk2f(c2));
      }
      int i0, i1;
      upk2(d, i0, i1);
      d0 = (uint32_t)i0;
      d1 = (uint32_t)i1;
#endif
      if constexpr (CLAMP) {
        d0 = (uint32_t)max((int)d0, dmin);
        d1 = (uint32_t)max((int)d1, dmin);
      }
    };
#ifndef ITA_NO_PRELOAD
    tmem_walk_pre<2>(pre, addr,
#else
    tmem_walk<2>(addr,
#endif
                 [&](auto, const uint32_t(&v)[32]) {
#pragma unroll
                   for (int i = 0; i < 32; i += 2) {
                     uint32_t d0, d1;
                     dist2(v[i], v[i + 1], d0, d1);
                     sum += shr_clamp(one, d0) + shr_clamp(one, d1);
                   }
                 },
                 NoRelease{});
#ifndef ITA_NO_PRELOAD
    tmem_ld32(addr(0), pre);
#endif
    sum = xchg_sum<NPART>(pc, HALF, row, sum);
    const uint32_t inv8 = ((((255u * 65536u) / sum) >> 8) << 1) << odd;
#ifndef ITA_NO_PRELOAD
    tmem_walk_pre<2>(pre, addr,
#else
    tmem_walk<2>(addr,
#endif
                 [&](auto gg, const uint32_t(&v)[32]) {
                   constexpr int g = decltype(gg)::value;
#pragma unroll
                   for (int j = 0; j < 32; j += 16) {
                     uint32_t w[4];
#pragma unroll
                     for (int q = 0; q < 4; ++q) {
                       uint32_t d[4];
                       dist2(v[j + q * 4], v[j + q * 4 + 1], d[0], d[1]);
                       dist2(v[j + q * 4 + 2], v[j + q * 4 + 3], d[2], d[3]);
                       w[q] = pack_u8x4((int)shr_clamp(inv8, d[0]), (int)shr_clamp(inv8, d[1]), (int)shr_clamp(inv8, d[2]),
                                        (int)shr_clamp(inv8, d[3]));
                     }
                     store16(g, j, make_uint4(w[0], w[1], w[2], w[3]));
                   }
                 },
                 release);
    };   // passes
    if (sat) passes(std::true_type{});
    else passes(std::false_type{});
  } else {
    auto dist = [&](uint32_t v) -> uint32_t { return (uint32_t)(mp - min(max(rq_off_bits<MODE>(v, rq), 0), 255)); };
#ifndef ITA_NO_PRELOAD
    tmem_walk_pre<2>(pre, addr,
#else
    tmem_walk<2>(addr,
#endif
                 [&](auto, const uint32_t(&v)[32]) {
#pragma unroll
                   for (int i = 0; i < 32; ++i) sum += shr_clamp(256u, dist(v[i]));
                 },
                 NoRelease{});
#ifndef ITA_NO_PRELOAD
    tmem_ld32(addr(0), pre);
#endif
    sum = xchg_sum<NPART>(pc, HALF, row, sum);
    const uint32_t inv8 = ((255u * 65536u) / sum) >> 8;
#ifndef ITA_NO_PRELOAD
    tmem_walk_pre<2>(pre, addr,
#else
    tmem_walk<2>(addr,
#endif
                 [&](auto gg, const uint32_t(&v)[32]) {
                   constexpr int g = decltype(gg)::value;
#pragma unroll
                   for (int j = 0; j < 32; j += 16) {
                     uint32_t w[4], wa[4];
#pragma unroll
                     for (int q = 0; q < 4; ++q) {
                       int o[4], a[4];
#pragma unroll
                       for (int i = 0; i < 4; ++i) {
                         const uint32_t d = dist(v[j + q * 4 + i]);
                         o[i] = (int)shr_clamp(inv8, d);
                         a[i] = mp - 128 - (int)d;
                       }
                       w[q] = pack_u8x4(o[0], o[1], o[2], o[3]);
                       if constexpr (DUMP) wa[q] = pack_s8x4(a[0], a[1], a[2], a[3]);
                     }
                     const uint4 ov = make_uint4(w[0], w[1], w[2], w[3]);
                     store16(g, j, ov);
                     if constexpr (DUMP) {
                       if (gAsm) *reinterpret_cast<uint4 *>(gAsm + HALF * 64 + g * 32 + j) = ov;
                       if (gA) *reinterpret_cast<uint4 *>(gA + HALF * 64 + g * 32 + j) = make_uint4(wa[0], wa[1], wa[2], wa[3]);
                     }
                   }
                 },
                 release);
  }
}

// Recalled golden-model streaming softmax (the stream16 variant): 16-column groups in
// column order, running max m and denominator, eps = 1/32 with round-half-up; then out = inv >> rhu((m - a) / 32).
// The row is walked in column order: part p continues from the state part p - 1 publishes (NPART - 1 hand-overs, each a
// named barrier of the NPART warps that share the rows), and everybody reads the final state of the last part.
template <int MODE, bool DUMP, int NPART = 2, class R>
__device__ __forceinline__ void itamax_stream_half(int HALF, uint32_t t, const FRq &rq, int8_t *k_buf, int row, int8_t *gA, uint8_t *gAsm,
                                                   const PairCtx &pc, R &&release) {
  auto addr = [&](int g) { return t + (uint32_t)g * 32u; };
  auto biased = [&](uint32_t acc) -> uint32_t {   // sat8(requant) + 128 in [0, 255]
    if constexpr (rqm_offset(MODE)) {
      return (uint32_t)min(max(rq_off_bits<MODE>(acc, rq), 0), 255);
    } else {
      return (uint32_t)(min(max(rq_apply<MODE, 0>((int)acc, rq, 0), -128), 127) + 128);
    }
  };
  uint32_t m = 0, tot = 0;
#pragma unroll
  for (int turn = 0; turn < NPART; ++turn) {
    if (HALF == turn) {
      if (turn > 0) {                // the part before has published its running state
        m = (uint32_t)pc.xch[(turn - 1) * kS + row].x;
        tot = (uint32_t)pc.xch[(turn - 1) * kS + row].y;
      }
      tmem_walk<2>(addr,
                   [&](auto, const uint32_t(&v)[32]) {
#pragma unroll
                     for (int h = 0; h < 2; ++h) {
                       uint32_t b[16];
                       uint32_t cm = 0;
#pragma unroll
                       for (int i = 0; i < 16; ++i) {
                         b[i] = biased(v[h * 16 + i]);
                         cm = max(cm, b[i]);
                       }
                       if (cm > m) {
                         tot = shr_clamp(tot, (cm - m + 16) >> 5);
                         m = cm;
                       }
#pragma unroll
                       for (int i = 0; i < 16; ++i) tot += 256u >> ((m - b[i] + 16) >> 5);
                     }
                   },
                   NoRelease{});
      pc.xch[turn * kS + row] = make_int2((int)m, (int)tot);
    }
    pair_sync<NPART>(pc);
  }
  m = (uint32_t)pc.xch[(NPART - 1) * kS + row].x;
  tot = (uint32_t)pc.xch[(NPART - 1) * kS + row].y;
  const uint32_t inv = (255u * 256u) / (tot < 1 ? 1 : tot);
  tmem_walk<2>(addr,
               [&](auto gg, const uint32_t(&v)[32]) {
                 constexpr int g = decltype(gg)::value;
#pragma unroll
                 for (int j = 0; j < 32; j += 16) {
                   uint32_t w[4], wa[4];
#pragma unroll
                   for (int q = 0; q < 4; ++q) {
                     int o[4], a[4];
#pragma unroll
                     for (int i = 0; i < 4; ++i) {
                       const uint32_t b = biased(v[j + q * 4 + i]);
                       o[i] = (int)(inv >> ((m - b + 16) >> 5));
                       a[i] = (int)b - 128;
                     }
                     w[q] = pack_u8x4(o[0], o[1], o[2], o[3]);
                     if constexpr (DUMP) wa[q] = pack_s8x4(a[0], a[1], a[2], a[3]);
                   }
                   const uint4 ov = make_uint4(w[0], w[1], w[2], w[3]);
                   *reinterpret_cast<uint4 *>(k_buf + ((HALF * 64 + g * 32 + j) >> 4) * (kS * 16) + row * 16) = ov;
                   if constexpr (DUMP) {
                     if (gAsm) *reinterpret_cast<uint4 *>(gAsm + HALF * 64 + g * 32 + j) = ov;
                     if (gA) *reinterpret_cast<uint4 *>(gA + HALF * 64 + g * 32 + j) = make_uint4(wa[0], wa[1], wa[2], wa[3]);
                   }
                 }
               },
               release);
}

// ---------------------------------------------------------------------------------------------------- epilogue warps
// One warpgroup (WG = 0: "A", 1: "B") of one slot: drains its sub-block of every chunk, in chunk order.  The chunk
// schedule is unrolled at compile time: a lone warp issues one dependent instruction every ~5 cycles, so run-time chunk
// bookkeeping (tried: a run-time loop over three code paths) costs more than the instruction-cache misses it saves.
template <int E, int P, int MODE, int SOFTMAX, int ALT, bool DUMP>
__device__ __forceinline__ void epilogue_loop(const FusedParams &prm, const SlotCtx &sc, const PairCtx &pc, int WG, int quad, int lane, long long iters,
                                              long long frame0, long long frame_stride) {
  using C = Cfg<E, P>;
  const int row = quad * 32 + lane;
  const uint32_t t_lane = sc.tmem + ((uint32_t)(quad * 32) << 16);
  int8_t *const q_buf = reinterpret_cast<int8_t *>(sc.smem) + C::OFF_Q, *const k_buf = reinterpret_cast<int8_t *>(sc.smem) + C::OFF_K,
               *const v_buf = reinterpret_cast<int8_t *>(sc.smem) + C::OFF_V;
  bool ok = true;
  uint32_t n = 0;                                   // running chunk number: region n & 1, its (n >> 1)-th use
  for (long long it = 0; it < iters && ok; ++it) {
    const long long f = frame0 + it * frame_stride;
    static_for<0, C::NCH>([&](auto cc) {
      constexpr int c = decltype(cc)::value;
      // both warpgroups run this one copy of the code (WG is a run-time value: a second, per-warpgroup copy of the unrolled
      // schedule doubles the instruction footprint, and the kernel is instruction-fetch heavy -- see the ncu summary)
      constexpr int kind = C::kind_of(c), SW = C::sub_w(c);
      const int stage = WG ? C::stage_of(c, 1) : C::stage_of(c, 0), col0 = WG ? C::stage_col0(c, 1) : C::stage_col0(c, 0);
      if (!ok) return;
      const uint32_t r = n & 1u, use = n >> 1;
      ++n;
      ITA_TR(prm, threadIdx.x >> 5, it, c, 0);
      ok = mbar_wait(sc.bars + BAR_FULL + r, use & 1, prm.abort_flag);
      if (!ok) return;
      ITA_TR(prm, threadIdx.x >> 5, it, c, 1);
      tc_fence_after();
      const uint32_t taddr = t_lane + r * kRegionCols + WG * SW;
      // hand the region back as soon as this warp's last load has landed
      auto release = make_release(
          [&] {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(sc.bars + BAR_EMPTY + r);
            ITA_TR(prm, threadIdx.x >> 5, it, c, 2);
          },
          [] {});
      if constexpr (kind == CK_S) {
        // ---- logits: this warpgroup's 64 key columns -> ITAmax -> Asm (u8) into the K buffer
        int8_t *gA = nullptr;
        uint8_t *gAsm = nullptr;
        if constexpr (DUMP) {
          if (prm.dA) gA = prm.dA + (f * kS + row) * kS;
          if (prm.dAsm) gAsm = prm.dAsm + (f * kS + row) * kS;
        }
        if constexpr (SOFTMAX == ITA_SOFTMAX_REPO_GLOBAL && rqm_offset(MODE))
          itamax_repo_half_fp<MODE, DUMP>(!rqm_torch_fp(MODE) || prm.torch_fast != 0, WG, taddr, prm.rq[3], k_buf, row, gA, gAsm, pc, release);
        else if constexpr (SOFTMAX == ITA_SOFTMAX_REPO_GLOBAL)
          itamax_repo_half<MODE, ALT, DUMP>(WG, taddr, prm.rq[3], k_buf, row, gA, gAsm, pc, release);
        else
          itamax_stream_half<MODE, DUMP>(WG, taddr, prm.rq[3], k_buf, row, gA, gAsm, pc, release);
      } else {
        int8_t *panel = stage == STG_Q ? q_buf : stage == STG_K ? k_buf : stage == STG_V ? v_buf : stage == STG_AV ? q_buf : nullptr;
        int8_t *g = nullptr;
        if constexpr (kind == CK_O) {
          g = prm.y + (f * kS + row) * E + col0;
        } else if constexpr (DUMP) {
          int8_t *base = stage == STG_Q ? prm.dQ : stage == STG_K ? prm.dK : stage == STG_V ? prm.dV : prm.dO;
          if (base) g = base + (f * kS + row) * P + col0;
        }
        tmem_walk<SW / 32>([&](int gi) { return taddr + (uint32_t)gi * 32u; },
                           [&](auto gg, const uint32_t(&v)[32]) {
                             constexpr int gi = decltype(gg)::value;
                             requant_store32<MODE, ALT>(v, prm.rq[stage], panel, row, col0 + gi * 32, g ? g + gi * 32 : nullptr);
                           },
                           release);
      }
      if constexpr (kind != CK_O) {
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(sc.bars + BAR_CHUNK + c);
      }
      ITA_TR(prm, threadIdx.x >> 5, it, c, 3);
    });
  }
}

// ---------------------------------------------------------------------------------------------------- the kernel
template <int E, int P, int MODE, int SOFTMAX, int ALT, bool DUMP>
__global__ void __launch_bounds__((Cfg<E, P>::THREADS), 1) k_ita_fused(const __grid_constant__ CUtensorMap xmap, const FusedParams prm) {
  using C = Cfg<E, P>;
  constexpr int kSlots = C::SLOTS, kWarpMma = C::WARP_MMA, kWarpLoad = C::WARP_LOAD;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + C::OFF_BAR);
  uint64_t *bar_w = bars;                          // weights + constant tile landed
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
        mbar_init(b + BAR_EMPTY + r, kEpiWarpsPerSlot);
      }
      for (int c = 0; c + 1 < C::NCH; ++c) mbar_init(b + BAR_CHUNK + c, kEpiWarpsPerSlot);
      mbar_init(b + BAR_TURN, 1);
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
          load_x_tile<E>(smem + C::OFF_SLOT + s * C::SLOT_BYTES + C::OFF_X, &xmap, (long long)blockIdx.x * kSlots + s,
                         slot_bars + s * kBarsPerSlot + BAR_X);
        // the tile after that goes to L2 now: its shared-memory buffer is freed late in the frame (C_XFREE), and the load
        // that follows should not have to go to HBM
        if (iters[s] > 1) l2_prefetch(prm.x + ((long long)blockIdx.x * kSlots + s + frame_stride) * (kS * E), kS * E);
      }
      bool ok = true;
      for (long long it = 0; it + 1 < mx && ok; ++it)
        for (int s = 0; s < kSlots && ok; ++s)
          if (it + 1 < iters[s]) {
            // every MMA of frame `it` that reads this slot's X tile (or the K columns it shares storage with) is complete
            ok = mbar_wait(slot_bars + s * kBarsPerSlot + BAR_XFREE, (uint32_t)(it & 1), prm.abort_flag);
            if (ok)
              load_x_tile<E>(smem + C::OFF_SLOT + s * C::SLOT_BYTES + C::OFF_X, &xmap,
                             (long long)blockIdx.x * kSlots + s + (it + 1) * frame_stride, slot_bars + s * kBarsPerSlot + BAR_X);
            if (ok && it + 2 < iters[s]) l2_prefetch(prm.x + ((long long)blockIdx.x * kSlots + s + (it + 2) * frame_stride) * (kS * E), kS * E);
          }
    }
    __syncwarp();
  } else if (warp >= kWarpMma) {
    // ================================================================= MMA issue for one slot
    // The whole warp walks the chunk schedule (barrier waits and operand addresses stay warp-uniform, so they live
    // in the uniform datapath); one elected lane issues the tcgen05.mma / commit instructions.
    const int slot = __shfl_sync(0xffffffffu, warp - kWarpMma, 0);
    const long long iters = iters_of(slot);
    if (iters > 0) {
      const bool leader = elect_one();
      uint64_t *b = slot_bars + slot * kBarsPerSlot;
      bool ok = mbar_wait(bar_w, 0, prm.abort_flag);
      // operand buffer addresses in 16-byte units (the descriptor address field)
      const uint32_t w16 = smem_u32(smem + C::OFF_W) >> 4;
      const uint32_t ca16 = smem_u32(smem + C::OFF_CA) >> 4, dqk16 = smem_u32(smem + C::OFF_DQK) >> 4, dav16 = smem_u32(smem + C::OFF_DAV) >> 4;
      const uint32_t slot16 = smem_u32(smem + C::OFF_SLOT + slot * C::SLOT_BYTES) >> 4;
      const uint32_t x16 = slot16 + (C::OFF_X >> 4), q16 = slot16 + (C::OFF_Q >> 4), k16 = slot16 + (C::OFF_K >> 4),
                     v16 = slot16 + (C::OFF_V >> 4);
      const uint32_t d_base = tmem_base + slot * kTmemColsPerSlot;
      uint32_t n = 0;                                 // running chunk number: region n & 1, its (n >> 1)-th use
      int written = 0;                                // positions of the current frame whose 'chunk written' barrier has been passed
      for (long long it = 0; it < iters && ok; ++it) {
        const uint32_t ph = (uint32_t)(it & 1);
        static_for<0, C::NCH>([&](auto cc) {
          constexpr int c = decltype(cc)::value;
          constexpr int kind = C::kind_of(c), col0 = C::col0_of(c), NC = C::ncols_of(c);
          if (!ok) return;
          const uint32_t r = n & 1u, use = n >> 1;
          ++n;
          ITA_TR(prm, warp, it, c, 0);
          // Keep the two slots out of phase: a slot starts a frame only after the other slot has issued chunk C_TURN of its
          // previous turn (slot 1: of the same iteration, slot 0: of the iteration before).  Left alone the slots fall into
          // lockstep (measured), where both sit in the same MMA-latency bubbles and the same epilogue bursts; out of phase
          // one slot's arithmetic fills the other's waits (+11 % measured).
          if constexpr (c == 0 && kSlots == 2) {
            uint64_t *other = slot_bars + (slot ^ 1) * kBarsPerSlot + BAR_TURN;
            if (slot == 1) ok = mbar_wait(other, ph, prm.abort_flag);
            else if (it > 0 && it - 1 < iters_of(1)) ok = mbar_wait(other, ph ^ 1u, prm.abort_flag);
            if (!ok) return;
          }
          if (use > 0) ok = mbar_wait(b + BAR_EMPTY + r, (use - 1) & 1, prm.abort_flag);
          ITA_TR(prm, warp, it, c, 1);
          // The digit K steps of the logits, the context and the output projection read only constants: they go first
          // (overwriting the region) and run while this thread still waits for the chunk's operands.
          constexpr int DSG = rqm_offset(MODE) ? 0 : 1;
          if (ok && leader) {
            const uint32_t d_early = d_base + r * kRegionCols;
            if constexpr (kind == CK_S && rqm_offset(MODE)) {
              constexpr uint32_t IDD = make_idesc_i8(kS, kS, 0, 0, 0, 0);
              tc_fence_after();
              umma_i8(d_early, make_smem_desc_rel(ca16, 0, 128, 0), make_smem_desc_rel(dqk16, 0, 128, 0), IDD, 0);
              umma_i8(d_early, make_smem_desc_rel(ca16, 256, 128, 0), make_smem_desc_rel(dqk16, 256, 128, 0), IDD, 1);
              if constexpr (rqm_qk_steps(MODE) == 3)
                umma_i8(d_early, make_smem_desc_rel(ca16, 512, 128, 0), make_smem_desc_rel(dqk16, 512, 128, 0), IDD, 1);
            } else if constexpr (kind == CK_AV && rqm_offset(MODE)) {
              constexpr uint32_t IDD = make_idesc_i8(kS, NC, 0, 0, 0, 0);
              tc_fence_after();
              umma_i8(d_early, make_smem_desc_rel(ca16, 0, 128, 0), make_smem_desc_rel(dav16, 0, 128, 0), IDD, 0);
            } else if constexpr (kind == CK_O) {
              constexpr uint32_t IDD = make_idesc_i8(kS, E, DSG, DSG, 0, 0);
              tc_fence_after();
              umma_i8(d_early, make_smem_desc_rel(ca16, 0, 128, 0), make_smem_desc_rel(w16, C::OFF_WO + (P / 16) * E * 16, E * 16, 128), IDD, 0);
              if constexpr (rqm_offset(MODE))
                umma_i8(d_early, make_smem_desc_rel(ca16, 256, 128, 0), make_smem_desc_rel(w16, C::OFF_WO + (P / 16 + 2) * E * 16, E * 16, 128), IDD, 1);
            }
          }
          __syncwarp();
          if constexpr (c == 0) {
            ok = ok && mbar_wait(b + BAR_X, ph, prm.abort_flag);
            written = 0;
          }
          if (!ok) return;
          tc_fence_after();
          ITA_TR(prm, warp, it, c, 2);
          const uint32_t d_tmem = d_base + r * kRegionCols;
          // wait until the epilogue of position `pos` (and so of every position before it) has written its operand buffer
          auto need = [&](int pos) {
            while (ok && written <= pos) {
              ok = mbar_wait(b + BAR_CHUNK + written, ph, prm.abort_flag);
              ++written;
              tc_fence_after();
            }
            return ok;
          };
          {
            if constexpr (kind == CK_PROJ) {
              // projections: A = X [S rows, K = E] K-major; B = rows [col0, +NC) of [Wq | Wk | Wv]^T [3P rows, K = E + 32]
              // K-major; the last K step multiplies the constant tile with the bias (+ offset) digits
              constexpr uint32_t ID = make_idesc_i8(kS, NC, 1, 1, 0, 0), IDD = make_idesc_i8(kS, NC, DSG, DSG, 0, 0);
              constexpr uint32_t wt = col0 * 16;
              if (leader) {
#pragma unroll
                for (int k = 0; k < E / 32; ++k)
                  umma_i8(d_tmem, make_smem_desc_rel(x16, k * 2 * kS * 16, kS * 16, 128),
                          make_smem_desc_rel(w16, wt + k * 2 * C::PW * 16, C::PW * 16, 128), ID, k > 0);
                umma_i8(d_tmem, make_smem_desc_rel(ca16, 0, 128, 0), make_smem_desc_rel(w16, wt + (E / 16) * C::PW * 16, C::PW * 16, 128),
                        IDD, 1);
              }
            } else if constexpr (kind == CK_S) {
              // logits: A = Q [S, K = P] K-major; B = K [S rows (keys), K = P] K-major (RQM_FP: on top of the offset steps).
              // A K step is issued as soon as the projection chunk with its K columns is written (Q comes before K).
              constexpr uint32_t ID = make_idesc_i8(kS, kS, 1, 1, 0, 0);
              static_for<0, P / 32>([&](auto kk) {
                constexpr int k = decltype(kk)::value;
                if (!need(C::dep_S(k))) return;
                if (leader)
                  umma_i8(d_tmem, make_smem_desc_rel(q16, k * 2 * kS * 16, kS * 16, 128),
                          make_smem_desc_rel(k16, k * 2 * kS * 16, kS * 16, 128), ID, (k > 0 || rqm_offset(MODE)) ? 1u : 0u);
              });
            } else if constexpr (kind == CK_AV) {
              // context: A = Asm [S, K = S] u8 K-major (in the K buffer); B = columns [col0, +NC) of V [K = S rows, N = P]
              // MN-major (RQM_FP: on top of the offset step)
              constexpr uint32_t ID = make_idesc_i8(kS, NC, 0, 1, 0, 1);
              if (need(C::dep_AV(c - C::C_AV0)) && leader) {
#pragma unroll
                for (int k = 0; k < kS / 32; ++k)
                  umma_i8(d_tmem, make_smem_desc_rel(k16, k * 2 * kS * 16, kS * 16, 128),
                          make_smem_desc_rel(v16, (col0 / 16) * kS * 16 + k * 32 * 16, 128, kS * 16), ID, (k > 0 || rqm_offset(MODE)) ? 1u : 0u);
              }
            } else {
              // output projection: A = O [S, K = P] K-major (in the Q buffer); B = Wo^T [E rows, K = P + 64], on top of the
              // digit steps issued above (one for the integer forms, two for RQM_FP); a K step goes out as soon as the A.V
              // chunk with its context columns is written
              constexpr uint32_t ID = make_idesc_i8(kS, E, 1, 1, 0, 0);
              constexpr uint32_t wo = C::OFF_WO;
              static_for<0, P / 32>([&](auto kk) {
                constexpr int k = decltype(kk)::value;
                if (!need(C::dep_O(k))) return;
                if (leader)
                  umma_i8(d_tmem, make_smem_desc_rel(q16, k * 2 * kS * 16, kS * 16, 128), make_smem_desc_rel(w16, wo + k * 2 * E * 16, E * 16, 128),
                          ID, 1);
              });
            }
          }
          if (!ok) return;
          if (leader) {
            umma_commit(b + BAR_FULL + r);
            if constexpr (c == C::C_TURN) mbar_arrive(b + BAR_TURN);
            // the X buffer may be overwritten once its last reader is complete: the last projection, or -- when X shares
            // storage with the K columns beyond the softmax output -- the logits MMA
            if constexpr (c == C::C_XFREE) umma_commit(b + BAR_XFREE);
          }
          __syncwarp();
          ITA_TR(prm, warp, it, c, 3);
        });
      }
    }
  } else {
    // ================================================================= epilogue warps
    const int slot = warp / kEpiWarpsPerSlot;
    const int quad = warp & 3;
    const int wg = (warp >> 2) & 1;
    SlotCtx sc;
    sc.bars = slot_bars + slot * kBarsPerSlot;
    sc.smem = smem + C::OFF_SLOT + slot * C::SLOT_BYTES;
    sc.tmem = tmem_base + slot * kTmemColsPerSlot;
    PairCtx pc;
    pc.bar_id = 1 + slot * 4 + quad;
    pc.xch = reinterpret_cast<int2 *>(smem + C::OFF_XCHG) + slot * 2 * kS;
    const long long iters = iters_of(slot);
    if (iters > 0) {
      const long long frame0 = (long long)blockIdx.x * kSlots + slot;
      epilogue_loop<E, P, MODE, SOFTMAX, ALT, DUMP>(prm, sc, pc, wg, quad, lane, iters, frame0, frame_stride);
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

// ---------------------------------------------------------------------------------------------------- host-side planning (fused.cu)
long long digit_capacity(int ncols);
bool plan_fp_stage(const Rq &r, long long bnd, long long bmax, long long cap, FRq *out);
bool plan_torch_stage(const Rq &r, long long bnd, long long bmax, long long cap, int form, FRq *out);
bool bias_digits(int32_t b, int8_t *g);                      // kBiasK signed radix-127 digits
bool offset_digits(long long v, int n, int8_t *g);           // n unsigned radix-255 digits
void const_tile(int8_t *dst, int ncols, const int8_t *row);  // 8-row constant tile in the panel layout
int make_x_tensor_map(CUtensorMap *map, const void *x, long long rows, int row_bytes);

// ---------------------------------------------------------------------------------------------------- instantiation
typedef void (*KernelFn)(const CUtensorMap, const FusedParams);

// One translation unit per (E, P) shape defines these (ITA_FUSED_DEFINE_SHAPE) so that the shapes compile in parallel.
struct FusedShape {
  int E, P, smem_bytes, slots, threads;
  // mode: RQM_*; softmax: ITA_SOFTMAX_*; alt: INT_HI element form (0 / 1 / 2); dump: intermediates to global
  KernelFn (*pick)(int mode, int softmax, int alt, bool dump);
};

#ifdef ITA_FUSED_ALL_ALT
#define ITA_FUSED_ALT1(E, P, SM) if (mode == RQM_INT_HI && softmax == SM && alt == 1 && !dump) return k_ita_fused<E, P, RQM_INT_HI, SM, 1, false>;
#else
#define ITA_FUSED_ALT1(E, P, SM)
#endif

#define ITA_FUSED_PICK_SM(E, P, SM)                                                                                   \
  if (mode == RQM_INT_HI && softmax == SM && alt == 2 && !dump) return k_ita_fused<E, P, RQM_INT_HI, SM, 2, false>;   \
  ITA_FUSED_ALT1(E, P, SM)                                                                                            \
  if (mode == RQM_INT_HI && softmax == SM) return dump ? k_ita_fused<E, P, RQM_INT_HI, SM, 0, true> : k_ita_fused<E, P, RQM_INT_HI, SM, 0, false>;     \
  if (mode == RQM_INT_GEN && softmax == SM) return dump ? k_ita_fused<E, P, RQM_INT_GEN, SM, 0, true> : k_ita_fused<E, P, RQM_INT_GEN, SM, 0, false>; \
  if (mode == RQM_TORCH_FP && softmax == SM) return dump ? k_ita_fused<E, P, RQM_TORCH_FP, SM, 0, true> : k_ita_fused<E, P, RQM_TORCH_FP, SM, 0, false>; \
  if (mode == RQM_TORCH_FP2 && softmax == SM) return dump ? k_ita_fused<E, P, RQM_TORCH_FP2, SM, 0, true> : k_ita_fused<E, P, RQM_TORCH_FP2, SM, 0, false>; \
  if (mode == RQM_FP && softmax == SM) return dump ? k_ita_fused<E, P, RQM_FP, SM, 0, true> : k_ita_fused<E, P, RQM_FP, SM, 0, false>;           \
  if (mode == RQM_TORCH && softmax == SM) return dump ? k_ita_fused<E, P, RQM_TORCH, SM, 0, true> : k_ita_fused<E, P, RQM_TORCH, SM, 0, false>;

#define ITA_FUSED_DEFINE_SHAPE(E, P)                                              \
  static KernelFn pick_##E##_##P(int mode, int softmax, int alt, bool dump) {     \
    ITA_FUSED_PICK_SM(E, P, 0)                                                    \
    ITA_FUSED_PICK_SM(E, P, 1)                                                    \
    return nullptr;                                                               \
  }                                                                               \
  extern const FusedShape fused_shape_##E##_##P = {E, P, Cfg<E, P>::SMEM_BYTES, Cfg<E, P>::SLOTS, Cfg<E, P>::THREADS, pick_##E##_##P};

}  // namespace ita
