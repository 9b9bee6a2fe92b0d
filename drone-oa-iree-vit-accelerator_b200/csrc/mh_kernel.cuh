// mh_kernel.cuh -- multi-head ITA attention with S = 256 keys (BASELINE config 3: H = 4, S = 256, E = 128, P = 64) on the
// same tcgen05 skeleton as fused_kernel.cuh, as two persistent sm_100a kernels:
//
//   k_ita_mh       per (frame, head): Q/K/V projections, Q.K^T (128 x 256 logits tile per MMA), ITAmax over 256 keys,
//                  A.V; nothing but the requantised context O [B,H,S,P] leaves the SM
//   k_ita_mh_out   per (frame, row tile): the output projection over all heads and the head reduction -- either the
//                  concatenated form of the repo's module (one accumulator over all heads, one requant;
//                  models/ITA/QAT/layers.py:110-123) or the golden model's per-head step6_O followed by step7_Osum
//
// Why two kernels: tensor memory.  A logits tile of one row tile is 128 lanes x 256 columns (half the SM's 512 columns) and
// has to stay resident for the three ITAmax passes; accumulating the output projection across heads in TMEM would take the
// other 256 columns (2 row tiles x 128), leaving nothing for the projections and A.V.  The context is 64 KB per frame and
// is re-read by the second kernel from L2/HBM (8 % of k_ita_mh's own time at config 3).
//
// k_ita_mh, one CTA per SM, frames strided over CTAs, an "item" = (frame, head):
//   shared memory   X [256 x E] (whole frame), Q [256 x P], K and V [256 x P] double-buffered (the next head's projections
//                   run while this head's attention is in progress), Asm [128 x 256] u8 (one row tile), per-head weight
//                   block [Wq | Wk | Wv]^T with the digit K columns + the head's offset tiles, double-buffered and streamed
//                   from L2 (30 KB per head), the constant digit A tile.  All operands in the no-swizzle panel layout of
//                   fused_kernel.cuh, with 256-row panels for X/Q/K/V.
//   tensor memory   RS = 256 columns (logits tile), RP = 192 (Q | K | V projection accumulators of one row tile), RA = 64
//                   (context chunk)
//   schedule        6 positions per iteration; S / AV belong to the item in its attention phase ("own"), P' are the
//                   projections of the NEXT item, placed where the own chain would otherwise bubble (A.V can only start
//                   when ITAmax has written the last probability):
//                       0: S t0   1: P' t0   2: AV t0   3: S t1   4: P' t1   5: AV t1
//                   One MMA-issuing warp walks the positions in order; 16 epilogue warps (4 TMEM lane quadrants x 4
//                   column parts) drain every position in the same order, so "the epilogue of position p is complete"
//                   (one mbarrier per position) orders every shared-memory hazard: Q is overwritten by the next head only
//                   after the logits MMA that read it, Asm of tile 1 only after A.V of tile 0, and so on.
//   roles           warps 0-15 epilogue (quad = w & 3, part = w >> 2), warp 16 MMA issue, warp 17 TMA loader + TMEM alloc
//
// Reference semantics: models/ITA/QAT/layers.py:101-127 (multi-head split of the projections, per-head attention),
// tests/export_and_validation_W_B.py:518-542 (golden model: per-head steps, step7_Osum), SURVEY.md App. A.3.
#pragma once
#include "fused_kernel.cuh"

namespace ita {

constexpr int kMhS = 256;                 // keys / rows per frame
constexpr int kMhTiles = kMhS / kS;       // row tiles of 128
constexpr int kMhParts = 4;               // column parts of the epilogue (warps per TMEM lane quadrant)
constexpr int kMhEpiWarps = 4 * kMhParts;
constexpr int kMhThreads = (kMhEpiWarps + 2) * 32;

// mbarriers of k_ita_mh
enum { MB_CA = 0, MB_W = 1 /* +buf */, MB_WFREE = 3 /* +buf */, MB_X = 5, MB_XFREE = 6, MB_FULL = 7 /* +region */, MB_EMPTY = 10 /* +region */,
       MB_POS = 13 /* +position */, MB_COUNT = 19 };
enum { RG_S = 0, RG_P = 1, RG_A = 2 };
constexpr int kMhPos = 6;

struct MhParams {
  const int8_t *wblocks;     // device: [H] weight blocks (MhCfg::WB_BYTES each)
  const int8_t *consts;      // device: the constant digit A tile (padded to 1024 B) | [H] offset tiles (DQK 768 B | DAV 256 B)
  const FRq *rq;             // device: [H][6] per-head stage constants (Q K V QK AV O)
  const int8_t *x;           // [B, 256, E] (for the L2 prefetch; the tiles come through the tensor map)
  int8_t *o;                 // [B, H, 256, P] requantised context
  long long batch;
  int H;
  int *abort_flag;
  int torch_fast;
  int8_t *dQ, *dK, *dV, *dA;   // debug dumps (natural layouts), any may be null
  uint8_t *dAsm;
  long long *trace;          // developer builds (-DITA_TRACE): [warp][iteration < 8][position][4] SM clock stamps of CTA 0
};

template <int E, int P>
struct MhCfg {
  static_assert(P == 64, "the chunk plan below is written for P = 64 (Q | K fill one 128-column chunk, V one 64-column chunk)");
  static_assert(E % 32 == 0 && E <= 128, "E a multiple of 32, at most 128");
  static constexpr int PW = 3 * P;
  static constexpr int WB_BYTES = PW * (E + kBiasK);              // [Wq | Wk | Wv]^T with the bias digit columns
  static constexpr int kMaxH = 8;
  static constexpr int OT_BYTES = kConstQKBytes + kConstAVBytes;  // a head's offset tiles: DQK | DAV (resident for every head)
  static constexpr int OT_DAV = kConstQKBytes;
  static constexpr int PR = kMhS;                                 // rows of the X/Q/K/V panels
  static constexpr int OFF_CA = 0;
  static constexpr int OFF_OT = 1024;                             // [kMaxH] offset tiles
  static constexpr int OFF_W = OFF_OT + kMaxH * OT_BYTES;         // 2 blocks
  static constexpr int OFF_X = OFF_W + 2 * WB_BYTES;
  static constexpr int OFF_Q = OFF_X + kMhS * E;
  static constexpr int OFF_K = OFF_Q + kMhS * P;                  // 2 buffers
  static constexpr int OFF_V = OFF_K + 2 * kMhS * P;              // 2 buffers
  static constexpr int OFF_ASM = OFF_V + 2 * kMhS * P;            // [128 x 256] u8
  static constexpr int OFF_XCHG = OFF_ASM + kS * kMhS;            // [parts][128] int2
  static constexpr int OFF_RQ = OFF_XCHG + kMhParts * kS * 8;     // [kMaxH][6] stage constants (a copy of MhParams::rq)
  static constexpr int OFF_BAR = OFF_RQ + kMaxH * 6 * (int)sizeof(FRq);
  static constexpr int SMEM_BYTES = OFF_BAR + 256;
  static_assert(WB_BYTES % 128 == 0 && OFF_X % 128 == 0, "alignment");
  static_assert(MB_COUNT * 8 + 8 <= 256, "barrier area");
  static_assert(SMEM_BYTES <= kSmemLimit, "does not fit shared memory");
  // TMEM columns of the four regions
  static constexpr uint32_t region_col(int r) { return r == RG_S ? 0u : r == RG_P ? 256u : 448u; }
  static constexpr int region_of(int pos) { return (pos == 0 || pos == 3) ? RG_S : (pos == 1 || pos == 4) ? RG_P : RG_A; }
  static constexpr int tile_of(int pos) { return pos / 3; }
  static constexpr bool is_next(int pos) { return pos == 1 || pos == 4; }
};

template <int E>
__device__ __forceinline__ void mh_load_x(void *dst, const CUtensorMap *map, long long frame, uint64_t *bar) {
  mbar_arrive_expect_tx(bar, kMhS * E);
#pragma unroll
  for (int c = 0; c < E / 16; ++c)
#pragma unroll
    for (int t = 0; t < kMhTiles; ++t)      // the tensor map's box is 16 B x 128 rows
      tma_load_2d(reinterpret_cast<uint8_t *>(dst) + c * kMhS * 16 + t * kS * 16, map, c * 16, (int)(frame * kMhS + t * kS), bar);
}

template <int E, int P, int MODE, int SOFTMAX, bool DUMP>
__global__ void __launch_bounds__(kMhThreads, 1) k_ita_mh(const __grid_constant__ CUtensorMap xmap, const MhParams prm) {
  using C = MhCfg<E, P>;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + C::OFF_BAR);
  uint32_t *s_tmem = reinterpret_cast<uint32_t *>(bars + MB_COUNT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int H = prm.H;
  // frames of this CTA: blockIdx.x, + gridDim.x, ...
  const long long nfr = ((long long)blockIdx.x < prm.batch) ? (prm.batch - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const long long n_items = nfr * H;

  if (threadIdx.x == 0) {
    mbar_init(bars + MB_CA, 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(bars + MB_W + b, 1);
      mbar_init(bars + MB_WFREE + b, 1);
    }
    mbar_init(bars + MB_X, 1);
    mbar_init(bars + MB_XFREE, 1);
    for (int r = 0; r < 3; ++r) {
      mbar_init(bars + MB_FULL + r, 1);
      mbar_init(bars + MB_EMPTY + r, kMhEpiWarps);
    }
    for (int p = 0; p < kMhPos; ++p) mbar_init(bars + MB_POS + p, kMhEpiWarps);
    fence_mbar_init();
  }
  if (warp == kMhEpiWarps + 1) tmem_alloc<512>(s_tmem);
  {   // the per-head stage constants are read at the start of every position: keep them in shared memory
    const uint32_t *src = reinterpret_cast<const uint32_t *>(prm.rq);
    uint32_t *dst = reinterpret_cast<uint32_t *>(smem + C::OFF_RQ);
    for (int i = threadIdx.x; i < H * 6 * (int)(sizeof(FRq) / 4); i += blockDim.x) dst[i] = src[i];
  }
  const FRq *const rq_tab = reinterpret_cast<const FRq *>(smem + C::OFF_RQ);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == kMhEpiWarps + 1) {
    // ================================================================= loader
    if (lane == 0 && n_items > 0) {
      tma_prefetch_desc(&xmap);
      mbar_arrive_expect_tx(bars + MB_CA, 1024 + H * C::OT_BYTES);
      bulk_g2s(smem + C::OFF_CA, prm.consts, 1024 + H * C::OT_BYTES, bars + MB_CA);
      bool ok = true;
      for (long long j = 0; j < n_items && ok; ++j) {
        const long long k = j / H;
        const int h = (int)(j - k * H);
        if (h == 0) {
          if (k > 0) ok = mbar_wait(bars + MB_XFREE, (uint32_t)((k - 1) & 1), prm.abort_flag);
          if (!ok) break;
          const long long f = blockIdx.x + k * (long long)gridDim.x;
          mh_load_x<E>(smem + C::OFF_X, &xmap, f, bars + MB_X);
          if (k + 1 < nfr) l2_prefetch(prm.x + (f + gridDim.x) * (long long)(kMhS * E), kMhS * E);
        }
        const int b = (int)(j & 1);
        if (j >= 2) ok = mbar_wait(bars + MB_WFREE + b, (uint32_t)(((j >> 1) - 1) & 1), prm.abort_flag);
        if (!ok) break;
        mbar_arrive_expect_tx(bars + MB_W + b, C::WB_BYTES);
        bulk_g2s(smem + C::OFF_W + b * C::WB_BYTES, prm.wblocks + (size_t)h * C::WB_BYTES, C::WB_BYTES, bars + MB_W + b);
      }
    }
    __syncwarp();
  } else if (warp == kMhEpiWarps) {
    // ================================================================= MMA issue
    if (n_items > 0) {
      const bool leader = elect_one();
      bool ok = mbar_wait(bars + MB_CA, 0, prm.abort_flag);
      const uint32_t ca16 = smem_u32(smem + C::OFF_CA) >> 4, x16 = smem_u32(smem + C::OFF_X) >> 4, q16 = smem_u32(smem + C::OFF_Q) >> 4,
                     asm16 = smem_u32(smem + C::OFF_ASM) >> 4;
      constexpr int DSG = rqm_offset(MODE) ? 0 : 1;       // digit operands: unsigned for the offset forms
      constexpr uint32_t PS = C::PR * 16;                 // panel stride of the 256-row buffers
      uint32_t use[3] = {0, 0, 0};                        // uses of each TMEM region so far
      for (long long it = 0; it <= n_items && ok; ++it) {
        const bool own = it >= 1, nxt = it < n_items;
        const uint32_t php = (uint32_t)((it - 1) & 1);    // parity of "completion number `it`" of a position barrier
        const int ob = (int)((it - 1) & 1), nb = (int)(it & 1);          // K/V/W buffer of the own / next item
        const int nh = nxt ? (int)(it % H) : 0;
        const uint32_t kown16 = (smem_u32(smem + C::OFF_K) >> 4) + ob * (kMhS * P / 16), vown16 = (smem_u32(smem + C::OFF_V) >> 4) + ob * (kMhS * P / 16);
        const uint32_t w16 = (smem_u32(smem + C::OFF_W) >> 4) + nb * (C::WB_BYTES / 16);
        const uint32_t ot16 = (smem_u32(smem + C::OFF_OT) >> 4) + (own ? (uint32_t)((it - 1) % H) : 0u) * (C::OT_BYTES / 16);   // the own head's offset tiles
        static_for<0, kMhPos>([&](auto pp) {
          constexpr int pos = decltype(pp)::value;
          constexpr int r = C::region_of(pos), t = C::tile_of(pos);
          if (!ok) return;
          if (C::is_next(pos) ? !nxt : !own) return;
          ITA_TR(prm, warp, it, pos, 0);
          if (use[r] > 0) ok = mbar_wait(bars + MB_EMPTY + r, (use[r] - 1) & 1, prm.abort_flag);
          if (!ok) return;
          ITA_TR(prm, warp, it, pos, 1);
          tc_fence_after();
          const uint32_t d = tmem_base + C::region_col(r);
          if constexpr (pos == 0 || pos == 3) {
            // ---- logits of row tile t: A = Q rows [128 t, +128) (K = P), B = K (all 256 keys); offset digit steps first
            if (rqm_offset(MODE)) {
              constexpr uint32_t IDD = make_idesc_i8(kS, kMhS, 0, 0, 0, 0);
              if (leader) {
                umma_i8(d, make_smem_desc_rel(ca16, 0, 128, 0), make_smem_desc_rel(ot16, 0, 128, 0), IDD, 0);
                umma_i8(d, make_smem_desc_rel(ca16, 256, 128, 0), make_smem_desc_rel(ot16, 256, 128, 0), IDD, 1);
                if constexpr (rqm_qk_steps(MODE) == 3)
                  umma_i8(d, make_smem_desc_rel(ca16, 512, 128, 0), make_smem_desc_rel(ot16, 512, 128, 0), IDD, 1);
              }
              __syncwarp();
            }
            if constexpr (pos == 0) {        // Q, K, V of the own item: positions 1 and 4 of the iteration before
              ok = mbar_wait(bars + MB_POS + 4, php, prm.abort_flag);
              if (!ok) return;
              tc_fence_after();
            }
            constexpr uint32_t ID = make_idesc_i8(kS, kMhS, 1, 1, 0, 0);
            if (leader) {
#pragma unroll
              for (int k = 0; k < P / 32; ++k)
                umma_i8(d, make_smem_desc_rel(q16, t * kS * 16 + k * 2 * PS, PS, 128), make_smem_desc_rel(kown16, k * 2 * PS, PS, 128), ID,
                        (k > 0 || rqm_offset(MODE)) ? 1u : 0u);
            }
          } else if constexpr (pos == 2 || pos == 5) {
            // ---- context of row tile t: A = Asm [128, K = 256] u8, B = V [K = 256 keys, N = P] MN-major
            if (rqm_offset(MODE)) {
              constexpr uint32_t IDD = make_idesc_i8(kS, P, 0, 0, 0, 0);
              if (leader) umma_i8(d, make_smem_desc_rel(ca16, 0, 128, 0), make_smem_desc_rel(ot16, C::OT_DAV, 128, 0), IDD, 0);
              __syncwarp();
            }
            // Asm of this tile: the ITAmax epilogue of position pos - 2 (V of the own item was waited for at position 0)
            ok = mbar_wait(bars + MB_POS + (pos - 2), php, prm.abort_flag);
            if (!ok) return;
            tc_fence_after();
            constexpr uint32_t ID = make_idesc_i8(kS, P, 0, 1, 0, 1);
            if (leader) {
#pragma unroll
              for (int k = 0; k < kMhS / 32; ++k)
                umma_i8(d, make_smem_desc_rel(asm16, k * 2 * kS * 16, kS * 16, 128), make_smem_desc_rel(vown16, k * 32 * 16, 128, PS), ID,
                        (k > 0 || rqm_offset(MODE)) ? 1u : 0u);
            }
          } else {
            // ---- projections of the next item, row tile t: Q | K into columns [0, 128) of RP, V into [128, 192)
            if constexpr (pos == 1) {
              ok = mbar_wait(bars + MB_W + nb, (uint32_t)((it >> 1) & 1), prm.abort_flag);
              if (ok && nh == 0) ok = mbar_wait(bars + MB_X, (uint32_t)((it / H) & 1), prm.abort_flag);
              if (!ok) return;
              tc_fence_after();
            }
            if (leader) {
              static_for<0, 2>([&](auto cc) {
                constexpr int c = decltype(cc)::value;
                constexpr int col0 = c == 0 ? 0 : 2 * P, NC = c == 0 ? 2 * P : P;
                constexpr uint32_t ID = make_idesc_i8(kS, NC, 1, 1, 0, 0), IDD = make_idesc_i8(kS, NC, DSG, DSG, 0, 0);
#pragma unroll
                for (int k = 0; k < E / 32; ++k)
                  umma_i8(d + col0, make_smem_desc_rel(x16, t * kS * 16 + k * 2 * PS, PS, 128),
                          make_smem_desc_rel(w16, col0 * 16 + k * 2 * C::PW * 16, C::PW * 16, 128), ID, k > 0);
                umma_i8(d + col0, make_smem_desc_rel(ca16, 0, 128, 0), make_smem_desc_rel(w16, col0 * 16 + (E / 16) * C::PW * 16, C::PW * 16, 128),
                        IDD, 1);
              });
            }
          }
          ITA_TR(prm, warp, it, pos, 2);
          if (leader) {
            umma_commit(bars + MB_FULL + r);
            if constexpr (pos == 4) {
              umma_commit(bars + MB_WFREE + nb);                   // the item's last projection has read its weight block
              if (nh == H - 1) umma_commit(bars + MB_XFREE);       // the frame's last projection has read X
            }
          }
          __syncwarp();
          ++use[r];
          ITA_TR(prm, warp, it, pos, 3);
        });
      }
    }
  } else {
    // ================================================================= epilogue warps
    const int quad = warp & 3, part = warp >> 2;
    const int rowin = quad * 32 + lane;                      // row inside a 128-row tile = TMEM lane
    const uint32_t t_lane = tmem_base + ((uint32_t)(quad * 32) << 16);
    PairCtx pc;
    pc.bar_id = 1 + quad;
    pc.xch = reinterpret_cast<int2 *>(smem + C::OFF_XCHG);
    int8_t *const q_buf = reinterpret_cast<int8_t *>(smem + C::OFF_Q);
    int8_t *const asm_buf = reinterpret_cast<int8_t *>(smem + C::OFF_ASM);
    bool ok = true;
    uint32_t use[3] = {0, 0, 0};
    for (long long it = 0; it <= n_items && ok; ++it) {
      const bool own = it >= 1, nxt = it < n_items;
      const int nb = (int)(it & 1);
      const long long jo = it - 1, jn = it;
      const long long fo = own ? blockIdx.x + (jo / H) * (long long)gridDim.x : 0, fn = nxt ? blockIdx.x + (jn / H) * (long long)gridDim.x : 0;
      const int ho = own ? (int)(jo % H) : 0, hn = nxt ? (int)(jn % H) : 0;
      int8_t *const k_buf = reinterpret_cast<int8_t *>(smem + C::OFF_K) + nb * (kMhS * P);
      int8_t *const v_buf = reinterpret_cast<int8_t *>(smem + C::OFF_V) + nb * (kMhS * P);
      static_for<0, kMhPos>([&](auto pp) {
        constexpr int pos = decltype(pp)::value;
        constexpr int r = C::region_of(pos), t = C::tile_of(pos);
        if (!ok) return;
        if (C::is_next(pos) ? !nxt : !own) return;
        ITA_TR(prm, warp, it, pos, 0);
        ok = mbar_wait(bars + MB_FULL + r, use[r] & 1, prm.abort_flag);
        ++use[r];
        if (!ok) return;
        ITA_TR(prm, warp, it, pos, 1);
        tc_fence_after();
        auto release = make_release(
            [&] {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(bars + MB_EMPTY + r);
              ITA_TR(prm, warp, it, pos, 2);
            },
            [] {});
        const int row = t * kS + rowin;                      // row inside the frame
        if constexpr (pos == 0 || pos == 3) {
          // ---- logits tile: this part's 64 key columns -> ITAmax -> Asm (u8)
          const FRq rq = rq_tab[ho * 6 + STG_S];
          const uint32_t taddr = t_lane + C::region_col(RG_S) + (uint32_t)part * 64u;
          int8_t *gA = nullptr;
          uint8_t *gAsm = nullptr;
          if constexpr (DUMP) {
            const long long base = ((fo * H + ho) * kMhS + row) * (long long)kMhS;
            if (prm.dA) gA = prm.dA + base;
            if (prm.dAsm) gAsm = prm.dAsm + base;
          }
          if constexpr (SOFTMAX == ITA_SOFTMAX_REPO_GLOBAL && rqm_offset(MODE))
            itamax_repo_half_fp<MODE, DUMP, kMhParts>(!rqm_torch_fp(MODE) || prm.torch_fast != 0, part, taddr, rq, asm_buf, rowin, gA, gAsm, pc, release);
          else if constexpr (SOFTMAX == ITA_SOFTMAX_REPO_GLOBAL)
            itamax_repo_half<MODE, 0, DUMP, kMhParts>(part, taddr, rq, asm_buf, rowin, gA, gAsm, pc, release);
          else
            itamax_stream_half<MODE, DUMP, kMhParts>(part, taddr, rq, asm_buf, rowin, gA, gAsm, pc, release);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(bars + MB_POS + pos);
        } else if constexpr (pos == 2 || pos == 5) {
          // ---- context chunk: 16 columns per part -> requant -> global O [B, H, 256, P]
          const FRq rq = rq_tab[ho * 6 + STG_AV];
          uint32_t v[16];
          tmem_ld16(t_lane + C::region_col(r) + (uint32_t)part * 16u, v);
          tmem_ld_wait();
          tmem_pin16(v);
          release.issue();
          int8_t *g = prm.o + ((fo * H + ho) * kMhS + row) * (long long)P + part * 16;
          requant_store32<MODE, 0, false, 16, C::PR>(v, rq, nullptr, row, 0, g);
        } else {
          // ---- projection accumulators of the next item's row tile: 32 of the Q | K columns (parts 0, 1 -> Q columns [0, 32),
          //      [32, 64); parts 2, 3 -> K) and 16 of the V columns per part, loaded together
          const int stage = part < 2 ? STG_Q : STG_K, col = (part & 1) * 32;
          const FRq rq = rq_tab[hn * 6 + stage];
          const FRq rqv = rq_tab[hn * 6 + STG_V];
          uint32_t v[32], v2[16];
          tmem_ld32(t_lane + C::region_col(r) + (uint32_t)part * 32u, v);
          tmem_ld16(t_lane + C::region_col(r) + 2u * P + (uint32_t)part * 16u, v2);
          tmem_wait32(v);                                      // tcgen05.wait::ld covers both loads
          tmem_pin16(v2);                                      // ... and nothing that reads v2 may move above it
          release.issue();
          int8_t *g = nullptr, *g2 = nullptr;
          if constexpr (DUMP) {
            int8_t *base = part < 2 ? prm.dQ : prm.dK;
            if (base) g = base + ((fn * H + hn) * kMhS + row) * (long long)P + col;
            if (prm.dV) g2 = prm.dV + ((fn * H + hn) * kMhS + row) * (long long)P + part * 16;
          }
          requant_store32<MODE, 0, false, 32, C::PR>(v, rq, part < 2 ? q_buf : k_buf, row, col, g);
          requant_store32<MODE, 0, false, 16, C::PR>(v2, rqv, v_buf, row, part * 16, g2);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(bars + MB_POS + pos);
        }
        ITA_TR(prm, warp, it, pos, 3);
      });
    }
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == kMhEpiWarps + 1) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------------------- output projection
// k_ita_mh_out: item = (frame, row tile).  A operand: the context rows of every head, [128 x P] per head (TMA from
// O [B, H, 256, P]); B: Wo^T [E rows, K = H*P] plus digit blocks [E x 64] (one for the concatenated form, one per head for
// the per-head form).
//   concatenated form   one accumulator [128 x E] over every head, double-buffered in TMEM, requantised once
//   per-head form       (HEADS_PYITA: step6_O per head, then step7_Osum) an item is two "units" of E/2 output columns; a unit
//                       has one [128 x E/2] accumulator per head (H * E/2 <= 256 columns), two units alternate in TMEM.  The
//                       epilogue requantises each head (biased bytes y + 128, saturated by the pack), adds the heads in 16-bit
//                       lanes (two elements per register) and maps the sum through a shared-memory table of the step7_Osum
//                       requantisation (the sum has at most 255 * H + 1 values, so the table is exact for every rounding mode).
// Roles: warps 0-7 epilogue (quad = w & 3, half = w >> 2), warp 8 MMA issue, warp 9 loader + TMEM alloc.
constexpr int kMoThreads = 10 * 32;
constexpr int kMoStages = 3;              // context tiles in flight
enum { MO_W = 0, MO_IN = 1 /* +stage */, MO_INFREE = 4 /* +stage */, MO_FULL = 7 /* +acc */, MO_EMPTY = 9 /* +acc */, MO_COUNT = 11 };

struct MoParams {
  const int8_t *wo;          // device: Wo^T panels [E x H*P] | digit blocks [ND][E x 64] | constant A tile
  const FRq *rq;             // device: [H] stage O constants (concatenated form: entry 0)
  Rq rq_sum;                 // step7_Osum (per-head form)
  int8_t *y;                 // [B, 256, E]
  int8_t *dOut;              // per-head form, debug: [B, H, 256, E]
  long long batch;
  int H;
  int *abort_flag;
};

template <int E, int P>
struct MoCfg {
  static constexpr int kMaxH = 4;
  static constexpr int kLutBytes = 1024;                         // step7_Osum table: index = sum of the biased head outputs
  static constexpr int wo_bytes(int H) { return E * H * P; }
  static constexpr int dig_bytes = E * kBiasKO;
  static constexpr int w_bytes(int H, bool per_head) { return wo_bytes(H) + (per_head ? H : 1) * dig_bytes + kConstABytes; }
  static constexpr int in_bytes(int H) { return kS * H * P; }
  static constexpr int OFF_W = 0;
  static constexpr int off_in(int H, bool per_head) { return ((w_bytes(H, per_head) + 127) / 128) * 128; }
  static constexpr int off_lut(int H, bool per_head) { return off_in(H, per_head) + kMoStages * in_bytes(H); }
  static constexpr int off_bar(int H, bool per_head) { return off_lut(H, per_head) + kLutBytes; }
  static constexpr int smem_bytes(int H, bool per_head) { return off_bar(H, per_head) + 128; }
};

template <int E, int P, int MODE, bool PER_HEAD, bool DUMP>
__global__ void __launch_bounds__(kMoThreads, 1) k_ita_mh_out(const __grid_constant__ CUtensorMap omap, const MoParams prm) {
  using C = MoCfg<E, P>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int H = prm.H;
  const int off_in = C::off_in(H, PER_HEAD), in_bytes = C::in_bytes(H);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + C::off_bar(H, PER_HEAD));
  uint32_t *s_tmem = reinterpret_cast<uint32_t *>(bars + MO_COUNT);
  int8_t *const lut = reinterpret_cast<int8_t *>(smem + C::off_lut(H, PER_HEAD));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long total = prm.batch * kMhTiles;            // items
  const long long n_items = ((long long)blockIdx.x < total) ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  constexpr int NU = PER_HEAD ? 2 : 1;                     // units per item (per-head form: two column halves)
  constexpr int UW = E / NU;                               // output columns of a unit
  constexpr uint32_t ACC_COLS = 256;                       // TMEM columns of one accumulator set
  if (threadIdx.x == 0) {
    mbar_init(bars + MO_W, 1);
    for (int s = 0; s < kMoStages; ++s) {
      mbar_init(bars + MO_IN + s, 1);
      mbar_init(bars + MO_INFREE + s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(bars + MO_FULL + a, 1);
      mbar_init(bars + MO_EMPTY + a, 8);
    }
    fence_mbar_init();
  }
  if (warp == 9) tmem_alloc<512>(s_tmem);
  if constexpr (PER_HEAD) {
    const Rq rs = prm.rq_sum;
    for (int i = threadIdx.x; i < C::kLutBytes; i += blockDim.x) lut[i] = (int8_t)requant_general(i - 128 * H, rs);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  const int wbytes = C::w_bytes(H, PER_HEAD);
  const int off_dig = C::wo_bytes(H), off_ca = off_dig + (PER_HEAD ? H : 1) * C::dig_bytes;

  if (warp == 9) {
    if (lane == 0 && n_items > 0) {
      tma_prefetch_desc(&omap);
      mbar_arrive_expect_tx(bars + MO_W, wbytes);
      bulk_g2s(smem + C::OFF_W, prm.wo, wbytes, bars + MO_W);
      bool ok = true;
      for (long long i = 0; i < n_items && ok; ++i) {
        const int s = (int)(i % kMoStages);
        const long long u = i / kMoStages;
        if (u > 0) ok = mbar_wait(bars + MO_INFREE + s, (uint32_t)((u - 1) & 1), prm.abort_flag);
        if (!ok) break;
        const long long item = blockIdx.x + i * (long long)gridDim.x;
        const long long f = item / kMhTiles;
        const int t = (int)(item % kMhTiles);
        uint8_t *dst = smem + off_in + s * in_bytes;
        mbar_arrive_expect_tx(bars + MO_IN + s, in_bytes);
        for (int h = 0; h < H; ++h)
#pragma unroll
          for (int c = 0; c < P / 16; ++c)
            tma_load_2d(dst + (h * (P / 16) + c) * kS * 16, &omap, c * 16, (int)(((f * H + h) * kMhS) + t * kS), bars + MO_IN + s);
      }
    }
    __syncwarp();
  } else if (warp == 8) {
    if (n_items > 0) {
      const bool leader = elect_one();
      bool ok = mbar_wait(bars + MO_W, 0, prm.abort_flag);
      const uint32_t w16 = smem_u32(smem + C::OFF_W) >> 4;
      constexpr int DSG = rqm_offset(MODE) ? 0 : 1;
      constexpr uint32_t ID = make_idesc_i8(kS, UW, 1, 1, 0, 0), IDD = make_idesc_i8(kS, UW, DSG, DSG, 0, 0);
      long long un = 0;                                      // running unit number: accumulator set un & 1, its (un >> 1)-th use
      for (long long i = 0; i < n_items && ok; ++i) {
        const int s = (int)(i % kMoStages);
        const long long u = i / kMoStages;
        ok = mbar_wait(bars + MO_IN + s, (uint32_t)(u & 1), prm.abort_flag);
        const uint32_t in16 = (smem_u32(smem + off_in) >> 4) + s * (in_bytes / 16);
#pragma unroll
        for (int c = 0; c < NU; ++c, ++un) {
          const int a = (int)(un & 1);
          const long long ua = un >> 1;
          if (ok && ua > 0) ok = mbar_wait(bars + MO_EMPTY + a, (uint32_t)((ua - 1) & 1), prm.abort_flag);
          if (!ok) break;
          tc_fence_after();
          if (leader) {
            const uint32_t n0 = (uint32_t)(c * UW * 16);        // first output column of the unit, as a row offset of the B panels
            for (int h = 0; h < H; ++h) {
              const uint32_t d = tmem_base + a * ACC_COLS + (PER_HEAD ? h * UW : 0);
              if (PER_HEAD || h == 0) {
                // digit steps: bias (+ offset) of this accumulator; the offset forms use both 32-column steps
                const uint32_t dg = off_dig + (PER_HEAD ? h : 0) * C::dig_bytes + n0;
                umma_i8(d, make_smem_desc_rel(w16, off_ca, 128, 0), make_smem_desc_rel(w16, dg, E * 16, 128), IDD, 0);
                if (rqm_offset(MODE))
                  umma_i8(d, make_smem_desc_rel(w16, off_ca + 256, 128, 0), make_smem_desc_rel(w16, dg + 2 * E * 16, E * 16, 128), IDD, 1);
              }
#pragma unroll
              for (int k = 0; k < P / 32; ++k)
                umma_i8(d, make_smem_desc_rel(in16, (h * (P / 16) + 2 * k) * kS * 16, kS * 16, 128),
                        make_smem_desc_rel(w16, (h * (P / 16) + 2 * k) * E * 16 + n0, E * 16, 128), ID, 1);
            }
            umma_commit(bars + MO_FULL + a);
            if (c == NU - 1) umma_commit(bars + MO_INFREE + s);
          }
          __syncwarp();
        }
      }
    }
  } else {
    const int quad = warp & 3, half = warp >> 2;
    const int rowin = quad * 32 + lane;
    const uint32_t t_lane = tmem_base + ((uint32_t)(quad * 32) << 16);
    constexpr int HW = UW / 2;                              // columns per warp and unit
    bool ok = true;
    long long un = 0;
    for (long long i = 0; i < n_items && ok; ++i) {
      const long long item = blockIdx.x + i * (long long)gridDim.x;
      const long long f = item / kMhTiles;
      const int row = (int)(item % kMhTiles) * kS + rowin;
#pragma unroll
      for (int c = 0; c < NU; ++c, ++un) {
        const int a = (int)(un & 1);
        const long long ua = un >> 1;
        ok = mbar_wait(bars + MO_FULL + a, (uint32_t)(ua & 1), prm.abort_flag);
        if (!ok) break;
        tc_fence_after();
        const int col = c * UW + half * HW;                    // first output column of this warp
        int8_t *g = prm.y + (f * kMhS + row) * (long long)E + col;
        auto release = make_release(
            [&] {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(bars + MO_EMPTY + a);
            },
            [] {});
        if constexpr (!PER_HEAD) {
          const FRq rq = prm.rq[0];
          tmem_walk<HW / 32>([&](int gi) { return t_lane + (uint32_t)(a * ACC_COLS + half * HW + gi * 32); },
                             [&](auto gg, const uint32_t(&v)[32]) {
                               constexpr int gi = decltype(gg)::value;
                               requant_store32<MODE, 0>(v, rq, nullptr, rowin, 0, g + gi * 32);
                             },
                             release);
        } else {
          static_assert(!PER_HEAD || HW == 32, "per-head form: 32 columns per warp and unit");
          uint32_t lo[8], hi[8];                               // 16-bit lane sums of the biased head outputs: elements (4q, 4q+2) / (4q+1, 4q+3)
#pragma unroll
          for (int q = 0; q < 8; ++q) lo[q] = hi[q] = 0;
          for (int h = 0; h < H; ++h) {
            const FRq rq = prm.rq[h];
            uint32_t v[32];
            tmem_ld32(t_lane + (uint32_t)(a * ACC_COLS + h * UW + half * HW), v);
            tmem_wait32(v);
            if (h == H - 1) release.issue();
            uint32_t w[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              int y[4];
              uint32_t ub;                                     // y + 128 per byte
              if constexpr (rqm_offset(MODE)) {
                rq_off_bits2<MODE>(v[q * 4], v[q * 4 + 1], rq, y[0], y[1]);
                rq_off_bits2<MODE>(v[q * 4 + 2], v[q * 4 + 3], rq, y[2], y[3]);
                ub = pack_u8x4(y[0], y[1], y[2], y[3]);
              } else {
#pragma unroll
                for (int e = 0; e < 4; ++e) y[e] = rq_apply<MODE, 0>((int)v[q * 4 + e], rq, e);
                ub = pack_s8x4(y[0], y[1], y[2], y[3]) ^ 0x80808080u;
              }
              lo[q] += ub & 0x00FF00FFu;
              hi[q] += (ub >> 8) & 0x00FF00FFu;
              w[q] = ub ^ 0x80808080u;
            }
            if constexpr (DUMP) {
              if (prm.dOut) {
                int8_t *go = prm.dOut + ((f * H + h) * kMhS + row) * (long long)E + col;
                *reinterpret_cast<uint4 *>(go) = make_uint4(w[0], w[1], w[2], w[3]);
                *reinterpret_cast<uint4 *>(go + 16) = make_uint4(w[4], w[5], w[6], w[7]);
              }
            }
          }
          uint32_t o[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const uint32_t b0 = (uint8_t)lut[lo[q] & 0xFFFFu], b2 = (uint8_t)lut[lo[q] >> 16], b1 = (uint8_t)lut[hi[q] & 0xFFFFu],
                           b3 = (uint8_t)lut[hi[q] >> 16];
            o[q] = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
          }
          *reinterpret_cast<uint4 *>(g) = make_uint4(o[0], o[1], o[2], o[3]);
          *reinterpret_cast<uint4 *>(g + 16) = make_uint4(o[4], o[5], o[6], o[7]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

typedef void (*MhKernelFn)(const CUtensorMap, const MhParams);
typedef void (*MoKernelFn)(const CUtensorMap, const MoParams);

}  // namespace ita
