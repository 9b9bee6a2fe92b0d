// ffn_kernel.cuh -- the ITA feed-forward block  F2 = rq_fc2( relu(rq_fc1(X.W1 + b1)) . W2 + b2 )  as one persistent sm_100a
// kernel on the same building blocks as the attention kernel (fused_kernel.cuh): int8 operands in shared memory in the
// panel layout, s32 accumulators in tensor memory, tcgen05.mma kind::i8, biases (and the accumulator offsets of the
// FMA-pipe requant forms) added on the tensor pipe by digit K steps, requantisation in the epilogue warps, the post-ReLU
// activations F1 [128 x F] never leaving shared memory.
//
// Reference semantics: ITAFeedForward_QAT.forward between its quant and dequant stubs (models/ITA/QAT/layers.py:61-75) ==
// golden-model feedforward_layer (tests/export_and_validation_W_B.py:544-553); ReLU requant parameters :278-301 (the
// converted model's quantised ReLU keeps the scale, so that requant is the identity -- checked on the host).
//
// Pipeline.  The block is light (8.4 Mop and 40 k requantised elements per frame at E=64, F=256), so the kernel keeps up to
// FOUR frames in flight per SM ("slots"), each with ONE epilogue warpgroup (4 warps = the 4 TMEM lane quadrants), one
// MMA-issuing warp and a ring of 2 accumulator regions x 64 columns.  A frame is F/64 fc1 chunks followed by E/64 fc2
// chunks, 64 accumulator columns each; the fc2 MMAs wait for the F1 buffer to be complete (mbarrier armed by the fc1
// epilogues).  Weights ([W1^T | digits], [W2^T | digits], the constant A tile) are bulk-copied once per CTA.
#pragma once
#include "fused_kernel.cuh"

namespace ita {

constexpr int kFfnMaxSlots = 4;
enum { FB_X = 0, FB_XFREE = 1, FB_FULL = 2 /* + region */, FB_EMPTY = 4 /* + region */, FB_F1 = 6, kFfnBarsPerSlot = 8 };

struct FfnParams {
  const int8_t *wpanels;     // device: [W1^T | bias digits] | [W2^T | bias digits] | constant A tile
  int8_t *y;                 // [B,S,E]
  int8_t *f1;                // optional [B,S,F] (post-ReLU), may be null
  long long batch;
  int *abort_flag;
  FRq rq[2];                 // fc1, fc2
};

template <int E, int F>
struct FfnCfg {
  static constexpr int S = kS;
  static constexpr int NC1 = F / 64, NC2 = E / 64, NCH = NC1 + NC2;
  static constexpr int W1_BYTES = F * (E + kBiasK);                      // W1^T [F rows, K = E + 32]
  static constexpr int W2_BYTES = E * (F + kBiasKO);                     // W2^T [E rows, K = F + 64]
  static constexpr int OFF_W1 = 0, OFF_W2 = W1_BYTES, OFF_CA = W1_BYTES + W2_BYTES;
  static constexpr int CA_BYTES = 8 * 64;
  static constexpr int WPANEL_BYTES = OFF_CA + CA_BYTES;
  static constexpr int X_BYTES = S * E, F1_BYTES = S * F;
  static constexpr int OFF_X = 0, OFF_F1 = X_BYTES;
  static constexpr int SLOT_BYTES = X_BYTES + F1_BYTES;
  static constexpr int OFF_SLOT = ((WPANEL_BYTES + 127) / 128) * 128;
  static constexpr int smem_for(int slots) { return OFF_SLOT + slots * SLOT_BYTES + 512; }
  static constexpr int SLOTS = smem_for(4) <= kSmemLimit ? 4 : smem_for(3) <= kSmemLimit ? 3 : smem_for(2) <= kSmemLimit ? 2 : 1;
  static constexpr int WARP_MMA = SLOTS * 4, WARP_LOAD = WARP_MMA + SLOTS, THREADS = (WARP_LOAD + 1) * 32;
  static constexpr int OFF_BAR = OFF_SLOT + SLOTS * SLOT_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 512;
  static constexpr int TMEM_COLS_PER_SLOT = 128;                         // 2 regions x 64 columns
  static_assert(E % 64 == 0 && E <= 128 && F % 64 == 0 && F <= 512, "E in {64, 128}, F a multiple of 64");
  static_assert(SMEM_BYTES <= kSmemLimit, "shape does not fit shared memory");
  static_assert((1 + kFfnMaxSlots * kFfnBarsPerSlot) * 8 + 8 <= 512, "barrier area");
  static_assert(WPANEL_BYTES % 16 == 0 && SLOT_BYTES % 128 == 0, "alignment");
};

template <int E, int F, int MODE, bool DUMP>
__global__ void __launch_bounds__((FfnCfg<E, F>::THREADS), 1) k_ita_ffn(const __grid_constant__ CUtensorMap xmap, const FfnParams prm) {
  using C = FfnCfg<E, F>;
  constexpr int kSlots = C::SLOTS;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + C::OFF_BAR);
  uint64_t *bar_w = bars;
  uint64_t *slot_bars = bars + 1;
  uint32_t *s_tmem = reinterpret_cast<uint32_t *>(bars + 1 + kSlots * kFfnBarsPerSlot);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long nframes = prm.batch;
  const long long frame_stride = (long long)gridDim.x * kSlots;
  auto iters_of = [&](int slot) -> long long {
    const long long first = (long long)blockIdx.x * kSlots + slot;
    return first >= nframes ? 0 : (nframes - first + frame_stride - 1) / frame_stride;
  };

  if (threadIdx.x == 0) {
    mbar_init(bar_w, 1);
    for (int s = 0; s < kSlots; ++s) {
      uint64_t *b = slot_bars + s * kFfnBarsPerSlot;
      mbar_init(b + FB_X, 1);
      mbar_init(b + FB_XFREE, 1);
      for (int r = 0; r < 2; ++r) {
        mbar_init(b + FB_FULL + r, 1);
        mbar_init(b + FB_EMPTY + r, 4);
      }
      mbar_init(b + FB_F1, 4 * C::NC1);
    }
    fence_mbar_init();
  }
  if (warp == C::WARP_LOAD) tmem_alloc<512>(s_tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == C::WARP_LOAD) {
    // ================================================================= loader: weights once, then the X tiles
    if (lane == 0) {
      tma_prefetch_desc(&xmap);
      mbar_arrive_expect_tx(bar_w, C::WPANEL_BYTES);
      bulk_g2s(smem, prm.wpanels, C::WPANEL_BYTES, bar_w);
      long long iters[kSlots], mx = 0;
      for (int s = 0; s < kSlots; ++s) {
        iters[s] = iters_of(s);
        mx = iters[s] > mx ? iters[s] : mx;
        if (iters[s] > 0)
          load_x_tile<E>(smem + C::OFF_SLOT + s * C::SLOT_BYTES + C::OFF_X, &xmap, (long long)blockIdx.x * kSlots + s,
                         slot_bars + s * kFfnBarsPerSlot + FB_X);
      }
      bool ok = true;
      for (long long it = 0; it + 1 < mx && ok; ++it)
        for (int s = 0; s < kSlots && ok; ++s)
          if (it + 1 < iters[s]) {
            ok = mbar_wait(slot_bars + s * kFfnBarsPerSlot + FB_XFREE, (uint32_t)(it & 1), prm.abort_flag);
            if (ok)
              load_x_tile<E>(smem + C::OFF_SLOT + s * C::SLOT_BYTES + C::OFF_X, &xmap,
                             (long long)blockIdx.x * kSlots + s + (it + 1) * frame_stride, slot_bars + s * kFfnBarsPerSlot + FB_X);
          }
    }
    __syncwarp();
  } else if (warp >= C::WARP_MMA) {
    // ================================================================= MMA issue for one slot
    const int slot = __shfl_sync(0xffffffffu, warp - C::WARP_MMA, 0);
    const long long iters = iters_of(slot);
    if (iters > 0) {
      const bool leader = elect_one();
      uint64_t *b = slot_bars + slot * kFfnBarsPerSlot;
      bool ok = mbar_wait(bar_w, 0, prm.abort_flag);
      const uint32_t w16 = smem_u32(smem) >> 4, ca16 = smem_u32(smem + C::OFF_CA) >> 4;
      const uint32_t slot16 = smem_u32(smem + C::OFF_SLOT + slot * C::SLOT_BYTES) >> 4;
      const uint32_t x16 = slot16 + (C::OFF_X >> 4), f16 = slot16 + (C::OFF_F1 >> 4);
      const uint32_t d_base = tmem_base + slot * C::TMEM_COLS_PER_SLOT;
      constexpr int DSG = rqm_offset(MODE) ? 0 : 1;                       // digit operands: u8 (offset forms) or s8
      constexpr uint32_t ID = make_idesc_i8(kS, 64, 1, 1, 0, 0), IDD = make_idesc_i8(kS, 64, DSG, DSG, 0, 0);
      uint32_t n = 0;
      for (long long it = 0; it < iters && ok; ++it) {
        const uint32_t ph = (uint32_t)(it & 1);
        static_for<0, C::NCH>([&](auto cc) {
          constexpr int c = decltype(cc)::value;
          if (!ok) return;
          const uint32_t r = n & 1u, use = n >> 1;
          ++n;
          if (use > 0) ok = mbar_wait(b + FB_EMPTY + r, (use - 1) & 1, prm.abort_flag);
          if (!ok) return;
          const uint32_t d = d_base + r * 64;
          if constexpr (c < C::NC1) {
            if constexpr (c == 0) ok = mbar_wait(b + FB_X, ph, prm.abort_flag);
            if (!ok) return;
            tc_fence_after();
            if (leader) {
              constexpr uint32_t wt = C::OFF_W1 + c * 64 * 16;
#pragma unroll
              for (int k = 0; k < E / 32; ++k)
                umma_i8(d, make_smem_desc_rel(x16, k * 2 * kS * 16, kS * 16, 128), make_smem_desc_rel(w16, wt + k * 2 * F * 16, F * 16, 128), ID,
                        k > 0);
              umma_i8(d, make_smem_desc_rel(ca16, 0, 128, 0), make_smem_desc_rel(w16, wt + (E / 16) * F * 16, F * 16, 128), IDD, 1);
              umma_commit(b + FB_FULL + r);
              if constexpr (c == C::NC1 - 1) umma_commit(b + FB_XFREE);
            }
          } else {
            constexpr int j = c - C::NC1;
            constexpr uint32_t wt = C::OFF_W2 + j * 64 * 16;
            if (leader) {      // digit steps first: they read constants only
              tc_fence_after();
              umma_i8(d, make_smem_desc_rel(ca16, 0, 128, 0), make_smem_desc_rel(w16, wt + (F / 16) * E * 16, E * 16, 128), IDD, 0);
              if constexpr (rqm_offset(MODE))
                umma_i8(d, make_smem_desc_rel(ca16, 256, 128, 0), make_smem_desc_rel(w16, wt + (F / 16 + 2) * E * 16, E * 16, 128), IDD, 1);
            }
            __syncwarp();
            if constexpr (j == 0) ok = mbar_wait(b + FB_F1, ph, prm.abort_flag);
            if (!ok) return;
            tc_fence_after();
            if (leader) {
#pragma unroll
              for (int k = 0; k < F / 32; ++k)
                umma_i8(d, make_smem_desc_rel(f16, k * 2 * kS * 16, kS * 16, 128), make_smem_desc_rel(w16, wt + k * 2 * E * 16, E * 16, 128), ID, 1);
              umma_commit(b + FB_FULL + r);
            }
          }
          __syncwarp();
        });
      }
    }
  } else {
    // ================================================================= epilogue: one warpgroup per slot
    const int slot = warp >> 2, quad = warp & 3;
    const long long iters = iters_of(slot);
    if (iters > 0) {
      uint64_t *b = slot_bars + slot * kFfnBarsPerSlot;
      const int row = quad * 32 + lane;
      const uint32_t t_lane = tmem_base + slot * C::TMEM_COLS_PER_SLOT + ((uint32_t)(quad * 32) << 16);
      int8_t *const f1_buf = reinterpret_cast<int8_t *>(smem + C::OFF_SLOT + slot * C::SLOT_BYTES + C::OFF_F1);
      bool ok = true;
      uint32_t n = 0;
      for (long long it = 0; it < iters && ok; ++it) {
        const long long f = (long long)blockIdx.x * kSlots + slot + it * frame_stride;
        static_for<0, C::NCH>([&](auto cc) {
          constexpr int c = decltype(cc)::value;
          if (!ok) return;
          const uint32_t r = n & 1u, use = n >> 1;
          ++n;
          ok = mbar_wait(b + FB_FULL + r, use & 1, prm.abort_flag);
          if (!ok) return;
          tc_fence_after();
          const uint32_t taddr = t_lane + r * 64;
          auto release = make_release(
              [&] {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(b + FB_EMPTY + r);
              },
              [] {});
          if constexpr (c < C::NC1) {
            int8_t *g = nullptr;
            if constexpr (DUMP)
              if (prm.f1) g = prm.f1 + (f * kS + row) * F + c * 64;
            tmem_walk<2>([&](int gi) { return taddr + (uint32_t)gi * 32u; },
                         [&](auto gg, const uint32_t(&v)[32]) {
                           constexpr int gi = decltype(gg)::value;
                           requant_store32<MODE, 0, true>(v, prm.rq[0], f1_buf, row, c * 64 + gi * 32, g ? g + gi * 32 : nullptr);
                         },
                         release);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(b + FB_F1);
          } else {
            int8_t *g = prm.y + (f * kS + row) * E + (c - C::NC1) * 64;
            tmem_walk<2>([&](int gi) { return taddr + (uint32_t)gi * 32u; },
                         [&](auto gg, const uint32_t(&v)[32]) {
                           constexpr int gi = decltype(gg)::value;
                           requant_store32<MODE, 0, false>(v, prm.rq[1], nullptr, row, 0, g + gi * 32);
                         },
                         release);
          }
        });
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == C::WARP_LOAD) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

typedef void (*FfnKernelFn)(const CUtensorMap, const FfnParams);
struct FfnShape {
  int E, F, smem_bytes, slots, threads;
  FfnKernelFn (*pick)(int mode, bool dump);
};

#define ITA_FFN_PICK(E, F, M) \
  if (mode == M) return dump ? k_ita_ffn<E, F, M, true> : k_ita_ffn<E, F, M, false>;
#define ITA_FFN_DEFINE_SHAPE(E, F)                                                                               \
  static FfnKernelFn ffn_pick_##E##_##F(int mode, bool dump) {                                                   \
    ITA_FFN_PICK(E, F, RQM_INT_GEN) ITA_FFN_PICK(E, F, RQM_TORCH) ITA_FFN_PICK(E, F, RQM_FP) ITA_FFN_PICK(E, F, RQM_TORCH_FP) \
    ITA_FFN_PICK(E, F, RQM_TORCH_FP2) return nullptr;                                                            \
  }                                                                                                              \
  extern const FfnShape ffn_shape_##E##_##F = {E, F, FfnCfg<E, F>::SMEM_BYTES, FfnCfg<E, F>::SLOTS, FfnCfg<E, F>::THREADS, ffn_pick_##E##_##F};

}  // namespace ita
