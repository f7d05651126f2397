// Flash-style non-causal attention, head_dim 64, on tcgen05 / TMEM / TMA (sm_100a).
//
// Replaces F.scaled_dot_product_attention at transformer_blocks.py:198-201 (frame + global
// attention of the info-sharing transformer, 12 heads) and encoders/dinov2.py:136 (DINOv2,
// 16 heads).  Q, K, V are read in place from the packed qkv projection output
// [tokens, 3*D] (q | k | v, each heads x 64: the layout `reshape(B,N,3,H,hd)` implies), O is
// written as [tokens, D] (= `x.transpose(1,2).reshape(B,N,C)`).
//
// One CTA = one (sequence, head, 256-query tile) = two 128-query tiles A and B that ping-pong on
// the tensor core while sharing every K/V tile in shared memory:
//   warp0      TMA loader (Q_A, Q_B once; 4-stage K and V rings)
//   warp1/2    MMA issuer of tile A / tile B (one elected thread each; a tcgen05.mma that waits for
//              room in the tensor-core queue stalls the scheduler partition it is issued from, so
//              the two tiles' issue streams sit on different partitions), warp3 idle
//   warps 4-7  softmax of tile A, warps 8-11 softmax of tile B (thread = one query row)
//   S_t(j) = Q_t K_j^T      SS MMA 128x128x64 -> TMEM S_t                       (t = A, B)
//   P_t(j) = exp2(...)      softmax warps: S row read ONCE into registers, bf16 P -> TMEM P_t
//   O_t   += P_t(j) V_j     TS MMA 128x64x128 (A operand from TMEM, V MN-major from smem),
//                           accumulated in TMEM over the whole K/V loop
// S is loaded from TMEM once
// (128 registers per row) and O never leaves TMEM inside the loop.  The running max is only
// advanced when a row's new maximum exceeds the reference by more than 2^8 (exp2 domain), so the
// O rescale (TMEM load-scale-store by the row's own thread) is rare; P <= 256 stays exact enough
// in bf16 and the final O / l is mathematically unchanged.  S_t is released to the tensor core as
// soon as it is in registers (P_t has its own TMEM columns), so Q K^T of the next K/V tile runs
// under the exp pass of the current one.
#pragma once
#include <type_traits>

#include "common.cuh"

namespace mab {

constexpr int ATT_BQ = 128;    // rows of one Q tile (one UMMA M)
constexpr int ATT_QT = 2;      // Q tiles per CTA (long sequences); short sequences use 1, two CTAs per SM
constexpr int ATT_BKV = 128;
constexpr int ATT_D = 64;
constexpr int ATT_KV_STAGES = 4;
#ifndef ATT_TURN_LATE
#define ATT_TURN_LATE 1   // 1: the softmax turn barrier sits after the row max (0: before it)
#endif
// Which of the 16 element pairs of every 32 S columns take their exponentials from a polynomial on
// the FMA pipe instead of the MUFU (bit k = pair k).  Round 1 (one thread per row, exp-pass turns) was bound by
// the SM's 16 MUFU lanes and gained 20 % from 5 spread pairs (0x1249).  With two threads per row and no turns
// the share no longer matters: none / 2 / 5 of 16 pairs = 817.9 / 818.4 / 815.9 TFLOP/s at 66 305 tokens,
// 441 / 436 / 405 on 8x16x1037 (profiles/r2_attn_variants.txt).  No single unit bounds the loop any more:
// reading every S row from TMEM TWICE costs nothing (822 vs 821), dropping the row-max -> exp dependency and
// the half-row exchange gains 1 %, dropping the exponentials altogether gives 1060 -- the softmax warps run
// a latency-bound chain (ncu: issue slots 66 % busy, no pipe above 50 %).  Default: no polynomial -- every
// exponential is the MUFU's ex2.approx, fewest instructions per element (what counts under the 1000 W cap
// of the many-view regime), one approximation less; the polynomial path stays selectable for experiments.
#ifndef ATT_POLY_MASK
#define ATT_POLY_MASK 0        // long sequences (two Q tiles per CTA)
#endif
#ifndef ATT_POLY_MASK_SHORT
#define ATT_POLY_MASK_SHORT 0  // short sequences (one Q tile per CTA, two CTAs per SM)
#endif
#ifndef ATT_NO_TURNS
#define ATT_NO_TURNS 1   // measured with two threads per row (66 305 tokens): turns 686, no turns 816 TFLOP/s
#endif
// QT = 2: 4 control warps (one per scheduler partition) + 8 softmax warps, 4-stage K/V rings, 512 TMEM columns.
// QT = 1: 2 control warps + 4 softmax warps, 2-stage rings, 256 TMEM columns -> two CTAs per SM whose
//         prologues / epilogues / MMA latencies overlap each other (short sequences: 9 K/V tiles per CTA).
// ATT_SPLIT threads share one query row (each owns 128 / ATT_SPLIT of the S columns of a K/V tile).  With
// one thread per row a scheduler partition holds ONE softmax warp per Q tile, and a single in-order warp
// has to cover its own MUFU / FMA latencies (~400 issued instructions per row and K/V tile at 0.46 IPC).
// Two threads per row put two warps of the same tile on every partition: the halves of a row run
// concurrently and only exchange the row max (through shared memory) and, once per CTA, the row sum.
// Measured (profiles/r2_attn_variants.txt, 66 305 / 8289 tokens): 1 thread per row 716 / 599 TFLOP/s,
// 2 threads with the exp-pass turns 686 / 576, 2 threads without turns 816 / 656.
#ifndef ATT_SPLIT
#define ATT_SPLIT 2
#endif
template <int QT> struct AttCfg {
  static constexpr int CTRL_WARPS = QT == 2 ? 4 : 2;
  static constexpr int SOFTMAX_THREADS = QT * 128 * ATT_SPLIT;
  static constexpr int THREADS = 32 * CTRL_WARPS + SOFTMAX_THREADS;
  static constexpr int STAGES = QT == 2 ? 4 : 2;
  static constexpr int TMEM_COLS = QT == 2 ? 512 : 256;
  static constexpr int XCH_BYTES = 2 * QT * 2 * 128 * 4;  // row-max / row-sum exchange, double-buffered
  static constexpr int SMEM_BYTES = (QT + 2 * STAGES) * 128 * 64 * 2 + 512 + XCH_BYTES + 1024;
};
constexpr int ATT_THREADS = AttCfg<2>::THREADS;
constexpr int ATT_TILE_BYTES = 128 * 64 * 2;  // 16 KB: Q, K and V tiles alike
constexpr int ATT_SMEM_BYTES = AttCfg<2>::SMEM_BYTES;

struct AttnParams {
  int num_seqs, heads;
  int q_len, kv_len;                // per-sequence lengths
  int q_seq_stride, kv_seq_stride;  // rows between consecutive sequences
  int q_col0, k_col0, v_col0;       // column of head 0 inside the Q / K / V tensor maps
  __nv_bfloat16* out;               // [rows, ldo]
  int ldo;
  float scale_log2e;                // softmax scale * log2(e)
  float* lse;                       // optional [rows, heads] natural-log sum-exp or null
  // Ring (context-parallel) attention: the online-softmax state of every query row is carried
  // between the per-block launches.  carry_o [rows, heads*64] fp32 un-normalised O, carry_ml
  // [rows, heads, 2] fp32 (running max of the raw scores, running sum).
  float* carry_o;
  float* carry_ml;
  int carry_in;                     // 1: start from the carried state instead of (-inf, 0, 0)
  int carry_out;                    // 1: store the state, skip the normalised bf16 output
};

// packed fp32x2 helpers (FFMA2 / FADD2 on sm_100a) and the 3-input max (FMNMX3)
__device__ __forceinline__ uint64_t pack_f2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack_f2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma_f2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t add_f2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ uint32_t cvt_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// LEAN: no carried online-softmax state and no log-sum-exp output (every single-GPU launch of the path):
// the carry prologue / epilogue and their live values are compiled out of that instantiation
template <int QT, bool LEAN = false>
__global__ void __launch_bounds__(AttCfg<QT>::THREADS, QT == 2 ? 1 : 2)
attention_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, AttnParams p_in) {
  AttnParams p = p_in;
  if constexpr (LEAN) { p.carry_in = 0; p.carry_out = 0; p.lse = nullptr; p.carry_o = nullptr; p.carry_ml = nullptr; }
  constexpr int KS = AttCfg<QT>::STAGES;
  constexpr int POLY_MASK = QT == 2 ? ATT_POLY_MASK : ATT_POLY_MASK_SHORT;
  constexpr int CTRL = AttCfg<QT>::CTRL_WARPS;
  constexpr int SPLIT = ATT_SPLIT;           // threads per query row
  constexpr int NC = ATT_BKV / SPLIT;        // S columns of one K/V tile per thread
  constexpr int OC = ATT_D / SPLIT;          // O columns per thread
  constexpr int TILE_THREADS = ATT_BQ * SPLIT;
  static_assert(SPLIT == 1 || SPLIT == 2, "one or two threads per query row");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sQ = smem;                                   // [QT] tiles
  uint8_t* sK = smem + QT * ATT_TILE_BYTES;         // [KS] tiles
  uint8_t* sV = sK + KS * ATT_TILE_BYTES;    // [KS] tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + KS * ATT_TILE_BYTES);
  uint64_t* q_full = bars;                         // [1]
  uint64_t* k_full = bars + 1;                     // [KS]
  uint64_t* k_empty = k_full + KS;      // [KS]
  uint64_t* v_full = k_empty + KS;      // [KS]
  uint64_t* v_empty = v_full + KS;      // [KS]
  uint64_t* s_full = v_empty + KS;      // [QT]  S_t(j) is in TMEM
  uint64_t* s_free = s_full + QT;              // [QT]  S_t(j) is in registers: S_t may be overwritten
  uint64_t* p_full = s_free + QT;              // [QT]  P_t(j) is in TMEM (and O_t rescaled)
  uint64_t* pv_done = p_full + QT;             // [QT]  O_t += P_t(j) V_j has completed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + QT);
  float* xch = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);  // [2][QT][2][128]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q_tile = blockIdx.x, head = blockIdx.y, seq = blockIdx.z;
  const int q_row0 = seq * p.q_seq_stride + q_tile * (QT * ATT_BQ);
  const int kv_row0 = seq * p.kv_seq_stride;
  const int nkv = (p.kv_len + ATT_BKV - 1) / ATT_BKV;
  // the second Q tile of the last CTA of a sequence may hold no query at all: skip it entirely
  const int ntile = (QT == 2 && q_tile * (QT * ATT_BQ) + ATT_BQ < p.q_len) ? 2 : 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(q_full, 1);
    for (int s = 0; s < KS; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], ntile);  // one commit per tile's MMA issuer
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], ntile);
    }
    for (int t = 0; t < QT; ++t) {
      mbar_init(&s_full[t], 1);
      mbar_init(&s_free[t], 4 * SPLIT);  // one arrive per softmax warp of the tile
      mbar_init(&p_full[t], 4 * SPLIT);
      mbar_init(&pv_done[t], 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, AttCfg<QT>::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();     // prologue above overlapped the previous kernel's tail (PDL)
  pdl_trigger();
  // broadcast through a shuffle so that the compiler knows the TMEM base is warp-uniform: otherwise every
  // tcgen05.mma is wrapped in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop (~100 cycles per instruction)
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  // TMEM columns (QT = 2): S_A @0, S_B @128, O_A @256, O_B @320, P_A @384, P_B @448 (bf16 pairs: 64
  // columns); (QT = 1): S @0, O @128, P @192
  auto TM_S = [&](int t) { return tmem_base + t * 128; };
  auto TM_O = [&](int t) { return tmem_base + QT * 128 + t * 64; };
  auto TM_P = [&](int t) { return tmem_base + QT * 192 + t * 64; };

  if (warp == 0) {
    // ===================================== loader =========================================
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, ntile * ATT_TILE_BYTES);
      for (int t = 0; t < ntile; ++t)
        tma_load_2d(sQ + t * ATT_TILE_BYTES, &tmQ, q_full, p.q_col0 + head * ATT_D, q_row0 + t * ATT_BQ);
      for (int j = 0; j < nkv; ++j) {
        const int st = j % KS;
        const uint32_t ph = (j / KS) & 1;
        mbar_wait(&k_empty[st], ph ^ 1);
        mbar_arrive_expect_tx(&k_full[st], ATT_TILE_BYTES);
        tma_load_2d(sK + st * ATT_TILE_BYTES, &tmK, &k_full[st], p.k_col0 + head * ATT_D,
                    kv_row0 + j * ATT_BKV);
        mbar_wait(&v_empty[st], ph ^ 1);
        mbar_arrive_expect_tx(&v_full[st], ATT_TILE_BYTES);
        tma_load_2d(sV + st * ATT_TILE_BYTES, &tmV, &v_full[st], p.v_col0 + head * ATT_D,
                    kv_row0 + j * ATT_BKV);
      }
    }
  } else if (warp >= 1 && warp <= QT) {
    // ============================== MMA issuer of tile (warp - 1) ==============================
    const int t = warp - 1;
    constexpr uint32_t IDESC_QK = make_idesc_bf16(ATT_BQ, ATT_BKV, 0, 0);
    constexpr uint32_t IDESC_PV = make_idesc_bf16(ATT_BQ, ATT_D, 0, 1);  // B (=V) is MN-major
    auto issue_qk = [&](int j) {  // S_t(j) = Q_t K_j^T
      if (elect_one()) {
        const int st = j % KS;
        const uint32_t a = smem_u32(sQ + t * ATT_TILE_BYTES), b = smem_u32(sK + st * ATT_TILE_BYTES);
#pragma unroll
        for (int k = 0; k < ATT_D / 16; ++k)
          umma_ss(TM_S(t), make_sw128_desc(a + k * 32), make_sw128_desc(b + k * 32), IDESC_QK, k != 0);
        umma_commit(&s_full[t]);
        umma_commit(&k_empty[st]);
      }
      __syncwarp();
    };
    if (t < ntile) {
      mbar_wait(q_full, 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      issue_qk(0);
      for (int j = 0; j < nkv; ++j) {
        const int st = j % KS;
        if (j + 1 < nkv) {
          mbar_wait(&k_full[(j + 1) % KS], ((j + 1) / KS) & 1);
          mbar_wait(&s_free[t], j & 1);  // S_t(j) has been read: the next Q K^T may overwrite it
          tc_fence_after();
          issue_qk(j + 1);
        }
        mbar_wait(&v_full[st], (j / KS) & 1);
        mbar_wait(&p_full[t], j & 1);    // P_t(j) is in TMEM, O_t carries the right scale
        tc_fence_after();
        if (elect_one()) {
          const uint32_t vb = smem_u32(sV + st * ATT_TILE_BYTES);
#pragma unroll
          for (int k = 0; k < ATT_BKV / 16; ++k)
            umma_ts(TM_O(t), TM_P(t) + k * 8, make_sw128_desc(vb + k * 2048, 1024, 1024), IDESC_PV,
                    (j > 0 || p.carry_in || k != 0) ? 1u : 0u);
          umma_commit(&pv_done[t]);
          umma_commit(&v_empty[st]);
        }
        __syncwarp();
      }
    }
  } else if (warp >= CTRL) {
    // ===================================== softmax ========================================
    const int sw = warp - CTRL;                         // softmax warp index: (tile, column half, quarter)
    const int t = sw / (4 * SPLIT);                     // Q tile of this warp
    const int hf = SPLIT == 2 ? (sw >> 2) & 1 : 0;      // which half of the S / O columns this thread owns
    const int q = warp & 3;                             // TMEM lane quarter (warp id % 4)
    const int row = q * 32 + lane;                      // query row within the tile
    const uint32_t lane_addr = uint32_t(q * 32) << 16;
    const bool row_ok = q_tile * (QT * ATT_BQ) + t * ATT_BQ + row < p.q_len;
    const long long grow = (long long)q_row0 + t * ATT_BQ + row;
    // row-max exchange slot of (buffer, tile, half, row); the two warps that share 32 rows meet on
    // named barrier 3 + 4 t + q (64 threads)
    auto xslot = [&](int buf, int h) { return xch + ((buf * QT + t) * 2 + h) * 128 + row; };
    const int pair_bar = 3 + 4 * t + q;
    // A warp whose 32 rows all lie beyond the sequence (1037 = 8 * 128 + 13: three of the four warps
    // of every sequence's last tile) only keeps the barrier protocol going: no TMEM traffic, no
    // exponentials, so the MUFU of the SM goes to the co-resident CTA.  Short-sequence kernel only:
    // the turn barriers of the two-tile kernel count every softmax thread, and the mere presence of
    // this branch costs its hot loop 17 % (840 -> 700 TFLOP/s measured).
    const bool idle_warp = QT == 1 && q_tile * ATT_BQ + q * 32 >= p.q_len;
    if (t < ntile && idle_warp) {
      for (int j = 0; j < nkv; ++j) {
        mbar_wait(&s_full[t], j & 1);  // one arrive per phase: never ahead of the tensor core
        if (lane == 0) mbar_arrive(&s_free[t]);
        if (j > 0) mbar_wait(&pv_done[t], (j - 1) & 1);
        if (lane == 0) mbar_arrive(&p_full[t]);
      }
    } else if (t < ntile) {
      const uint32_t ts = TM_S(t) + lane_addr + hf * NC, to = TM_O(t) + lane_addr + hf * OC,
                     tp = TM_P(t) + lane_addr + hf * (NC / 2);
      const float sc = p.scale_log2e;
      const uint64_t sc2 = pack_f2(sc, sc);
      float m_ref = -INFINITY, l_run = 0.f;  // reference max of the exponentials, running sum
      if (p.carry_in) {  // resume from the carried state: O_t <- carry_o
        uint32_t o[OC];
        if (row_ok) {
          const float4* src = reinterpret_cast<const float4*>(p.carry_o + grow * p.ldo + head * ATT_D + hf * OC);
#pragma unroll
          for (int d = 0; d < OC / 4; ++d) {
            const float4 v = src[d];
            o[4 * d] = __float_as_uint(v.x); o[4 * d + 1] = __float_as_uint(v.y);
            o[4 * d + 2] = __float_as_uint(v.z); o[4 * d + 3] = __float_as_uint(v.w);
          }
        } else {
#pragma unroll
          for (int d = 0; d < OC; ++d) o[d] = 0u;
        }
#pragma unroll
        for (int c = 0; c < OC; c += 16) MAB_TMEM_ST_16(to + c, (o + c));
        if (row_ok) {
          const float2 ml = *reinterpret_cast<const float2*>(p.carry_ml + (grow * p.heads + head) * 2);
          m_ref = ml.x; l_run = hf == 0 ? ml.y : 0.f;  // the halves' sums are added at the end
        }
        tmem_wait_st();
      }

      // The exp pass is bound by the SM's 16 MUFU lanes: the two tiles take turns on it (named
      // barriers 1 = "A's turn", 2 = "B's turn") so that one tile's TMEM traffic, barrier waits and
      // MMA latency always sit under the other tile's exp pass.
      const bool take_turns = ntile == 2 && !ATT_NO_TURNS;
      if (take_turns && t == 1) bar_arrive_named(1, 2 * TILE_THREADS);  // tile A goes first
      // One K/V tile of the softmax.  Full tiles and the partial last tile of a sequence are two
      // separate instantiations, so that the hot (full) body stays one straight-line block.
      auto kv_tile = [&](const int j, auto full_tag) {
        constexpr bool full = decltype(full_tag)::value;
        // columns >= valid (of this thread's NC) are padding / the next sequence
        const int valid = p.kv_len - j * ATT_BKV - hf * NC;
        mbar_wait(&s_full[t], j & 1);
        tc_fence_after();
        // ---- the whole S row into registers, then hand S_t back to the tensor core ----
        uint32_t v[NC];
#pragma unroll
        for (int c = 0; c < NC; c += 32) MAB_TMEM_LD_32(ts + c, (v + c));
        tmem_wait_ld();
#ifdef ATT_LD_TWICE   // experiment: the same S row read from TMEM a second time (is the loop bound by the TMEM read port?)
#pragma unroll
        for (int c = 0; c < NC; c += 32) MAB_TMEM_LD_32(ts + c, (v + c));
        tmem_wait_ld();
#endif
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_free[t]);
#if !ATT_TURN_LATE
        if (take_turns) bar_sync_named(1 + t, 2 * TILE_THREADS);
#endif
        if (!full || (QT == 2 && valid < NC)) {  // partial last tile: columns >= valid become -inf (exp2 -> 0), common path unmasked
#pragma unroll
          for (int i = 0; i < NC; ++i)
            if (i >= valid) v[i] = 0xff800000u;
        }
        // ---- row max (four independent chains) ----
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int i = 0; i < NC; i += 8) {
          mx0 = max3(mx0, __uint_as_float(v[i]), __uint_as_float(v[i + 1]));
          mx1 = max3(mx1, __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
          mx2 = max3(mx2, __uint_as_float(v[i + 4]), __uint_as_float(v[i + 5]));
          mx3 = max3(mx3, __uint_as_float(v[i + 6]), __uint_as_float(v[i + 7]));
        }
        float mx = fmaxf(max3(mx0, mx1, mx2), mx3);
        if (SPLIT == 2) {
          // the other half of the row: through shared memory, double-buffered by the parity of j (the
          // slot of tile j + 2 is only written after the partner has passed the barrier of tile j + 1)
          *xslot(j & 1, hf) = mx;
          bar_sync_named(pair_bar, 64);
          mx = fmaxf(mx, *xslot(j & 1, hf ^ 1));
        }
        // mx is finite: every tile has >= 1 valid column (in one of the halves)
        // advance the reference max only when it is exceeded by more than 2^8 in the exp2 domain
        const bool grow_ref = (mx - m_ref) * sc > 8.0f;  // true on the first tile (m_ref = -inf)
        const float m_new = grow_ref ? mx : m_ref;
        const float alpha = grow_ref ? fast_exp2((m_ref - m_new) * sc) : 1.0f;  // exp2(-inf) = 0
        const bool rescale = __any_sync(0xffffffffu, grow_ref) && (j > 0 || p.carry_in);
        float nmb = -m_new * sc;
#if ATT_TURN_LATE
        // the turn covers the exp pass only: the row max above (ALU pipe) runs under the other tile's
        // exponentials, so the MUFU never waits for it.  nmb feeds every exponential: passing it
        // through the barrier keeps the exp pass behind it and the max in front of it.
        if (take_turns) bar_sync_named_dep(1 + t, 2 * TILE_THREADS, nmb);
#endif
        const uint64_t nmb2 = pack_f2(nmb, nmb);
        // ---- P = exp2(S*c - m*c) -> bf16 (registers), row sum ----
        // the pairs selected by ATT_POLY_MASK are evaluated on the FMA/ALU pipes
        // (round-to-nearest split x = n + f, |f| <= 0.5, degree-3 minimax 2^f, exponent add: relative
        // error 7.5e-5, far below the bf16 rounding of P) instead of the 16-lane MUFU, which is the
        // unit that bounds this loop.
        uint32_t pk[NC / 2];
        uint64_t sum_a = pack_f2(0.f, 0.f), sum_b = sum_a;
        const uint64_t magic2 = pack_f2(12582912.f, 12582912.f), nmagic2 = pack_f2(-12582912.f, -12582912.f);
        const uint64_t neg1_2 = pack_f2(-1.f, -1.f);
        const uint64_t pc3 = pack_f2(0.0551716685f, 0.0551716685f), pc2 = pack_f2(0.2426111251f, 0.2426111251f);
        const uint64_t pc1 = pack_f2(0.6932609677f, 0.6932609677f), pc0 = pack_f2(0.9999280572f, 0.9999280572f);
        auto exp2_poly2 = [&](float xa, float xb, float& ea, float& eb) {
          const uint64_t x2 = pack_f2(fmaxf(xa, -125.f), fmaxf(xb, -125.f));
          const uint64_t r2 = add_f2(x2, magic2);          // low mantissa bits = round(x)
          const uint64_t f2 = fma_f2(add_f2(r2, nmagic2), neg1_2, x2);
          uint64_t p2 = fma_f2(pc3, f2, pc2);
          p2 = fma_f2(p2, f2, pc1);
          p2 = fma_f2(p2, f2, pc0);
          float ra, rb, pa, pb;
          unpack_f2(r2, ra, rb);
          unpack_f2(p2, pa, pb);
          ea = __int_as_float(__float_as_int(pa) + (__float_as_int(ra) << 23));
          eb = __int_as_float(__float_as_int(pb) + (__float_as_int(rb) << 23));
        };
        if constexpr (full) {
#pragma unroll
          for (int i = 0; i < NC; i += 2) {
            float x0, x1, e0, e1;
            unpack_f2(fma_f2(pack_f2(__uint_as_float(v[i]), __uint_as_float(v[i + 1])), sc2, nmb2), x0, x1);
#ifdef ATT_NO_EXP   // experiment (wrong results): no exponential at all -- is the loop bound by the softmax arithmetic?
            e0 = x0; e1 = x1;
#else
            if ((POLY_MASK >> ((i >> 1) & 15)) & 1) {
              exp2_poly2(x0, x1, e0, e1);
            } else {
              e0 = fast_exp2(x0); e1 = fast_exp2(x1);
            }
#endif
            if (i & 2) sum_b = add_f2(sum_b, pack_f2(e0, e1));
            else sum_a = add_f2(sum_a, pack_f2(e0, e1));
            pk[i >> 1] = cvt_bf16x2(e0, e1);
          }
        } else {
          // partial last K/V tile (1037 = 8 * 128 + 13): only the 32-column chunks that hold a valid
          // column are evaluated (masked columns are -inf -> 0), the others are P = 0 outright
#pragma unroll
          for (int c = 0; c < NC; c += 32) {
            if (c < valid) {
#pragma unroll
              for (int i = c; i < c + 32; i += 2) {
                float x0, x1;
                unpack_f2(fma_f2(pack_f2(__uint_as_float(v[i]), __uint_as_float(v[i + 1])), sc2, nmb2), x0, x1);
                const float e0 = fast_exp2(x0), e1 = fast_exp2(x1);
                sum_a = add_f2(sum_a, pack_f2(e0, e1));
                pk[i >> 1] = cvt_bf16x2(e0, e1);
              }
            } else {
#pragma unroll
              for (int i = c; i < c + 32; i += 2) pk[i >> 1] = 0u;
            }
          }
        }
        float s0, s1;
        unpack_f2(add_f2(sum_a, sum_b), s0, s1);
        const float row_sum = s0 + s1;  // depends on every exponential of the pass
#if ATT_TURN_LATE
        if (take_turns && !(t == 1 && j == nkv - 1)) bar_arrive_named_dep(1 + (t ^ 1), 2 * TILE_THREADS, row_sum);
#else
        if (take_turns && !(t == 1 && j == nkv - 1)) bar_arrive_named(1 + (t ^ 1), 2 * TILE_THREADS);
#endif
        // ---- P_t and O_t are free once P_t(j-1) V has completed ----
#ifndef ATT_NO_PV_WAIT   // experiment (race, wrong results): is the wait for P V(j-1) on the critical path?
        if (j > 0) {
          mbar_wait(&pv_done[t], (j - 1) & 1);
          tc_fence_after();
        }
#endif
        if (rescale) {  // warp-uniform; rows that keep their reference multiply by 1
          const uint64_t a2 = pack_f2(alpha, alpha);
#pragma unroll
          for (int h = 0; h < OC / 32; ++h) {
            uint32_t o[32];
            MAB_TMEM_LD_32(to + h * 32, o);
            tmem_wait_ld();
#pragma unroll
            for (int d = 0; d < 32; d += 2) {
              float y0, y1;
              unpack_f2(fma_f2(pack_f2(__uint_as_float(o[d]), __uint_as_float(o[d + 1])), a2, pack_f2(0.f, 0.f)),
                        y0, y1);
              o[d] = __float_as_uint(y0); o[d + 1] = __float_as_uint(y1);
            }
            MAB_TMEM_ST_16(to + h * 32, o);
            MAB_TMEM_ST_16(to + h * 32 + 16, (o + 16));
          }
        }
#pragma unroll
        for (int c = 0; c < NC / 2; c += 16) MAB_TMEM_ST_16(tp + c, (pk + c));
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[t]);
        l_run = fmaf(l_run, alpha, row_sum);
        m_ref = m_new;
      };
      if constexpr (QT == 2) {
        // long sequences: one instantiation, the (single) partial tile is masked at run time and goes
        // through the common pass -- a second copy of the body costs the hot loop its schedule
        // (measured 836 -> 740 TFLOP/s at 66 305 tokens)
        for (int j = 0; j < nkv; ++j) kv_tile(j, std::true_type{});
      } else {
        // short sequences (1037 tokens = 9 tiles): the partial tile is 1/9 of the work
        const int nfull = p.kv_len / ATT_BKV;
        for (int j = 0; j < nfull; ++j) kv_tile(j, std::true_type{});
        if (nfull < nkv) kv_tile(nfull, std::false_type{});
      }

      // ---- epilogue: O_t out of TMEM once the last P V has completed ----
      if (SPLIT == 2) {  // the row sum of the two halves (m_ref is the same in both by construction)
        *xslot(nkv & 1, hf) = l_run;
        bar_sync_named(pair_bar, 64);
        l_run += *xslot(nkv & 1, hf ^ 1);
      }
      mbar_wait(&pv_done[t], (nkv - 1) & 1);
      tc_fence_after();
      uint32_t o[OC];
#pragma unroll
      for (int c = 0; c < OC; c += 32) MAB_TMEM_LD_32(to + c, (o + c));
      tmem_wait_ld();
      if (row_ok && p.carry_out) {
        if (hf == 0)
          *reinterpret_cast<float2*>(p.carry_ml + (grow * p.heads + head) * 2) = make_float2(m_ref, l_run);
        float4* dst = reinterpret_cast<float4*>(p.carry_o + grow * p.ldo + head * ATT_D + hf * OC);
#pragma unroll
        for (int d = 0; d < OC / 4; ++d)
          dst[d] = make_float4(__uint_as_float(o[4 * d]), __uint_as_float(o[4 * d + 1]),
                               __uint_as_float(o[4 * d + 2]), __uint_as_float(o[4 * d + 3]));
      } else if (row_ok) {
        const float inv = 1.0f / l_run;
        uint4* dst = reinterpret_cast<uint4*>(p.out + grow * p.ldo + head * ATT_D + hf * OC);
#pragma unroll
        for (int d = 0; d < OC; d += 8)
          dst[d >> 3] = make_uint4(
              pack_bf16(__uint_as_float(o[d]) * inv, __uint_as_float(o[d + 1]) * inv),
              pack_bf16(__uint_as_float(o[d + 2]) * inv, __uint_as_float(o[d + 3]) * inv),
              pack_bf16(__uint_as_float(o[d + 4]) * inv, __uint_as_float(o[d + 5]) * inv),
              pack_bf16(__uint_as_float(o[d + 6]) * inv, __uint_as_float(o[d + 7]) * inv));
        if (p.lse && hf == 0)
          p.lse[grow * p.heads + head] = m_ref * p.scale_log2e * 0.6931471805599453f + logf(l_run);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, AttCfg<QT>::TMEM_COLS);
  }
}

}  // namespace mab
