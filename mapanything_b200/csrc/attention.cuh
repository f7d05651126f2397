// Flash-style non-causal attention, head_dim 64, on tcgen05 / TMEM / TMA (sm_100a).
//
// Replaces F.scaled_dot_product_attention at transformer_blocks.py:198-201 (frame + global
// attention of the info-sharing transformer, 12 heads) and encoders/dinov2.py:136 (DINOv2,
// 16 heads).  Q, K, V are read in place from the packed qkv projection output
// [tokens, 3*D] (q | k | v, each heads x 64: the layout `reshape(B,N,3,H,hd)` implies), O is
// written as [tokens, D] (= `x.transpose(1,2).reshape(B,N,C)`).
//
// One CTA = one (sequence, head, 128-query tile).  warp0: TMA loader (Q once, K/V rings),
// warp1: MMA issuer, warps 2..5: softmax (thread = query row).
//   S_j = Q K_j^T           SS MMA  128x128x64 -> TMEM S[j&1]
//   P_j = exp2(...)         softmax warps, bf16, written back over S[j&1] (TMEM)
//   O_j = P_j V_j           TS MMA  128x64x128 (A from TMEM, V MN-major from smem) -> TMEM O[j&1]
//   O   = O*alpha_j + O_j   accumulated in registers by the softmax warps (no TMEM read-modify-write)
#pragma once
#include "common.cuh"

namespace mab {

constexpr int ATT_BQ = 128;
constexpr int ATT_BKV = 128;
constexpr int ATT_D = 64;
constexpr int ATT_KV_STAGES = 4;
constexpr int ATT_THREADS = 192;
constexpr int ATT_TILE_BYTES = 128 * 64 * 2;  // 16 KB: Q, K and V tiles alike
constexpr int ATT_SMEM_BYTES = (1 + 2 * ATT_KV_STAGES) * ATT_TILE_BYTES + 512 + 1024;

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

__global__ void __launch_bounds__(ATT_THREADS, 1)
attention_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + ATT_TILE_BYTES;
  uint8_t* sV = sK + ATT_KV_STAGES * ATT_TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + ATT_KV_STAGES * ATT_TILE_BYTES);
  uint64_t* q_full = bars;                         // [1]
  uint64_t* k_full = bars + 1;                     // [KS]
  uint64_t* k_empty = k_full + ATT_KV_STAGES;      // [KS]
  uint64_t* v_full = k_empty + ATT_KV_STAGES;      // [KS]
  uint64_t* v_empty = v_full + ATT_KV_STAGES;      // [KS]
  uint64_t* s_full = v_empty + ATT_KV_STAGES;      // [2]
  uint64_t* p_full = s_full + 2;                   // [2]
  uint64_t* o_full = p_full + 2;                   // [2]
  uint64_t* o_empty = o_full + 2;                  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q_tile = blockIdx.x, head = blockIdx.y, seq = blockIdx.z;
  const int q_row0 = seq * p.q_seq_stride + q_tile * ATT_BQ;
  const int kv_row0 = seq * p.kv_seq_stride;
  const int nkv = (p.kv_len + ATT_BKV - 1) / ATT_BKV;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(q_full, 1);
    for (int s = 0; s < ATT_KV_STAGES; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], 1);
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&p_full[s], 4);
      mbar_init(&o_full[s], 1);
      mbar_init(&o_empty[s], 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t TM_S = tmem_base;         // S0 @ +0, S1 @ +128 ; P_j overlays the first 64 cols
  const uint32_t TM_O = tmem_base + 256;   // O0 @ +256, O1 @ +320

  if (warp == 0) {
    // ===================================== loader =========================================
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, ATT_TILE_BYTES);
      tma_load_2d(sQ, &tmQ, q_full, p.q_col0 + head * ATT_D, q_row0);
      for (int j = 0; j < nkv; ++j) {
        const int st = j % ATT_KV_STAGES;
        const uint32_t ph = (j / ATT_KV_STAGES) & 1;
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
  } else if (warp == 1) {
    // ===================================== MMA issuer =====================================
    constexpr uint32_t IDESC_QK = make_idesc_bf16(ATT_BQ, ATT_BKV, 0, 0);
    constexpr uint32_t IDESC_PV = make_idesc_bf16(ATT_BQ, ATT_D, 0, 1);  // B (=V) is MN-major
    auto issue_qk = [&](int j) {
      const int st = j % ATT_KV_STAGES;
      mbar_wait(&k_full[st], (j / ATT_KV_STAGES) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t a = smem_u32(sQ), b = smem_u32(sK + st * ATT_TILE_BYTES);
#pragma unroll
        for (int k = 0; k < ATT_D / 16; ++k)
          umma_ss(TM_S + (j & 1) * 128, make_sw128_desc(a + k * 32), make_sw128_desc(b + k * 32),
                  IDESC_QK, k != 0);
        umma_commit(&k_empty[st]);
        umma_commit(&s_full[j & 1]);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0);
    issue_qk(0);
    if (nkv > 1) issue_qk(1);
    for (int j = 0; j < nkv; ++j) {
      const int st = j % ATT_KV_STAGES;
      const int b = j & 1;
      const uint32_t use = (j >> 1) & 1;
      mbar_wait(&p_full[b], use);          // P_j is in TMEM
      mbar_wait(&v_full[st], (j / ATT_KV_STAGES) & 1);
      mbar_wait(&o_empty[b], use ^ 1);     // O_{j-2} has been consumed
      tc_fence_after();
      if (lane == 0) {
        const uint32_t vb = smem_u32(sV + st * ATT_TILE_BYTES);
#pragma unroll
        for (int k = 0; k < ATT_BKV / 16; ++k)
          umma_ts(TM_O + b * 64, TM_S + b * 128 + k * 8, make_sw128_desc(vb + k * 2048, 1024, 1024),
                  IDESC_PV, k != 0);
        umma_commit(&v_empty[st]);
        umma_commit(&o_full[b]);
      }
      __syncwarp();
      if (j + 2 < nkv) issue_qk(j + 2);    // overwrites S[b] (and P_j) after PV(j) in issue order
    }
  } else {
    // ===================================== softmax ========================================
    const int q = warp & 3;
    const int row = q * 32 + lane;                      // query row within the tile
    const uint32_t lane_addr = uint32_t(q * 32) << 16;
    float m_run = -INFINITY, l_run = 0.f, alpha_pending = 1.f;
    float o_acc[ATT_D];
#pragma unroll
    for (int d = 0; d < ATT_D; ++d) o_acc[d] = 0.f;
    const bool row_ok = q_tile * ATT_BQ + row < p.q_len;
    if (p.carry_in && row_ok) {
      const long long grow = (long long)q_row0 + row;
      const float2 ml = *reinterpret_cast<const float2*>(p.carry_ml + (grow * p.heads + head) * 2);
      m_run = ml.x; l_run = ml.y;
      const float4* src = reinterpret_cast<const float4*>(p.carry_o + grow * p.ldo + head * ATT_D);
#pragma unroll
      for (int d = 0; d < ATT_D / 4; ++d) {
        const float4 v = src[d];
        o_acc[4 * d] = v.x; o_acc[4 * d + 1] = v.y; o_acc[4 * d + 2] = v.z; o_acc[4 * d + 3] = v.w;
      }
    }

    auto consume_o = [&](int j, float alpha) {
      const int b = j & 1;
      mbar_wait(&o_full[b], (j >> 1) & 1);
      tc_fence_after();
      uint32_t v0[32], v1[32];
      MAB_TMEM_LD_32(TM_O + lane_addr + b * 64, v0);
      MAB_TMEM_LD_32(TM_O + lane_addr + b * 64 + 32, v1);
      tmem_wait_ld();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&o_empty[b]);
#pragma unroll
      for (int d = 0; d < 32; ++d) {
        o_acc[d] = fmaf(o_acc[d], alpha, __uint_as_float(v0[d]));
        o_acc[d + 32] = fmaf(o_acc[d + 32], alpha, __uint_as_float(v1[d]));
      }
    };

    for (int j = 0; j < nkv; ++j) {
      const int b = j & 1;
      const uint32_t ts = TM_S + lane_addr + b * 128;
      const int valid = p.kv_len - j * ATT_BKV;  // columns >= valid are padding / next sequence
      mbar_wait(&s_full[b], (j >> 1) & 1);
      tc_fence_after();
      // pass 1: row max
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        MAB_TMEM_LD_32(ts + c * 32, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (c * 32 + i < valid) mx = fmaxf(mx, __uint_as_float(v[i]));
      }
      const float m_new = fmaxf(m_run, mx);  // finite: every tile has >= 1 valid column
      const float alpha = fast_exp2((m_run - m_new) * p.scale_log2e);  // 0 on the first tile
      const float mb = m_new * p.scale_log2e;
      float sum = 0.f;
      // pass 2: P = exp2(S*c - m*c) -> bf16 -> TMEM (overlaying already-consumed S columns)
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32], pk[16];
        MAB_TMEM_LD_32(ts + c * 32, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float e0 = fast_exp2(fmaf(__uint_as_float(v[i]), p.scale_log2e, -mb));
          float e1 = fast_exp2(fmaf(__uint_as_float(v[i + 1]), p.scale_log2e, -mb));
          if (c * 32 + i >= valid) e0 = 0.f;
          if (c * 32 + i + 1 >= valid) e1 = 0.f;
          const uint32_t pr = pack_bf16(e0, e1);
          // the row sum uses the bf16-rounded probabilities that the PV MMA actually sees
          sum += bf16_lo(pr) + bf16_hi(pr);
          pk[i >> 1] = pr;
        }
        MAB_TMEM_ST_16(ts + c * 16, pk);
      }
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[b]);
      l_run = l_run * alpha + sum;
      m_run = m_new;
      // fold the previous tile's PV product while the tensor core works on this one
      if (j > 0) consume_o(j - 1, alpha_pending);
      alpha_pending = alpha;
    }
    consume_o(nkv - 1, alpha_pending);

    if (row_ok && p.carry_out) {
      const long long grow = (long long)q_row0 + row;
      *reinterpret_cast<float2*>(p.carry_ml + (grow * p.heads + head) * 2) = make_float2(m_run, l_run);
      float4* dst = reinterpret_cast<float4*>(p.carry_o + grow * p.ldo + head * ATT_D);
#pragma unroll
      for (int d = 0; d < ATT_D / 4; ++d)
        dst[d] = make_float4(o_acc[4 * d], o_acc[4 * d + 1], o_acc[4 * d + 2], o_acc[4 * d + 3]);
    } else if (row_ok) {
      const float inv = 1.0f / l_run;
      const long long grow = (long long)q_row0 + row;
      uint4* dst = reinterpret_cast<uint4*>(p.out + grow * p.ldo + head * ATT_D);
#pragma unroll
      for (int d = 0; d < ATT_D; d += 8)
        dst[d >> 3] = make_uint4(pack_bf16(o_acc[d] * inv, o_acc[d + 1] * inv),
                                 pack_bf16(o_acc[d + 2] * inv, o_acc[d + 3] * inv),
                                 pack_bf16(o_acc[d + 4] * inv, o_acc[d + 5] * inv),
                                 pack_bf16(o_acc[d + 6] * inv, o_acc[d + 7] * inv));
      if (p.lse)
        p.lse[grow * p.heads + head] = m_run * p.scale_log2e * 0.6931471805599453f + logf(l_run);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace mab
