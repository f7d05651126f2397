// tcgen05 / TMEM / TMA GEMM family of libmab200 (sm_100a).
//
//   C[M,N] = A[M,K] * B[N,K]^T     bf16 operands (K-major, 128B-swizzled smem tiles), fp32 TMEM accum
//
// Two A-operand producers share one kernel:
//   MODE_GEMM : A is a row-major [M,K] bf16 matrix (every nn.Linear and 1x1 conv of the path:
//               transformer_blocks.py:75-78,146-154, dinov2 layers/attention.py:48-50, mlp.py:29-31)
//   MODE_CONV3: implicit-GEMM 3x3 / stride 1 / pad 1 convolution over an NHWC bf16 activation
//               (dpt_block.py:34-69,128-146, dpt.py:271-277).  A-tile = 8x16 output pixels; for
//               each of the 9 taps the TMA box is shifted by (dy-1, dx-1) and out-of-image
//               coordinates are zero-filled by the TMA unit, so no im2col buffer ever exists.
//
// Persistent, warp-specialised: warp0 = TMA producer, warp1 = MMA issuer (+TMEM owner),
// warps 2..9 = epilogue.  smem ring of STAGES {A 128x64, B BNx64}; TMEM holds two BN-column fp32
// accumulators so the epilogue of tile i overlaps the main loop of tile i+1.
#pragma once
#include "common.cuh"

namespace mab {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_EPI_WARPS = 8;   // two per TMEM lane quarter, each owns half of the tile's columns
constexpr int GEMM_THREADS = 64 + 32 * GEMM_EPI_WARPS;
constexpr int CONV_TH = 8;   // conv tile: 8 rows x 16 cols of output pixels = 128 GEMM rows
constexpr int CONV_TW = 16;

enum { MODE_GEMM = 0, MODE_CONV3 = 1 };
enum { ACT_NONE = 0, ACT_GELU = 1, ACT_RELU = 2 };
enum { EPI_GENERIC = 0, EPI_REGRESS = 1,
       EPI_RESID = 2,    // in-place fp32 residual update only (proj / fc2 of every transformer block)
       EPI_PLAIN = 3 };  // bias (+ GELU) -> bf16 only (qkv / fc1 of every transformer block)
// EPI_RESID / EPI_PLAIN carry exactly the epilogue their layers need: 135 / ~150 registers and no spills,
// where the generic kernel (every operand combination of the dense head and the tests behind a switch)
// sits on the 168-register cap of a 10-warp CTA with 140-420 B of spills

struct GemmEpilogue {
  const float* bias;             // [N] or null
  const float* gamma;            // [N] LayerScale or null
  const float* res_f32;          // fp32 residual, same indexing as out (may alias out_f32)
  const __nv_bfloat16* res1;     // bf16 residuals (RCU skip / fusion add), same indexing as out
  const __nv_bfloat16* res2;
  float* out_f32;                // any subset of the three outputs
  __nv_bfloat16* out_bf16;
  __nv_bfloat16* out_relu;       // relu(value) (input of the next ResidualConvUnit)
  int ldo;                       // output row pitch in elements
  int act;                       // ACT_*
  int split_n;                   // > 0: out_bf16 row = [hi | lo | hi] with hi at col, lo at col +
                                 // split_n, hi again at col + 2*split_n (bf16x3 operand of the next
                                 // GEMM: x ~ hi + lo, |x - hi - lo| <= 2^-17 |x|)
  int act_post;                  // 1: the activation is applied AFTER the residual adds
                                 // (ResidualBlock: act(conv2(.) + shortcut), dense_rep_encoder.py:31-52)
  const float* res2_f32;         // second fp32 residual and fp32 relu(value) copy: the fp32 / TF32 dense
  float* out_relu_f32;           // head (the reference's autocast-off region, model.py:1774)
  int round_tf32;                // 1: out_f32 / out_relu_f32 are stored rounded to the nearest TF32 value
                                 // (they are the A operands of the next kind::tf32 contraction, tf32_rn)
};

// Fused tail of the dense head: ReLU -> 1x1 conv 128->6 -> adaptor -> pointmap (EPI_REGRESS)
struct RegressEpilogue {
  const float* bias2;        // [128]   dpt_regressor_head.conv2.0.bias
  const float* w3;           // [6,128] dpt_regressor_head.conv2.2.weight
  const float* b3;           // [6]
  const float* pose;         // [views,8]: t(3) q(4, unit) pad  (pose head output after adaptor)
  const float* scale;        // [1] metric scale
  float* pts3d;              // [views,H,W,3]
  float* pts3d_cam;          // [views,H,W,3]
  float* rays;               // [views,H,W,3]
  float* depth;              // [views,H,W]   (depth_along_ray, x scale)
  float* conf;               // [views,H,W]
  float* mask_logits;        // [views,H,W]
  unsigned char* mask;       // [views,H,W]   non_ambiguous_mask (0/1)
  float* logits6;            // optional [views,H,W,6] raw regressor output (tests) or null
};

struct GemmShape {
  int M, N, K;          // MODE_GEMM: A is [M,K].  MODE_CONV3: K = 9*Cpad, M unused
  int batch, H, W;      // MODE_CONV3: activation is [batch,H,W,C]
  int cblocks;          // MODE_CONV3: Cpad/64
  int creal;            // MODE_CONV3: un-padded input channels (algorithmic FLOP accounting)
};

__device__ __forceinline__ float gelu_erf(float x) {
  // exact-erf GELU (nn.GELU default, transformer_blocks.py:68) as x * sigmoid(2u) with
  // erf(x/sqrt2) ~ tanh(u), u = x (a0 + a1 x^2 + a2 x^4) on |x| <= 5 (minimax fit: |gelu error|
  // <= 2.6e-5 absolute, below the bf16 rounding of the stored activation); the sigmoid form has no
  // cancellation in the negative tail.  2 MUFU + 7 FP32 ops per element, branch-free (rcp.approx:
  // the IEEE rcp.rn compiles to a guarded slow-path call that serialises the whole chunk).
#ifdef MAB_GELU_AS
  const float z = fabsf(x) * 0.70710678118654752f;
  const float t = __frcp_rn(fmaf(0.3275911f, z, 1.0f));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  const float e_ = 1.0f - p * t * __expf(-z * z);
  return 0.5f * x * (1.0f + copysignf(e_, x));
#endif
#ifdef MAB_GELU_TANH
  {
    // Experiment, NOT the default: the fit below through ONE MUFU op instead of two, x * sigmoid(2u) =
    // 0.5 x (1 + tanh(u)).  3 us faster per fc1 GEMM (profiles/r2_gemm_variants.txt: 62.9 -> 59.8 us,
    // 44.7 -> 41.5 us) but tanh.approx.f32 is only 2^-11 RELATIVE to tanh, and 1 + tanh(u) cancels in the
    // negative tail: up to 7e-4 absolute on activations that should be ~1e-3 -- as much noise as the bf16
    // rounding of the whole hidden activation, for 0.7 % of the step.  Parity first.
    const float xt = fminf(fmaxf(x, -5.0f), 5.0f);
    const float xx = xt * xt;
    float pl = fmaf(xx, -3.51902e-04f, 3.70080e-02f);
    pl = fmaf(pl, xx, 7.97505e-01f);
    float th;
    asm("tanh.approx.f32 %0, %1;" : "=f"(th) : "f"(xt * pl));
    const float hx = 0.5f * x;
    return fmaf(hx, th, hx);
  }
#endif
  const float xc = fminf(fmaxf(x, -5.0f), 5.0f);
  const float x2 = xc * xc;
  // -2 * log2(e) * (a0 + a1 x^2 + a2 x^4)
  float pnl = fmaf(x2, 1.01537568e-03f, -1.06782574e-01f);
  pnl = fmaf(pnl, x2, -2.30111381e+00f);
  const float e = fast_exp2(xc * pnl);
  return x * fast_rcp(1.0f + e);
}

template <int BN, int STAGES>
struct GemmSmem {
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  static constexpr int B_BYTES = (BN / 2) * GEMM_BK * 2;  // this CTA's half of the pair's B tile
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // per epilogue warp: 32 rows x 32 floats, float4 columns XOR-swizzled by (row & 7): both the
  // row-per-lane writes and the 4-rows-per-instruction reads are bank-conflict free
  static constexpr int EPI_BYTES = GEMM_EPI_WARPS * 32 * 32 * 4;
  static constexpr int BAR_BYTES = 256;
  static constexpr int TOTAL = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024;  // +align slack
};

// ---- store half of the generic epilogue, specialised on the operands that exist -----------------
enum {
  EF_OUTBF16 = 1, EF_OUTF32 = 2, EF_OUTRELU = 4, EF_RESF32 = 8, EF_RES1 = 16, EF_RES2 = 32,
  EF_GAMMA = 64, EF_GELU = 128, EF_RELU = 256, EF_RES2F = 512, EF_OUTRELUF = 1024,
  EF_GENERIC = 1 << 20  // generic: test pointers at run time
};
__device__ __forceinline__ int epilogue_flags(const GemmEpilogue& ep) {
  if (ep.split_n || ep.act_post) return EF_GENERIC;
  return (ep.out_bf16 ? EF_OUTBF16 : 0) | (ep.out_f32 ? EF_OUTF32 : 0) | (ep.out_relu ? EF_OUTRELU : 0) |
         (ep.res_f32 ? EF_RESF32 : 0) | (ep.res1 ? EF_RES1 : 0) | (ep.res2 ? EF_RES2 : 0) |
         (ep.gamma ? EF_GAMMA : 0) | (ep.act == ACT_GELU ? EF_GELU : 0) | (ep.act == ACT_RELU ? EF_RELU : 0) |
         (ep.res2_f32 ? EF_RES2F : 0) | (ep.out_relu_f32 ? EF_OUTRELUF : 0);
}

template <int F, bool RED = false>
__device__ __forceinline__ void epilogue_chunks(const GemmEpilogue& ep, int N, uint32_t t_row, float* stg,
                                                int lane, int sub_r, int sub_c4, const long long (&ro)[8],
                                                uint32_t row_mask, int c_begin, int c_end, int n0) {
  constexpr bool G = (F & EF_GENERIC) != 0;
  const bool has_of = G ? ep.out_f32 != nullptr : (F & EF_OUTF32) != 0;
  const bool has_ob = G ? ep.out_bf16 != nullptr : (F & EF_OUTBF16) != 0;
  const bool has_or = G ? ep.out_relu != nullptr : (F & EF_OUTRELU) != 0;
  const bool has_rf = G ? ep.res_f32 != nullptr : (F & EF_RESF32) != 0;
  const bool has_r1 = G ? ep.res1 != nullptr : (F & EF_RES1) != 0;
  const bool has_r2 = G ? ep.res2 != nullptr : (F & EF_RES2) != 0;
  const bool has_g = G ? ep.gamma != nullptr : (F & EF_GAMMA) != 0;
  const bool has_r2f = G ? ep.res2_f32 != nullptr : (F & EF_RES2F) != 0;
  const bool has_orf = G ? ep.out_relu_f32 != nullptr : (F & EF_OUTRELUF) != 0;
  const bool gelu = G ? ep.act == ACT_GELU : (F & EF_GELU) != 0;
  const bool relu = G ? ep.act == ACT_RELU : (F & EF_RELU) != 0;
  const uint64_t keep = l2_policy_evict_last();
  // In-place fp32 residual (x += gamma * (A B^T + bias): out_f32 == res_f32 and nothing else written -- the
  // proj and fc2 GEMMs of every transformer block): the add is performed by the L2 with a fire-and-forget
  // vector reduction (red.global.add.v4.f32), so the epilogue never waits for a residual load.  One fp32 add
  // per element either way (not an accumulation: deterministic).  The path lives in its OWN kernel
  // instantiation (EPI_RESID): inside the generic kernel its mere presence cost every other epilogue
  // registers (in the step: fc1 67 -> 74 us, qkv 43.7 -> 45.4 us, while proj went 33.3 -> 27.5 us).
  const bool red_inplace = RED;
#pragma unroll 1
  for (int c0 = c_begin; c0 < c_end; c0 += 32) {
    const int ncol0 = n0 + c0;
    if (ncol0 >= N) break;
    const int col = ncol0 + sub_c4 * 4;
    const uint32_t mask = col < N ? row_mask : 0u;  // N is a multiple of 4 for every layer of the path
    // every global load of the chunk is issued first: they do not depend on the accumulator, so the TMEM
    // load and the shared-memory transpose below run under their latency (one or two warps per scheduler
    // have to cover their own latencies)
    float4 b4 = make_float4(0, 0, 0, 0), g4 = make_float4(1, 1, 1, 1);
    if (mask && ep.bias) b4 = __ldg(reinterpret_cast<const float4*>(ep.bias + col));
    if (mask && has_g) g4 = __ldg(reinterpret_cast<const float4*>(ep.gamma + col));
    float4 a[8], rf[8];
    uint2 r1[8], r2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const bool ok = (mask >> i) & 1;
      const long long off = ro[i] + col;
      if (has_rf && !red_inplace) rf[i] = ok ? ld_f4_l2hint(ep.res_f32 + off, keep) : make_float4(0, 0, 0, 0);
      if (has_r1) r1[i] = ok ? *reinterpret_cast<const uint2*>(ep.res1 + off) : make_uint2(0, 0);
      if (has_r2) r2[i] = ok ? *reinterpret_cast<const uint2*>(ep.res2 + off) : make_uint2(0, 0);
    }
    uint32_t v[32];
    MAB_TMEM_LD_32(t_row + c0, v);
    tmem_wait_ld();
    __syncwarp();  // previous chunk's readers are done with stg
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<float4*>(stg + lane * 32 + ((j ^ (lane & 7)) << 2)) =
          make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                      __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = 4 * i + sub_r;
      a[i] = *reinterpret_cast<const float4*>(stg + r * 32 + ((sub_c4 ^ (r & 7)) << 2));
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (!((mask >> i) & 1)) continue;
      float4 x = a[i];
      x.x += b4.x; x.y += b4.y; x.z += b4.z; x.w += b4.w;
      const bool post = G && ep.act_post;
      if (gelu && !post) {
        x.x = gelu_erf(x.x); x.y = gelu_erf(x.y); x.z = gelu_erf(x.z); x.w = gelu_erf(x.w);
      } else if (relu && !post) {
        x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f);
      }
      if (has_g) { x.x *= g4.x; x.y *= g4.y; x.z *= g4.z; x.w *= g4.w; }
      if (red_inplace) {
        red_add_f4(ep.out_f32 + ro[i] + col, x);
        continue;
      }
      if (has_rf) { x.x += rf[i].x; x.y += rf[i].y; x.z += rf[i].z; x.w += rf[i].w; }
      if (has_r1) {
        x.x += bf16_lo(r1[i].x); x.y += bf16_hi(r1[i].x); x.z += bf16_lo(r1[i].y); x.w += bf16_hi(r1[i].y);
      }
      if (has_r2) {
        x.x += bf16_lo(r2[i].x); x.y += bf16_hi(r2[i].x); x.z += bf16_lo(r2[i].y); x.w += bf16_hi(r2[i].y);
      }
      if (has_r2f) {  // loaded here, not with the prefetched operands: 32 more live registers otherwise
        const float4 e = *reinterpret_cast<const float4*>(ep.res2_f32 + ro[i] + col);
        x.x += e.x; x.y += e.y; x.z += e.z; x.w += e.w;
      }
      if (post && gelu) {
        x.x = gelu_erf(x.x); x.y = gelu_erf(x.y); x.z = gelu_erf(x.z); x.w = gelu_erf(x.w);
      } else if (post && relu) {
        x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f);
      }
      const long long off = ro[i] + col;
      if (ep.round_tf32) x = tf32_rn(x);
      if (has_of) {
        if (has_rf) st_f4_l2hint(ep.out_f32 + off, x, keep);  // the residual stream stays in L2
        else *reinterpret_cast<float4*>(ep.out_f32 + off) = x;
      }
      if (has_ob) {
        const uint2 hi = make_uint2(pack_bf16(x.x, x.y), pack_bf16(x.z, x.w));
        *reinterpret_cast<uint2*>(ep.out_bf16 + off) = hi;
        if (G && ep.split_n) {
          const uint2 lo = make_uint2(pack_bf16(x.x - bf16_lo(hi.x), x.y - bf16_hi(hi.x)),
                                      pack_bf16(x.z - bf16_lo(hi.y), x.w - bf16_hi(hi.y)));
          *reinterpret_cast<uint2*>(ep.out_bf16 + off + ep.split_n) = lo;
          *reinterpret_cast<uint2*>(ep.out_bf16 + off + 2 * ep.split_n) = hi;
        }
      }
      if (has_or)
        *reinterpret_cast<uint2*>(ep.out_relu + off) =
            make_uint2(pack_bf16(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f)),
                       pack_bf16(fmaxf(x.z, 0.f), fmaxf(x.w, 0.f)));
      if (has_orf)
        *reinterpret_cast<float4*>(ep.out_relu_f32 + off) =
            make_float4(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f), fmaxf(x.z, 0.f), fmaxf(x.w, 0.f));
    }
  }
}

// CTA pairs (cta_group::2): the two CTAs of a cluster (two SMs of one TPC) compute ONE 256 x BN
// tile.  Each CTA stages its own 128 A rows and its own BN/2 B rows; the leader CTA issues
// tcgen05.mma.cta_group::2 (UMMA M = 256) which reads both CTAs' shared memory and writes both
// CTAs' TMEM (128 lanes each).  Per CTA and k-block that is 16 + BN/4 KB of TMA traffic and of
// operand reads instead of 16 + BN/2 KB: a single-CTA 128x256 tile is bound by shared-memory
// bandwidth (TMA writes + UMMA operand reads ~192 B/clk against 128 B/clk), not by the tensor core
// (measured: 44 % of peak in the main loop, independent of L2 traffic).
// TF = 1: fp32 operands consumed as TF32 (kind::tf32) for the reference's fp32/TF32 regions
// (geometric-input encoders, model.py:1377): a k-block is 32 fp32 = the same 128-byte swizzled rows,
// so the pipeline, descriptors and byte counts are those of the bf16 kernel.
template <int BN, int STAGES, int MODE, int EPI, int CL, int TF = 0>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            GemmShape shp, GemmEpilogue ep, RegressEpilogue rg) {
  using SM = GemmSmem<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* epi_smem = smem + STAGES * SM::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + SM::EPI_BYTES);
  uint64_t* full_bar = bars;                 // [STAGES]
  uint64_t* empty_bar = bars + STAGES;       // [STAGES]
  uint64_t* tfull_bar = bars + 2 * STAGES;   // [2]
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // ---- tile bookkeeping -------------------------------------------------------------------
  int tiles_x = 0, tiles_y = 0, num_m;
  if (MODE == MODE_CONV3) {
    tiles_x = (shp.W + CONV_TW - 1) / CONV_TW;
    tiles_y = (shp.H + CONV_TH - 1) / CONV_TH;
    num_m = shp.batch * tiles_x * tiles_y;
  } else {
    num_m = (shp.M + GEMM_BM - 1) / GEMM_BM;
  }
  const int num_n = (shp.N + BN - 1) / BN;
  const int crank = CL > 1 ? (int)cluster_ctarank() : 0;
  const int num_mg = (num_m + CL - 1) / CL;   // groups of CL adjacent M tiles
  const int num_tiles = num_mg * num_n;       // per cluster
  const int cluster_id = blockIdx.x / CL, num_clusters = gridDim.x / CL;
  // tile index -> (m_blk, n_blk); m_blk may be >= num_m for the last group (loads zero-fill,
  // stores are predicated off) so that every CTA of a cluster runs the same number of tiles
  auto tile_m = [&](int tile) { return (tile % num_mg) * CL + crank; };
  auto tile_n = [&](int tile) { return tile / num_mg; };
  constexpr int BKE = TF ? GEMM_BK / 2 : GEMM_BK;  // elements per 128-byte k-block row
  const int num_kb = (MODE == MODE_CONV3) ? 9 * shp.cblocks : (shp.K + BKE - 1) / BKE;
  constexpr uint32_t TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128
                                 : (2 * BN <= 256) ? 256 : 512;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);   // leader only: its producer arms it with BOTH CTAs' bytes
      mbar_init(&empty_bar[s], 1);  // the leader's MMA commit arrives in both CTAs
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);                     // the leader's MMA commit arrives in both CTAs
      mbar_init(&tempty_bar[s], 2 * GEMM_EPI_WARPS);   // leader only: epilogue warps of both CTAs
    }
    fence_barrier_init();
  }
  static_assert(CL == 2, "the kernel is written for CTA pairs");
  if (warp == 1) tmem_alloc2(tmem_slot, TMEM_COLS);
  tc_fence_before();
  cluster_sync_all();  // peers' barriers are initialised and TMEM is allocated in both CTAs
  tc_fence_after();
  // broadcast through a shuffle so that the compiler knows the TMEM base is warp-uniform: otherwise every
  // tcgen05.mma is wrapped in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop (~100 cycles per instruction)
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  // everything above overlapped the tail of the previous kernel in the stream (PDL)
  pdl_wait();
  pdl_trigger();

  if (warp == 0) {
    // ================================ TMA producer ==========================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int m_blk = tile_m(tile), n_blk = tile_n(tile);
        int cn = 0, cy = 0, cx = 0;
        if (MODE == MODE_CONV3) {
          cx = (m_blk % tiles_x) * CONV_TW;
          cy = ((m_blk / tiles_x) % tiles_y) * CONV_TH;
          cn = m_blk / (tiles_x * tiles_y);
        }
        for (int kb = 0; kb < num_kb; ++kb) {
#ifdef MAB_GEMM_NOLOAD
          continue;  // experiment: MMA rate without any operand traffic
#endif
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * SM::STAGE_BYTES;
          uint8_t* sb = sa + SM::A_BYTES;
          // both CTAs' loads complete on the LEADER's full barrier, armed once with all the bytes
          if (crank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * SM::STAGE_BYTES);
          const uint32_t fb = mapa_u32(smem_u32(&full_bar[stage]), 0);
          if (MODE == MODE_CONV3) {
            const int tap = kb / shp.cblocks, cb = kb - tap * shp.cblocks;
            const int dy = tap / 3 - 1, dx = tap % 3 - 1;
            tma_load_4d_cg2(sa, &tmA, fb, cb * BKE, cx + dx, cy + dy, cn);
          } else {
            tma_load_2d_cg2(sa, &tmA, fb, kb * BKE, m_blk * GEMM_BM);
          }
          tma_load_2d_cg2(sb, &tmB, fb, kb * BKE, n_blk * BN + crank * (BN / 2));
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1 && crank == 0) {
    // ================================ MMA issuer (leader CTA) ================================
    constexpr uint32_t IDESC = TF ? make_idesc_tf32(2 * GEMM_BM, BN) : make_idesc_bf16(2 * GEMM_BM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++it) {
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1);  // epilogue has drained this accumulator
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
#if !defined(MAB_GEMM_NOLOAD) && !defined(MAB_GEMM_NOWAIT)
        mbar_wait(&full_bar[stage], phase);
#endif
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * SM::STAGE_BYTES);
          const uint32_t sb = sa + SM::A_BYTES;
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k) {
            // +32 bytes per UMMA_K step (16 bf16 or 8 tf32) inside the 128B swizzle atom
            if (TF)
              umma_ss2_tf32(d_tmem, make_sw128_desc(sa + k * 32), make_sw128_desc(sb + k * 32), IDESC,
                            (kb | k) != 0);
            else
              umma_ss2(d_tmem, make_sw128_desc(sa + k * 32), make_sw128_desc(sb + k * 32), IDESC,
                       (kb | k) != 0);
          }
          umma_commit2(&empty_bar[stage]);                      // frees the smem slot in both CTAs
          if (kb == num_kb - 1) umma_commit2(&tfull_bar[acc]);  // accumulator complete, both CTAs
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 2) {
    // ================================ epilogue (warps 2..9) ==================================
    const int q = warp & 3;             // TMEM lane quarter this warp may access (warp id % 4)
    const int chalf = (warp - 2) >> 2;  // which half of the tile's columns this warp drains
    float* stg = reinterpret_cast<float*>(epi_smem) + (warp - 2) * 32 * 32;
    const int epi_flags = epilogue_flags(ep);
    int it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++it) {
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      const int m_blk = tile_m(tile), n_blk = tile_n(tile);
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16) + acc * BN;

      int cn = 0, cy = 0, cx = 0;
      if (MODE == MODE_CONV3) {
        cx = (m_blk % tiles_x) * CONV_TW;
        cy = ((m_blk / tiles_x) % tiles_y) * CONV_TH;
        cn = m_blk / (tiles_x * tiles_y);
      }
      // row (0..127 within the tile) -> element offset of the output row, or -1 if out of range
      auto row_offset = [&](int r) -> long long {
        if (MODE == MODE_CONV3) {
          const int y = cy + r / CONV_TW, x = cx + r % CONV_TW;
          if (y >= shp.H || x >= shp.W || cn >= shp.batch) return -1;
          return ((long long)(cn * shp.H + y) * shp.W + x);
        } else {
          const int m = m_blk * GEMM_BM + r;
          return m < shp.M ? (long long)m : -1;
        }
      };

      if (EPI == EPI_REGRESS) {
        // thread = one output pixel holding all BN(=128) channels of conv2.0 (column-half 1 warps idle)
        if (chalf == 1) {
          if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty_bar[acc]), 0));
          continue;
        }
        float lg[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) lg[k] = __ldg(rg.b3 + k);
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
          uint32_t v[32];
          MAB_TMEM_LD_32(t_row + c0, v);
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float a = fmaxf(__uint_as_float(v[j]) + __ldg(rg.bias2 + c0 + j), 0.f);
#pragma unroll
            for (int k = 0; k < 6; ++k) lg[k] = fmaf(a, __ldg(rg.w3 + k * BN + c0 + j), lg[k]);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty_bar[acc]), 0));
        const long long pix = row_offset(q * 32 + lane);
        if (pix >= 0) {
          // adaptors.py:432-466 (unit-sphere rays), :256-280 (exp depth), :1044-1073 (1+exp conf),
          // :1126-1131 (sigmoid mask) ; geometry.py:855-907 ; model.py:1907-1923
          const float* ps = rg.pose + cn * 8;
          const float s = __ldg(rg.scale);
          const float nrm = fmaxf(sqrtf(lg[0] * lg[0] + lg[1] * lg[1] + lg[2] * lg[2]), 1e-8f);
          const float rx = lg[0] / nrm, ry = lg[1] / nrm, rz = lg[2] / nrm;
          const float d = expf(lg[3]);
          const float cf = 1.0f + expf(lg[4]);
          const float qx = ps[3], qy = ps[4], qz = ps[5], qw = ps[6];
          const float px = d * rx, py = d * ry, pz = d * rz;
          const float r00 = 1 - 2 * (qy * qy + qz * qz), r01 = 2 * (qx * qy - qw * qz),
                      r02 = 2 * (qx * qz + qw * qy);
          const float r10 = 2 * (qx * qy + qw * qz), r11 = 1 - 2 * (qx * qx + qz * qz),
                      r12 = 2 * (qy * qz - qw * qx);
          const float r20 = 2 * (qx * qz - qw * qy), r21 = 2 * (qy * qz + qw * qx),
                      r22 = 1 - 2 * (qx * qx + qy * qy);
          const float wx = r00 * px + r01 * py + r02 * pz + ps[0];
          const float wy = r10 * px + r11 * py + r12 * pz + ps[1];
          const float wz = r20 * px + r21 * py + r22 * pz + ps[2];
          rg.pts3d[pix * 3 + 0] = wx * s; rg.pts3d[pix * 3 + 1] = wy * s; rg.pts3d[pix * 3 + 2] = wz * s;
          rg.pts3d_cam[pix * 3 + 0] = px * s; rg.pts3d_cam[pix * 3 + 1] = py * s; rg.pts3d_cam[pix * 3 + 2] = pz * s;
          rg.rays[pix * 3 + 0] = rx; rg.rays[pix * 3 + 1] = ry; rg.rays[pix * 3 + 2] = rz;
          rg.depth[pix] = d * s;
          rg.conf[pix] = cf;
          rg.mask_logits[pix] = lg[5];
          // sigmoid(x) > 0.5 evaluated exactly as the reference does in fp32
          rg.mask[pix] = (1.0f / (1.0f + expf(-lg[5])) > 0.5f) ? 1 : 0;
          if (rg.logits6) {
#pragma unroll
            for (int k = 0; k < 6; ++k) rg.logits6[pix * 6 + k] = lg[k];
          }
        }
        continue;
      }

      // ---- generic epilogue: TMEM -> regs -> swizzled smem transpose -> coalesced global ------
      // pass 1 thread = row `lane` of this warp's 32-row slab, 32 columns per chunk
      // pass 2 thread = (row 4*i + lane/8, columns 4*(lane%8)..+3): 8 lanes cover 128 contiguous B
      // The store half is specialised at compile time on which optional operands exist (EF_*): a
      // warp's chunk is ~100 instructions instead of ~450 of null-pointer predication.
      const int sub_r = lane >> 3, sub_c4 = lane & 7;
      long long ro[8];
      uint32_t row_mask = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        ro[i] = row_offset(q * 32 + 4 * i + sub_r);
        if (ro[i] >= 0) row_mask |= 1u << i;
        ro[i] = (ro[i] < 0 ? 0 : ro[i]) * ep.ldo;
      }
      const int c_begin = chalf * (BN / 2), c_end = (chalf + 1) * (BN / 2);
      const int n0 = n_blk * BN;
#define MAB_EPI_CASE(F) \
  case F: epilogue_chunks<F>(ep, shp.N, t_row, stg, lane, sub_r, sub_c4, ro, row_mask, c_begin, c_end, n0); break;
      if constexpr (EPI == EPI_PLAIN) {  // the launcher guarantees: bf16 output only, bias, optional GELU
        if (ep.act == ACT_GELU)
          epilogue_chunks<EF_OUTBF16 | EF_GELU>(ep, shp.N, t_row, stg, lane, sub_r, sub_c4, ro,
                                                                     row_mask, c_begin, c_end, n0);
        else
          epilogue_chunks<EF_OUTBF16>(ep, shp.N, t_row, stg, lane, sub_r, sub_c4, ro, row_mask,
                                                           c_begin, c_end, n0);
      } else if constexpr (EPI == EPI_RESID) {  // the launcher guarantees out_f32 == res_f32, fp32 only, no activation
        if (ep.gamma)
          epilogue_chunks<EF_OUTF32 | EF_RESF32 | EF_GAMMA, true>(ep, shp.N, t_row, stg, lane, sub_r,
                                                                                sub_c4, ro, row_mask, c_begin, c_end, n0);
        else
          epilogue_chunks<EF_OUTF32 | EF_RESF32, true>(ep, shp.N, t_row, stg, lane, sub_r, sub_c4, ro,
                                                                     row_mask, c_begin, c_end, n0);
      } else if constexpr (TF) {  // fp32 operands: geometric-input encoders and the fp32 / TF32 dense head
        switch (epi_flags) {
          MAB_EPI_CASE(EF_OUTF32)
          MAB_EPI_CASE(EF_OUTF32 | EF_RELU)
          MAB_EPI_CASE(EF_OUTF32 | EF_GELU)
          MAB_EPI_CASE(EF_OUTF32 | EF_OUTRELUF)
          MAB_EPI_CASE(EF_OUTF32 | EF_RESF32)
          MAB_EPI_CASE(EF_OUTF32 | EF_RESF32 | EF_RES2F | EF_OUTRELUF)
          default:
            epilogue_chunks<EF_GENERIC>(ep, shp.N, t_row, stg, lane, sub_r, sub_c4, ro, row_mask, c_begin, c_end, n0);
        }
      } else {
        switch (epi_flags) {
          MAB_EPI_CASE(EF_OUTBF16)
          MAB_EPI_CASE(EF_OUTBF16 | EF_GELU)
          MAB_EPI_CASE(EF_OUTBF16 | EF_RELU)
          MAB_EPI_CASE(EF_OUTF32)
          MAB_EPI_CASE(EF_OUTF32 | EF_RESF32)
          MAB_EPI_CASE(EF_OUTF32 | EF_RESF32 | EF_GAMMA)
          MAB_EPI_CASE(EF_OUTBF16 | EF_OUTRELU)
          MAB_EPI_CASE(EF_OUTBF16 | EF_RES1)
          MAB_EPI_CASE(EF_OUTBF16 | EF_RES1 | EF_OUTRELU)
          MAB_EPI_CASE(EF_OUTBF16 | EF_RES1 | EF_RES2 | EF_OUTRELU)
          default:
            epilogue_chunks<EF_GENERIC>(ep, shp.N, t_row, stg, lane, sub_r, sub_c4, ro, row_mask, c_begin, c_end, n0);
        }
      }
#undef MAB_EPI_CASE
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty_bar[acc]), 0));
    }
  }

  tc_fence_before();
  // no CTA may exit while a peer can still multicast into its smem or arrive on its barriers
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, TMEM_COLS);
  }
}

}  // namespace mab
