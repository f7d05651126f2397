// Host-side helpers: error handling, TMA tensor-map encoding, kernel launchers.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "attention.cuh"
#include "elementwise.cuh"
#include "gemm.cuh"
#include "postprocess.cuh"
#include "export.cuh"
#include "image.cuh"

namespace mab {

inline char* err_buf() {
  static thread_local char buf[512] = "";
  return buf;
}
inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}
#define MAB_CUDA(expr)                                                                    \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess)                                                                \
      return ::mab::fail((int)_e, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                         __FILE__, __LINE__);                                             \
  } while (0)
#define MAB_TRY(expr)         \
  do {                        \
    int _r = (expr);          \
    if (_r != 0) return _r;   \
  } while (0)
#define MAB_ARG(cond, msg) \
  do {                     \
    if (!(cond)) return ::mab::fail(-1, "argument error: %s (%s)", msg, #cond); \
  } while (0)

// ---- launch accounting + optional per-launch CUDA-event profiler (bench.py's roofline leg) -------
enum { CAT_GEMM = 0, CAT_CONV3 = 1, CAT_CONV3_REGRESS = 2, CAT_ATTENTION = 3, CAT_LAYERNORM = 4,
       CAT_ELEMENTWISE = 5, CAT_POSTPROCESS = 6, CAT_GEMM_TF32 = 7, CAT_CONV3_TF32 = 8, CAT_COUNT = 9 };
struct ProfRec { int cat; double flops, bytes; cudaEvent_t e0, e1; };
struct ProfState {
  bool on = false;
  long long launches = 0;             // every kernel launch of the library, profiling or not
  std::vector<ProfRec> recs;
  std::vector<cudaEvent_t> pool;
  size_t used = 0;
  cudaEvent_t get() {
    if (used == pool.size()) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      pool.push_back(e);
    }
    return pool[used++];
  }
};
inline ProfState& prof() {
  static ProfState p;
  return p;
}
// RAII around ONE kernel launch: counts it; when profiling, brackets it with events on its stream.
struct ProfScope {
  cudaStream_t st;
  bool rec;
  ProfRec r;
  ProfScope(int cat, double flops, double bytes, cudaStream_t s) : st(s) {
    ProfState& p = prof();
    ++p.launches;
    rec = p.on;
    if (rec) {
      r.cat = cat; r.flops = flops; r.bytes = bytes;
      r.e0 = p.get(); r.e1 = p.get();
      cudaEventRecord(r.e0, st);
    }
  }
  ~ProfScope() {
    if (rec) {
      cudaEventRecord(r.e1, st);
      prof().recs.push_back(r);
    }
  }
};
#define MAB_LAUNCH(cat, flops, bytes, st, ...)       \
  do {                                               \
    ::mab::ProfScope _ps(cat, flops, bytes, st);     \
    __VA_ARGS__;                                     \
  } while (0)

// Launch with the programmatic-stream-serialization attribute (PDL): only for kernels that call
// pdl_wait() before their first global-memory access.  MAB_NO_PDL=1 in the environment turns the
// attribute off (plain stream order) for debugging.
inline bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("MAB_NO_PDL");
    on = (e && e[0] == '1') ? 0 : 1;
  }
  return on == 1;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              int cluster, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (cluster > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

// ---- cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda link dependency)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 row-major [rows, cols] (row pitch ld elements), box = [box_rows, 64 cols], 128B swizzle
inline int make_tmap_2d(CUtensorMap* m, const void* ptr, long long rows, long long cols,
                        long long ld, int box_rows, bool f32 = false) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return fail(-2, "cuTensorMapEncodeTiled entry point not available");
  const int es = f32 ? 4 : 2;  // fp32 operands (TF32 GEMM): 32 elements per 128-byte box row
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || ((ld * es) & 15))
    return fail(-1, "TMA operand must be 16B aligned with a 16B-multiple row pitch (ld=%lld)", ld);
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * es};
  cuuint32_t box[2] = {(cuuint32_t)(128 / es), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(ptr), dims, strides, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(-3, "cuTensorMapEncodeTiled(2d) failed: %d", (int)r);
  return 0;
}

// bf16 NHWC [batch,H,W,C], box = [1, CONV_TH, CONV_TW, 64], 128B swizzle
inline int make_tmap_nhwc(CUtensorMap* m, const void* ptr, int batch, int H, int W, int C,
                          bool f32 = false) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return fail(-2, "cuTensorMapEncodeTiled entry point not available");
  const cuuint64_t es = f32 ? 4 : 2;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || ((C * es) & 15))
    return fail(-1, "NHWC TMA operand must be 16B aligned with a 16B-multiple pixel pitch (C=%d)", C);
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)batch};
  cuuint64_t strides[3] = {(cuuint64_t)C * es, (cuuint64_t)W * C * es, (cuuint64_t)H * W * C * es};
  cuuint32_t box[4] = {(cuuint32_t)(128 / es), CONV_TW, CONV_TH, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                  const_cast<void*>(ptr), dims, strides, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(-3, "cuTensorMapEncodeTiled(4d) failed: %d", (int)r);
  return 0;
}

// Per-device caches: one process may hold handles on several GPUs (function attributes and the SM
// count belong to the current device's context).
constexpr int MAB_MAX_DEVICES = 64;
inline int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev >= 0 && dev < MAB_MAX_DEVICES ? dev : 0;
}
inline int sm_count() {
  static int n[MAB_MAX_DEVICES] = {0};
  const int dev = current_device();
  if (!n[dev]) {
    cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev);
    if (n[dev] <= 0) n[dev] = 148;
  }
  return n[dev];
}
inline int ew_grid(long long total, int block = 256) {
  long long g = (total + block - 1) / block;
  const long long cap = (long long)sm_count() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// ---- GEMM / conv launchers ---------------------------------------------------------------------
constexpr int GEMM_CL = 2;  // CTA pair (cta_group::2): one 256 x BN tile per cluster (gemm.cuh)

template <int BN, int STAGES, int MODE, int EPI, int CL, int TF = 0>
inline int launch_gemm_t(const CUtensorMap& tA, const CUtensorMap& tB, const GemmShape& shp,
                         const GemmEpilogue& ep, const RegressEpilogue& rg, cudaStream_t st) {
  using SM = GemmSmem<BN, STAGES>;
  auto kern = gemm_kernel<BN, STAGES, MODE, EPI, CL, TF>;
  static bool attr_done[MAB_MAX_DEVICES] = {false};
  if (!attr_done[current_device()]) {
    MAB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::TOTAL));
    attr_done[current_device()] = true;
  }
  int num_m;
  if (MODE == MODE_CONV3)
    num_m = shp.batch * ((shp.W + CONV_TW - 1) / CONV_TW) * ((shp.H + CONV_TH - 1) / CONV_TH);
  else
    num_m = (shp.M + GEMM_BM - 1) / GEMM_BM;
  const int groups = ((num_m + CL - 1) / CL) * ((shp.N + BN - 1) / BN);  // tiles per cluster slot
  if (groups <= 0) return 0;
  const int max_clusters = sm_count() / CL;
  const int grid = (groups < max_clusters ? groups : max_clusters) * CL;
  // ALGORITHMIC FLOPs: the un-padded K of the layer (creal for a conv; shp.creal > 0 for a linear layer whose
  // K is padded (patch embed 588 -> 640) or tripled (bf16x3 pose head: three bf16 products per fp32 product))
  double flops;
  if (MODE == MODE_CONV3)
    flops = 2.0 * shp.batch * shp.H * shp.W * (double)shp.N * 9.0 * shp.creal;
  else
    flops = 2.0 * shp.M * (double)shp.N * (shp.creal > 0 ? shp.creal : shp.K);
  // kind::tf32 launches have their own categories: half the tensor rate, not comparable with the bf16 peak
  const int cat = MODE == MODE_CONV3 ? (EPI == EPI_REGRESS ? CAT_CONV3_REGRESS : (TF ? CAT_CONV3_TF32 : CAT_CONV3))
                                     : (TF ? CAT_GEMM_TF32 : CAT_GEMM);
  ProfScope ps(cat, flops, 0.0, st);
  MAB_CUDA(launch_pdl(kern, dim3(grid), dim3(GEMM_THREADS), SM::TOTAL, st, CL, tA, tB, shp, ep, rg));
  return 0;
}

// C[M,N] = epilogue(A[M,K] B[N,K]^T); lda/ldb row pitches in elements
inline int launch_gemm(const void* A, long long lda, const void* B, long long ldb, int M, int N,
                       int K, const GemmEpilogue& ep, cudaStream_t st, int k_algorithmic = 0) {
  MAB_ARG(M > 0 && N > 0 && K > 0, "empty GEMM");
  MAB_ARG(N % 4 == 0, "N must be a multiple of 4");
  CUtensorMap tA, tB;
  GemmShape shp{M, N, K, 0, 0, 0, 0, k_algorithmic};
  RegressEpilogue rg{};
  MAB_TRY(make_tmap_2d(&tA, A, M, K, lda, GEMM_BM));
  // x += gamma * (A W^T + b) in place and nothing else: the residual-update kernel (red.global.add epilogue)
  const bool resid = ep.res_f32 && ep.out_f32 == ep.res_f32 && !ep.out_bf16 && !ep.out_relu && !ep.res1 &&
                     !ep.res2 && !ep.res2_f32 && !ep.out_relu_f32 && ep.act == ACT_NONE && !ep.split_n &&
                     !ep.act_post && !ep.round_tf32 && N > 128 && !getenv("MAB_NO_EPI_RED");
  // bias (+ GELU) -> bf16 and nothing else: the plain kernel (qkv, fc1)
  const bool plain = ep.out_bf16 && !ep.out_f32 && !ep.out_relu && !ep.res_f32 && !ep.res1 && !ep.res2 &&
                     !ep.res2_f32 && !ep.out_relu_f32 && !ep.gamma && (ep.act == ACT_NONE || ep.act == ACT_GELU) &&
                     !ep.split_n && !ep.act_post && !ep.round_tf32 && N > 128 && !getenv("MAB_NO_EPI_PLAIN");
  if (N > 128) {
    // tile width by wave quantisation: the persistent grid has sm_count/2 clusters, each 256 x BN
    // tile costs ~BN (x1.12 at BN = 192, measured: the B half-tile amortises less of the A traffic); e.g. the
    // info-sharing width 768 gives 99 tiles of 256 (2 waves on 74 clusters, 67 % full) but 132
    // tiles of 192 (2 waves, 89 % full)
    const int num_mg = ((M + GEMM_BM - 1) / GEMM_BM + GEMM_CL - 1) / GEMM_CL, clusters = sm_count() / GEMM_CL;
    auto cost = [&](int bn, double unit) {
      const int tiles = num_mg * ((N + bn - 1) / bn);
      return ((tiles + clusters - 1) / clusters) * bn * unit;
    };
    if (cost(192, 1.12) < cost(256, 1.0)) {
      MAB_TRY(make_tmap_2d(&tB, B, N, K, ldb, 192 / GEMM_CL));
      if (resid) return launch_gemm_t<192, 6, MODE_GEMM, EPI_RESID, GEMM_CL>(tA, tB, shp, ep, rg, st);
      if (plain) return launch_gemm_t<192, 6, MODE_GEMM, EPI_PLAIN, GEMM_CL>(tA, tB, shp, ep, rg, st);
      return launch_gemm_t<192, 6, MODE_GEMM, EPI_GENERIC, GEMM_CL>(tA, tB, shp, ep, rg, st);
    }
    MAB_TRY(make_tmap_2d(&tB, B, N, K, ldb, 256 / GEMM_CL));
    if (resid) return launch_gemm_t<256, 6, MODE_GEMM, EPI_RESID, GEMM_CL>(tA, tB, shp, ep, rg, st);
    if (plain) return launch_gemm_t<256, 6, MODE_GEMM, EPI_PLAIN, GEMM_CL>(tA, tB, shp, ep, rg, st);
    return launch_gemm_t<256, 6, MODE_GEMM, EPI_GENERIC, GEMM_CL>(tA, tB, shp, ep, rg, st);
  }
  MAB_TRY(make_tmap_2d(&tB, B, N, K, ldb, 128 / GEMM_CL));
  return launch_gemm_t<128, 8, MODE_GEMM, EPI_GENERIC, GEMM_CL>(tA, tB, shp, ep, rg, st);
}

// 3x3/s1/p1 conv, x NHWC bf16 [batch,H,W,C], w packed [Cout, 9*Cpad] (Cpad = ceil64(C))
inline int launch_conv3(const void* x, int batch, int H, int W, int C, const void* w, int Cout,
                        const GemmEpilogue& ep, cudaStream_t st) {
  const int cblocks = (C + 63) / 64;
  CUtensorMap tA, tB;
  GemmShape shp{0, Cout, 9 * cblocks * 64, batch, H, W, cblocks, C};
  RegressEpilogue rg{};
  MAB_TRY(make_tmap_nhwc(&tA, x, batch, H, W, C));
  const int m_pairs = (batch * ((W + CONV_TW - 1) / CONV_TW) * ((H + CONV_TH - 1) / CONV_TH) + GEMM_CL - 1) / GEMM_CL;
  const bool narrow = m_pairs * ((Cout + 255) / 256) * 2 <= sm_count() / GEMM_CL;  // see launch_conv3_tf32
  if (Cout > 128 && !narrow) {
    MAB_TRY(make_tmap_2d(&tB, w, Cout, 9LL * cblocks * 64, 9LL * cblocks * 64, 256 / GEMM_CL));
    return launch_gemm_t<256, 6, MODE_CONV3, EPI_GENERIC, GEMM_CL>(tA, tB, shp, ep, rg, st);
  }
  MAB_TRY(make_tmap_2d(&tB, w, Cout, 9LL * cblocks * 64, 9LL * cblocks * 64, 128 / GEMM_CL));
  return launch_gemm_t<128, 8, MODE_CONV3, EPI_GENERIC, GEMM_CL>(tA, tB, shp, ep, rg, st);
}

// ---- TF32 variants for the reference's fp32 regions: fp32 operands, same tiles --------------------
// C[M,N] = epilogue(A[M,K] B[N,K]^T), A/B fp32 row-major, K a multiple of 4
inline int launch_gemm_tf32(const float* A, long long lda, const float* B, long long ldb, int M, int N,
                            int K, const GemmEpilogue& ep, cudaStream_t st) {
  MAB_ARG(M > 0 && N > 0 && K > 0 && N % 4 == 0, "bad TF32 GEMM shape");
  CUtensorMap tA, tB;
  GemmShape shp{M, N, K, 0, 0, 0, 0, 0};
  RegressEpilogue rg{};
  MAB_TRY(make_tmap_2d(&tA, A, M, K, lda, GEMM_BM, true));
  if (N <= 128) {
    MAB_TRY(make_tmap_2d(&tB, B, N, K, ldb, 128 / GEMM_CL, true));
    return launch_gemm_t<128, 8, MODE_GEMM, EPI_GENERIC, GEMM_CL, 1>(tA, tB, shp, ep, rg, st);
  }
  MAB_TRY(make_tmap_2d(&tB, B, N, K, ldb, 256 / GEMM_CL, true));
  return launch_gemm_t<256, 6, MODE_GEMM, EPI_GENERIC, GEMM_CL, 1>(tA, tB, shp, ep, rg, st);
}
// 3x3/s1/p1 conv, x NHWC fp32 [batch,H,W,C], w packed fp32 [Cout, 9*Cpad] (Cpad = ceil32(C))
inline int launch_conv3_tf32(const float* x, int batch, int H, int W, int C, int Creal, const float* w,
                             int Cout, const GemmEpilogue& ep, cudaStream_t st) {
  const int cblocks = (C + 31) / 32;
  CUtensorMap tA, tB;
  GemmShape shp{0, Cout, 9 * cblocks * 32, batch, H, W, cblocks, Creal};
  RegressEpilogue rg{};
  MAB_TRY(make_tmap_nhwc(&tA, x, batch, H, W, C, true));
  // small feature maps (28x37 and 14x19 levels of the DPT head at 8 views): 256-wide tiles would occupy
  // fewer than half of the 74 clusters for the whole K loop; 128-wide tiles double the clusters at work
  const int m_pairs = (batch * ((W + CONV_TW - 1) / CONV_TW) * ((H + CONV_TH - 1) / CONV_TH) + GEMM_CL - 1) / GEMM_CL;
  const bool narrow = m_pairs * ((Cout + 255) / 256) * 2 <= sm_count() / GEMM_CL;
  if (Cout <= 128 || narrow) {
    MAB_TRY(make_tmap_2d(&tB, w, Cout, 9LL * cblocks * 32, 9LL * cblocks * 32, 128 / GEMM_CL, true));
    return launch_gemm_t<128, 8, MODE_CONV3, EPI_GENERIC, GEMM_CL, 1>(tA, tB, shp, ep, rg, st);
  }
  MAB_TRY(make_tmap_2d(&tB, w, Cout, 9LL * cblocks * 32, 9LL * cblocks * 32, 256 / GEMM_CL, true));
  return launch_gemm_t<256, 6, MODE_CONV3, EPI_GENERIC, GEMM_CL, 1>(tA, tB, shp, ep, rg, st);
}

// last conv of the dense head with the fused 1x1 + adaptor + pointmap epilogue (C = Cout = 128)
inline int launch_conv3_regress(const void* x, int batch, int H, int W, const void* w,
                                const RegressEpilogue& rg, cudaStream_t st) {
  CUtensorMap tA, tB;
  GemmShape shp{0, 128, 9 * 2 * 64, batch, H, W, 2, 128};
  GemmEpilogue ep{};
  MAB_TRY(make_tmap_nhwc(&tA, x, batch, H, W, 128));
  MAB_TRY(make_tmap_2d(&tB, w, 128, 9LL * 128, 9LL * 128, 128 / GEMM_CL));
  return launch_gemm_t<128, 8, MODE_CONV3, EPI_REGRESS, GEMM_CL>(tA, tB, shp, ep, rg, st);
}

// the same with fp32 activations / weights (kind::tf32): the reference's precision for this conv
inline int launch_conv3_regress_tf32(const float* x, int batch, int H, int W, const float* w,
                                     const RegressEpilogue& rg, cudaStream_t st) {
  CUtensorMap tA, tB;
  GemmShape shp{0, 128, 9 * 4 * 32, batch, H, W, 4, 128};
  GemmEpilogue ep{};
  MAB_TRY(make_tmap_nhwc(&tA, x, batch, H, W, 128, true));
  MAB_TRY(make_tmap_2d(&tB, w, 128, 9LL * 128, 9LL * 128, 128 / GEMM_CL, true));
  return launch_gemm_t<128, 8, MODE_CONV3, EPI_REGRESS, GEMM_CL, 1>(tA, tB, shp, ep, rg, st);
}

// ---- attention launcher ------------------------------------------------------------------------
// K/V length up to which the one-Q-tile kernel (two CTAs per SM) is used; MAB_ATT_SHORT_MAX in the
// environment or mab_set_attention_short_max() (tests drive both kernels at small sizes with it)
inline int& att_short_max() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("MAB_ATT_SHORT_MAX");
    v = e ? atoi(e) : 6000;  // measured crossover between 4 and 8 views of global attention
  }
  return v;
}
// q_src / kv_src: packed [rows, 3*heads*64] bf16 buffers (q | k | v)
inline int launch_attention(const void* q_src, long long q_rows_total, const void* kv_src,
                            long long kv_rows_total, int heads, int num_seqs, int q_len, int kv_len,
                            int q_seq_stride, int kv_seq_stride, float scale, void* out, float* lse,
                            cudaStream_t st, float* carry_o = nullptr, float* carry_ml = nullptr,
                            int carry_in = 0, int carry_out = 0, int kv_packed = 0) {
  MAB_ARG(!(carry_in || carry_out) || (carry_o && carry_ml), "carry buffers missing");
  MAB_ARG(heads > 0 && num_seqs > 0 && q_len > 0 && kv_len > 0, "empty attention");
  const int D = heads * 64;
  CUtensorMap tQ, tKV;
  MAB_TRY(make_tmap_2d(&tQ, q_src, q_rows_total, 3LL * D, 3LL * D, ATT_BQ));
  // kv_packed: kv_src is a [rows, 2*heads*64] K|V buffer (gathered from the other ranks)
  const long long kv_ld = kv_packed ? 2LL * D : 3LL * D;
  MAB_TRY(make_tmap_2d(&tKV, kv_src, kv_rows_total, kv_ld, kv_ld, ATT_BKV));
  AttnParams p;
  p.num_seqs = num_seqs; p.heads = heads; p.q_len = q_len; p.kv_len = kv_len;
  p.q_seq_stride = q_seq_stride; p.kv_seq_stride = kv_seq_stride;
  p.q_col0 = 0; p.k_col0 = kv_packed ? 0 : D; p.v_col0 = kv_packed ? D : 2 * D;
  p.out = reinterpret_cast<__nv_bfloat16*>(out); p.ldo = D;
  p.scale_log2e = scale * 1.4426950408889634f;
  p.lse = lse;
  p.carry_o = carry_o; p.carry_ml = carry_ml; p.carry_in = carry_in; p.carry_out = carry_out;
  static bool attr_done[MAB_MAX_DEVICES] = {false};
  if (!attr_done[current_device()]) {
    MAB_CUDA(cudaFuncSetAttribute(attention_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  AttCfg<2>::SMEM_BYTES));
    MAB_CUDA(cudaFuncSetAttribute(attention_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  AttCfg<1>::SMEM_BYTES));
    MAB_CUDA(cudaFuncSetAttribute(attention_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  AttCfg<1>::SMEM_BYTES));
    attr_done[current_device()] = true;
  }
  ProfScope ps(CAT_ATTENTION, 4.0 * q_len * (double)kv_len * 64.0 * heads * num_seqs, 0.0, st);
  static const bool no_lean = getenv("MAB_NO_ATT_LEAN") != nullptr;
  // one-tile kernel only: the lean instantiation is 2.5 % faster there (465 vs 453 TFLOP/s on 8x16x1037), while
  // the same change costs the two-tile kernel its schedule (821 -> 724 TFLOP/s at 66 305 tokens)
  const bool lean = !carry_in && !carry_out && !lse && !no_lean;
  // short sequences (frame / encoder attention: 9 K/V tiles): one Q tile per CTA, two CTAs per SM;
  // long ones (global attention): two Q tiles per CTA sharing every K/V tile in shared memory
  if (kv_len <= att_short_max()) {
    dim3 grid((q_len + ATT_BQ - 1) / ATT_BQ, heads, num_seqs);
    MAB_CUDA(launch_pdl(lean ? attention_kernel<1, true> : attention_kernel<1, false>, grid, dim3(AttCfg<1>::THREADS),
                        AttCfg<1>::SMEM_BYTES, st, 1, tQ, tKV, tKV, p));
    return 0;
  }
  dim3 grid((q_len + ATT_QT * ATT_BQ - 1) / (ATT_QT * ATT_BQ), heads, num_seqs);
  MAB_CUDA(launch_pdl(attention_kernel<2, false>, grid, dim3(AttCfg<2>::THREADS), AttCfg<2>::SMEM_BYTES, st, 1, tQ,
                      tKV, tKV, p));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

inline int launch_layernorm(const float* x, int rows_out, int C, int group_out, int group_in,
                            int skip, const float* w, const float* b, const float* w2,
                            const float* b2, float eps, void* out_bf16, float* out_f32,
                            float* out_mid, cudaStream_t st, void* out_split = nullptr, int round_f32 = 0) {
  if (rows_out <= 0) return 0;
  auto os = reinterpret_cast<__nv_bfloat16*>(out_split);
  const int block = 256, rows_per_block = block / 32;
  const int grid = (rows_out + rows_per_block - 1) / rows_per_block;
  auto ob = reinterpret_cast<__nv_bfloat16*>(out_bf16);
  // algorithmic bytes: fp32 row in, bf16 and/or fp32 row(s) out
  const double bytes = (double)rows_out * C * (4.0 + (out_bf16 ? 2.0 : 0.0) + (out_f32 ? 4.0 : 0.0) +
                                               (out_mid ? 4.0 : 0.0) + (out_split ? 6.0 : 0.0));
  ProfScope ps(CAT_LAYERNORM, 0.0, bytes, st);
  static const bool no_lean = getenv("MAB_NO_LN_LEAN") != nullptr;
  const bool lean = ob && !out_f32 && !out_mid && !os && !w2 && group_out == group_in && skip == 0 && !no_lean;
  if (lean && C == 768)
    MAB_CUDA(launch_pdl(layernorm_kernel<6, true>, dim3(grid), dim3(block), 0, st, 1, x, rows_out, group_out,
                        group_in, skip, w, b, w2, b2, eps, ob, out_f32, out_mid, os, round_f32));
  else if (lean && C == 1024)
    MAB_CUDA(launch_pdl(layernorm_kernel<8, true>, dim3(grid), dim3(block), 0, st, 1, x, rows_out, group_out,
                        group_in, skip, w, b, w2, b2, eps, ob, out_f32, out_mid, os, round_f32));
  else if (C == 768)
    MAB_CUDA(launch_pdl(layernorm_kernel<6>, dim3(grid), dim3(block), 0, st, 1, x, rows_out, group_out,
                        group_in, skip, w, b, w2, b2, eps, ob, out_f32, out_mid, os, round_f32));
  else if (C == 1024)
    MAB_CUDA(launch_pdl(layernorm_kernel<8>, dim3(grid), dim3(block), 0, st, 1, x, rows_out, group_out,
                        group_in, skip, w, b, w2, b2, eps, ob, out_f32, out_mid, os, round_f32));
  else if (C == 128)
    MAB_CUDA(launch_pdl(layernorm_kernel<1>, dim3(grid), dim3(block), 0, st, 1, x, rows_out, group_out,
                        group_in, skip, w, b, w2, b2, eps, ob, out_f32, out_mid, os, round_f32));
  else
    return fail(-1, "layernorm: unsupported width %d (supported 128, 768, 1024)", C);
  MAB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace mab
