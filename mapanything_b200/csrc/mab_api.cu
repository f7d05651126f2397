// libmab200: C ABI (include/mab200.h) + host runtime that sequences the sm_100a kernels of the
// MapAnything inference hot path.  Part 1: weight packing kernels and the per-kernel entry points.
#include "../../include/mab200.h"

#include <dlfcn.h>

#include <map>
#include <mutex>
#include <vector>

#include "host_util.cuh"

using namespace mab;
typedef __nv_bfloat16 bf16;

// ------------------------------------------------------------------------------------------------
// weight packing kernels (run once in mab_finalize_weights)
// ------------------------------------------------------------------------------------------------
namespace {

// fp32 [rows, cols] -> bf16 [rows, ldo] zero padded
__device__ __forceinline__ void store_as(bf16* p, float v) { *p = __float2bfloat16(v); }
__device__ __forceinline__ void store_as(float* p, float v) { *p = tf32_rn(v); }  // kind::tf32 operand

template <typename T>
__global__ void pack_matrix_kernel(const float* __restrict__ w, int rows, int cols, int ldo,
                                   T* __restrict__ out) {
  const long long total = (long long)rows * ldo;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = i % ldo;
    const long long r = i / ldo;
    store_as(out + i, c < cols ? w[r * cols + c] : 0.f);
  }
}

// fp32 [rows, cols] -> bf16 [rows, 3*cols] = [hi | hi | lo]: B operand of a bf16x3 GEMM whose A rows
// are [hi | lo | hi], so that A.B^T = a_hi.w_hi + a_lo.w_hi + a_hi.w_lo ~ fp32 a.w (error ~2^-16)
__global__ void pack_matrix_split_kernel(const float* __restrict__ w, int rows, int cols,
                                         bf16* __restrict__ out) {
  const long long total = (long long)rows * cols;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = i % cols;
    const long long r = i / cols;
    const float x = w[i];
    const bf16 hi = __float2bfloat16(x);
    const bf16 lo = __float2bfloat16(x - __bfloat162float(hi));
    bf16* o = out + r * 3 * cols;
    o[c] = hi; o[cols + c] = hi; o[2 * cols + c] = lo;
  }
}

// conv weight [Cout, Cin, 3, 3] fp32 -> [Cout, 9, Cpad] bf16 (tap = ky*3+kx), zero padded channels
template <typename T>
__global__ void pack_conv3_kernel(const float* __restrict__ w, int Cout, int Cin, int Cpad,
                                  T* __restrict__ out) {
  const long long total = (long long)Cout * 9 * Cpad;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = i % Cpad;
    const int tap = (i / Cpad) % 9;
    const int co = i / ((long long)Cpad * 9);
    store_as(out + i, c < Cin ? w[((long long)co * Cin + c) * 9 + tap] : 0.f);
  }
}

// ConvTranspose2d weight [Cin, Cout, k, k] -> GEMM B [k*k*Cout, Cin] with row (i*k+j)*Cout + co
template <typename T>
__global__ void pack_convT_kernel(const float* __restrict__ w, int Cin, int Cout, int k,
                                  T* __restrict__ out) {
  const long long total = (long long)k * k * Cout * Cin;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int ci = idx % Cin;
    const int row = idx / Cin;
    const int co = row % Cout, ij = row / Cout;
    store_as(out + idx, w[((long long)ci * Cout + co) * k * k + ij]);
  }
}

// bias [Cout] -> [k*k*Cout] repeated
__global__ void tile_bias_kernel(const float* __restrict__ b, int Cout, int reps, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < Cout * reps) out[i] = b[i % Cout];
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// per-kernel C entry points
// ------------------------------------------------------------------------------------------------
extern "C" const char* mab_last_error(void) { return err_buf(); }
extern "C" int mab_version(void) { return 100; }

extern "C" long long mab_launch_count(void) { return prof().launches; }
extern "C" int mab_prof_enable(int on) {
  ProfState& p = prof();
  p.on = on != 0;
  p.recs.clear();
  p.used = 0;
  return 0;
}
// Synchronises the device, then sums the recorded launches of category `cat` (CAT_*; include/mab200.h).
extern "C" int mab_prof_query(int cat, long long* launches, double* ms, double* flops, double* bytes) {
  MAB_ARG(cat >= 0 && cat < CAT_COUNT && launches && ms && flops && bytes, "bad category");
  MAB_CUDA(cudaDeviceSynchronize());
  *launches = 0; *ms = 0; *flops = 0; *bytes = 0;
  for (const ProfRec& r : prof().recs) {
    if (r.cat != cat) continue;
    float t = 0.f;
    MAB_CUDA(cudaEventElapsedTime(&t, r.e0, r.e1));
    ++*launches; *ms += t; *flops += r.flops; *bytes += r.bytes;
  }
  return 0;
}

// Per-launch records of the profiler, in launch order: category, device milliseconds, algorithmic FLOPs and
// bytes of up to `max` launches; returns the number of records held (synchronises the device).
extern "C" long long mab_prof_records(int* cat, double* ms, double* flops, double* bytes, long long max) {
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  long long n = 0;
  for (const ProfRec& r : prof().recs) {
    if (n < max) {
      float t = 0.f;
      if (cudaEventElapsedTime(&t, r.e0, r.e1) != cudaSuccess) return -1;
      if (cat) cat[n] = r.cat;
      if (ms) ms[n] = t;
      if (flops) flops[n] = r.flops;
      if (bytes) bytes[n] = r.bytes;
    }
    ++n;
  }
  return n;
}

// ---- a1 pre-processing entry points (inference.py:222-311) ---------------------------------------
extern "C" int mab_rays_from_intrinsics(const void* K, int views, int H, int W, void* rays, void* stream) {
  MAB_ARG(K && rays && views > 0 && H > 0 && W > 0, "bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, st,
             rays_from_intrinsics_kernel<<<ew_grid((long long)views * H * W), 256, 0, st>>>(
                 (const float*)K, views, H, W, (float*)rays));
  MAB_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int mab_normalize_rays(const void* rays_in, long long pixels, void* rays_out, void* stream) {
  MAB_ARG(rays_in && rays_out && pixels > 0, "bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, st,
             normalize_rays_kernel<<<ew_grid(pixels), 256, 0, st>>>((const float*)rays_in, pixels,
                                                                    (float*)rays_out));
  MAB_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int mab_depth_z_to_along_ray(const void* depth_z, const void* rays, long long pixels, void* out,
                                        void* stream) {
  MAB_ARG(depth_z && rays && out && pixels > 0, "bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, st,
             depth_z_to_along_ray_kernel<<<ew_grid(pixels), 256, 0, st>>>((const float*)depth_z,
                                                                          (const float*)rays, pixels,
                                                                          (float*)out));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int mab_layernorm(const float* x, int rows, int C, const float* w, const float* b,
                             float eps, void* out_bf16, float* out_f32, void* stream) {
  MAB_ARG(x && w && b && (out_bf16 || out_f32), "null argument");
  return launch_layernorm(x, rows, C, rows > 0 ? rows : 1, rows > 0 ? rows : 1, 0, w, b, nullptr,
                          nullptr, eps, out_bf16, out_f32, nullptr, (cudaStream_t)stream);
}

extern "C" int mab_gemm(const void* A, const void* B, int M, int N, int K, const float* bias,
                        const float* gamma, const float* res_f32, int act, float* out_f32,
                        void* out_bf16, void* stream) {
  MAB_ARG(A && B && (out_f32 || out_bf16), "null argument");
  GemmEpilogue ep{};
  ep.bias = bias; ep.gamma = gamma; ep.res_f32 = res_f32; ep.out_f32 = out_f32;
  ep.out_bf16 = reinterpret_cast<bf16*>(out_bf16); ep.ldo = N; ep.act = act;
  return launch_gemm(A, K, B, K, M, N, K, ep, (cudaStream_t)stream);
}

// K/V length up to which attention uses the one-Q-tile kernel (< 0: back to the default)
extern "C" int mab_set_attention_short_max(int kv_len) {
  att_short_max() = kv_len < 0 ? 6000 : kv_len;
  return 0;
}

extern "C" int mab_attention(const void* q_src, int q_rows_total, const void* kv_src,
                             int kv_rows_total, int heads, int num_seqs, int q_len, int kv_len,
                             int q_seq_stride, int kv_seq_stride, float scale, void* out, float* lse,
                             void* stream) {
  MAB_ARG(q_src && kv_src && out, "null argument");
  MAB_ARG(q_seq_stride >= 0 && kv_seq_stride >= 0 &&
          (long long)(num_seqs - 1) * q_seq_stride + q_len <= q_rows_total &&
          (long long)(num_seqs - 1) * kv_seq_stride + kv_len <= kv_rows_total,
          "sequences reach past the end of the buffers");
  return launch_attention(q_src, q_rows_total, kv_src, kv_rows_total, heads, num_seqs, q_len, kv_len,
                          q_seq_stride, kv_seq_stride, scale, out, lse, (cudaStream_t)stream);
}

extern "C" int mab_attention_carry(const void* q_src, int q_rows_total, const void* kv_src,
                                   int kv_rows_total, int heads, int num_seqs, int q_len, int kv_len,
                                   int q_seq_stride, int kv_seq_stride, float scale, void* out,
                                   float* carry_o, float* carry_ml, int carry_in, int carry_out,
                                   void* stream) {
  MAB_ARG(q_src && kv_src && (out || carry_out), "null argument");
  MAB_ARG(q_seq_stride >= 0 && kv_seq_stride >= 0 &&
          (long long)(num_seqs - 1) * q_seq_stride + q_len <= q_rows_total &&
          (long long)(num_seqs - 1) * kv_seq_stride + kv_len <= kv_rows_total,
          "sequences reach past the end of the buffers");
  return launch_attention(q_src, q_rows_total, kv_src, kv_rows_total, heads, num_seqs, q_len, kv_len,
                          q_seq_stride, kv_seq_stride, scale, out, nullptr, (cudaStream_t)stream,
                          carry_o, carry_ml, carry_in, carry_out);
}

extern "C" int mab_pack_conv3_weight(const float* w, int Cout, int Cin, void* out, void* stream) {
  MAB_ARG(w && out && Cout > 0 && Cin > 0, "bad arguments");
  const int Cpad = (Cin + 63) / 64 * 64;
  pack_conv3_kernel<<<ew_grid((long long)Cout * 9 * Cpad), 256, 0, (cudaStream_t)stream>>>(
      w, Cout, Cin, Cpad, reinterpret_cast<bf16*>(out));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int mab_conv3x3(const void* x, int batch, int H, int W, int C, const void* w_packed,
                           int Cout, const float* bias, const void* res1, const void* res2, int act,
                           void* out_bf16, void* out_relu, void* stream) {
  MAB_ARG(x && w_packed && (out_bf16 || out_relu), "null argument");
  MAB_ARG(batch > 0 && H > 0 && W > 0 && C > 0 && Cout > 0 && Cout % 4 == 0, "bad conv shape");
  GemmEpilogue ep{};
  ep.bias = bias; ep.res1 = reinterpret_cast<const bf16*>(res1);
  ep.res2 = reinterpret_cast<const bf16*>(res2); ep.out_bf16 = reinterpret_cast<bf16*>(out_bf16);
  ep.out_relu = reinterpret_cast<bf16*>(out_relu); ep.ldo = Cout; ep.act = act;
  return launch_conv3(x, batch, H, W, C, w_packed, Cout, ep, (cudaStream_t)stream);
}

// ---- TF32 variants (fp32 operands) used by the geometric-input encoders ---------------------------
extern "C" int mab_gemm_tf32(const float* A, const float* B, int M, int N, int K, const float* bias,
                             const float* res_f32, int act, int act_post, float* out_f32, void* stream) {
  MAB_ARG(A && B && out_f32, "null argument");
  GemmEpilogue ep{};
  ep.bias = bias; ep.res_f32 = res_f32; ep.out_f32 = out_f32; ep.ldo = N; ep.act = act; ep.act_post = act_post;
  return launch_gemm_tf32(A, K, B, K, M, N, K, ep, (cudaStream_t)stream);
}
extern "C" int mab_pack_conv3_weight_f32(const float* w, int Cout, int Cin, int Cpad, float* out, void* stream) {
  MAB_ARG(w && out && Cout > 0 && Cin > 0, "bad arguments");
  MAB_ARG(Cpad >= Cin && Cpad % 32 == 0, "Cpad must be a multiple of 32 and >= Cin");
  pack_conv3_kernel<<<ew_grid((long long)Cout * 9 * Cpad), 256, 0, (cudaStream_t)stream>>>(w, Cout, Cin, Cpad, out);
  MAB_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int mab_conv3x3_tf32(const float* x, int batch, int H, int W, int C, const float* w_packed,
                                int Cout, const float* bias, const float* res_f32, int act, int act_post,
                                float* out_f32, void* stream) {
  MAB_ARG(x && w_packed && out_f32, "null argument");
  MAB_ARG(batch > 0 && H > 0 && W > 0 && Cout > 0 && Cout % 4 == 0, "bad conv shape");
  MAB_ARG(C > 0 && C % 32 == 0, "C must be a multiple of 32 (pad the channels)");
  GemmEpilogue ep{};
  ep.bias = bias; ep.res_f32 = res_f32; ep.out_f32 = out_f32; ep.ldo = Cout; ep.act = act; ep.act_post = act_post;
  return launch_conv3_tf32(x, batch, H, W, C, C, w_packed, Cout, ep, (cudaStream_t)stream);
}

extern "C" int mab_resize_bilinear(const void* x, int batch, int Hi, int Wi, int C, int Hv, int Wv,
                                   int Ho, int Wo, void* out, void* out_relu, void* stream) {
  MAB_ARG(x && (out || out_relu), "null argument");
  MAB_ARG(C > 0 && C % 8 == 0, "C must be a multiple of 8");
  MAB_ARG(Hi > 0 && Wi > 0 && Hv >= Ho && Wv >= Wo, "resize: bad sizes");
  MAB_ARG(batch > 0 && batch <= 65535 && Ho > 0 && Ho <= 65535 && Wo > 0, "resize: bad sizes");
  const dim3 grid((Wo * (C / 8) + 255) / 256, Ho, batch);
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, (cudaStream_t)stream,
             resize_bilinear_ac_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const bf16*>(x), batch, Hi, Wi, C, Hv, Wv, Ho, Wo,
      reinterpret_cast<bf16*>(out), reinterpret_cast<bf16*>(out_relu)));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int mab_posembed_resize(const float* pos, int grid, int C, int h, int w, float* out,
                                   void* stream) {
  MAB_ARG(pos && out && grid > 0 && C > 0 && h > 0 && w > 0, "bad arguments");
  // vision_transformer.py:223-228: scale_factor = ((h + 0.1)/M, (w + 0.1)/M); ATen then uses
  // 1/scale_factor as the source step
  if (h == grid && w == grid) {  // vision_transformer.py:214-215: same grid -> returned unchanged
    MAB_CUDA(cudaMemcpyAsync(out, pos, (size_t)(1 + h * w) * C * sizeof(float),
                             cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
  }
  const float inv_sy = 1.0f / (float)((double)(h + 0.1) / grid);
  const float inv_sx = 1.0f / (float)((double)(w + 0.1) / grid);
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, (cudaStream_t)stream,
             posembed_bicubic_kernel<<<ew_grid((long long)(1 + h * w) * C), 256, 0, (cudaStream_t)stream>>>(
      pos, grid, C, h, w, inv_sy, inv_sx, out));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// RoPE2D in place (curope.cpp:49-69 `rope_2d(tokens, positions, base, fwd)`): tokens addressed as
// tokens + token * token_pitch + head * head_dim (elements), positions [n_tokens, 2] int64 (y, x).
extern "C" int mab_rope2d(void* tokens, int dtype, long long n_tokens, int heads, int head_dim,
                          long long token_pitch, const long long* positions, float base, float fwd,
                          void* stream) {
  MAB_ARG(tokens && positions, "null argument");
  MAB_ARG(dtype == 0 || dtype == 1, "dtype must be MAB_F32 or MAB_BF16");
  MAB_ARG(n_tokens >= 0 && heads > 0, "bad sizes");
  // curope requires D % 4 == 0 (kernels.cu:92); the vectorised kernel rotates 4 frequencies per thread
  MAB_ARG(head_dim > 0 && head_dim % 16 == 0, "head_dim must be a multiple of 16");
  MAB_ARG(token_pitch >= (long long)heads * head_dim && token_pitch % 4 == 0, "bad token pitch");
  MAB_ARG(reinterpret_cast<uintptr_t>(tokens) % 16 == 0, "tokens must be 16-byte aligned");
  if (n_tokens == 0) return 0;
  const long long threads = n_tokens * heads * (head_dim / 8);
  const double bytes = 2.0 * (double)n_tokens * heads * head_dim * (dtype == 0 ? 4 : 2) + 16.0 * n_tokens;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 0)
    MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, bytes, st,
               rope2d_kernel<float><<<ew_grid(threads), 256, 0, st>>>(
                   reinterpret_cast<float*>(tokens), n_tokens, heads, head_dim, token_pitch, positions, base, fwd));
  else
    MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, bytes, st,
               rope2d_kernel<bf16><<<ew_grid(threads), 256, 0, st>>>(
                   reinterpret_cast<bf16*>(tokens), n_tokens, heads, head_dim, token_pitch, positions, base, fwd));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// ---- image input path (csrc/image.cuh) -------------------------------------------------------
namespace {

double img_sinc(double x) {
  if (x == 0.0) return 1.0;
  x *= M_PI;
  return sin(x) / x;
}
double img_filter(int filter, double x) {  // Pillow Resample.c lanczos_filter / bicubic_filter (a = -0.5)
  if (filter == 0) return (-3.0 <= x && x < 3.0) ? img_sinc(x) * img_sinc(x / 3) : 0.0;
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
  if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
  return 0.0;
}
// Pillow precompute_coeffs + normalize_coeffs_8bpc for output samples [first, first + count) of a
// resize in_size -> out_size; bounds (xmin, n) pairs and ksize coefficients per sample
int img_coeffs(int in_size, int out_size, int filter, int first, int count, std::vector<int>* bounds,
               std::vector<int>* kk) {
  double scale = (double)((float)in_size - 0.0f) / out_size, filterscale = scale;
  if (filterscale < 1.0) filterscale = 1.0;
  const double support = (filter == 0 ? 3.0 : 2.0) * filterscale;
  const int ksize = (int)ceil(support) * 2 + 1;
  bounds->assign((size_t)count * 2, 0);
  kk->assign((size_t)count * ksize, 0);
  const double ss = 1.0 / filterscale;
  std::vector<double> w(ksize);
  for (int j = 0; j < count; ++j) {
    const int xx = first + j;
    const double center = 0.0 + (xx + 0.5) * scale;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x) {
      w[x] = img_filter(filter, (x + xmin - center + 0.5) * ss);
      ww += w[x];
    }
    for (int x = 0; x < xmax; ++x) {
      const double v = ww != 0.0 ? w[x] / ww : w[x];
      (*kk)[(size_t)j * ksize + x] = v < 0 ? (int)(-0.5 + v * (1 << IMG_PRECISION_BITS))
                                           : (int)(0.5 + v * (1 << IMG_PRECISION_BITS));
    }
    (*bounds)[2 * j] = xmin;
    (*bounds)[2 * j + 1] = xmax;
  }
  return ksize;
}

}  // namespace

// cropping.py:230-244 (rescale so that the image contains the crop; LANCZOS when shrinking, BICUBIC
// otherwise) and :441-447 (centred crop), evaluated in double like numpy does
extern "C" int mab_image_plan_make(int in_w, int in_h, int target_w, int target_h, mab_image_plan* plan) {
  MAB_ARG(plan && in_w > 0 && in_h > 0 && target_w > 0 && target_h > 0, "bad sizes");
  const double sw = (double)target_w / in_w, sh = (double)target_h / in_h;
  const double scale_final = (sw > sh ? sw : sh) + 1e-8;
  plan->in_w = in_w; plan->in_h = in_h;
  plan->resize_w = (int)floor(in_w * scale_final);
  plan->resize_h = (int)floor(in_h * scale_final);
  plan->filter = scale_final < 1 ? 0 : 1;
  plan->out_w = target_w; plan->out_h = target_h;
  // Python floor division (the crop offsets are >= 0 by construction of scale_final)
  plan->crop_left = (int)floor((plan->resize_w - target_w) / 2.0);
  plan->crop_top = (int)floor((plan->resize_h - target_h) / 2.0);
  MAB_ARG(plan->crop_left >= 0 && plan->crop_top >= 0, "target larger than the rescaled image");
  return 0;
}

extern "C" size_t mab_preprocess_workspace_bytes(const mab_image_plan* p) {
  if (!p || p->out_w <= 0 || p->out_h <= 0) return 0;
  auto ks = [](int in, int out, int filter) {
    double fs = (double)in / out; if (fs < 1.0) fs = 1.0;
    return (size_t)((int)ceil((filter == 0 ? 3.0 : 2.0) * fs) * 2 + 1);
  };
  const size_t tabs = (size_t)p->out_w * (2 + ks(p->in_w, p->resize_w, p->filter)) +
                      (size_t)p->out_h * (2 + ks(p->in_h, p->resize_h, p->filter));
  return tabs * sizeof(int) + (size_t)p->in_h * p->out_w * 3 + 1024;
}

extern "C" int mab_preprocess_image(const unsigned char* rgb, const mab_image_plan* p, const float* mean,
                                    const float* std_, float* out, unsigned char* out_u8, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  MAB_ARG(rgb && p && mean && std_ && out && workspace, "null argument");
  MAB_ARG(p->in_w > 0 && p->in_h > 0 && p->resize_w > 0 && p->resize_h > 0 && p->out_w > 0 && p->out_h > 0 &&
          (p->filter == 0 || p->filter == 1), "bad image plan");
  MAB_ARG(p->crop_left >= 0 && p->crop_top >= 0 && p->crop_left + p->out_w <= p->resize_w &&
          p->crop_top + p->out_h <= p->resize_h, "crop outside the resized image");
  MAB_ARG(workspace_bytes >= mab_preprocess_workspace_bytes(p), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  // The coefficient tables depend on the plan only: computed once per plan and kept in PINNED host memory
  // (one cudaHostAlloc per distinct plan, at its first use), so that the uploads below are truly
  // asynchronous on the caller's stream -- copies from pageable vectors stage synchronously.
  struct Tables { int* pinned = nullptr; size_t nhb = 0, nhk = 0, nvb = 0, nvk = 0; int hks = 0, vks = 0, y0 = 0, y1 = 0; };
  static std::map<std::vector<int>, Tables> cache;
  static std::mutex cache_mu;
  const bool need_h = p->resize_w != p->in_w, need_v = p->resize_h != p->in_h;
  Tables tb;
  {
    std::lock_guard<std::mutex> lock(cache_mu);
    const std::vector<int> key = {p->in_w, p->in_h, p->resize_w, p->resize_h, p->filter, p->crop_left, p->crop_top,
                                  p->out_w, p->out_h};
    auto it = cache.find(key);
    if (it == cache.end()) {
      std::vector<int> hb, hk, vb, vk;
      Tables t;
      // a pass is skipped when that dimension is unchanged (ImagingResample need_horizontal / need_vertical)
      if (need_h) t.hks = img_coeffs(p->in_w, p->resize_w, p->filter, p->crop_left, p->out_w, &hb, &hk);
      else {  // identity taps: one coefficient of 1.0 per column
        t.hks = 1; hb.resize((size_t)p->out_w * 2); hk.assign(p->out_w, 1 << IMG_PRECISION_BITS);
        for (int x = 0; x < p->out_w; ++x) { hb[2 * x] = p->crop_left + x; hb[2 * x + 1] = 1; }
      }
      t.y0 = p->crop_top; t.y1 = p->crop_top + p->out_h;  // source rows the vertical pass reads
      if (need_v) {
        t.vks = img_coeffs(p->in_h, p->resize_h, p->filter, p->crop_top, p->out_h, &vb, &vk);
        t.y0 = vb[0]; t.y1 = vb[2 * (p->out_h - 1)] + vb[2 * (p->out_h - 1) + 1];
        for (int y = 0; y < p->out_h; ++y) {  // bounds are monotone, but do not rely on it
          if (vb[2 * y] < t.y0) t.y0 = vb[2 * y];
          if (vb[2 * y] + vb[2 * y + 1] > t.y1) t.y1 = vb[2 * y] + vb[2 * y + 1];
        }
      }
      t.nhb = hb.size(); t.nhk = hk.size(); t.nvb = vb.size(); t.nvk = vk.size();
      const size_t total = t.nhb + t.nhk + t.nvb + t.nvk;
      MAB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&t.pinned), (total > 0 ? total : 1) * sizeof(int), cudaHostAllocPortable));
      int* q = t.pinned;
      memcpy(q, hb.data(), t.nhb * sizeof(int)); q += t.nhb;
      memcpy(q, hk.data(), t.nhk * sizeof(int)); q += t.nhk;
      memcpy(q, vb.data(), t.nvb * sizeof(int)); q += t.nvb;
      memcpy(q, vk.data(), t.nvk * sizeof(int));
      it = cache.emplace(key, t).first;
    }
    tb = it->second;
  }
  const int hks = tb.hks, vks = tb.vks, y0 = tb.y0, rows = tb.y1 - tb.y0;
  uint8_t* w8 = reinterpret_cast<uint8_t*>(workspace);
  int* d_hb = reinterpret_cast<int*>(w8);
  int* d_hk = d_hb + tb.nhb;
  int* d_vb = d_hk + tb.nhk;
  int* d_vk = d_vb + tb.nvb;
  unsigned char* tmp = reinterpret_cast<unsigned char*>(d_vk + tb.nvk);
  // one upload: hb | hk | vb | vk are contiguous on both sides
  MAB_CUDA(cudaMemcpyAsync(d_hb, tb.pinned, (tb.nhb + tb.nhk + tb.nvb + tb.nvk) * sizeof(int), cudaMemcpyHostToDevice, st));
  const double in_bytes = (double)rows * p->in_w * 3, tmp_bytes = (double)rows * p->out_w * 3;
  dim3 gh((p->out_w + 127) / 128, rows), gv((p->out_w + 127) / 128, p->out_h);
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, in_bytes + tmp_bytes, st,
             image_resample_h_kernel<<<gh, 128, 0, st>>>(rgb, p->in_w, y0, rows, p->out_w, d_hb, d_hk, hks, tmp));
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, tmp_bytes + (double)p->out_w * p->out_h * 12.0, st,
             image_resample_v_norm_kernel<<<gv, 128, 0, st>>>(tmp, y0, p->out_w, p->out_h, d_vb, d_vk, vks, mean[0],
                                                             mean[1], mean[2], std_[0], std_[1], std_[2], out, out_u8));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// recover_pinhole_intrinsics_from_ray_directions (geometry.py:304-447, regression branch): least squares
// of x = cx + fx * (dx / dz) on the stride-max(1, H//50) x max(1, W//50) grid, one block per view
extern "C" int mab_recover_intrinsics(const float* rays, int views, int H, int W, float* K_out, void* stream) {
  MAB_ARG(rays && K_out && views > 0 && H > 0 && W > 0, "bad arguments");
  MAB_ARG((long long)H * W <= 1000000, "the reference switches to its geometric branch above 1e6 pixels");
  cudaStream_t st = (cudaStream_t)stream;
  MAB_LAUNCH(CAT_POSTPROCESS, 0.0, 0.0, st,
             intrinsics_pose_kernel<<<views, 256, 0, st>>>(rays, H, W, nullptr, nullptr, K_out, nullptr, nullptr,
                                                          nullptr));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// depth maps of preprocess_inputs (image.py:560-567 -> cropping.py:246-253,345-349)
extern "C" int mab_resize_nearest_crop(const float* src, int in_h, int in_w, int resize_w, int resize_h,
                                       int crop_left, int crop_top, int out_w, int out_h, float* out,
                                       void* stream) {
  MAB_ARG(src && out && in_h > 0 && in_w > 0 && resize_w > 0 && resize_h > 0 && out_w > 0 && out_h > 0, "bad sizes");
  MAB_ARG(crop_left >= 0 && crop_top >= 0 && crop_left + out_w <= resize_w && crop_top + out_h <= resize_h,
          "crop outside the resized map");
  cudaStream_t st = (cudaStream_t)stream;
  const double ifx = 1.0 / ((double)resize_w / in_w), ify = 1.0 / ((double)resize_h / in_h);
  dim3 grid((out_w + 127) / 128, out_h);
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, (double)out_w * out_h * 8.0, st,
             resize_nearest_crop_kernel<<<grid, 128, 0, st>>>(src, in_h, in_w, ifx, ify, crop_left, crop_top, out_w,
                                                             out_h, out));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// geometry.py:24-114 + the demos' mask-filtered cloud (see csrc/export.cuh).
extern "C" size_t mab_export_workspace_bytes(int views, int H, int W) {
  if (views <= 0 || H <= 0 || W <= 0) return 0;
  const long long tiles = (long long)views * (((long long)H * W + EXP_TILE - 1) / EXP_TILE);
  return (size_t)tiles * (sizeof(int) + sizeof(long long)) + 512;
}

extern "C" int mab_export_points(const float* depth_z, const float* intrinsics, const float* camera_poses,
                                 const unsigned char* mask, const float* img_no_norm, int views, int H,
                                 int W, float* world_points, float* out_xyz, unsigned char* out_rgb,
                                 long long capacity, long long* counts, void* workspace,
                                 size_t workspace_bytes, void* stream) {
  MAB_ARG(depth_z && intrinsics && counts && workspace, "null argument");
  MAB_ARG(views > 0 && H > 0 && W > 0 && capacity >= 0, "bad sizes");
  MAB_ARG(!out_rgb || img_no_norm, "out_rgb needs img_no_norm");
  MAB_ARG((long long)H * W < (1LL << 31), "view too large");
  MAB_ARG(workspace_bytes >= mab_export_workspace_bytes(views, H, W), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const int npix = H * W, tpv = (npix + EXP_TILE - 1) / EXP_TILE;
  const long long tiles = (long long)views * tpv;
  MAB_ARG(tiles < (1LL << 31), "too many tiles");
  long long* tile_offsets = reinterpret_cast<long long*>(workspace);  // 8-byte aligned first
  int* tile_counts = reinterpret_cast<int*>(tile_offsets + tiles);
  const double px = (double)views * npix;
  MAB_LAUNCH(CAT_POSTPROCESS, 0.0, px * 5.0, st,
             export_count_kernel<<<(int)tiles, EXP_THREADS, 0, st>>>(depth_z, mask, npix, tpv, tile_counts));
  MAB_LAUNCH(CAT_POSTPROCESS, 0.0, (double)tiles * 12.0, st,
             export_scan_kernel<<<1, 1024, 0, st>>>(tile_counts, (int)tiles, tpv, views, tile_offsets, counts));
  MAB_LAUNCH(CAT_POSTPROCESS, 0.0, px * (5.0 + (world_points ? 12.0 : 0.0) + 27.0), st,
             export_scatter_kernel<<<(int)tiles, EXP_THREADS, 0, st>>>(
                 depth_z, intrinsics, camera_poses, mask, img_no_norm, W, npix, tpv, tile_offsets, capacity,
                 world_points, out_xyz, out_rgb));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// ---- allocation-free variant of mab_export_points (no counterpart in the reference) --------------
// The tile counts / offsets live in a file-scope device array instead of a caller workspace, so the
// call needs no cudaMalloc anywhere (graph capture, serving loops).  Every GPU that loads the library
// holds the array until exit.  In HOST code `g_export_static_scratch` is only a stand-in for the
// symbol: the device address comes from cudaGetSymbolAddress, once per device with that device
// current, and is what the kernels get.
static __device__ __align__(256) unsigned char
    g_export_static_scratch[(size_t)MAB_EXPORT_STATIC_MAX_TILES * (sizeof(int) + sizeof(long long))];

namespace {
constexpr int kMaxDevices = 64;
int export_static_scratch_address(void** out) {
  static void* addr[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  MAB_CUDA(cudaGetDevice(&dev));
  MAB_ARG(dev >= 0 && dev < kMaxDevices, "device ordinal out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    MAB_CUDA(cudaGetSymbolAddress(&p, g_export_static_scratch));
    MAB_ARG(p != nullptr, "cudaGetSymbolAddress returned a null address");
    addr[dev] = p;
  }
  *out = addr[dev];
  return 0;
}
}  // namespace

extern "C" int mab_export_points_static(const float* depth_z, const float* intrinsics,
                                        const float* camera_poses, const unsigned char* mask,
                                        const float* img_no_norm, int views, int H, int W,
                                        float* world_points, float* out_xyz, unsigned char* out_rgb,
                                        long long capacity, long long* counts, void* stream) {
  MAB_ARG(depth_z && intrinsics && counts, "null argument");
  MAB_ARG(views > 0 && H > 0 && W > 0 && capacity >= 0, "bad sizes");
  MAB_ARG(!out_rgb || img_no_norm, "out_rgb needs img_no_norm");
  MAB_ARG((long long)H * W < (1LL << 31), "view too large");
  const int npix = H * W, tpv = (npix + EXP_TILE - 1) / EXP_TILE;
  const long long tiles = (long long)views * tpv;
  if (tiles > (long long)MAB_EXPORT_STATIC_MAX_TILES)
    return fail(-1, "argument error: mab_export_points_static: %lld tiles of %d pixels exceed the static "
                "scratch of MAB_EXPORT_STATIC_MAX_TILES = %d tiles; use mab_export_points with a workspace",
                tiles, EXP_TILE, (int)MAB_EXPORT_STATIC_MAX_TILES);
  void* scratch = nullptr;
  MAB_TRY(export_static_scratch_address(&scratch));
  cudaStream_t st = (cudaStream_t)stream;
  // same carving as mab_export_points; every entry the scan and scatter passes read is written by
  // the count / scan pass of THIS call first, so nothing has to start at zero
  long long* tile_offsets = reinterpret_cast<long long*>(scratch);
  int* tile_counts = reinterpret_cast<int*>(tile_offsets + tiles);
  prof().launches += 3;  // counted, never event-bracketed: the profiler's event pool allocates
  export_count_kernel<<<(int)tiles, EXP_THREADS, 0, st>>>(depth_z, mask, npix, tpv, tile_counts);
  export_scan_kernel<<<1, 1024, 0, st>>>(tile_counts, (int)tiles, tpv, views, tile_offsets, counts);
  export_scatter_kernel<<<(int)tiles, EXP_THREADS, 0, st>>>(depth_z, intrinsics, camera_poses, mask,
                                                            img_no_norm, W, npix, tpv, tile_offsets,
                                                            capacity, world_points, out_xyz, out_rgb);
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// workspace: views*H*W*(3*4 + 1 + 4) bytes  (normals fp32x3, normal mask u8, angle fp32)
extern "C" int mab_edge_mask(const float* pts3d, float* pts3d_cam, const unsigned char* fmask,
                             int views, int H, int W, float normal_tol_deg, float depth_rtol,
                             void* workspace, unsigned char* mask_out, void* stream) {
  MAB_ARG(pts3d && pts3d_cam && fmask && workspace && mask_out, "null argument");
  MAB_ARG(views > 0 && H > 0 && W > 0, "bad sizes");
  cudaStream_t st = (cudaStream_t)stream;
  const long long npix = (long long)views * H * W;
  float* normals = reinterpret_cast<float*>(workspace);
  float* ang = normals + npix * 3;
  unsigned char* nmask = reinterpret_cast<unsigned char*>(ang + npix);
  MAB_LAUNCH(CAT_POSTPROCESS, 0.0, 0.0, st,
             normals_kernel<<<ew_grid(npix), 256, 0, st>>>(pts3d, fmask, views, H, W, normals, nmask));
  MAB_LAUNCH(CAT_POSTPROCESS, 0.0, 0.0, st,
             normal_angle_kernel<<<ew_grid(npix), 256, 0, st>>>(normals, nmask, views, H, W, ang));
  const double tol = (double)normal_tol_deg * 3.14159265358979323846 / 180.0;
  MAB_LAUNCH(CAT_POSTPROCESS, 0.0, 0.0, st,
             edge_mask_apply_kernel<<<ew_grid(npix), 256, 0, st>>>(ang, fmask, views, H, W, 1, tol, depth_rtol,
                                                        nullptr, pts3d_cam, nullptr, nullptr, mask_out));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// ================================================================================================
// Part 2: model handle, weights, workspace plan
// ================================================================================================
namespace {

struct RawTensor {
  float* p = nullptr;
  std::vector<int64_t> shape;
  long long numel = 0;
};

std::string canonical_key(const std::string& k) {
  // aliases of the reference state dict (model.py:398-400, dpt_block.py:71-78, dpt.py:176-178)
  std::string s = k;
  auto starts = [&](const char* p) { return s.rfind(p, 0) == 0; };
  if (starts("dense_head.0.")) s = "dpt_feature_head." + s.substr(13);
  else if (starts("dense_head.1.")) s = "dpt_regressor_head." + s.substr(13);
  const std::string a = "dpt_feature_head.scratch.layer_rn.";
  const std::string b = "dpt_feature_head.input_process.";
  if (s.rfind(a, 0) == 0 && s.size() > a.size() + 1) {
    const int i = s[a.size()] - '0';
    s = "dpt_feature_head.scratch.layer" + std::to_string(i + 1) + "_rn" + s.substr(a.size() + 1);
  } else if (s.rfind(b, 0) == 0 && s.size() > b.size() + 3 && s.substr(b.size() + 1, 3) == ".1.") {
    const int i = s[b.size()] - '0';
    s = "dpt_feature_head.scratch.layer" + std::to_string(i + 1) + "_rn" + s.substr(b.size() + 2);
  }
  return s;
}

struct Bump {  // bump allocator over the caller's workspace (256-byte aligned)
  uint8_t* base = nullptr;
  size_t off = 0;
  template <typename T>
  T* take(long long count) {
    off = (off + 255) & ~size_t(255);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += (size_t)count * sizeof(T);
    return p;
  }
};

struct Plan {  // every workspace pointer of one forward
  int V = 0, Vtot = 0, H = 0, W = 0, h = 0, w = 0, T = 0;
  int Ve = 0;  // views per encoder pass (cfg.enc_chunk): the encoder's activations are sized for Ve views
  // The encoder runs on `lanes` streams at once, each pass of Vl views on the lane whose turn it is, with
  // its own activations: persistent GEMM / attention kernels leave SMs idle in their last wave (M = 8 views
  // x 1037 rows is 33 CTA-pair tiles per column of tiles on 74 clusters: 89 % full) and while their last
  // tiles drain; the other lane's kernels fill those SMs.
  int lanes = 1, Vl = 0;
  struct EncBufs { bf16* im2col; float* patch; float* xs; bf16* ln; bf16* qkv; bf16* att; bf16* hid;
                   float* enc_out; float* geo_sum; float* geo_vec; float* geo_small; } lane[2];
  int has_scale = 0, has_ref = 0;
  long long Lenc = 0, L = 0;  // encoder rows V*(T+1); info-sharing rows V*T (+1)
  // encoder
  bf16* im2col; float* patch; float* pos; float* xs; bf16* ln; bf16* qkv; bf16* att; bf16* hid;
  float* enc_out; bf16* fused; float* geo_sum; float* geo_vec; float* geo_small;
  // info sharing
  float* ys; int* view_rows; bf16* tap[3]; bf16* fin; float* scale_tok; float* carry_o; float* carry_ml;
  // heads (small)
  float* scale_tmp; float* pose_pool; float* pose_tmp; float* pose_raw; bf16* ph_in; bf16* ph_a; bf16* ph_b; bf16* ph_c; bf16* ph_d;
  // dense head chunk buffers: bf16, or fp32 when cfg.head_tf32 (typed by stage_heads_range<T>)
  int vc = 0;
  void* d_a; void* d_s; void* d_u; void* lvl[4]; void* lvl_relu[4]; void* t0; void* t1; void* t2; void* t3;
  void* path; void* feat8; void* r1; void* r2;
  // fp32 copies of the dense head's inputs (cfg.head_tf32): the reference hands the head fp32 features
  float* fused32 = nullptr; float* tap32[3] = {nullptr, nullptr, nullptr}; float* fin32 = nullptr;
  size_t total = 0;
};

}  // namespace

struct mab_handle {
  mab_config cfg;
  int device = 0;
  bool finalized = false;
  std::map<std::string, RawTensor> raw;       // fp32 device copies, canonical keys
  std::map<std::string, bf16*> packed;        // bf16 GEMM / conv operands
  std::map<std::string, float*> packed32;     // fp32 (TF32) GEMM / conv operands, padded layouts
  std::map<std::string, float*> extra;        // derived fp32 tensors (tiled biases)
  std::vector<void*> owned;
  std::vector<std::string> packed_keys;        // raw tensors that only fed a pack_* kernel: freed by finalize
  std::vector<int> view_pe_rows;               // mab_set_view_pe_rows: one table row per view of the shard
  // view-sharded forward inside the library (mab_set_comm / mab_forward_sharded): NCCL entry points resolved
  // with dlopen from the library the caller names (no link-time dependency), the caller's communicator
  struct Nccl {
    void* lib = nullptr; void* comm = nullptr; int rank = 0, world = 1;
    int (*GroupStart)() = nullptr; int (*GroupEnd)() = nullptr;
    int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
  } nccl;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_kv_ready = nullptr, ev_kv_done = nullptr, ev_kv_far = nullptr;
  cudaStream_t lane_stream = nullptr;          // second encoder lane (the first is the caller's stream)
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  Plan plan;

  const float* F(const std::string& k) const {
    auto it = raw.find(k);
    return it == raw.end() ? nullptr : it->second.p;
  }
  const bf16* P(const std::string& k) const {
    auto it = packed.find(k);
    return it == packed.end() ? nullptr : it->second;
  }
  const float* P32(const std::string& k) const {
    auto it = packed32.find(k);
    return it == packed32.end() ? nullptr : it->second;
  }
};

namespace {

void make_plan(const mab_handle* hd, int V, int Vtot, int H, int W, int has_scale, int has_ref,
               void* ws, Plan* pl) {
  const mab_config& c = hd->cfg;
  Plan& p = *pl;
  p.V = V; p.Vtot = Vtot; p.H = H; p.W = W; p.h = H / c.patch; p.w = W / c.patch; p.T = p.h * p.w;
  p.has_scale = has_scale; p.has_ref = has_ref;
  p.Lenc = (long long)V * (p.T + 1);
  p.L = (long long)V * p.T + (has_scale ? 1 : 0);
  const long long VT = (long long)V * p.T;
  const int D = c.enc_dim, A = c.aat_dim;
  Bump b; b.base = reinterpret_cast<uint8_t*>(ws);
  // The encoder (a3-a6) is per-view independent and runs Ve views at a time (mab_stage_encode), so its
  // activations cost Ve, not V, views of memory; what stays resident per view is the fused token (bf16 +
  // fp32 for the fp32 head), the info-sharing stream, its taps and the info-sharing activations, which
  // share the ln | qkv | att | hid buffers with the encoder (each sized for the larger of the two uses).
  p.Ve = (c.enc_chunk > 0 && c.enc_chunk < V) ? c.enc_chunk : V;
  p.lanes = (p.Ve >= 2 && c.enc_lanes == 2) ? 2 : 1;
  p.Vl = (p.Ve + p.lanes - 1) / p.lanes;
  const long long VTe = (long long)p.Vl * p.T, Le = (long long)p.Vl * (p.T + 1);
  auto mx = [](long long a, long long b2) { return a > b2 ? a : b2; };
  p.pos = b.take<float>((long long)(p.T + 1) * D);
  for (int l = 0; l < p.lanes; ++l) {
    Plan::EncBufs& e = p.lane[l];
    // lane 0's ln | qkv | att | hid also serve the info-sharing blocks (all local tokens at width A)
    const long long rows_a = l == 0 ? VT + 1 : 0;
    e.im2col = b.take<bf16>(VTe * 640);
    e.patch = b.take<float>(VTe * D);
    e.xs = b.take<float>(Le * D);
    e.ln = b.take<bf16>(mx(Le * D, rows_a * A));
    e.qkv = b.take<bf16>(mx(Le * 3 * D, rows_a * 3 * A));
    e.att = b.take<bf16>(mx(Le * D, rows_a * A));
    e.hid = b.take<bf16>(mx(Le * 4 * D, rows_a * 4 * A));
    e.enc_out = b.take<float>(VTe * D);
    e.geo_sum = b.take<float>(VTe * D);          // sum of the masked dense geometric features
    e.geo_vec = b.take<float>((long long)p.Vl * D); // sum of the masked per-view (global) features
    e.geo_small = b.take<float>((long long)p.Vl * 4096 + 64);  // MLP scratch + masks + scalar inputs
  }
  if (p.lanes == 1) p.lane[1] = p.lane[0];
  p.im2col = p.lane[0].im2col; p.patch = p.lane[0].patch; p.xs = p.lane[0].xs; p.ln = p.lane[0].ln;
  p.qkv = p.lane[0].qkv; p.att = p.lane[0].att; p.hid = p.lane[0].hid; p.enc_out = p.lane[0].enc_out;
  p.geo_sum = p.lane[0].geo_sum; p.geo_vec = p.lane[0].geo_vec; p.geo_small = p.lane[0].geo_small;
  p.fused = b.take<bf16>((VT + 1) * D);
  p.ys = b.take<float>((VT + 1) * A);
  p.view_rows = b.take<int>(V);  // view_pos_table row of every view (use_pe_for_non_reference_views)
  const bool tf = c.head_tf32 != 0;
  p.tap[0] = p.tap[1] = p.tap[2] = p.fin = nullptr;
  p.fused32 = p.tap32[0] = p.tap32[1] = p.tap32[2] = p.fin32 = nullptr;
  if (tf) {
    p.fused32 = c.aat_tap2 > 0 ? nullptr : b.take<float>(VT * D);  // three taps: the head takes no encoder features
    p.tap32[0] = b.take<float>(VT * A);
    p.tap32[1] = b.take<float>(VT * A);
    p.tap32[2] = c.aat_tap2 > 0 ? b.take<float>(VT * A) : nullptr;
    p.fin32 = b.take<float>(VT * A);
  } else {
    p.tap[0] = b.take<bf16>(VT * A);
    p.tap[1] = b.take<bf16>(VT * A);
    p.tap[2] = c.aat_tap2 > 0 ? b.take<bf16>(VT * A) : nullptr;  // third intermediate feature (48-layer configs)
    p.fin = b.take<bf16>(VT * A);
  }
  p.scale_tok = b.take<float>(A);
  p.carry_o = Vtot > V ? b.take<float>((VT + 1) * A) : nullptr;   // ring attention state
  p.carry_ml = Vtot > V ? b.take<float>((VT + 1) * (A / 64) * 2) : nullptr;
  p.scale_tmp = b.take<float>(4 * 256);
  p.pose_pool = b.take<float>((long long)V * 784);
  p.pose_tmp = b.take<float>((long long)V * 784 * 2);
  p.pose_raw = b.take<float>((long long)V * 8);
  // pose + dense head, per chunk of vc views
  p.vc = c.head_chunk < 1 ? 1 : (c.head_chunk > V ? V : c.head_chunk);
  const long long vc = p.vc, T = p.T;
  p.ph_in = b.take<bf16>(vc * T * 3 * A);
  p.ph_a = b.take<bf16>(vc * T * 3 * 784);
  p.ph_b = b.take<bf16>(vc * T * 3 * 784);
  p.ph_c = b.take<bf16>(vc * T * 3 * 784);
  p.ph_d = b.take<bf16>(vc * T * 3 * 784);
  const int h = p.h, w = p.w;
  const long long es = tf ? 4 : 2;  // bytes per activation element of the dense head
  auto act = [&](long long elems) { return (void*)b.take<uint8_t>(elems * es); };
  p.d_a = act(vc * T * 768);
  p.d_s = act(vc * T * 6912 / 1);  // >= T*1536 (convT) and >= Ho*Wo*9*768 (im2col s2)
  p.d_u = act(vc * 16 * T * 96);   // 4h x 4w x 96 == 2h x 2w x 192 * 2
  const long long lvl_pix[4] = {16LL * T, 4LL * T, T, (long long)((h + 1) / 2) * ((w + 1) / 2)};
  for (int i = 0; i < 4; ++i) {
    p.lvl[i] = act(vc * lvl_pix[i] * 256);
    p.lvl_relu[i] = act(vc * lvl_pix[i] * 256);
  }
  const long long big = vc * 16 * T * 256;  // 4h x 4w x 256
  p.t0 = act(big); p.t1 = act(big); p.t2 = act(big); p.t3 = act(big);
  p.path = act(big);
  p.feat8 = act(vc * 64 * T * 256);
  p.r1 = act(vc * 64 * T * 128);
  p.r2 = act(vc * (long long)H * W * 128);
  p.total = b.off + 256;
}

}  // namespace

extern "C" int mab_create(const mab_config* cfg, int device, mab_handle** out) {
  MAB_ARG(cfg && out, "null argument");
  MAB_ARG(cfg->patch == 14 && cfg->enc_dim == 1024 && (cfg->aat_dim == 768 || cfg->aat_dim == 1024),
          "unsupported architecture");
  MAB_ARG(cfg->enc_heads * 64 == cfg->enc_dim && cfg->aat_heads * 64 == cfg->aat_dim, "head_dim must be 64");
  MAB_ARG(cfg->enc_depth > 0 && cfg->enc_pos_grid > 0 && cfg->aat_depth > 0 && cfg->head_chunk > 0,
          "depths, position grid and head_chunk must be positive");
  MAB_ARG(cfg->aat_tap0 >= 0 && cfg->aat_tap0 < cfg->aat_tap1 && cfg->aat_tap1 < cfg->aat_depth,
          "intermediate feature indices must satisfy 0 <= tap0 < tap1 < depth");
  MAB_ARG(cfg->aat_tap2 <= 0 || (cfg->aat_tap1 < cfg->aat_tap2 && cfg->aat_tap2 < cfg->aat_depth),
          "a third intermediate feature index must satisfy tap1 < tap2 < depth");
  MAB_ARG(!cfg->aat_nonref_view_pe || cfg->aat_ref_view_pe,
          "view PE of the non-reference views needs the reference-view PE (aat.py:64)");
  MAB_ARG(device >= 0, "bad device");
  MAB_CUDA(cudaSetDevice(device));
  mab_handle* h = new mab_handle();
  h->cfg = *cfg;
  h->device = device;
  // set-up time only: the second encoder lane's stream and the two events that fork / join it
  cudaError_t e1 = cudaStreamCreateWithFlags(&h->lane_stream, cudaStreamNonBlocking);
  cudaError_t e2 = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming);
  cudaError_t e3 = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming);
  if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) {
    delete h;
    return fail((int)(e1 != cudaSuccess ? e1 : e2 != cudaSuccess ? e2 : e3), "could not create the encoder lane stream / events");
  }
  *out = h;
  return 0;
}

extern "C" int mab_destroy(mab_handle* h) {
  if (!h) return 0;
  for (auto& kv : h->raw) cudaFree(kv.second.p);
  for (void* p : h->owned) cudaFree(p);
  if (h->comm_stream) cudaStreamDestroy(h->comm_stream);
  if (h->ev_kv_ready) cudaEventDestroy(h->ev_kv_ready);
  if (h->ev_kv_done) cudaEventDestroy(h->ev_kv_done);
  if (h->ev_kv_far) cudaEventDestroy(h->ev_kv_far);
  if (h->nccl.lib) dlclose(h->nccl.lib);   // the communicator stays the caller's
  if (h->lane_stream) cudaStreamDestroy(h->lane_stream);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  delete h;
  return 0;
}

// view_pos_table row of every view this rank holds, for the next mab_stage_encode / mab_forward of a
// handle with aat_nonref_view_pe (host array, copied; -1 = no encoding for that view)
extern "C" int mab_set_view_pe_rows(mab_handle* h, const int* rows, int n) {
  MAB_ARG(h && rows && n > 0, "bad arguments");
  MAB_ARG(h->cfg.aat_nonref_view_pe, "the handle was created without aat_nonref_view_pe");
  for (int i = 0; i < n; ++i) MAB_ARG(rows[i] >= -1, "rows must be >= -1");
  h->view_pe_rows.assign(rows, rows + n);
  return 0;
}

extern "C" int mab_set_weight(mab_handle* h, const char* ref_key, const void* ptr, int dtype,
                              int ndim, const int64_t* shape) {
  MAB_ARG(h && ref_key && ptr && shape, "null argument");
  MAB_ARG(dtype == MAB_F32, "weights must be fp32");
  MAB_ARG(!h->finalized, "weights are frozen after mab_finalize_weights");
  MAB_ARG(ndim >= 1 && ndim <= 8, "ndim must be in 1..8");
  long long n = 1;
  for (int i = 0; i < ndim; ++i) {
    MAB_ARG(shape[i] > 0 && shape[i] < (1LL << 31), "every dimension must be positive");
    n *= shape[i];
  }
  MAB_CUDA(cudaSetDevice(h->device));
  const std::string key = canonical_key(ref_key);
  RawTensor& t = h->raw[key];
  if (t.p && t.numel != n) { cudaFree(t.p); t.p = nullptr; }
  if (!t.p) MAB_CUDA(cudaMalloc(&t.p, (size_t)n * sizeof(float)));
  t.numel = n;
  t.shape.assign(shape, shape + ndim);
  MAB_CUDA(cudaMemcpy(t.p, ptr, (size_t)n * sizeof(float), cudaMemcpyDefault));
  return 0;
}

// ================================================================================================
// Part 3: weight packing
// ================================================================================================
namespace {

int pack_linear(mab_handle* h, const std::string& key, int ld_pad, cudaStream_t st) {
  auto it = h->raw.find(key);
  if (it == h->raw.end()) return fail(-4, "missing weight %s", key.c_str());
  const RawTensor& t = it->second;
  const int rows = (int)t.shape[0];
  const int cols = (int)(t.numel / rows);
  const int ld = ld_pad > 0 ? ld_pad : cols;
  bf16* out;
  MAB_CUDA(cudaMalloc(&out, (size_t)rows * ld * sizeof(bf16)));
  h->owned.push_back(out);
  pack_matrix_kernel<bf16><<<ew_grid((long long)rows * ld), 256, 0, st>>>(t.p, rows, cols, ld, out);
  MAB_CUDA(cudaGetLastError());
  h->packed[key] = out;
  h->packed_keys.push_back(key);
  return 0;
}

int pack_linear_split(mab_handle* h, const std::string& key, cudaStream_t st) {
  auto it = h->raw.find(key);
  if (it == h->raw.end()) return fail(-4, "missing weight %s", key.c_str());
  const RawTensor& t = it->second;
  const int rows = (int)t.shape[0];
  const int cols = (int)(t.numel / rows);
  bf16* out;
  MAB_CUDA(cudaMalloc(&out, (size_t)rows * 3 * cols * sizeof(bf16)));
  h->owned.push_back(out);
  pack_matrix_split_kernel<<<ew_grid((long long)rows * cols), 256, 0, st>>>(t.p, rows, cols, out);
  MAB_CUDA(cudaGetLastError());
  h->packed[key] = out;
  h->packed_keys.push_back(key);
  return 0;
}

int pack_conv3(mab_handle* h, const std::string& key, cudaStream_t st) {
  auto it = h->raw.find(key);
  if (it == h->raw.end()) return fail(-4, "missing weight %s", key.c_str());
  const RawTensor& t = it->second;
  const int Cout = (int)t.shape[0], Cin = (int)t.shape[1];
  const int Cpad = (Cin + 63) / 64 * 64;
  bf16* out;
  MAB_CUDA(cudaMalloc(&out, (size_t)Cout * 9 * Cpad * sizeof(bf16)));
  h->owned.push_back(out);
  pack_conv3_kernel<bf16><<<ew_grid((long long)Cout * 9 * Cpad), 256, 0, st>>>(t.p, Cout, Cin, Cpad, out);
  MAB_CUDA(cudaGetLastError());
  h->packed[key] = out;
  h->packed_keys.push_back(key);
  return 0;
}

// fp32 operands of the TF32 kernels: [rows, ld] zero padded / [Cout, 9, ceil32(Cin) or cpad]
int pack_linear_f32(mab_handle* h, const std::string& key, int ld_pad, cudaStream_t st) {
  auto it = h->raw.find(key);
  if (it == h->raw.end()) return fail(-4, "missing weight %s", key.c_str());
  const RawTensor& t = it->second;
  const int rows = (int)t.shape[0];
  const int cols = (int)(t.numel / rows);
  const int ld = ld_pad > 0 ? ld_pad : cols;
  float* out;
  MAB_CUDA(cudaMalloc(&out, (size_t)rows * ld * sizeof(float)));
  h->owned.push_back(out);
  pack_matrix_kernel<float><<<ew_grid((long long)rows * ld), 256, 0, st>>>(t.p, rows, cols, ld, out);
  MAB_CUDA(cudaGetLastError());
  h->packed32[key] = out;
  h->packed_keys.push_back(key);
  return 0;
}
int pack_conv3_f32(mab_handle* h, const std::string& key, int cpad, cudaStream_t st) {
  auto it = h->raw.find(key);
  if (it == h->raw.end()) return fail(-4, "missing weight %s", key.c_str());
  const RawTensor& t = it->second;
  const int Cout = (int)t.shape[0], Cin = (int)t.shape[1];
  float* out;
  MAB_CUDA(cudaMalloc(&out, (size_t)Cout * 9 * cpad * sizeof(float)));
  h->owned.push_back(out);
  pack_conv3_kernel<float><<<ew_grid((long long)Cout * 9 * cpad), 256, 0, st>>>(t.p, Cout, Cin, cpad, out);
  MAB_CUDA(cudaGetLastError());
  h->packed32[key] = out;
  h->packed_keys.push_back(key);
  return 0;
}

int pack_convT(mab_handle* h, const std::string& wkey, const std::string& bkey, cudaStream_t st, bool f32 = false) {
  auto it = h->raw.find(wkey);
  auto ib = h->raw.find(bkey);
  if (it == h->raw.end() || ib == h->raw.end()) return fail(-4, "missing weight %s", wkey.c_str());
  const RawTensor& t = it->second;
  const int Cin = (int)t.shape[0], Cout = (int)t.shape[1], k = (int)t.shape[2];
  void* out;
  MAB_CUDA(cudaMalloc(&out, (size_t)k * k * Cout * Cin * (f32 ? sizeof(float) : sizeof(bf16))));
  h->owned.push_back(out);
  if (f32)
    pack_convT_kernel<float><<<ew_grid((long long)k * k * Cout * Cin), 256, 0, st>>>(t.p, Cin, Cout, k, (float*)out);
  else
    pack_convT_kernel<bf16><<<ew_grid((long long)k * k * Cout * Cin), 256, 0, st>>>(t.p, Cin, Cout, k, (bf16*)out);
  float* bt;
  MAB_CUDA(cudaMalloc(&bt, (size_t)k * k * Cout * sizeof(float)));
  h->owned.push_back(bt);
  tile_bias_kernel<<<(k * k * Cout + 255) / 256, 256, 0, st>>>(ib->second.p, Cout, k * k, bt);
  MAB_CUDA(cudaGetLastError());
  if (f32) h->packed32[wkey] = (float*)out;
  else h->packed[wkey] = (bf16*)out;
  h->packed_keys.push_back(wkey);
  h->extra[bkey + ".tiled"] = bt;
  return 0;
}

}  // namespace

extern "C" int mab_finalize_weights(mab_handle* h, void* stream) {
  MAB_ARG(h, "null handle");
  if (h->finalized) return 0;  // the weights are frozen: nothing new to pack
  MAB_CUDA(cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;
  const mab_config& c = h->cfg;
  MAB_TRY(pack_linear(h, "encoder.model.patch_embed.proj.weight", 640, st));
  for (int i = 0; i < c.enc_depth; ++i) {
    const std::string p = "encoder.model.blocks." + std::to_string(i);
    for (const char* s : {".attn.qkv.weight", ".attn.proj.weight", ".mlp.fc1.weight", ".mlp.fc2.weight"})
      MAB_TRY(pack_linear(h, p + s, 0, st));
  }
  // aat.py:121-124: nn.Identity when the trunk is as wide as the encoder
  if (c.aat_dim != c.enc_dim) MAB_TRY(pack_linear(h, "info_sharing.proj_embed.weight", 0, st));
  for (int i = 0; i < c.aat_depth; ++i) {
    const std::string p = "info_sharing.self_attention_blocks." + std::to_string(i);
    for (const char* s : {".attn.qkv.weight", ".attn.proj.weight", ".mlp.fc1.weight", ".mlp.fc2.weight"})
      MAB_TRY(pack_linear(h, p + s, 0, st));
  }
  const std::string ip = "dpt_feature_head.input_process.";
  const std::string sc = "dpt_feature_head.scratch.";
  if (c.head_tf32) {
    // dense head with fp32 activations and TF32 contractions, the reference's own precision for this
    // region (autocast off, model.py:1774; dpt.py:180-232,285-311): fp32 operands, channels already
    // multiples of 32 (96, 192, 384, 768, 256, 128)
    auto cin_of = [&](const std::string& k) { auto it = h->raw.find(k); return it == h->raw.end() ? 0 : (int)it->second.shape[1]; };
    auto conv32 = [&](const std::string& k) {
      const int cin = cin_of(k);
      if (cin <= 0 || cin % 32) return fail(-4, "missing weight %s (or input channels not a multiple of 32)", k.c_str());
      return pack_conv3_f32(h, k, cin, st);
    };
    for (int i = 0; i < 4; ++i) MAB_TRY(pack_linear_f32(h, ip + std::to_string(i) + ".0.0.weight", 0, st));
    MAB_TRY(pack_convT(h, ip + "0.0.1.weight", ip + "0.0.1.bias", st, true));
    MAB_TRY(pack_convT(h, ip + "1.0.1.weight", ip + "1.0.1.bias", st, true));
    MAB_TRY(conv32(ip + "3.0.1.weight"));
    for (int i = 1; i <= 4; ++i) {
      MAB_TRY(conv32(sc + "layer" + std::to_string(i) + "_rn.weight"));
      const std::string r = sc + "refinenet" + std::to_string(i);
      for (const char* u : {".resConfUnit1", ".resConfUnit2"}) {
        if (i == 4 && std::string(u) == ".resConfUnit1") continue;  // dpt.py:82-83
        MAB_TRY(conv32(r + u + ".conv1.weight"));
        MAB_TRY(conv32(r + u + ".conv2.weight"));
      }
      MAB_TRY(pack_linear_f32(h, r + ".out_conv.weight", 0, st));
    }
    MAB_TRY(conv32("dpt_regressor_head.conv1.weight"));
    MAB_TRY(conv32("dpt_regressor_head.conv2.0.weight"));
  } else {
    for (int i = 0; i < 4; ++i) MAB_TRY(pack_linear(h, ip + std::to_string(i) + ".0.0.weight", 0, st));
    MAB_TRY(pack_convT(h, ip + "0.0.1.weight", ip + "0.0.1.bias", st));
    MAB_TRY(pack_convT(h, ip + "1.0.1.weight", ip + "1.0.1.bias", st));
    MAB_TRY(pack_conv3(h, ip + "3.0.1.weight", st));
    for (int i = 1; i <= 4; ++i) {
      MAB_TRY(pack_conv3(h, sc + "layer" + std::to_string(i) + "_rn.weight", st));
      const std::string r = sc + "refinenet" + std::to_string(i);
      for (const char* u : {".resConfUnit1", ".resConfUnit2"}) {
        if (i == 4 && std::string(u) == ".resConfUnit1") continue;  // dpt.py:82-83
        MAB_TRY(pack_conv3(h, r + u + ".conv1.weight", st));
        MAB_TRY(pack_conv3(h, r + u + ".conv2.weight", st));
      }
      MAB_TRY(pack_linear(h, r + ".out_conv.weight", 0, st));
    }
    MAB_TRY(pack_conv3(h, "dpt_regressor_head.conv1.weight", st));
    MAB_TRY(pack_conv3(h, "dpt_regressor_head.conv2.0.weight", st));
  }
  // pose head: bf16x3 operands (the reference runs it in fp32/TF32, model.py:1774; its output is
  // normalised to a unit quaternion, which plain bf16 operands do not resolve to 0.1 deg)
  MAB_TRY(pack_linear_split(h, "pose_head.proj.weight", st));
  for (int b = 0; b < 2; ++b)
    for (int k = 1; k <= 3; ++k)
      MAB_TRY(pack_linear_split(h, "pose_head.res_conv." + std::to_string(b) + ".res_conv" + std::to_string(k) + ".weight", st));
  // geometric-input encoders (model.py:176-210): dense conv stacks as bf16 implicit GEMMs
  for (const char* enc : {"ray_dirs_encoder", "depth_encoder"}) {
    const std::string e = enc;
    if (!h->F(e + ".conv_in.weight")) continue;
    // fp32 in the reference (autocast off, TF32 on GPU; model.py:1377): kept as fp32/TF32 operands
    MAB_TRY(pack_conv3_f32(h, e + ".conv_in.weight", e == "depth_encoder" ? 256 : 640, st));
    for (int i = 0; i < 2; ++i) {
      const std::string r = e + ".encoder." + std::to_string(i);
      MAB_TRY(pack_conv3_f32(h, r + ".conv1.weight", i == 0 ? 640 : 768, st));
      MAB_TRY(pack_conv3_f32(h, r + ".conv2.weight", i == 0 ? 768 : 1024, st));
      MAB_TRY(pack_linear_f32(h, r + ".shortcut.weight", i == 0 ? 640 : 0, st));
    }
    MAB_TRY(pack_linear_f32(h, e + ".encoder.2.weight", 0, st));
  }
  // required fp32 tensors that are used as-is
  for (const char* k : {"encoder.model.cls_token", "encoder.model.pos_embed", "encoder.model.norm.weight",
                        "fusion_norm_layer.weight", "scale_token", "info_sharing.norm.weight", "dpt_regressor_head.conv2.2.weight",
                        "pose_head.more_mlps.0.weight", "pose_head.fc_t.weight", "pose_head.fc_rot.weight",
                        "scale_head.proj.weight", "scale_head.output_proj.weight"})
    if (!h->F(k)) return fail(-4, "missing weight %s", k);
  // the buffer only exists in checkpoints with the reference-view PE (aat.py:160-172)
  if (h->cfg.aat_ref_view_pe && !h->F("info_sharing.view_pos_table"))
    return fail(-4, "missing weight %s", "info_sharing.view_pos_table");
  // fc_t | fc_rot fused into one [7,784] matrix + [7] bias for the pose tail
  {
    float* w7; float* b7;
    MAB_CUDA(cudaMalloc(&w7, 7 * 784 * sizeof(float)));
    MAB_CUDA(cudaMalloc(&b7, 8 * sizeof(float)));
    h->owned.push_back(w7); h->owned.push_back(b7);
    MAB_CUDA(cudaMemcpyAsync(w7, h->F("pose_head.fc_t.weight"), 3 * 784 * 4, cudaMemcpyDeviceToDevice, st));
    MAB_CUDA(cudaMemcpyAsync(w7 + 3 * 784, h->F("pose_head.fc_rot.weight"), 4 * 784 * 4, cudaMemcpyDeviceToDevice, st));
    MAB_CUDA(cudaMemcpyAsync(b7, h->F("pose_head.fc_t.bias"), 3 * 4, cudaMemcpyDeviceToDevice, st));
    MAB_CUDA(cudaMemcpyAsync(b7 + 3, h->F("pose_head.fc_rot.bias"), 4 * 4, cudaMemcpyDeviceToDevice, st));
    h->extra["pose_head.fc7.weight"] = w7;
    h->extra["pose_head.fc7.bias"] = b7;
  }
  MAB_CUDA(cudaStreamSynchronize(st));  // one-time setup, not on the inference path
  // the fp32 staging copies of every packed GEMM / conv operand are dead now (> 2 GB for the production
  // model); biases, norms, LayerScale vectors, tokens, tables and the small fp32 heads stay (read through F())
  for (const std::string& k : h->packed_keys) {
    auto it = h->raw.find(k);
    if (it != h->raw.end() && it->second.p) {
      cudaFree(it->second.p);
      it->second.p = nullptr;
    }
  }
  h->packed_keys.clear();
  h->finalized = true;
  return 0;
}

// test/diagnostic hook: device pointer of a named intermediate of the last forward
extern "C" int mab_debug_buffer(mab_handle* h, const char* name, void** ptr, long long* numel,
                                int* dtype) {
  MAB_ARG(h && name && ptr && numel && dtype && h->plan.V > 0, "no forward has run");
  const Plan& p = h->plan;
  const long long VT = (long long)p.V * p.T;
  const int D = h->cfg.enc_dim, A = h->cfg.aat_dim;
  const std::string n = name;
  if (n == "enc_tokens") { *ptr = p.xs; *numel = (long long)p.Vl * (p.T + 1) * D; *dtype = MAB_F32; }  // lane 0, its last pass
  else if (n == "fused" && p.fused32) { *ptr = p.fused32; *numel = VT * D; *dtype = MAB_F32; }
  else if (n == "fused") { *ptr = p.fused; *numel = VT * D; *dtype = MAB_BF16; }
  else if (n == "aat_stream") { *ptr = p.ys; *numel = p.L * A; *dtype = MAB_F32; }
  else if (n == "tap0" && p.tap32[0]) { *ptr = p.tap32[0]; *numel = VT * A; *dtype = MAB_F32; }
  else if (n == "tap1" && p.tap32[1]) { *ptr = p.tap32[1]; *numel = VT * A; *dtype = MAB_F32; }
  else if (n == "tap2" && p.tap32[2]) { *ptr = p.tap32[2]; *numel = VT * A; *dtype = MAB_F32; }
  else if (n == "final" && p.fin32) { *ptr = p.fin32; *numel = VT * A; *dtype = MAB_F32; }
  else if (n == "tap0" && p.tap[0]) { *ptr = p.tap[0]; *numel = VT * A; *dtype = MAB_BF16; }
  else if (n == "tap1" && p.tap[1]) { *ptr = p.tap[1]; *numel = VT * A; *dtype = MAB_BF16; }
  else if (n == "tap2" && p.tap[2]) { *ptr = p.tap[2]; *numel = VT * A; *dtype = MAB_BF16; }
  else if (n == "final" && p.fin) { *ptr = p.fin; *numel = VT * A; *dtype = MAB_BF16; }
  else if (n == "scale_token") { *ptr = p.scale_tok; *numel = A; *dtype = MAB_F32; }
  else if (n == "pose_raw") { *ptr = p.pose_raw; *numel = (long long)p.V * 7; *dtype = MAB_F32; }
  else if (n == "pose_pool") { *ptr = p.pose_pool; *numel = (long long)p.V * 784; *dtype = MAB_F32; }
  else if (n == "feat8") { *ptr = p.feat8; *numel = (long long)p.vc * 64 * p.T * 256; *dtype = h->cfg.head_tf32 ? MAB_F32 : MAB_BF16; }
  else return fail(-1, "unknown debug buffer %s", name);
  return 0;
}

extern "C" size_t mab_workspace_bytes(mab_handle* h, int views_local, int views_total, int H, int W) {
  if (!h || views_local <= 0 || H % h->cfg.patch || W % h->cfg.patch) return 0;
  Plan p;
  make_plan(h, views_local, views_total > views_local ? views_total : views_local, H, W, 1, 1, nullptr, &p);
  return p.total;
}

// ================================================================================================
// Part 4: forward stages
// ================================================================================================
namespace {

struct Ctx {
  mab_handle* h;
  cudaStream_t st;
  // y = epilogue(A[M,K] W[N,K]^T); W, bias looked up by reference key
  int linear(const void* A, int M, int K, long long lda, const std::string& wkey, int N,
             const float* bias, int act, const float* gamma, const float* res_f32, float* out_f32,
             bf16* out_bf16, const bf16* res1 = nullptr, long long ldw = 0, int k_algorithmic = 0) {
    const bf16* W = h->P(wkey);
    if (!W) return fail(-4, "weight %s was not packed", wkey.c_str());
    GemmEpilogue ep{};
    ep.bias = bias; ep.gamma = gamma; ep.res_f32 = res_f32; ep.out_f32 = out_f32;
    ep.out_bf16 = out_bf16; ep.res1 = res1; ep.ldo = N; ep.act = act;
    return launch_gemm(A, lda, W, ldw > 0 ? ldw : K, M, N, K, ep, st, k_algorithmic);
  }
  // bf16x3 linear: A rows are [hi | lo | hi] (pitch 3K), W rows [hi | hi | lo]; the output is
  // written in the same split form (pitch 3N) so that layers chain; the residual is a split value
  int linear_x3(const bf16* A, int M, int K, const std::string& wkey, int N, const float* bias, int act,
                const bf16* res_split, bf16* out_split) {
    const bf16* W = h->P(wkey);
    if (!W) return fail(-4, "weight %s was not packed", wkey.c_str());
    GemmEpilogue ep{};
    ep.bias = bias; ep.act = act; ep.out_bf16 = out_split; ep.ldo = 3 * N; ep.split_n = N;
    ep.res1 = res_split; ep.res2 = res_split ? res_split + N : nullptr;
    return launch_gemm(A, 3LL * K, W, 3LL * K, M, N, 3 * K, ep, st, K);
  }
  int conv3(const bf16* x, int batch, int H, int W, int C, const std::string& wkey, int Cout,
            const float* bias, int act, const bf16* res1, const bf16* res2, bf16* out, bf16* out_relu) {
    const bf16* Wp = h->P(wkey);
    if (!Wp) return fail(-4, "weight %s was not packed", wkey.c_str());
    GemmEpilogue ep{};
    ep.bias = bias; ep.res1 = res1; ep.res2 = res2; ep.out_bf16 = out; ep.out_relu = out_relu;
    ep.ldo = Cout; ep.act = act;
    return launch_conv3(x, batch, H, W, C, Wp, Cout, ep, st);
  }
};

// ---- dense-head operators, typed by the activation element: bf16 (bf16 operands) or float (fp32
// activations, TF32 contractions -- the reference's precision for this region, model.py:1774) ----------
int head_linear(Ctx& c, const bf16* A, int M, int K, const std::string& wkey, int N, const float* bias, bf16* out) {
  return c.linear(A, M, K, K, wkey, N, bias, ACT_NONE, nullptr, nullptr, nullptr, out);
}
int head_linear(Ctx& c, const float* A, int M, int K, const std::string& wkey, int N, const float* bias, float* out) {
  const float* W = c.h->P32(wkey);
  if (!W) return fail(-4, "weight %s was not packed (fp32)", wkey.c_str());
  GemmEpilogue ep{};
  ep.bias = bias; ep.out_f32 = out; ep.ldo = N; ep.round_tf32 = 1;
  return launch_gemm_tf32(A, K, W, K, M, N, K, ep, c.st);
}
int head_conv3(Ctx& c, const bf16* x, int batch, int H, int W, int C, const std::string& wkey, int Cout,
               const float* bias, int act, const bf16* res1, const bf16* res2, bf16* out, bf16* out_relu) {
  return c.conv3(x, batch, H, W, C, wkey, Cout, bias, act, res1, res2, out, out_relu);
}
int head_conv3(Ctx& c, const float* x, int batch, int H, int W, int C, const std::string& wkey, int Cout,
               const float* bias, int act, const float* res1, const float* res2, float* out, float* out_relu) {
  const float* Wp = c.h->P32(wkey);
  if (!Wp) return fail(-4, "weight %s was not packed (fp32)", wkey.c_str());
  GemmEpilogue ep{};
  ep.bias = bias; ep.res_f32 = res1; ep.res2_f32 = res2; ep.out_f32 = out; ep.out_relu_f32 = out_relu;
  ep.ldo = Cout; ep.act = act; ep.round_tf32 = 1;
  return launch_conv3_tf32(x, batch, H, W, C, C, Wp, Cout, ep, c.st);
}
// the two layout kernels move 16-byte units: an fp32 tensor is a bf16 tensor of twice the channels
int head_pixel_shuffle(Ctx& c, const void* in, int es, int views, int h, int w, int k, int C, void* out) {
  const int cb = C * es / 2;  // channels in bf16 units
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
             pixel_shuffle_kernel<<<ew_grid((long long)views * h * w * k * k * (cb / 8)), 256, 0, c.st>>>(
                 reinterpret_cast<const bf16*>(in), views, h, w, k, cb, reinterpret_cast<bf16*>(out)));
  MAB_CUDA(cudaGetLastError());
  return 0;
}
int head_im2col_s2(Ctx& c, const void* in, int es, int views, int H, int W, int C, int Ho, int Wo, void* out) {
  const int cb = C * es / 2;
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
             im2col3x3_s2_kernel<<<ew_grid((long long)views * Ho * Wo * 9 * (cb / 8)), 256, 0, c.st>>>(
                 reinterpret_cast<const bf16*>(in), views, H, W, cb, Ho, Wo, reinterpret_cast<bf16*>(out)));
  MAB_CUDA(cudaGetLastError());
  return 0;
}
int head_resize(Ctx& c, const bf16* x, int batch, int Hi, int Wi, int C, int Hv, int Wv, int Ho, int Wo, bf16* out) {
  return mab_resize_bilinear(x, batch, Hi, Wi, C, Hv, Wv, Ho, Wo, out, nullptr, c.st);
}
int head_resize(Ctx& c, const float* x, int batch, int Hi, int Wi, int C, int Hv, int Wv, int Ho, int Wo, float* out) {
  MAB_ARG(batch > 0 && batch <= 65535 && Ho > 0 && Ho <= 65535 && Wo > 0 && C % 4 == 0, "resize: bad sizes");
  const dim3 grid((Wo * (C / 4) + 255) / 256, (Ho + RESIZE_ROWS - 1) / RESIZE_ROWS, batch);
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, (double)batch * C * 4.0 * ((double)Ho * Wo + (double)Hi * Wi), c.st,
             resize_bilinear_ac_f32_kernel<<<grid, 256, 0, c.st>>>(x, batch, Hi, Wi, C, Hv, Wv, Ho, Wo, out, nullptr));
  MAB_CUDA(cudaGetLastError());
  return 0;
}
int head_regress(Ctx& c, const bf16* x, int batch, int H, int W, const RegressEpilogue& rg) {
  const bf16* w2 = c.h->P("dpt_regressor_head.conv2.0.weight");
  if (!w2) return fail(-4, "dpt_regressor_head.conv2.0.weight was not packed");
  return launch_conv3_regress(x, batch, H, W, w2, rg, c.st);
}
int head_regress(Ctx& c, const float* x, int batch, int H, int W, const RegressEpilogue& rg) {
  const float* w2 = c.h->P32("dpt_regressor_head.conv2.0.weight");
  if (!w2) return fail(-4, "dpt_regressor_head.conv2.0.weight was not packed (fp32)");
  return launch_conv3_regress_tf32(x, batch, H, W, w2, rg, c.st);
}

// one pre-LN transformer block over `rows` tokens of width D (block.py:93-118,
// transformer_blocks.py:467-468): x += g1*proj(attn(LN1 x)); x += g2*fc2(gelu(fc1(LN2 x)))
int block_qkv(Ctx& c, const std::string& p, float* x, long long rows, int D, bf16* ln, bf16* qkv) {
  mab_handle* h = c.h;
  MAB_TRY(launch_layernorm(x, (int)rows, D, (int)rows, (int)rows, 0, h->F(p + ".norm1.weight"),
                           h->F(p + ".norm1.bias"), nullptr, nullptr, 1e-6f, ln, nullptr, nullptr, c.st));
  return c.linear(ln, (int)rows, D, D, p + ".attn.qkv.weight", 3 * D, h->F(p + ".attn.qkv.bias"),
                  ACT_NONE, nullptr, nullptr, nullptr, qkv);
}
int block_tail(Ctx& c, const std::string& p, float* x, long long rows, int D, bf16* att, bf16* ln,
               bf16* hid, const float* g1, const float* g2) {
  mab_handle* h = c.h;
  MAB_TRY(c.linear(att, (int)rows, D, D, p + ".attn.proj.weight", D, h->F(p + ".attn.proj.bias"),
                   ACT_NONE, g1, x, x, nullptr));
  MAB_TRY(launch_layernorm(x, (int)rows, D, (int)rows, (int)rows, 0, h->F(p + ".norm2.weight"),
                           h->F(p + ".norm2.bias"), nullptr, nullptr, 1e-6f, ln, nullptr, nullptr, c.st));
  MAB_TRY(c.linear(ln, (int)rows, D, D, p + ".mlp.fc1.weight", 4 * D, h->F(p + ".mlp.fc1.bias"),
                   ACT_GELU, nullptr, nullptr, nullptr, hid));
  return c.linear(hid, (int)rows, 4 * D, 4 * D, p + ".mlp.fc2.weight", D, h->F(p + ".mlp.fc2.bias"),
                  ACT_NONE, g2, x, x, nullptr);
}

}  // namespace

namespace {

// DenseRepresentationEncoder (dense_rep_encoder.py:234-287) on the token-major PixelUnshuffle of
// the input: conv3x3 -> 2 x ResidualBlock(conv3x3, GELU, conv3x3, + 1x1 shortcut, GELU) -> conv1x1
// -> LayerNorm; the result is accumulated into p.geo_sum scaled by the per-view mask.
// fp32 activations, TF32 tensor-core contractions (the reference's precision for this region).
// `scratch` holds 3840 floats per token: R0 640 | R1 640 | R2 768 | R3 768 | R5 1024, where the
// second block's shortcut/output reuses R0+R1 (dead by then).
int dense_rep_encode(Ctx& c, const std::string& e, float* scratch, int cin_pad, int cin_real, const float* mask,
                     bool first) {
  mab_handle* h = c.h;
  Plan& p = h->plan;
  const int V = p.V, hh = p.h, ww = p.w;
  const long long VT = (long long)V * p.T;
  float* unsh = scratch;                       // R0 [VT, cin_pad] (written by the caller)
  float* a0 = scratch + VT * 640;              // R1 conv_in output, 588 channels padded to 640
  float* x1 = a0 + VT * 640;                   // R2 shortcut of block 0, then its output (in place)
  float* b0 = x1 + VT * 768;                   // R3 conv1 output of block 0
  float* b1 = b0 + VT * 768;                   // R5 conv1 output of block 1, later the 1x1 output
  float* x2 = scratch;                         // R0+R1: shortcut of block 1, then its output
  float* y = b1;
  MAB_CUDA(cudaMemsetAsync(a0, 0, (size_t)VT * 640 * sizeof(float), c.st));  // zero the 52 pad channels
  auto conv = [&](const float* x, int C, int Creal, const std::string& key, int Cout, const float* bias,
                  int act, const float* res, int act_post, float* out, int ldo) {
    const float* W = h->P32(key);
    if (!W) return fail(-4, "weight %s was not packed", key.c_str());
    GemmEpilogue ep{};
    ep.bias = bias; ep.out_f32 = out; ep.ldo = ldo; ep.act = act; ep.res_f32 = res; ep.act_post = act_post;
    ep.round_tf32 = 1;
    return launch_conv3_tf32(x, V, hh, ww, C, Creal, W, Cout, ep, c.st);
  };
  auto lin = [&](const float* x, int K, int lda, const std::string& key, int N, const float* bias, float* out) {
    const float* W = h->P32(key);
    if (!W) return fail(-4, "weight %s was not packed", key.c_str());
    GemmEpilogue ep{};
    ep.bias = bias; ep.out_f32 = out; ep.ldo = N; ep.round_tf32 = 1;
    return launch_gemm_tf32(x, lda, W, K, (int)VT, N, K, ep, c.st);
  };
  MAB_TRY(conv(unsh, cin_pad, cin_real, e + ".conv_in.weight", 588, h->F(e + ".conv_in.bias"), ACT_NONE,
               nullptr, 0, a0, 640));
  const float* x = a0;
  int cin = 640, creal = 588;
  float* xo[2] = {x1, x2};
  float* bb[2] = {b0, b1};
  const int cout[2] = {768, 1024};
  for (int i = 0; i < 2; ++i) {
    const std::string r = e + ".encoder." + std::to_string(i);
    MAB_TRY(conv(x, cin, creal, r + ".conv1.weight", cout[i], h->F(r + ".conv1.bias"), ACT_GELU, nullptr, 0,
                 bb[i], cout[i]));
    // block 1 writes its shortcut over R0+R1 = block 0's input, so conv1 above must read x first
    MAB_TRY(lin(x, cin, cin, r + ".shortcut.weight", cout[i], h->F(r + ".shortcut.bias"), xo[i]));
    MAB_TRY(conv(bb[i], cout[i], cout[i], r + ".conv2.weight", cout[i], h->F(r + ".conv2.bias"), ACT_GELU,
                 xo[i], 1, xo[i], cout[i]));
    x = xo[i]; cin = creal = cout[i];
  }
  MAB_TRY(lin(x, 1024, 1024, e + ".encoder.2.weight", 1024, h->F(e + ".encoder.2.bias"), y));
  MAB_TRY(launch_layernorm(y, (int)VT, 1024, (int)VT, (int)VT, 0, h->F(e + ".norm_layer.weight"),
                           h->F(e + ".norm_layer.bias"), nullptr, nullptr, 1e-6f, nullptr, y, nullptr, c.st));
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
             accum_masked_rows_kernel<<<ew_grid(VT * 256), 256, 0, c.st>>>(p.geo_sum, y, mask, p.T, 1024,
                                                                          VT * 256, first ? 1 : 0));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// GlobalRepresentationEncoder (global_rep_encoder.py:54-67,83-105): Linear-GELU x3, Linear,
// LayerNorm; accumulated into p.geo_vec scaled by the per-view mask.  x [V, cin] fp32.
int global_rep_encode(Ctx& c, const std::string& e, const float* x, int cin, const float* mask, bool first,
                      float* t0, float* t1) {
  mab_handle* h = c.h;
  Plan& p = h->plan;
  const int V = p.V;
  auto lin = [&](const char* name, const float* in, int K, int N, int act, float* out) {
    const std::string k = e + "." + name;
    const long long warps = (long long)N * ((V + 7) / 8);
    MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
               small_linear_kernel<<<(int)((warps * 32 + 255) / 256), 256, 0, c.st>>>(
                   in, h->F(k + ".weight"), h->F(k + ".bias"), V, K, N, act, out, K, N));
  };
  if (!h->F(e + ".encoder.1.weight")) return fail(-4, "missing weights of %s", e.c_str());
  lin("encoder.0.0.0.0", x, cin, 128, 1, t0);
  lin("encoder.0.0.1", t0, 128, 256, 1, t1);
  lin("encoder.0.1", t1, 256, 512, 1, t0);
  lin("encoder.1", t0, 512, 1024, 0, t1);
  MAB_TRY(launch_layernorm(t1, V, 1024, V, V, 0, h->F(e + ".norm_layer.weight"), h->F(e + ".norm_layer.bias"),
                           nullptr, nullptr, 1e-6f, nullptr, t1, nullptr, c.st));
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
             accum_masked_rows_kernel<<<ew_grid((long long)V * 256), 256, 0, c.st>>>(
                 p.geo_vec, t1, mask, 1, 1024, (long long)V * 256, first ? 1 : 0));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// a6: model.py:1377-1436.  p.enc_out holds the encoder-normed patch tokens (fp32); writes p.fused.
int encode_geometric_inputs(Ctx& c, const mab_inputs* in) {
  mab_handle* h = c.h;
  Plan& p = h->plan;
  const int V = p.V, T = p.T, D = h->cfg.enc_dim;
  const long long VT = (long long)V * T;
  // small per-view scratch: [0,V) depth norm | masks 5 x V | in_scale V | in_trans 3V | in_tscale V | MLP 2 x V*1024
  float* s = p.geo_small;
  float* depth_norm = s; float* m_ray = s + V; float* m_depth = s + 2 * V; float* m_scale = s + 3 * V;
  float* m_cam = s + 4 * V; float* m_tscale = s + 5 * V; float* in_scale = s + 6 * V;
  float* in_trans = s + 7 * V; float* in_tscale = s + 10 * V;
  float* t0 = s + 16 * V; float* t1 = t0 + (long long)V * 1024;
  const bool has_rays = in->ray_dirs && in->ray_mask, has_depth = in->depth_along_ray && in->depth_mask;
  const bool has_cam = in->cam_quats && in->cam_trans && in->cam_mask;
  if (has_depth)
    MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
               depth_norm_kernel<<<V, 256, 0, c.st>>>(in->depth_along_ray, (long long)p.H * p.W, depth_norm));
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
             geo_prepare_kernel<<<(V + 127) / 128, 128, 0, c.st>>>(
                 V, has_depth ? depth_norm : nullptr, has_cam ? in->cam_trans : nullptr, in->pose_trans_norm,
                 has_rays ? in->ray_mask : nullptr, has_depth ? in->depth_mask : nullptr,
                 has_cam ? in->cam_mask : nullptr, in->depth_metric_mask, in->pose_metric_mask, in_scale,
                 in_trans, in_tscale, m_ray, m_depth, m_scale, m_cam, m_tscale));
  MAB_CUDA(cudaGetLastError());
  // the encoder's activation buffers (ln | qkv | att | hid, contiguous in the plan) are free here
  bool first_dense = true, first_vec = true;
  float* scratch = reinterpret_cast<float*>(p.ln);
  // ln | qkv | att | hid are contiguous in the plan and hold at least (T+1) * 9 * D bf16 per view of this
  // pass = 18 KB per token, against 15 KB per token of scratch
  if ((size_t)VT * 3840 * sizeof(float) > (size_t)p.Lenc * 9 * D * sizeof(bf16))
    return fail(-5, "geometric encoder scratch overflows the encoder activation buffers");
  for (int which = 0; which < 2; ++which) {
    if ((which == 0 && !has_rays) || (which == 1 && !has_depth)) continue;
    const int cpad = which == 0 ? 640 : 256;
    if (which == 0)
      MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
                 patch_im2col_kernel<<<ew_grid(VT * cpad), 256, 0, c.st>>>(in->ray_dirs, V, p.H, p.W, 3, 1, cpad,
                                                                          nullptr, 0, nullptr, scratch));
    else
      MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
                 patch_im2col_kernel<<<ew_grid(VT * cpad), 256, 0, c.st>>>(in->depth_along_ray, V, p.H, p.W, 1, 0,
                                                                          cpad, depth_norm, 1, nullptr, scratch));
    MAB_TRY(dense_rep_encode(c, which == 0 ? "ray_dirs_encoder" : "depth_encoder", scratch, cpad,
                             which == 0 ? 588 : 196, which == 0 ? m_ray : m_depth, first_dense));
    first_dense = false;
  }
  if (has_depth) {
    MAB_TRY(global_rep_encode(c, "depth_scale_encoder", in_scale, 1, m_scale, first_vec, t0, t1));
    first_vec = false;
  }
  if (has_cam) {
    MAB_TRY(global_rep_encode(c, "cam_rot_encoder", in->cam_quats, 4, m_cam, first_vec, t0, t1));
    MAB_TRY(global_rep_encode(c, "cam_trans_encoder", in_trans, 3, m_cam, false, t0, t1));
    MAB_TRY(global_rep_encode(c, "cam_trans_scale_encoder", in_tscale, 1, m_tscale, false, t0, t1));
    first_vec = false;
  }
  MAB_LAUNCH(CAT_LAYERNORM, 0.0, (double)VT * D * 6.0, c.st,
             fuse_norm_kernel<8><<<(int)((VT + 7) / 8), 256, 0, c.st>>>(
                 p.enc_out, first_dense ? nullptr : p.geo_sum, first_vec ? nullptr : p.geo_vec, (int)VT, T,
                 h->F("fusion_norm_layer.weight"), h->F("fusion_norm_layer.bias"), 1e-6f, p.fused,
                 c.h->cfg.aat_dim == D ? p.ys : p.fused32, c.h->cfg.aat_dim == D ? 0 : 1));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

// a2-a6 for the views of one encoder pass (h->plan describes the pass: V views, outputs at p.fused /
// p.fused32 / p.ys): patch embed, the DINOv2 blocks, encoder norm, geometric encoders and fusion norm
static int encode_views(Ctx& c, const mab_inputs* in, bool multimodal, bool identity_embed) {
  mab_handle* h = c.h;
  const mab_config& cf = h->cfg;
  Plan& p = h->plan;
  const int V = p.V, T = p.T, D = cf.enc_dim;
  const long long VT = (long long)V * T;
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
             patch_im2col_kernel<<<ew_grid(VT * 640), 256, 0, c.st>>>(in->images, V, p.H, p.W, 3, 0, 640,
                                                                      nullptr, 0, p.im2col, nullptr));
  MAB_TRY(c.linear(p.im2col, (int)VT, 640, 640, "encoder.model.patch_embed.proj.weight", D,
                   h->F("encoder.model.patch_embed.proj.bias"), ACT_NONE, nullptr, nullptr, p.patch,
                   nullptr, nullptr, 640, 588));
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
             assemble_tokens_kernel<<<ew_grid(p.Lenc * D / 4), 256, 0, c.st>>>(
      p.patch, h->F("encoder.model.cls_token"), p.pos, V, T, D, p.xs));
  MAB_CUDA(cudaGetLastError());
  for (int i = 0; i < cf.enc_depth; ++i) {
    const std::string b = "encoder.model.blocks." + std::to_string(i);
    MAB_TRY(block_qkv(c, b, p.xs, p.Lenc, D, p.ln, p.qkv));
    MAB_TRY(launch_attention(p.qkv, p.Lenc, p.qkv, p.Lenc, cf.enc_heads, V, T + 1, T + 1, T + 1,
                             T + 1, 0.125f, p.att, nullptr, c.st));
    MAB_TRY(block_tail(c, b, p.xs, p.Lenc, D, p.att, p.ln, p.hid, h->F(b + ".ls1.gamma"),
                       h->F(b + ".ls2.gamma")));
  }
  if (!multimodal) {
    // encoder norm on the patch tokens only (drop cls), immediately followed by the fusion norm
    // (image-only: the geometric features are exactly zero, model.py:931-968 / SURVEY 8a a6)
    MAB_TRY(launch_layernorm(p.xs, (int)VT, D, T, T + 1, 1, h->F("encoder.model.norm.weight"),
                             h->F("encoder.model.norm.bias"), h->F("fusion_norm_layer.weight"),
                             h->F("fusion_norm_layer.bias"), 1e-6f, p.fused, identity_embed ? p.ys : p.fused32,
                             nullptr, c.st, nullptr, identity_embed ? 0 : 1));
  } else {
    MAB_TRY(launch_layernorm(p.xs, (int)VT, D, T, T + 1, 1, h->F("encoder.model.norm.weight"),
                             h->F("encoder.model.norm.bias"), nullptr, nullptr, 1e-6f, nullptr, p.enc_out,
                             nullptr, c.st));
    MAB_TRY(encode_geometric_inputs(c, in));
  }
  return 0;
}

// a2-a7: patch embed, 24 DINOv2 blocks, encoder norm (+ fusion norm), proj_embed, ref-view PE
extern "C" int mab_stage_encode(mab_handle* h, const mab_inputs* in, int views_total, void* workspace,
                                size_t workspace_bytes, int has_scale_token, void* stream) {
  MAB_ARG(h && in && workspace && h->finalized, "handle not ready");
  MAB_ARG(in->views > 0 && in->H > 0 && in->W > 0 && in->H % 14 == 0 && in->W % 14 == 0 && in->images,
          "bad inputs");
  MAB_CUDA(cudaSetDevice(h->device));
  const mab_config& cf = h->cfg;
  Ctx c{h, (cudaStream_t)stream};
  // has_scale_token bit0: this rank owns the scale token; bit1: this rank owns view 0
  make_plan(h, in->views, views_total > in->views ? views_total : in->views, in->H, in->W,
            has_scale_token & 1, (has_scale_token >> 1) & 1, workspace, &h->plan);
  Plan& p = h->plan;
  if (p.total > workspace_bytes)
    return fail(-5, "workspace too small: need %zu bytes, got %zu", p.total, workspace_bytes);
  const int V = p.V, T = p.T, D = cf.enc_dim;
  const long long VT = (long long)V * T;
  const bool multimodal = in->ray_dirs || in->depth_along_ray || in->cam_quats;
  const bool identity_embed = cf.aat_dim == D;  // no proj_embed: the fused tokens ARE the residual stream
  if (multimodal && !h->P32("ray_dirs_encoder.conv_in.weight"))
    return fail(-4, "geometric inputs given but the checkpoint has no geometric encoders");

  MAB_TRY(mab_posembed_resize(h->F("encoder.model.pos_embed"), cf.enc_pos_grid, D, p.h, p.w, p.pos, c.st));
  // a3-a6 are per-view independent: passes of p.Vl views, alternating between the encoder lanes (streams),
  // each pass writing its rows of the fused tokens (bf16, + fp32 for the fp32 head / identity-embed stream)
  const Plan whole = h->plan;
  int rc = 0;
  const bool two = whole.lanes == 2;
  if (two) {  // fork: the second lane starts after everything already queued on the caller's stream
    MAB_CUDA(cudaEventRecord(h->ev_fork, c.st));
    MAB_CUDA(cudaStreamWaitEvent(h->lane_stream, h->ev_fork, 0));
  }
  int pass = 0;
  for (int v0 = 0; v0 < whole.V && rc == 0; v0 += whole.Vl, ++pass) {
    const int vc = whole.V - v0 < whole.Vl ? whole.V - v0 : whole.Vl;
    const int l = two ? pass & 1 : 0;
    mab_inputs sub = *in;
    const long long px = (long long)whole.H * whole.W;
    sub.views = vc;
    sub.images = in->images + (long long)v0 * 3 * px;
    if (in->ray_dirs) sub.ray_dirs = in->ray_dirs + (long long)v0 * px * 3;
    if (in->depth_along_ray) sub.depth_along_ray = in->depth_along_ray + (long long)v0 * px;
    if (in->cam_quats) sub.cam_quats = in->cam_quats + (long long)v0 * 4;
    if (in->cam_trans) sub.cam_trans = in->cam_trans + (long long)v0 * 3;
    if (in->ray_mask) sub.ray_mask = in->ray_mask + v0;
    if (in->depth_mask) sub.depth_mask = in->depth_mask + v0;
    if (in->cam_mask) sub.cam_mask = in->cam_mask + v0;
    if (in->depth_metric_mask) sub.depth_metric_mask = in->depth_metric_mask + v0;
    if (in->pose_metric_mask) sub.pose_metric_mask = in->pose_metric_mask + v0;
    // the helpers read h->plan: present the pass as a scene of vc views on this lane's activations whose
    // outputs start at view v0 (kernel arguments are captured at launch, so the plan may change under
    // kernels that are still queued)
    Plan& q = h->plan;
    q = whole;
    q.V = vc;
    q.Lenc = (long long)vc * (whole.T + 1);
    const Plan::EncBufs& e = whole.lane[l];
    q.im2col = e.im2col; q.patch = e.patch; q.xs = e.xs; q.ln = e.ln; q.qkv = e.qkv; q.att = e.att; q.hid = e.hid;
    q.enc_out = e.enc_out; q.geo_sum = e.geo_sum; q.geo_vec = e.geo_vec; q.geo_small = e.geo_small;
    q.fused = whole.fused + (long long)v0 * whole.T * D;
    if (whole.fused32) q.fused32 = whole.fused32 + (long long)v0 * whole.T * D;
    if (identity_embed) q.ys = whole.ys + (long long)v0 * whole.T * D;
    Ctx cl{h, l == 0 ? c.st : h->lane_stream};
    rc = encode_views(cl, &sub, multimodal, identity_embed);
  }
  if (two && rc == 0) {  // join
    MAB_CUDA(cudaEventRecord(h->ev_join, h->lane_stream));
    MAB_CUDA(cudaStreamWaitEvent(c.st, h->ev_join, 0));
  }
  h->plan = whole;
  if (rc) return rc;
  if (identity_embed) {
    // the fp32 fused tokens were written straight into the stream; the scale token joins them as is
    if (p.has_scale)
      MAB_CUDA(cudaMemcpyAsync(p.ys + VT * D, h->F("scale_token"), (size_t)D * sizeof(float),
                               cudaMemcpyDeviceToDevice, c.st));
  } else {
    if (p.has_scale) {
      MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
                 f32_to_bf16_kernel<<<4, 256, 0, c.st>>>(h->F("scale_token"), p.fused + VT * D, D));
      MAB_CUDA(cudaGetLastError());
    }
    // proj_embed (aat.py:121-124,600)
    MAB_TRY(c.linear(p.fused, (int)p.L, D, D, "info_sharing.proj_embed.weight", cf.aat_dim,
                     h->F("info_sharing.proj_embed.bias"), ACT_NONE, nullptr, nullptr, p.ys, nullptr));
  }
  // reference-view PE on view 0's tokens (aat.py:618-626)
  if (cf.aat_ref_view_pe && cf.aat_nonref_view_pe) {
    // use_pe_for_non_reference_views (aat.py:618-645): view 0 takes row 0 and every other view the row the
    // caller drew for it (torch.randint in the reference, or 1..V-1); rows were checked against the table
    MAB_ARG((int)h->view_pe_rows.size() == V, "mab_set_view_pe_rows must give one row per view of this shard");
    const RawTensor& tab = h->raw.at("info_sharing.view_pos_table");
    for (int r : h->view_pe_rows)
      MAB_ARG(r < (int)tab.shape[0], "view_pos_table row out of range (max_num_views_for_pe)");
    MAB_CUDA(cudaMemcpyAsync(p.view_rows, h->view_pe_rows.data(), (size_t)V * sizeof(int),
                             cudaMemcpyHostToDevice, c.st));
    MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
               add_view_rows_kernel<<<ew_grid(VT * cf.aat_dim), 256, 0, c.st>>>(p.ys, tab.p, p.view_rows, V, T,
                                                                              cf.aat_dim));
    MAB_CUDA(cudaGetLastError());
  } else if (p.has_ref && cf.aat_ref_view_pe) {  // distinguish_ref_and_non_ref_views (aat.py:618-626)
    MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
               add_row_vector_kernel<<<ew_grid((long long)T * cf.aat_dim), 256, 0, c.st>>>(
        p.ys, h->F("info_sharing.view_pos_table"), T, cf.aat_dim));
    MAB_CUDA(cudaGetLastError());
  }
  return 0;
}

// gat_ifr_24_layers*.yaml (model_type "global_attention"): every block attends over the whole scene;
// the alternating trunk does so in its even blocks only (aat.py:657-705)
static bool global_layer(const mab_config& cf, int layer) { return cf.aat_all_global || layer % 2 == 0; }

extern "C" int mab_stage_aat_qkv(mab_handle* h, int layer, void** qkv_out, int* rows_out,
                                 void* stream) {
  MAB_ARG(h && h->plan.V > 0, "mab_stage_encode must run first");
  MAB_ARG(layer >= 0 && layer < h->cfg.aat_depth, "layer out of range");
  Ctx c{h, (cudaStream_t)stream};
  Plan& p = h->plan;
  const long long rows = global_layer(h->cfg, layer) ? p.L : (long long)p.V * p.T;  // frame layers: scale token skips
  const std::string b = "info_sharing.self_attention_blocks." + std::to_string(layer);
  MAB_TRY(block_qkv(c, b, p.ys, rows, h->cfg.aat_dim, p.ln, p.qkv));
  if (qkv_out) *qkv_out = p.qkv;
  if (rows_out) *rows_out = (int)rows;
  return 0;
}

// softmax scale of the info-sharing attention: head_dim^-0.5, times the entropy scaling of q
// sqrt(1.4 * log(N) / log(444)) when configured (transformer_blocks.py:108-110,190-196; N = tokens of
// the attention call: the whole scene + scale token for a global layer, T for a frame layer),
// evaluated in fp32 like the reference's 0-dim tensors
static float aat_softmax_scale(const mab_config& cf, long long n_tokens) {
  if (!cf.aat_entropy_scaling) return 0.125f;
  const float f = sqrtf((1.4f * logf((float)n_tokens)) / logf(444.0f));
  return 0.125f * f;
}

// Attention of block `layer` against one K/V block.  Odd (frame) layers and single-GPU global
// layers take one call (flags = FIRST|LAST, kv = NULL).  View-sharded global layers call it once
// per ring step with the qkv rows of the rank whose block is currently held: the online-softmax
// state is carried in the workspace between the calls (attention.cuh).
extern "C" int mab_stage_aat_attend(mab_handle* h, int layer, const void* kv, int kv_rows, int flags,
                                    void* stream) {
  MAB_ARG(h && h->plan.V > 0, "mab_stage_encode must run first");
  MAB_ARG(layer >= 0 && layer < h->cfg.aat_depth, "layer out of range");
  const mab_config& cf = h->cfg;
  Ctx c{h, (cudaStream_t)stream};
  Plan& p = h->plan;
  const int T = p.T, V = p.V;
  const int first = flags & MAB_RING_FIRST, last = flags & MAB_RING_LAST;
  if (global_layer(cf, layer)) {  // global attention over every token of the scene (aat.py:657-661)
    const long long rows = p.L;
    const void* kvp = kv ? kv : p.qkv;
    const int kvr = kv ? kv_rows : (int)rows;
    MAB_ARG((first && last) || p.carry_o, "ring attention needs views_total > views at mab_stage_encode");
    // sharded scenes: every rank sees Vtot * T + 1 tokens in total (the scale token lives on one rank)
    const long long n_scene = (long long)p.Vtot * T + 1;
    return launch_attention(p.qkv, rows, kvp, kvr, cf.aat_heads, 1, (int)rows, kvr, 0, 0,
                            aat_softmax_scale(cf, n_scene), p.att, nullptr, c.st, p.carry_o, p.carry_ml,
                            first ? 0 : 1, last ? 0 : 1);
  }
  MAB_ARG(first && last && !kv, "frame attention is local to a view: one step");
  const long long rows = (long long)V * T;  // frame attention inside each view (aat.py:662-705)
  return launch_attention(p.qkv, rows, p.qkv, rows, cf.aat_heads, V, T, T, T, T, aat_softmax_scale(cf, T), p.att,
                          nullptr, c.st);
}

// K|V columns of this rank's current qkv rows -> dst [rows, 2*aat_dim] bf16 (contiguous), the block
// the other ranks gather for the global attention of an even layer.
extern "C" int mab_stage_aat_export_kv(mab_handle* h, void* dst, void* stream) {
  MAB_ARG(h && h->plan.V > 0 && dst, "mab_stage_aat_qkv must run first");
  Plan& p = h->plan;
  const size_t A = h->cfg.aat_dim;
  MAB_CUDA(cudaMemcpy2DAsync(dst, 2 * A * sizeof(bf16), p.qkv + A, 3 * A * sizeof(bf16), 2 * A * sizeof(bf16),
                             (size_t)p.L, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}

// Global attention of even block `layer` against a gathered K|V block [kv_rows, 2*aat_dim]; the
// online-softmax state is carried between calls exactly as in mab_stage_aat_attend.
extern "C" int mab_stage_aat_attend_kv(mab_handle* h, int layer, const void* kv, int kv_rows, int flags,
                                       void* stream) {
  MAB_ARG(h && h->plan.V > 0 && kv && kv_rows > 0, "mab_stage_encode must run first");
  MAB_ARG(layer >= 0 && layer < h->cfg.aat_depth, "layer out of range");
  MAB_ARG(global_layer(h->cfg, layer), "only the global layers attend across ranks");
  const mab_config& cf = h->cfg;
  Plan& p = h->plan;
  const int first = flags & MAB_RING_FIRST, last = flags & MAB_RING_LAST;
  MAB_ARG((first && last) || p.carry_o, "carried attention needs views_total > views at mab_stage_encode");
  return launch_attention(p.qkv, p.L, kv, kv_rows, cf.aat_heads, 1, (int)p.L, kv_rows, 0, 0,
                          aat_softmax_scale(cf, (long long)p.Vtot * p.T + 1), p.att, nullptr, (cudaStream_t)stream,
                          p.carry_o, p.carry_ml, first ? 0 : 1, last ? 0 : 1, 1);
}

// proj + residual, LN2, MLP + residual of block `layer`, intermediate taps and the final norm
extern "C" int mab_stage_aat_finish(mab_handle* h, int layer, void* stream) {
  MAB_ARG(h && h->plan.V > 0, "mab_stage_encode must run first");
  MAB_ARG(layer >= 0 && layer < h->cfg.aat_depth, "layer out of range");
  const mab_config& cf = h->cfg;
  Ctx c{h, (cudaStream_t)stream};
  Plan& p = h->plan;
  const int A = cf.aat_dim, T = p.T, V = p.V;
  const std::string b = "info_sharing.self_attention_blocks." + std::to_string(layer);
  const long long rows = global_layer(cf, layer) ? p.L : (long long)V * T;
  MAB_TRY(block_tail(c, b, p.ys, rows, A, p.att, p.ln, p.hid, nullptr, nullptr));
  // intermediate taps + final norm (aat.py:706-716), token-major == NHWC for the dense head
  const long long VT = (long long)V * T;
  const float* nw = h->F("info_sharing.norm.weight");
  const float* nb = h->F("info_sharing.norm.bias");
  int tap = -1;
  if (layer == cf.aat_tap0) tap = 0;
  else if (layer == cf.aat_tap1) tap = 1;
  else if (cf.aat_tap2 > 0 && layer == cf.aat_tap2) tap = 2;
  // bf16 for the bf16 head; fp32 for the fp32 / TF32 head (exactly one of the two buffers exists)
  if (tap >= 0)
    MAB_TRY(launch_layernorm(p.ys, (int)VT, A, (int)VT, (int)VT, 0, nw, nb, nullptr, nullptr, 1e-6f,
                             p.tap[tap], p.tap32[tap], nullptr, c.st, nullptr, 1));
  if (layer == cf.aat_depth - 1) {
    MAB_TRY(launch_layernorm(p.ys, (int)VT, A, (int)VT, (int)VT, 0, nw, nb, nullptr, nullptr, 1e-6f,
                             p.fin, p.fin32, nullptr, c.st, nullptr, 1));
    if (p.has_scale)
      MAB_TRY(launch_layernorm(p.ys + VT * A, 1, A, 1, 1, 0, nw, nb, nullptr, nullptr, 1e-6f, nullptr,
                               p.scale_tok, nullptr, c.st));
  }
  return 0;
}

// ================================================================================================
// Part 5: heads (a10-a14), whole forward, post-processing (a16)
// ================================================================================================
namespace {

int small_linear(cudaStream_t st, const float* x, const float* w, const float* b, int rows, int K,
                 int N, int act, float* y) {
  const long long warps = (long long)N * ((rows + 7) / 8);
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, st,
             small_linear_kernel<<<(int)((warps * 32 + 255) / 256), 256, 0, st>>>(x, w, b, rows, K, N, act, y, K, N));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

// ResidualConvUnit_custom (dpt_block.py:156-177) on `x` (with relu(x) precomputed in x_relu):
//   out = conv2(relu(conv1(relu(x)) + b1)) + b2 + x (+ extra)
template <typename T>
int rcu(Ctx& c, const std::string& p, const T* x, const T* x_relu, const T* extra, int batch,
        int H, int W, T* tmp_relu, T* out, T* out_relu) {
  mab_handle* h = c.h;
  MAB_TRY(head_conv3(c, x_relu, batch, H, W, 256, p + ".conv1.weight", 256, h->F(p + ".conv1.bias"), ACT_RELU,
                     (const T*)nullptr, (const T*)nullptr, tmp_relu, (T*)nullptr));
  return head_conv3(c, tmp_relu, batch, H, W, 256, p + ".conv2.weight", 256, h->F(p + ".conv2.bias"), ACT_NONE,
                    x, extra, out, out_relu);
}

}  // namespace

// scale head (mlp_head.py:58-92, adaptors.py:190-212) on the final scale token -> metric scale [1]
extern "C" int mab_stage_scale_head(mab_handle* h, float* metric_scale_out, void* stream) {
  MAB_ARG(h && metric_scale_out && h->plan.V > 0 && h->plan.has_scale, "this rank has no scale token");
  Ctx c{h, (cudaStream_t)stream};
  Plan& p = h->plan;
  const int A = h->cfg.aat_dim;
  float* t0 = p.scale_tmp; float* t1 = p.scale_tmp + 256; float* t2 = p.scale_tmp + 512;
  MAB_TRY(small_linear(c.st, p.scale_tok, h->F("scale_head.proj.weight"), h->F("scale_head.proj.bias"), 1, A, 196, 0, t0));
  MAB_TRY(small_linear(c.st, t0, h->F("scale_head.mlp.0.0.weight"), h->F("scale_head.mlp.0.0.bias"), 1, 196, 196, 2, t1));
  MAB_TRY(small_linear(c.st, t1, h->F("scale_head.mlp.1.0.weight"), h->F("scale_head.mlp.1.0.bias"), 1, 196, 196, 2, t2));
  MAB_TRY(small_linear(c.st, t2, h->F("scale_head.output_proj.weight"), h->F("scale_head.output_proj.bias"), 1, 196, 1, 0, t0));
  MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st, scale_finalize_kernel<<<1, 32, 0, c.st>>>(t0, metric_scale_out));
  MAB_CUDA(cudaGetLastError());
  return 0;
}

namespace {
template <typename T>
int stage_heads_range_t(mab_handle* h, const mab_inputs* in, mab_outputs* out, const float* metric_scale_in,
                        int view_lo, int view_hi, bool with_scale, void* stream);
int stage_heads_range(mab_handle* h, const mab_inputs* in, mab_outputs* out, const float* metric_scale_in,
                      int view_lo, int view_hi, bool with_scale, void* stream);
// every array of mab_outputs except logits6 is written by the heads
bool outputs_complete(const mab_outputs* o) {
  return o->pts3d && o->pts3d_cam && o->ray_directions && o->depth_along_ray && o->conf && o->mask_logits &&
         o->non_ambiguous_mask && o->pose && o->metric_scale;
}
}

extern "C" int mab_stage_heads(mab_handle* h, const mab_inputs* in, mab_outputs* out,
                               const float* metric_scale_in, void* stream) {
  MAB_ARG(h && in && out && h->plan.V > 0, "mab_stage_encode must run first");
  MAB_ARG(outputs_complete(out), "only mab_outputs.logits6 may be NULL");
  return stage_heads_range(h, in, out, metric_scale_in, 0, h->plan.V, true, stream);
}

// Heads of the views [view_lo, view_hi) of this rank only (pointers in `out` address the rank's whole
// arrays).  out->metric_scale must already hold the scene's metric scale (mab_stage_scale_head or
// a broadcast).  Lets the caller post-process and copy out finished views while later ones compute.
extern "C" int mab_stage_heads_range(mab_handle* h, const mab_inputs* in, mab_outputs* out, int view_lo,
                                     int view_hi, void* stream) {
  MAB_ARG(h && in && out && h->plan.V > 0, "mab_stage_encode must run first");
  MAB_ARG(view_lo >= 0 && view_lo < view_hi && view_hi <= h->plan.V, "bad view range");
  MAB_ARG(outputs_complete(out), "only mab_outputs.logits6 may be NULL");
  return stage_heads_range(h, in, out, out->metric_scale, view_lo, view_hi, false, stream);
}

namespace {
int stage_heads_range(mab_handle* h, const mab_inputs* in, mab_outputs* out, const float* metric_scale_in,
                      int view_lo, int view_hi, bool with_scale, void* stream) {
  if (h->cfg.head_tf32)
    return stage_heads_range_t<float>(h, in, out, metric_scale_in, view_lo, view_hi, with_scale, stream);
  return stage_heads_range_t<bf16>(h, in, out, metric_scale_in, view_lo, view_hi, with_scale, stream);
}

template <typename T>
int stage_heads_range_t(mab_handle* h, const mab_inputs* in, mab_outputs* out, const float* metric_scale_in,
                        int view_lo, int view_hi, bool with_scale, void* stream) {
  constexpr bool TF = sizeof(T) == 4;
  constexpr int ES = (int)sizeof(T);
  const mab_config& cf = h->cfg;
  Ctx c{h, (cudaStream_t)stream};
  Plan& p = h->plan;
  const int V = p.V, T_ = p.T, A = cf.aat_dim, hh = p.h, ww = p.w, H = p.H, W = p.W;

  // ---- scale head: computed here by the rank that owns the scale token, unless the caller
  // already ran mab_stage_scale_head and broadcast its result (view-sharded run) ------------------
  if (!with_scale) {
    // the caller has placed the metric scale in out->metric_scale
  } else if (metric_scale_in) {
    if (metric_scale_in != out->metric_scale)
      MAB_CUDA(cudaMemcpyAsync(out->metric_scale, metric_scale_in, sizeof(float), cudaMemcpyDeviceToDevice, c.st));
  } else {
    MAB_ARG(p.has_scale, "rank without the scale token needs metric_scale_in");
    MAB_TRY(mab_stage_scale_head(h, out->metric_scale, stream));
  }

  // ---- dense head: DPT feature pyramid + regressor + fused adaptor/pointmap epilogue -----------
  const std::string ip = "dpt_feature_head.input_process.";
  const std::string sc = "dpt_feature_head.scratch.";
  const int h3 = (hh + 1) / 2, w3 = (ww + 1) / 2;
  for (int v0 = view_lo; v0 < view_hi; v0 += p.vc) {
    const int vc = (view_hi - v0 < p.vc) ? view_hi - v0 : p.vc;
    const int M = vc * T_;
    // model.py:1724-1768: [encoder features, 2 intermediate, final] or, with three intermediate
    // features (48-layer configs), [3 intermediate, final] and no encoder features
    const bool three = cf.aat_tap2 > 0;
    const int K0 = three ? A : cf.enc_dim;
    // head inputs: bf16 token-major buffers, or their fp32 twins for the fp32 / TF32 head
    auto head_in = [&](const bf16* b16, const float* f32, int dim) -> const T* {
      if (TF) return reinterpret_cast<const T*>(f32 + (long long)v0 * T_ * dim);
      return reinterpret_cast<const T*>(b16 + (long long)v0 * T_ * dim);
    };
    const T* f0 = three ? head_in(p.tap[0], p.tap32[0], A) : head_in(p.fused, p.fused32, cf.enc_dim);
    const T* f1 = head_in(p.tap[three ? 1 : 0], p.tap32[three ? 1 : 0], A);
    const T* f2 = head_in(p.tap[three ? 2 : 1], p.tap32[three ? 2 : 1], A);
    const T* f3 = head_in(p.fin, p.fin32, A);
    T* d_a = (T*)p.d_a; T* d_s = (T*)p.d_s; T* d_u = (T*)p.d_u; T* t0 = (T*)p.t0; T* t1 = (T*)p.t1;
    T* t2 = (T*)p.t2; T* t3 = (T*)p.t3; T* path = (T*)p.path; T* feat8 = (T*)p.feat8; T* r1 = (T*)p.r1; T* r2 = (T*)p.r2;
    T* lvl[4]; T* lvl_relu[4];
    for (int i = 0; i < 4; ++i) { lvl[i] = (T*)p.lvl[i]; lvl_relu[i] = (T*)p.lvl_relu[i]; }
    const T* nil = nullptr;
    // ---- pose head (pose_head.py:115-159, adaptors.py:714-732) in bf16x3 (~fp32) arithmetic:
    // final norm of this chunk's tokens as a split value, 7 token-wise GEMMs, mean over tokens
    {
      const std::string ph = "pose_head.";
      MAB_TRY(launch_layernorm(p.ys + (long long)v0 * T_ * A, M, A, M, M, 0, h->F("info_sharing.norm.weight"),
                               h->F("info_sharing.norm.bias"), nullptr, nullptr, 1e-6f, nullptr, nullptr,
                               nullptr, c.st, p.ph_in));
      MAB_TRY(c.linear_x3(p.ph_in, M, A, ph + "proj.weight", 784, h->F(ph + "proj.bias"), ACT_NONE, nullptr, p.ph_a));
      bf16* feat = p.ph_a; bf16* other = p.ph_d;
      for (int b = 0; b < 2; ++b) {
        const std::string r = ph + "res_conv." + std::to_string(b) + ".res_conv";
        MAB_TRY(c.linear_x3(feat, M, 784, r + "1.weight", 784, h->F(r + "1.bias"), ACT_RELU, nullptr, p.ph_b));
        MAB_TRY(c.linear_x3(p.ph_b, M, 784, r + "2.weight", 784, h->F(r + "2.bias"), ACT_RELU, nullptr, p.ph_c));
        MAB_TRY(c.linear_x3(p.ph_c, M, 784, r + "3.weight", 784, h->F(r + "3.bias"), ACT_RELU, feat, other));
        bf16* t = feat; feat = other; other = t;
      }
      dim3 blk(128, 8), grd((784 + 127) / 128, vc);
      MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
                 token_mean_kernel<<<grd, blk, 0, c.st>>>(feat, T_, 784, 3 * 784, 784, p.pose_pool + (long long)v0 * 784));
      MAB_CUDA(cudaGetLastError());
      float* pool = p.pose_pool + (long long)v0 * 784;
      float* m0 = p.pose_tmp; float* m1 = p.pose_tmp + (long long)V * 784;
      MAB_TRY(small_linear(c.st, pool, h->F(ph + "more_mlps.0.weight"), h->F(ph + "more_mlps.0.bias"), vc, 784, 784, 2, m0));
      MAB_TRY(small_linear(c.st, m0, h->F(ph + "more_mlps.2.weight"), h->F(ph + "more_mlps.2.bias"), vc, 784, 784, 2, m1));
      MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
                 small_linear_kernel<<<(int)(((long long)7 * ((vc + 7) / 8) * 32 + 255) / 256), 256, 0, c.st>>>(
                     m1, h->extra["pose_head.fc7.weight"], h->extra["pose_head.fc7.bias"], vc, 784, 7, 0,
                     p.pose_raw + (long long)v0 * 7, 784, 7));
      MAB_LAUNCH(CAT_ELEMENTWISE, 0.0, 0.0, c.st,
                 pose_finalize_kernel<<<(vc + 127) / 128, 128, 0, c.st>>>(p.pose_raw + (long long)v0 * 7, vc,
                                                                          out->pose + (long long)v0 * 8));
      MAB_CUDA(cudaGetLastError());
    }
    // layer 0: 1x1 (1024->96) -> convT 4x4/s4 -> 3x3 (96->256, no bias)           dpt.py:107-128
    MAB_TRY(head_linear(c, f0, M, K0, ip + "0.0.0.weight", 96, h->F(ip + "0.0.0.bias"), d_a));
    MAB_TRY(head_linear(c, d_a, M, 96, ip + "0.0.1.weight", 16 * 96, h->extra[ip + "0.0.1.bias.tiled"], d_s));
    MAB_TRY(head_pixel_shuffle(c, d_s, ES, vc, hh, ww, 4, 96, d_u));
    MAB_TRY(head_conv3(c, d_u, vc, 4 * hh, 4 * ww, 96, sc + "layer1_rn.weight", 256, nullptr, ACT_NONE, nil, nil, lvl[0], lvl_relu[0]));
    // layer 1: 1x1 (768->192) -> convT 2x2/s2 -> 3x3 (192->256)
    MAB_TRY(head_linear(c, f1, M, A, ip + "1.0.0.weight", 192, h->F(ip + "1.0.0.bias"), d_a));
    MAB_TRY(head_linear(c, d_a, M, 192, ip + "1.0.1.weight", 4 * 192, h->extra[ip + "1.0.1.bias.tiled"], d_s));
    MAB_TRY(head_pixel_shuffle(c, d_s, ES, vc, hh, ww, 2, 192, d_u));
    MAB_TRY(head_conv3(c, d_u, vc, 2 * hh, 2 * ww, 192, sc + "layer2_rn.weight", 256, nullptr, ACT_NONE, nil, nil, lvl[1], lvl_relu[1]));
    // layer 2: 1x1 (768->384) -> 3x3 (384->256)
    MAB_TRY(head_linear(c, f2, M, A, ip + "2.0.0.weight", 384, h->F(ip + "2.0.0.bias"), d_a));
    MAB_TRY(head_conv3(c, d_a, vc, hh, ww, 384, sc + "layer3_rn.weight", 256, nullptr, ACT_NONE, nil, nil, lvl[2], lvl_relu[2]));
    // layer 3: 1x1 (768->768) -> 3x3/s2 (im2col + GEMM) -> 3x3 (768->256)
    MAB_TRY(head_linear(c, f3, M, A, ip + "3.0.0.weight", 768, h->F(ip + "3.0.0.bias"), d_a));
    MAB_TRY(head_im2col_s2(c, d_a, ES, vc, hh, ww, 768, h3, w3, d_s));
    MAB_TRY(head_linear(c, d_s, vc * h3 * w3, 9 * 768, ip + "3.0.1.weight", 768, h->F(ip + "3.0.1.bias"), d_u));
    MAB_TRY(head_conv3(c, d_u, vc, h3, w3, 768, sc + "layer4_rn.weight", 256, nullptr, ACT_NONE, nil, nil, lvl[3], lvl_relu[3]));
    MAB_CUDA(cudaGetLastError());

    // refinenet4: RCU2 -> out_conv (1x1, commutes with the bilinear x2) -> x2 -> crop   dpt.py:213
    MAB_TRY(rcu<T>(c, sc + "refinenet4.resConfUnit2", lvl[3], lvl_relu[3], nil, vc, h3, w3, t0, t1, (T*)nullptr));
    MAB_TRY(head_linear(c, t1, vc * h3 * w3, 256, sc + "refinenet4.out_conv.weight", 256, h->F(sc + "refinenet4.out_conv.bias"), t2));
    MAB_TRY(head_resize(c, t2, vc, h3, w3, 256, 2 * h3, 2 * w3, hh, ww, path));
    // refinenet3/2/1: x = path + RCU1(skip); x = RCU2(x); out_conv; x2                dpt_block.py:225-255
    int lh = hh, lw = ww;
    for (int l = 2; l >= 0; --l) {
      const std::string r = sc + "refinenet" + std::to_string(l + 1);
      MAB_TRY(rcu<T>(c, r + ".resConfUnit1", lvl[l], lvl_relu[l], path, vc, lh, lw, t0, t1, t2));
      MAB_TRY(rcu<T>(c, r + ".resConfUnit2", t1, t2, nil, vc, lh, lw, t0, t3, (T*)nullptr));
      MAB_TRY(head_linear(c, t3, vc * lh * lw, 256, r + ".out_conv.weight", 256, h->F(r + ".out_conv.bias"), t0));
      T* dst = (l == 0) ? feat8 : path;
      MAB_TRY(head_resize(c, t0, vc, lh, lw, 256, 2 * lh, 2 * lw, 2 * lh, 2 * lw, dst));
      lh *= 2; lw *= 2;
    }
    // regressor: 3x3 (256->128) -> bilinear to HxW -> [3x3 (128->128) + ReLU + 1x1 (128->6) +
    // adaptor + pointmap] in one kernel                                               dpt.py:285-311
    MAB_TRY(head_conv3(c, feat8, vc, lh, lw, 256, "dpt_regressor_head.conv1.weight", 128, h->F("dpt_regressor_head.conv1.bias"), ACT_NONE, nil, nil, r1, (T*)nullptr));
    MAB_TRY(head_resize(c, r1, vc, lh, lw, 128, H, W, H, W, r2));
    RegressEpilogue rg{};
    const long long px0 = (long long)v0 * H * W;
    rg.bias2 = h->F("dpt_regressor_head.conv2.0.bias");
    rg.w3 = h->F("dpt_regressor_head.conv2.2.weight");
    rg.b3 = h->F("dpt_regressor_head.conv2.2.bias");
    rg.pose = out->pose + (long long)v0 * 8;
    rg.scale = out->metric_scale;
    rg.pts3d = out->pts3d + px0 * 3; rg.pts3d_cam = out->pts3d_cam + px0 * 3;
    rg.rays = out->ray_directions + px0 * 3; rg.depth = out->depth_along_ray + px0;
    rg.conf = out->conf + px0; rg.mask_logits = out->mask_logits + px0;
    rg.mask = out->non_ambiguous_mask + px0;
    rg.logits6 = out->logits6 ? out->logits6 + px0 * 6 : nullptr;
    MAB_TRY(head_regress(c, r2, vc, H, W, rg));
  }
  return 0;
}
}  // namespace

extern "C" int mab_forward(mab_handle* h, const mab_inputs* in, mab_outputs* out, void* workspace,
                           size_t workspace_bytes, void* stream) {
  MAB_ARG(h && in && out, "null argument");
  MAB_TRY(mab_stage_encode(h, in, in->views, workspace, workspace_bytes, 3, stream));
  for (int l = 0; l < h->cfg.aat_depth; ++l) {
    MAB_TRY(mab_stage_aat_qkv(h, l, nullptr, nullptr, stream));
    MAB_TRY(mab_stage_aat_attend(h, l, nullptr, 0, MAB_RING_FIRST | MAB_RING_LAST, stream));
    MAB_TRY(mab_stage_aat_finish(h, l, stream));
  }
  return mab_stage_heads(h, in, out, nullptr, stream);
}

// ---- view-sharded forward as ONE call: the K|V exchange of the global layers issued from here ----------
#define MAB_NCCL(expr)                                                                          \
  do {                                                                                          \
    int _r = (expr);                                                                            \
    if (_r != 0)                                                                                \
      return ::mab::fail(1000 + _r, "%s failed: %s", #expr,                                     \
                         h->nccl.GetErrorString ? h->nccl.GetErrorString(_r) : "NCCL error");  \
  } while (0)

extern "C" int mab_set_comm(mab_handle* h, void* nccl_comm, int rank, int world, const char* nccl_library) {
  MAB_ARG(h && nccl_comm && nccl_library, "null argument");
  MAB_ARG(world >= 1 && rank >= 0 && rank < world, "bad rank / world");
  MAB_CUDA(cudaSetDevice(h->device));
  mab_handle::Nccl& n = h->nccl;
  if (!n.lib) {
    n.lib = dlopen(nccl_library, RTLD_NOW | RTLD_GLOBAL);
    if (!n.lib) return fail(-6, "dlopen(%s) failed: %s", nccl_library, dlerror());
  }
  auto sym = [&](const char* name) { return dlsym(n.lib, name); };
  n.GroupStart = reinterpret_cast<decltype(n.GroupStart)>(sym("ncclGroupStart"));
  n.GroupEnd = reinterpret_cast<decltype(n.GroupEnd)>(sym("ncclGroupEnd"));
  n.Send = reinterpret_cast<decltype(n.Send)>(sym("ncclSend"));
  n.Recv = reinterpret_cast<decltype(n.Recv)>(sym("ncclRecv"));
  n.Broadcast = reinterpret_cast<decltype(n.Broadcast)>(sym("ncclBroadcast"));
  n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("ncclGetErrorString"));
  if (!n.GroupStart || !n.GroupEnd || !n.Send || !n.Recv || !n.Broadcast)
    return fail(-6, "%s does not export the NCCL point-to-point API", nccl_library);
  n.comm = nccl_comm; n.rank = rank; n.world = world;
  if (!h->comm_stream) {  // set-up time: the exchange stream and its two events
    MAB_CUDA(cudaStreamCreateWithFlags(&h->comm_stream, cudaStreamNonBlocking));
    MAB_CUDA(cudaEventCreateWithFlags(&h->ev_kv_ready, cudaEventDisableTiming));
    MAB_CUDA(cudaEventCreateWithFlags(&h->ev_kv_done, cudaEventDisableTiming));
    MAB_CUDA(cudaEventCreateWithFlags(&h->ev_kv_far, cudaEventDisableTiming));
  }
  return 0;
}

extern "C" size_t mab_kv_gather_bytes(mab_handle* h, int views_total, int H, int W) {
  if (!h || views_total <= 0 || H % h->cfg.patch || W % h->cfg.patch) return 0;
  const size_t T = (size_t)(H / h->cfg.patch) * (W / h->cfg.patch);
  return ((size_t)views_total * T + 1) * 2 * h->cfg.aat_dim * sizeof(bf16);
}

extern "C" int mab_forward_sharded(mab_handle* h, const mab_inputs* in, int views_total, mab_outputs* out,
                                   void* workspace, size_t workspace_bytes, void* kv_all, size_t kv_all_bytes,
                                   int with_heads, void* stream) {
  MAB_ARG(h && in && out && kv_all, "null argument");
  mab_handle::Nccl& n = h->nccl;
  MAB_ARG(n.comm, "mab_set_comm must be called first");
  MAB_ARG(views_total >= n.world && in->views > 0, "every rank needs a view");
  cudaStream_t st = (cudaStream_t)stream;
  // the shard table of mapanything_b200.parallel.shard_table: contiguous blocks, earlier ranks take the extra view
  const int base = views_total / n.world, extra = views_total % n.world;
  MAB_ARG(in->views == base + (n.rank < extra ? 1 : 0), "in->views does not match the shard table");
  const size_t T = (size_t)(in->H / h->cfg.patch) * (in->W / h->cfg.patch);
  const size_t row_bytes = 2 * (size_t)h->cfg.aat_dim * sizeof(bf16);
  std::vector<size_t> offs(n.world + 1, 0);  // byte offsets of the ranks' K|V blocks in kv_all
  for (int r = 0; r < n.world; ++r)
    offs[r + 1] = offs[r] + ((size_t)(base + (r < extra ? 1 : 0)) * T + (r == 0 ? 1 : 0)) * row_bytes;
  MAB_ARG(kv_all_bytes >= offs[n.world], "kv_all too small (mab_kv_gather_bytes)");
  uint8_t* kv = reinterpret_cast<uint8_t*>(kv_all);

  MAB_TRY(mab_stage_encode(h, in, views_total, workspace, workspace_bytes, n.rank == 0 ? 3 : 0, stream));
  for (int l = 0; l < h->cfg.aat_depth; ++l) {
    MAB_TRY(mab_stage_aat_qkv(h, l, nullptr, nullptr, stream));
    if (global_layer(h->cfg, l) && n.world > 1) {
      // Every rank exports its K|V rows, posts sends to / receives from all peers in ONE NCCL group on the
      // exchange stream, attends to its own block while the blocks travel, then to the remote rows in at
      // most two contiguous spans (before / after its own block); the online-softmax state is carried.
      // MAB_EXCHANGE_GROUPS=2 splits the exchange into a near group (ranks within world / 4) and a far group
      // and attends to the near blocks while the far ones travel (up to four remote spans): measured at 4 GPUs,
      // 32 views, 33.5 ms against 33.2 ms with one group -- the extra attention launches cost what the earlier
      // start buys, so one group is the default.
      static const bool one_group = !(getenv("MAB_EXCHANGE_GROUPS") && atoi(getenv("MAB_EXCHANGE_GROUPS")) == 2);
      const int reach = one_group ? n.world : (n.world / 4 > 1 ? n.world / 4 : 1);
      const int near_lo = n.rank - reach < 0 ? 0 : n.rank - reach;
      const int near_hi = n.rank + reach > n.world - 1 ? n.world - 1 : n.rank + reach;  // inclusive
      MAB_TRY(mab_stage_aat_export_kv(h, kv + offs[n.rank], stream));
      MAB_CUDA(cudaEventRecord(h->ev_kv_ready, st));
      MAB_CUDA(cudaStreamWaitEvent(h->comm_stream, h->ev_kv_ready, 0));
      auto exchange = [&](int lo, int hi, bool inside) -> int {  // peers in [lo, hi] (inside) or outside it
        bool any = false;
        for (int r = 0; r < n.world; ++r) any |= r != n.rank && ((r >= lo && r <= hi) == inside);
        if (!any) return 0;
        MAB_NCCL(n.GroupStart());
        for (int r = 0; r < n.world; ++r) {
          if (r == n.rank || ((r >= lo && r <= hi) != inside)) continue;
          MAB_NCCL(n.Send(kv + offs[n.rank], offs[n.rank + 1] - offs[n.rank], 0 /* ncclInt8 */, r, n.comm, h->comm_stream));
          MAB_NCCL(n.Recv(kv + offs[r], offs[r + 1] - offs[r], 0, r, n.comm, h->comm_stream));
        }
        MAB_NCCL(n.GroupEnd());
        return 0;
      };
      MAB_TRY(exchange(near_lo, near_hi, true));
      MAB_CUDA(cudaEventRecord(h->ev_kv_done, h->comm_stream));
      MAB_TRY(exchange(near_lo, near_hi, false));
      MAB_CUDA(cudaEventRecord(h->ev_kv_far, h->comm_stream));
      // remote spans in the order they become available: near-before, near-after, far-before, far-after
      struct Span { size_t lo, hi; int ev; } spans[4] = {
          {offs[near_lo], offs[n.rank], 0}, {offs[n.rank + 1], offs[near_hi + 1], 0},
          {0, offs[near_lo], 1}, {offs[near_hi + 1], offs[n.world], 1}};
      int last = -1;
      for (int i = 0; i < 4; ++i) if (spans[i].hi > spans[i].lo) last = i;
      MAB_TRY(mab_stage_aat_attend(h, l, nullptr, 0, MAB_RING_FIRST | (last < 0 ? MAB_RING_LAST : 0), stream));
      int waited = -1;
      for (int i = 0; i < 4; ++i) {
        if (spans[i].hi <= spans[i].lo) continue;
        if (spans[i].ev > waited) {
          MAB_CUDA(cudaStreamWaitEvent(st, spans[i].ev == 0 ? h->ev_kv_done : h->ev_kv_far, 0));
          waited = spans[i].ev;
        }
        MAB_TRY(mab_stage_aat_attend_kv(h, l, kv + spans[i].lo, (int)((spans[i].hi - spans[i].lo) / row_bytes),
                                        i == last ? MAB_RING_LAST : 0, stream));
      }
    } else {
      MAB_TRY(mab_stage_aat_attend(h, l, nullptr, 0, MAB_RING_FIRST | MAB_RING_LAST, stream));
    }
    MAB_TRY(mab_stage_aat_finish(h, l, stream));
  }
  // metric scale: computed by the rank that owns the scale token, broadcast to the others
  if (n.rank == 0) MAB_TRY(mab_stage_scale_head(h, out->metric_scale, stream));
  if (n.world > 1) {
    // the next layer's exchange must not start before this broadcast is queued behind the last one: same
    // communicator, one stream at a time -- the broadcast rides the exchange stream as well
    MAB_CUDA(cudaEventRecord(h->ev_kv_ready, st));
    MAB_CUDA(cudaStreamWaitEvent(h->comm_stream, h->ev_kv_ready, 0));
    MAB_NCCL(n.Broadcast(out->metric_scale, out->metric_scale, 1, 7 /* ncclFloat32 */, 0, n.comm, h->comm_stream));
    MAB_CUDA(cudaEventRecord(h->ev_kv_done, h->comm_stream));
    MAB_CUDA(cudaStreamWaitEvent(st, h->ev_kv_done, 0));
  }
  if (with_heads) return mab_stage_heads(h, in, out, out->metric_scale, stream);
  return 0;
}

// workspace: views*H*W*18 + views*8 + 512 bytes
extern "C" int mab_postprocess(mab_handle* h, const mab_inputs* in, mab_outputs* raw,
                               mab_post_outputs* post, const mab_post_options* opt, void* workspace,
                               size_t workspace_bytes, void* stream) {
  MAB_ARG(h && in && raw && post && opt && workspace, "null argument");
  cudaStream_t st = (cudaStream_t)stream;
  const int V = in->views, H = in->H, W = in->W;
  MAB_ARG(V > 0 && H > 0 && W > 0 && in->images, "bad inputs");
  MAB_ARG(outputs_complete(raw), "only mab_outputs.logits6 may be NULL");
  MAB_ARG(post->img_no_norm && post->depth_z && post->intrinsics && post->camera_poses && post->cam_trans &&
          post->cam_quats && post->mask, "every mab_post_outputs array is written");
  const long long npix = (long long)V * H * W;
  MAB_ARG(workspace_bytes >= (size_t)npix * 18 + (size_t)V * 8 + 512, "post-process workspace too small");
  const unsigned char* fmask = raw->non_ambiguous_mask;
  if (opt->apply_mask && opt->apply_confidence_mask) {
    // inference.py:440-452: torch.quantile(conf[b], p/100) per view, linear interpolation
    unsigned char* fm = reinterpret_cast<unsigned char*>(workspace) + ((size_t)npix * 17 + 255) / 256 * 256;
    float* stat = reinterpret_cast<float*>(fm + ((size_t)npix + 255) / 256 * 256);
    const long long n = (long long)H * W;
    const float q = (float)((double)opt->confidence_percentile / 100.0);
    MAB_ARG(q >= 0.f && q <= 1.f, "confidence_percentile must be in [0, 100]");
    const float rank = q * (float)(n - 1);
    const long long k_lo = (long long)floorf(rank), k_hi = (long long)ceilf(rank);
    MAB_LAUNCH(CAT_POSTPROCESS, 0.0, 0.0, st,
               conf_order_stat_kernel<<<dim3(2, V), 1024, 0, st>>>(raw->conf, n, k_lo, k_hi, stat));
    MAB_LAUNCH(CAT_POSTPROCESS, 0.0, 0.0, st,
               conf_mask_kernel<<<ew_grid(npix), 256, 0, st>>>(raw->conf, raw->non_ambiguous_mask, stat,
                                                              rank - (float)k_lo, n, npix, fm));
    fmask = fm;
  }
  MAB_LAUNCH(CAT_POSTPROCESS, 0.0, 0.0, st,
             denorm_image_kernel<<<ew_grid(npix), 256, 0, st>>>(in->images, V, H, W, 0.485f, 0.456f, 0.406f,
                                                     0.229f, 0.224f, 0.225f, post->img_no_norm));
  MAB_LAUNCH(CAT_POSTPROCESS, 0.0, 0.0, st,
             intrinsics_pose_kernel<<<V, 256, 0, st>>>(raw->ray_directions, H, W, raw->pose, raw->metric_scale,
                                            post->intrinsics, post->camera_poses, post->cam_trans,
                                            post->cam_quats));
  if (opt->apply_mask) {
    if (opt->mask_edges) {
      MAB_TRY(mab_edge_mask(raw->pts3d, raw->pts3d_cam, fmask, V, H, W, opt->edge_normal_threshold_deg,
                            opt->edge_depth_threshold, workspace, post->mask, st));
    } else {
      MAB_CUDA(cudaMemcpyAsync(post->mask, fmask, npix, cudaMemcpyDeviceToDevice, st));
    }
  }
  MAB_LAUNCH(CAT_POSTPROCESS, 0.0, 0.0, st,
             apply_mask_kernel<<<ew_grid(npix), 256, 0, st>>>(post->mask, npix, opt->apply_mask, raw->pts3d,
                                                   raw->pts3d_cam, raw->depth_along_ray, post->depth_z));
  MAB_CUDA(cudaGetLastError());
  return 0;
}
