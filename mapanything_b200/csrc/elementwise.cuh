// Bandwidth-bound kernels of libmab200: LayerNorm, patch im2col, token assembly, bicubic
// pos-embed resize, bilinear (align_corners) resize, pixel shuffle, strided im2col, pooling and
// the small fp32 head linears.  All are coalesced + vectorised; roofline = HBM GB/s.
#pragma once
#include "common.cuh"

namespace mab {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ----------------------------------------------------------------------------------------------
// LayerNorm over the last dim (C = 32*4*VPL), fp32 in, bf16 and/or fp32 out.  One warp per row.
// Row gather: output row o reads input row (o / group_out) * group_in + skip + o % group_out
// (drops the cls token of each view: vision_transformer.py:305-312).
// Optional second LayerNorm applied to the result of the first (image-only fusion norm,
// model.py:1423-1431 directly after the encoder norm).
// ----------------------------------------------------------------------------------------------
// LEAN: the LayerNorm in front of every qkv / fc1 GEMM (96 of the 102 launches of a step): contiguous rows,
// bf16 output only -- the optional outputs, the second norm and the row gather are compiled out
template <int VPL, bool LEAN = false>  // float4 per lane: C = 128 * VPL
__global__ void __launch_bounds__(256, LEAN ? 4 : 1) layernorm_kernel(const float* __restrict__ x, int rows_out, int group_out,
                                 int group_in, int skip, const float* __restrict__ w,
                                 const float* __restrict__ b, const float* __restrict__ w2_in,
                                 const float* __restrict__ b2, float eps,
                                 __nv_bfloat16* __restrict__ out_bf16, float* __restrict__ out_f32_in,
                                 float* __restrict__ out_mid_f32_in,
                                 __nv_bfloat16* __restrict__ out_split_in, int round_f32) {
  const float* w2 = LEAN ? nullptr : w2_in;
  float* out_f32 = LEAN ? nullptr : out_f32_in;
  float* out_mid_f32 = LEAN ? nullptr : out_mid_f32_in;
  __nv_bfloat16* out_split = LEAN ? nullptr : out_split_in;
  if (LEAN) { group_out = rows_out > 0 ? rows_out : 1; group_in = group_out; skip = 0; }
  constexpr int C = 128 * VPL;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  pdl_wait();
  pdl_trigger();
  if (warp >= rows_out) return;
  const long long in_row = (long long)(warp / group_out) * group_in + skip + warp % group_out;
  const float* xr = x + in_row * C;
  const uint64_t keep = l2_policy_evict_last();  // the residual stream is read again by the next epilogue
  float4 v[VPL];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    v[i] = ld_f4_l2hint(xr + 4 * (lane + 32 * i), keep);
    s += v[i].x + v[i].y + v[i].z + v[i].w;
  }
  auto normalise = [&](const float* ww, const float* bb) {
    const float mean = warp_sum(s) * (1.0f / C);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const float a = v[i].x - mean, b_ = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += a * a + b_ * b_ + c * c + d * d;
    }
    const float rstd = rsqrtf(warp_sum(q) * (1.0f / C) + eps);
    s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(ww) + lane + 32 * i);
      const float4 h = __ldg(reinterpret_cast<const float4*>(bb) + lane + 32 * i);
      v[i].x = (v[i].x - mean) * rstd * g.x + h.x;
      v[i].y = (v[i].y - mean) * rstd * g.y + h.y;
      v[i].z = (v[i].z - mean) * rstd * g.z + h.z;
      v[i].w = (v[i].w - mean) * rstd * g.w + h.w;
      s += v[i].x + v[i].y + v[i].z + v[i].w;
    }
  };
  normalise(w, b);
  if (out_mid_f32) {
#pragma unroll
    for (int i = 0; i < VPL; ++i)
      reinterpret_cast<float4*>(out_mid_f32 + (long long)warp * C)[lane + 32 * i] = v[i];
  }
  if (w2) normalise(w2, b2);
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    // round_f32: the fp32 row is an operand of a kind::tf32 contraction (dense-head input), see tf32_rn
    if (out_f32) reinterpret_cast<float4*>(out_f32 + (long long)warp * C)[lane + 32 * i] = round_f32 ? tf32_rn(v[i]) : v[i];
    if (out_bf16)
      reinterpret_cast<uint2*>(out_bf16 + (long long)warp * C)[lane + 32 * i] =
          make_uint2(pack_bf16(v[i].x, v[i].y), pack_bf16(v[i].z, v[i].w));
    if (out_split) {  // bf16x3 A operand: row = [hi | lo | hi], pitch 3*C (see GemmEpilogue::split_n)
      const uint2 hi = make_uint2(pack_bf16(v[i].x, v[i].y), pack_bf16(v[i].z, v[i].w));
      const uint2 lo = make_uint2(pack_bf16(v[i].x - bf16_lo(hi.x), v[i].y - bf16_hi(hi.x)),
                                  pack_bf16(v[i].z - bf16_lo(hi.y), v[i].w - bf16_hi(hi.y)));
      uint2* row = reinterpret_cast<uint2*>(out_split + (long long)warp * 3 * C);
      row[lane + 32 * i] = hi;
      row[C / 4 + lane + 32 * i] = lo;
      row[2 * (C / 4) + lane + 32 * i] = hi;
    }
  }
}

// ----------------------------------------------------------------------------------------------
// Patch im2col = PixelUnshuffle(14) in token-major form: [views, nch, H, W] (or NHWC) fp32 ->
// bf16 [views*h*w, KP] with k = c*196 + ky*14 + kx, zero padded to KP.  This is both the
// flattening of the 14x14/s14 patch-embed conv weight (patch_embed.py:65-80) and the channel
// order of nn.PixelUnshuffle(14) in DenseRepresentationEncoder (dense_rep_encoder.py:112-114).
// Optional depth transform (model.py:1021-1137): x / norm[view], then sign(x) * log1p(|x|)
// (normalize_depth_using_non_zero_pixels + apply_log_to_norm, geometry.py:1594-1626,1737-1750).
// ----------------------------------------------------------------------------------------------
__global__ void patch_im2col_kernel(const float* __restrict__ img, int views, int H, int W, int nch,
                                    int nhwc, int KP, const float* __restrict__ view_norm, int log_mode,
                                    __nv_bfloat16* __restrict__ out, float* __restrict__ out_f32) {
  const int h = H / 14, w = W / 14;
  const long long total = (long long)views * h * w * KP;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = i % KP;
    const long long row = i / KP;
    float val = 0.f;
    if (k < nch * 196) {
      const int c = k / 196, ky = (k % 196) / 14, kx = k % 14;
      const int px = row % w, py = (row / w) % h, v = row / (w * h);
      const int y = py * 14 + ky, x = px * 14 + kx;
      val = nhwc ? img[(((long long)v * H + y) * W + x) * nch + c]
                 : img[(((long long)v * nch + c) * H + y) * W + x];
      if (view_norm) val = val / view_norm[v];
      if (log_mode) val = copysignf(log1pf(fabsf(val)), val);
    }
    if (out) out[i] = __float2bfloat16(val);
    if (out_f32) out_f32[i] = tf32_rn(val);  // fp32 operand of the TF32 geometric-input encoders
  }
}

// ----------------------------------------------------------------------------------------------
// a1 pre-processing (inference.py:222-311), one thread per pixel
// ----------------------------------------------------------------------------------------------
// unit rays of a pinhole camera (get_rays_in_camera_frame, geometry.py:186-241): K [views,3,3]
__global__ void rays_from_intrinsics_kernel(const float* __restrict__ K, int views, int H, int W,
                                            float* __restrict__ rays) {
  const long long total = (long long)views * H * W;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int x = i % W, y = (i / W) % H, v = i / ((long long)W * H);
    const float* k = K + v * 9;
    const float xx = ((float)x - k[2]) / k[0], yy = ((float)y - k[5]) / k[4];
    const float n = sqrtf(xx * xx + yy * yy + 1.0f);
    rays[i * 3 + 0] = xx / n; rays[i * 3 + 1] = yy / n; rays[i * 3 + 2] = 1.0f / n;
  }
}

// r / (|r| + 1e-8)   (inference.py:251-256)
__global__ void normalize_rays_kernel(const float* __restrict__ in, long long npix, float* __restrict__ out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npix;
       i += (long long)gridDim.x * blockDim.x) {
    const float a = in[i * 3], b = in[i * 3 + 1], c = in[i * 3 + 2];
    const float n = sqrtf(a * a + b * b + c * c) + 1e-8f;
    out[i * 3] = a / n; out[i * 3 + 1] = b / n; out[i * 3 + 2] = c / n;
  }
}

// depth_along_ray = | depth_z * (r / r_z) |   (convert_z_depth_to_depth_along_ray, inference.py:263-274)
__global__ void depth_z_to_along_ray_kernel(const float* __restrict__ depth_z, const float* __restrict__ rays,
                                            long long npix, float* __restrict__ out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npix;
       i += (long long)gridDim.x * blockDim.x) {
    const float z = depth_z[i], rz = rays[i * 3 + 2];
    const float a = z * (rays[i * 3] / rz), b = z * (rays[i * 3 + 1] / rz), c = z * (rz / rz);
    out[i] = sqrtf(a * a + b * b + c * c);
  }
}

// mean of the non-zero (valid > 0) pixels of each view's depth map, clipped to >= 1e-8
// (normalize_depth_using_non_zero_pixels, geometry.py:1594-1626).  One block per view.
__global__ void depth_norm_kernel(const float* __restrict__ depth, long long npix, float* __restrict__ norm) {
  const int v = blockIdx.x;
  const float* d = depth + (long long)v * npix;
  float s = 0.f, n = 0.f;
  for (long long i = threadIdx.x; i < npix; i += blockDim.x) {
    const float x = d[i];
    if (x > 0.f) { s += x; n += 1.f; }
  }
  __shared__ float ss[256], sn[256];
  ss[threadIdx.x] = s; sn[threadIdx.x] = n;
  __syncthreads();
  for (int k = blockDim.x / 2; k > 0; k >>= 1) {
    if (threadIdx.x < k) { ss[threadIdx.x] += ss[threadIdx.x + k]; sn[threadIdx.x] += sn[threadIdx.x + k]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) norm[v] = fmaxf(ss[0] / (sn[0] + 1e-8f), 1e-8f);
}

// dst[row, :] (+)= scale[view(row)] * src[row, :]   (token-major [views*T, C] fp32; model.py:968,1137)
__global__ void accum_masked_rows_kernel(float* __restrict__ dst, const float* __restrict__ src,
                                         const float* __restrict__ scale, int T, int C, long long total4,
                                         int first) {
  const int c4n = C / 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / c4n;
    const float m = scale[row / T];
    float4 s4 = reinterpret_cast<const float4*>(src)[i];
    float4 d4 = first ? make_float4(0, 0, 0, 0) : reinterpret_cast<float4*>(dst)[i];
    d4.x += m * s4.x; d4.y += m * s4.y; d4.z += m * s4.z; d4.w += m * s4.w;
    reinterpret_cast<float4*>(dst)[i] = d4;
  }
}

// small per-view preparation of the geometric-input MLP encoders' inputs and masks
// (model.py:1141-1167,1170-1289): one thread per view.
//   in_scale[v]  = log(depth_norm[v] + 1e-8)          m_scale[v] = depth_mask & depth_metric_mask
//   in_trans[v]  = trans[v] / pose_norm               m_cam[v]   = cam_mask
//   in_tscale[v] = log(pose_norm + 1e-8)              m_tscale[v]= cam_mask & pose_metric_mask
__global__ void geo_prepare_kernel(int views, const float* __restrict__ depth_norm,
                                   const float* __restrict__ trans, float pose_norm,
                                   const unsigned char* __restrict__ ray_mask,
                                   const unsigned char* __restrict__ depth_mask,
                                   const unsigned char* __restrict__ cam_mask,
                                   const unsigned char* __restrict__ depth_metric,
                                   const unsigned char* __restrict__ pose_metric,
                                   float* __restrict__ in_scale, float* __restrict__ in_trans,
                                   float* __restrict__ in_tscale, float* __restrict__ m_ray,
                                   float* __restrict__ m_depth, float* __restrict__ m_scale,
                                   float* __restrict__ m_cam, float* __restrict__ m_tscale) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= views) return;
  const bool dm = depth_mask && depth_mask[v], cm = cam_mask && cam_mask[v];
  m_ray[v] = (ray_mask && ray_mask[v]) ? 1.f : 0.f;
  m_depth[v] = dm ? 1.f : 0.f;
  m_scale[v] = (dm && depth_metric && depth_metric[v]) ? 1.f : 0.f;
  m_cam[v] = cm ? 1.f : 0.f;
  m_tscale[v] = (cm && pose_metric && pose_metric[v]) ? 1.f : 0.f;
  in_scale[v] = depth_norm ? logf(depth_norm[v] + 1e-8f) : 0.f;
  const float pn = fmaxf(pose_norm, 1e-8f);
  for (int k = 0; k < 3; ++k) in_trans[v * 3 + k] = trans ? trans[v * 3 + k] / pn : 0.f;
  in_tscale[v] = logf(pn + 1e-8f);
}

// fused = LayerNorm(enc[row] + dense[row] + vec[view(row)])  -> bf16   (model.py:1423-1431)
// one warp per row, C = 128 * VPL.  out_f32 (optional): the same rows in fp32, the residual stream of
// an info-sharing trunk without proj_embed (width 1024, aat.py:121-124)
template <int VPL>
__global__ void fuse_norm_kernel(const float* __restrict__ enc, const float* __restrict__ dense,
                                 const float* __restrict__ vec, int rows, int T,
                                 const float* __restrict__ w, const float* __restrict__ b, float eps,
                                 __nv_bfloat16* __restrict__ out, float* __restrict__ out_f32, int round_f32) {
  constexpr int C = 128 * VPL;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float4* er = reinterpret_cast<const float4*>(enc + (long long)warp * C);
  const float4* dr = dense ? reinterpret_cast<const float4*>(dense + (long long)warp * C) : nullptr;
  const float4* vr = vec ? reinterpret_cast<const float4*>(vec + (long long)(warp / T) * C) : nullptr;
  float4 v[VPL];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    v[i] = er[lane + 32 * i];
    if (dr) { const float4 d = dr[lane + 32 * i]; v[i].x += d.x; v[i].y += d.y; v[i].z += d.z; v[i].w += d.w; }
    if (vr) { const float4 d = vr[lane + 32 * i]; v[i].x += d.x; v[i].y += d.y; v[i].z += d.z; v[i].w += d.w; }
    s += v[i].x + v[i].y + v[i].z + v[i].w;
  }
  const float mean = warp_sum(s) * (1.0f / C);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const float a = v[i].x - mean, b_ = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
    q += a * a + b_ * b_ + c * c + d * d;
  }
  const float rstd = rsqrtf(warp_sum(q) * (1.0f / C) + eps);
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const float4 g = __ldg(reinterpret_cast<const float4*>(w) + lane + 32 * i);
    const float4 h = __ldg(reinterpret_cast<const float4*>(b) + lane + 32 * i);
    const float4 y = make_float4((v[i].x - mean) * rstd * g.x + h.x, (v[i].y - mean) * rstd * g.y + h.y,
                                 (v[i].z - mean) * rstd * g.z + h.z, (v[i].w - mean) * rstd * g.w + h.w);
    reinterpret_cast<uint2*>(out + (long long)warp * C)[lane + 32 * i] =
        make_uint2(pack_bf16(y.x, y.y), pack_bf16(y.z, y.w));
    if (out_f32) reinterpret_cast<float4*>(out_f32 + (long long)warp * C)[lane + 32 * i] = round_f32 ? tf32_rn(y) : y;
  }
}

// tokens[v, 0] = cls + pos[0]; tokens[v, 1+t] = patch[v*T + t] + pos[1+t]
// (vision_transformer.py:244-251; patch already contains the conv bias)
__global__ void assemble_tokens_kernel(const float* __restrict__ patch, const float* __restrict__ cls,
                                       const float* __restrict__ pos, int views, int T, int C,
                                       float* __restrict__ tokens) {
  const long long total = (long long)views * (T + 1) * (C / 4);
  const int c4n = C / 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c4 = i % c4n;
    const int t = (i / c4n) % (T + 1);
    const int v = i / ((long long)c4n * (T + 1));
    float4 a = (t == 0) ? reinterpret_cast<const float4*>(cls)[c4]
                        : reinterpret_cast<const float4*>(patch)[((long long)v * T + t - 1) * c4n + c4];
    const float4 p = reinterpret_cast<const float4*>(pos)[(long long)t * c4n + c4];
    a.x += p.x; a.y += p.y; a.z += p.z; a.w += p.w;
    reinterpret_cast<float4*>(tokens)[i] = a;
  }
}

// Bicubic (A = -0.75, align_corners = False, explicit scale factor) resize of the DINOv2 position
// embedding grid, bit-compatible in structure with ATen's upsample_bicubic2d as called from
// vision_transformer.py:208-242: src = (dst + 0.5) / scale_factor - 0.5, border-clamped taps.
// pos_in [1 + M*M, C] -> pos_out [1 + h*w, C] (row 0 = class position, copied).
__device__ __forceinline__ void cubic_coeffs(float t, float* c) {
  const float A = -0.75f;
  auto cc1 = [&](float x) { return ((A + 2.f) * x - (A + 3.f)) * x * x + 1.f; };
  auto cc2 = [&](float x) { return ((A * x - 5.f * A) * x + 8.f * A) * x - 4.f * A; };
  c[0] = cc2(t + 1.f);
  c[1] = cc1(t);
  c[2] = cc1(1.f - t);
  c[3] = cc2(2.f - t);
}
__global__ void posembed_bicubic_kernel(const float* __restrict__ pos_in, int M, int C, int h, int w,
                                        float inv_sy, float inv_sx, float* __restrict__ pos_out) {
  const long long total = (long long)(1 + h * w) * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = i % C;
    const int t = i / C;
    if (t == 0) { pos_out[i] = pos_in[c]; continue; }
    const int oy = (t - 1) / w, ox = (t - 1) % w;
    const float sy = inv_sy * (oy + 0.5f) - 0.5f, sx = inv_sx * (ox + 0.5f) - 0.5f;
    const int iy = (int)floorf(sy), ix = (int)floorf(sx);
    float cy[4], cx[4];
    cubic_coeffs(sy - iy, cy);
    cubic_coeffs(sx - ix, cx);
    float acc = 0.f;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int yy = min(max(iy - 1 + a, 0), M - 1);
      float rowv = 0.f;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int xx = min(max(ix - 1 + b, 0), M - 1);
        rowv += pos_in[(long long)(1 + yy * M + xx) * C + c] * cx[b];
      }
      acc += rowv * cy[a];
    }
    pos_out[i] = acc;
  }
}

// x[r, :] += pe[:] for r < rows  (reference-view positional encoding, aat.py:618-626)
__global__ void add_row_vector_kernel(float* __restrict__ x, const float* __restrict__ pe, int rows,
                                      int C) {
  const long long total = (long long)rows * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x)
    x[i] += pe[i % C];
}

// x[v * T + t, :] += table[rows[v], :]  (view positional encoding of every view of the shard,
// use_pe_for_non_reference_views, aat.py:618-645; rows[v] < 0: none for that view)
__global__ void add_view_rows_kernel(float* __restrict__ x, const float* __restrict__ table,
                                     const int* __restrict__ rows, int views, int T, int C) {
  const long long per_view = (long long)T * C, total = per_view * views;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int r = rows[i / per_view];
    if (r >= 0) x[i] += table[(long long)r * C + i % C];
  }
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                   long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    y[i] = __float2bfloat16(x[i]);
}

// ----------------------------------------------------------------------------------------------
// Bilinear resize, align_corners=True, NHWC bf16, 8 channels per thread.
// Computes the (Hv x Wv) virtual output grid from (Hi x Wi) but stores only the top-left
// (Ho x Wo) crop (dpt.py:213 crops refinenet4's 2x up-sampled 28x38 to 28x37).
// dpt_block.py:251-253 (x2) and dpt.py:304 (224x296 -> 392x518).  Optional second output relu(x).
// ----------------------------------------------------------------------------------------------
__global__ void resize_bilinear_ac_kernel(const __nv_bfloat16* __restrict__ in, int views, int Hi,
                                          int Wi, int C, int Hv, int Wv, int Ho, int Wo,
                                          __nv_bfloat16* __restrict__ out,
                                          __nv_bfloat16* __restrict__ out_relu) {
  // grid = (ceil(Wo * C/8 / 256), Ho, views): the output row and view come from the block index, so the
  // per-thread index arithmetic is one 32-bit division (the flat grid-stride version spent its time in
  // four 64-bit divisions per 16-byte store)
  const int c8n = C / 8;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= Wo * c8n) return;
  const int ox = idx / c8n, c8 = idx - ox * c8n;
  const int oy = blockIdx.y, v = blockIdx.z;
  const float ry = Hv > 1 ? float(Hi - 1) / float(Hv - 1) : 0.f;
  const float rx = Wv > 1 ? float(Wi - 1) / float(Wv - 1) : 0.f;
  const float sy = ry * oy, sx = rx * ox;
  const int y0 = min((int)sy, Hi - 1), x0 = min((int)sx, Wi - 1);
  const int y1 = min(y0 + 1, Hi - 1), x1 = min(x0 + 1, Wi - 1);
  const float ly = sy - y0, lx = sx - x0, hy = 1.f - ly, hx = 1.f - lx;
  const uint4* base = reinterpret_cast<const uint4*>(in) + (long long)v * Hi * Wi * c8n + c8;
  const uint4 a = base[((long long)y0 * Wi + x0) * c8n], b = base[((long long)y0 * Wi + x1) * c8n];
  const uint4 c = base[((long long)y1 * Wi + x0) * c8n], d = base[((long long)y1 * Wi + x1) * c8n];
  const uint32_t* pa = &a.x; const uint32_t* pb = &b.x; const uint32_t* pc = &c.x; const uint32_t* pd = &d.x;
  uint32_t o[4], orl[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    // ATen: hy*(hx*a + lx*b) + ly*(hx*c + lx*d)
    const float lo = hy * (hx * bf16_lo(pa[k]) + lx * bf16_lo(pb[k])) +
                     ly * (hx * bf16_lo(pc[k]) + lx * bf16_lo(pd[k]));
    const float hi = hy * (hx * bf16_hi(pa[k]) + lx * bf16_hi(pb[k])) +
                     ly * (hx * bf16_hi(pc[k]) + lx * bf16_hi(pd[k]));
    o[k] = pack_bf16(lo, hi);
    orl[k] = pack_bf16(fmaxf(lo, 0.f), fmaxf(hi, 0.f));
  }
  const long long i = (((long long)v * Ho + oy) * Wo + ox) * c8n + c8;
  if (out) reinterpret_cast<uint4*>(out)[i] = make_uint4(o[0], o[1], o[2], o[3]);
  if (out_relu) reinterpret_cast<uint4*>(out_relu)[i] = make_uint4(orl[0], orl[1], orl[2], orl[3]);
}

// the same on fp32 NHWC activations (fp32 / TF32 dense head): thread = 4 channels of RESIZE_ROWS consecutive
// output rows of one output column -- the column's x taps and weights are shared, and the 4 * RESIZE_ROWS
// 16-byte loads are all issued before the first blend (the one-row version ran at 2.1 TB/s of stores:
// latency-bound, 395 us for the 224x296 -> 392x518 resize of 8 views)
constexpr int RESIZE_ROWS = 4;
__global__ void resize_bilinear_ac_f32_kernel(const float* __restrict__ in, int views, int Hi, int Wi, int C,
                                              int Hv, int Wv, int Ho, int Wo, float* __restrict__ out,
                                              float* __restrict__ out_relu) {
  const int c4n = C / 4;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= Wo * c4n) return;
  const int ox = idx / c4n, c4 = idx - ox * c4n;
  const int oy0 = blockIdx.y * RESIZE_ROWS, v = blockIdx.z;
  const float ry = Hv > 1 ? float(Hi - 1) / float(Hv - 1) : 0.f;
  const float rx = Wv > 1 ? float(Wi - 1) / float(Wv - 1) : 0.f;
  const float sx = rx * ox;
  const int x0 = min((int)sx, Wi - 1), x1 = min(x0 + 1, Wi - 1);
  const float lx = sx - x0, hx = 1.f - lx;
  const float4* base = reinterpret_cast<const float4*>(in) + (long long)v * Hi * Wi * c4n + c4;
  float4 a[RESIZE_ROWS], b[RESIZE_ROWS], c[RESIZE_ROWS], d[RESIZE_ROWS];
  float ly[RESIZE_ROWS];
#pragma unroll
  for (int r = 0; r < RESIZE_ROWS; ++r) {
    const int oy = min(oy0 + r, Ho - 1);
    const float sy = ry * oy;
    const int y0 = min((int)sy, Hi - 1), y1 = min(y0 + 1, Hi - 1);
    ly[r] = sy - y0;
    a[r] = __ldg(base + ((long long)y0 * Wi + x0) * c4n); b[r] = __ldg(base + ((long long)y0 * Wi + x1) * c4n);
    c[r] = __ldg(base + ((long long)y1 * Wi + x0) * c4n); d[r] = __ldg(base + ((long long)y1 * Wi + x1) * c4n);
  }
#pragma unroll
  for (int r = 0; r < RESIZE_ROWS; ++r) {
    const int oy = oy0 + r;
    if (oy >= Ho) break;
    const float hy = 1.f - ly[r];
    float4 o;  // ATen: hy*(hx*a + lx*b) + ly*(hx*c + lx*d)
    o.x = hy * (hx * a[r].x + lx * b[r].x) + ly[r] * (hx * c[r].x + lx * d[r].x);
    o.y = hy * (hx * a[r].y + lx * b[r].y) + ly[r] * (hx * c[r].y + lx * d[r].y);
    o.z = hy * (hx * a[r].z + lx * b[r].z) + ly[r] * (hx * c[r].z + lx * d[r].z);
    o.w = hy * (hx * a[r].w + lx * b[r].w) + ly[r] * (hx * c[r].w + lx * d[r].w);
    o = tf32_rn(o);  // operand of the next kind::tf32 contraction
    const long long i = (((long long)v * Ho + oy) * Wo + ox) * c4n + c4;
    if (out) __stcs(reinterpret_cast<float4*>(out) + i, o);   // streamed: read once by the next conv
    if (out_relu)
      __stcs(reinterpret_cast<float4*>(out_relu) + i,
             make_float4(fmaxf(o.x, 0.f), fmaxf(o.y, 0.f), fmaxf(o.z, 0.f), fmaxf(o.w, 0.f)));
  }
}

// ConvTranspose2d with kernel == stride (dpt.py:117-127,139-149) = GEMM + pixel shuffle.
// in: GEMM output [views*h*w, k*k*C] with column (i*k + j)*C + c  ->  NHWC [views, h*k, w*k, C]
__global__ void pixel_shuffle_kernel(const __nv_bfloat16* __restrict__ in, int views, int h, int w,
                                     int k, int C, __nv_bfloat16* __restrict__ out) {
  const int c8n = C / 8;
  const long long total = (long long)views * h * k * w * k * c8n;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c8 = i % c8n;
    const int ox = (i / c8n) % (w * k);
    const int oy = (i / ((long long)c8n * w * k)) % (h * k);
    const int v = i / ((long long)c8n * w * k * h * k);
    const int y = oy / k, ii = oy % k, x = ox / k, jj = ox % k;
    const long long src = (((long long)v * h + y) * w + x) * (k * k * c8n) + (ii * k + jj) * c8n + c8;
    reinterpret_cast<uint4*>(out)[i] = reinterpret_cast<const uint4*>(in)[src];
  }
}

// im2col for the 3x3 / stride 2 / pad 1 conv of act_4_postprocess (dpt.py:160-166):
// NHWC [views,H,W,C] -> [views*Ho*Wo, 9*C], column tap*C + c
__global__ void im2col3x3_s2_kernel(const __nv_bfloat16* __restrict__ in, int views, int H, int W,
                                    int C, int Ho, int Wo, __nv_bfloat16* __restrict__ out) {
  const int c8n = C / 8;
  const long long total = (long long)views * Ho * Wo * 9 * c8n;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c8 = i % c8n;
    const int tap = (i / c8n) % 9;
    const int ox = (i / ((long long)c8n * 9)) % Wo;
    const int oy = (i / ((long long)c8n * 9 * Wo)) % Ho;
    const int v = i / ((long long)c8n * 9 * Wo * Ho);
    const int y = oy * 2 - 1 + tap / 3, x = ox * 2 - 1 + tap % 3;
    uint4 val = make_uint4(0, 0, 0, 0);
    if (y >= 0 && y < H && x >= 0 && x < W)
      val = reinterpret_cast<const uint4*>(in)[(((long long)v * H + y) * W + x) * c8n + c8];
    reinterpret_cast<uint4*>(out)[i] = val;
  }
}

// mean over the T tokens of each view: bf16 [views*T, C] -> fp32 [views, C]  (pose_head.py:150)
// ld = row pitch in elements; lo_off > 0: x is a bf16x2 split value, the low part lo_off elements
// behind the high part
__global__ void token_mean_kernel(const __nv_bfloat16* __restrict__ x, int T, int C, int ld, int lo_off,
                                  float* __restrict__ out) {
  const int v = blockIdx.y;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;  // blockDim.x channels, blockDim.y row lanes
  __shared__ float part[8][128];
  float s = 0.f;
  if (c < C)
#pragma unroll 8  // independent loads in flight: the loop is DRAM-latency bound otherwise (67 -> ~10 us)
    for (int t = threadIdx.y; t < T; t += blockDim.y) {
      const __nv_bfloat16* r = x + ((long long)v * T + t) * ld + c;
      s += __bfloat162float(r[0]) + (lo_off ? __bfloat162float(r[lo_off]) : 0.f);
    }
  part[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    float tot = 0.f;
    for (int k = 0; k < blockDim.y; ++k) tot += part[k][threadIdx.x];
    out[(long long)v * C + c] = tot / T;
  }
}

// y[r, n] = act(sum_k x[r,k] * w[n,k] + b[n]) in fp32 for the tiny head MLPs
// (pose_head.py:152-157, mlp_head.py:80-90, global_rep_encoder.py:83-105).  One warp per (n, 8 rows).
// act: 0 none, 1 gelu (exact), 2 relu
__global__ void small_linear_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                    const float* __restrict__ b, int rows, int K, int N, int act,
                                    float* __restrict__ y, int ldx, int ldy) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int n = warp % N, rblk = warp / N;
  const int r0 = rblk * 8;
  if (r0 >= rows) return;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll 4  // independent loads in flight (the loop is latency-bound: one warp per output column)
  for (int k = lane; k < K; k += 32) {
    const float wv = __ldg(w + (long long)n * K + k);
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (r0 + i < rows) acc[i] = fmaf(x[(long long)(r0 + i) * ldx + k], wv, acc[i]);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float v = warp_sum(acc[i]);
    if (lane == 0 && r0 + i < rows) {
      v += b ? b[n] : 0.f;
      if (act == 1) v = 0.5f * v * (1.0f + erff(v * 0.70710678118654752f));
      else if (act == 2) v = fmaxf(v, 0.f);
      y[(long long)(r0 + i) * ldy + n] = v;
    }
  }
}

// pose adaptor (adaptors.py:714-732): raw [views, 7] (t | q) -> pose [views, 8] (t | q/|q| | 0)
__global__ void pose_finalize_kernel(const float* __restrict__ raw, int views, float* __restrict__ pose) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= views) return;
  const float* r = raw + v * 7;
  const float n = fmaxf(sqrtf(r[3] * r[3] + r[4] * r[4] + r[5] * r[5] + r[6] * r[6]), 1e-8f);
  float* p = pose + v * 8;
  p[0] = r[0]; p[1] = r[1]; p[2] = r[2];
  p[3] = r[3] / n; p[4] = r[4] / n; p[5] = r[5] / n; p[6] = r[6] / n; p[7] = 0.f;
}

// scale adaptor (adaptors.py:190-212): exp then clip to >= 1e-8
__global__ void scale_finalize_kernel(const float* __restrict__ raw, float* __restrict__ scale) {
  if (threadIdx.x == 0 && blockIdx.x == 0) scale[0] = fmaxf(expf(raw[0]), 1e-8f);
}

// ----------------------------------------------------------------------------------------------
// RoPE2D, in place (the reference tree's one native kernel: uniception/models/libs/croco/curope/
// kernels.cu:17-108, scalar loop curope.cpp:11-47; used by Attention.forward when a custom
// positional encoding is configured, transformer_blocks.py:179-184).  A head vector of D elements
// is [uY(Q) | vY(Q) | uX(Q) | vX(Q)], Q = D/4; for the y half (x half) the pair (u_d, v_d) is
// rotated by the angle fwd * pos_y(x) / base^(d/Q).  Tokens are addressed as
// base + token * pitch + head * D, so the q and k thirds of a packed [rows, 3*H*D] qkv buffer are
// rotated where they lie (no head split / transposes).  Thread = 4 consecutive d of one half of
// one (token, head): two 16-byte (fp32) or 8-byte (bf16) loads and stores, all lanes of a warp
// touch one contiguous run of the row.  Bandwidth bound: D*sizeof(T) read + written per head.
// ----------------------------------------------------------------------------------------------
template <typename T>
__global__ void rope2d_kernel(T* __restrict__ tokens, long long n_tokens, int heads, int D, long long pitch,
                              const long long* __restrict__ pos, float base, float fwd) {
  const int Q = D >> 2, gq = Q >> 2;           // gq groups of 4 per quarter
  const int per_head = 2 * gq;                 // threads per (token, head)
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long th = idx / per_head;         // token * heads + head
  if (th >= n_tokens * heads) return;
  const int sub = (int)(idx - th * per_head);
  const int half = sub / gq, g = sub - half * gq;
  const long long tok = th / heads;
  const int head = (int)(th - tok * heads);
  const float p = (float)pos[tok * 2 + half];
  T* u_ptr = tokens + tok * pitch + (long long)head * D + half * 2 * Q + 4 * g;
  T* v_ptr = u_ptr + Q;
  float u[4], v[4];
  if (sizeof(T) == 4) {
    const float4 a = *reinterpret_cast<const float4*>(u_ptr), b = *reinterpret_cast<const float4*>(v_ptr);
    u[0] = a.x; u[1] = a.y; u[2] = a.z; u[3] = a.w; v[0] = b.x; v[1] = b.y; v[2] = b.z; v[3] = b.w;
  } else {
    const uint2 a = *reinterpret_cast<const uint2*>(u_ptr), b = *reinterpret_cast<const uint2*>(v_ptr);
    u[0] = bf16_lo(a.x); u[1] = bf16_hi(a.x); u[2] = bf16_lo(a.y); u[3] = bf16_hi(a.y);
    v[0] = bf16_lo(b.x); v[1] = bf16_hi(b.x); v[2] = bf16_lo(b.y); v[3] = bf16_hi(b.y);
  }
  float ru[4], rv[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int d = 4 * g + i;
    const float ang = fwd * p / powf(base, (float)d / (float)Q);  // curope.cpp:36
    float s, c;
    sincosf(ang, &s, &c);
    ru[i] = u[i] * c - v[i] * s;
    rv[i] = v[i] * c + u[i] * s;
  }
  if (sizeof(T) == 4) {
    *reinterpret_cast<float4*>(u_ptr) = make_float4(ru[0], ru[1], ru[2], ru[3]);
    *reinterpret_cast<float4*>(v_ptr) = make_float4(rv[0], rv[1], rv[2], rv[3]);
  } else {
    *reinterpret_cast<uint2*>(u_ptr) = make_uint2(pack_bf16(ru[0], ru[1]), pack_bf16(ru[2], ru[3]));
    *reinterpret_cast<uint2*>(v_ptr) = make_uint2(pack_bf16(rv[0], rv[1]), pack_bf16(rv[2], rv[3]));
  }
}

}  // namespace mab
