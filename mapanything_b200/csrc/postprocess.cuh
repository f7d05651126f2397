// GPU replacement for postprocess_model_outputs_for_inference (mapanything/utils/inference.py:314-506),
// which the reference runs per view on the CPU in numpy.  Float ops are issued in the same order and
// without FMA contraction (__fmul_rn/__fadd_rn) so that the boolean masks agree with numpy's fp32
// arithmetic on identical inputs.
#pragma once
#include "common.cuh"

namespace mab {

// img_no_norm = clip(img * std + mean, 0, 1), NCHW -> NHWC   (image.py:93-131)
__global__ void denorm_image_kernel(const float* __restrict__ img, int views, int H, int W,
                                    float m0, float m1, float m2, float s0, float s1, float s2,
                                    float* __restrict__ out) {
  const long long total = (long long)views * H * W;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long v = i / ((long long)H * W), pix = i % ((long long)H * W);
    const float* src = img + v * 3 * H * W + pix;
    const float r = __fadd_rn(__fmul_rn(src[0], s0), m0);
    const float g = __fadd_rn(__fmul_rn(src[(long long)H * W], s1), m1);
    const float b = __fadd_rn(__fmul_rn(src[2LL * H * W], s2), m2);
    out[i * 3 + 0] = fminf(fmaxf(r, 0.f), 1.f);
    out[i * 3 + 1] = fminf(fmaxf(g, 0.f), 1.f);
    out[i * 3 + 2] = fminf(fmaxf(b, 0.f), 1.f);
  }
}

// Pinhole intrinsics by least squares on a sub-sampled ray grid (geometry.py:378-430) and the
// 4x4 cam2world matrix (inference.py:390-405).  One block per view.
__global__ void intrinsics_pose_kernel(const float* __restrict__ rays, int H, int W,
                                       const float* __restrict__ pose, const float* __restrict__ scale,
                                       float* __restrict__ K_out, float* __restrict__ pose_out,
                                       float* __restrict__ cam_trans, float* __restrict__ cam_quats) {
  const int v = blockIdx.x;
  const int sh = max(1, H / 50), sw = max(1, W / 50);
  const int nh = (H + sh - 1) / sh, nw = (W + sw - 1) / sw;
  double a[7] = {0, 0, 0, 0, 0, 0, 0};  // n, Srx, Srx2, Sx, Srx*x, | Sry.. handled below
  double sry = 0, sry2 = 0, sy = 0, sryy = 0;
  for (int i = threadIdx.x; i < nh * nw; i += blockDim.x) {
    const int y = (i / nw) * sh, x = (i % nw) * sw;
    const float* r = rays + (((long long)v * H + y) * W + x) * 3;
    const double rx = (double)(r[0] / r[2]), ry = (double)(r[1] / r[2]);
    a[0] += 1.0; a[1] += rx; a[2] += rx * rx; a[3] += x; a[4] += rx * x;
    sry += ry; sry2 += ry * ry; sy += y; sryy += ry * y;
  }
  __shared__ double red[9][256];
  double vals[9] = {a[0], a[1], a[2], a[3], a[4], sry, sry2, sy, sryy};
  for (int k = 0; k < 9; ++k) red[k][threadIdx.x] = vals[k];
  __syncthreads();
  for (int s = blockDim.x / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s)
      for (int k = 0; k < 9; ++k) red[k][threadIdx.x] += red[k][threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double n = red[0][0];
    // [n Sr; Sr Sr2] [c; f] = [Sx; Srx]
    const double detx = n * red[2][0] - red[1][0] * red[1][0];
    const double cx = (red[3][0] * red[2][0] - red[1][0] * red[4][0]) / detx;
    const double fx = (n * red[4][0] - red[1][0] * red[3][0]) / detx;
    const double dety = n * red[6][0] - red[5][0] * red[5][0];
    const double cy = (red[7][0] * red[6][0] - red[5][0] * red[8][0]) / dety;
    const double fy = (n * red[8][0] - red[5][0] * red[7][0]) / dety;
    float* K = K_out + v * 9;
    K[0] = (float)fx; K[1] = 0; K[2] = (float)cx; K[3] = 0; K[4] = (float)fy; K[5] = (float)cy;
    K[6] = 0; K[7] = 0; K[8] = 1.f;
    if (!pose) return;  // intrinsics only (mab_recover_intrinsics)
    const float* p = pose + v * 8;
    const float s = scale[0];
    // quaternion_to_rotation_matrix re-normalises (geometry.py:617)
    const float qn = sqrtf(p[3] * p[3] + p[4] * p[4] + p[5] * p[5] + p[6] * p[6]);
    const float x = p[3] / qn, y = p[4] / qn, z = p[5] / qn, w = p[6] / qn;
    float* P = pose_out + v * 16;
    P[0] = 1 - 2 * (y * y + z * z); P[1] = 2 * (x * y - w * z); P[2] = 2 * (x * z + w * y);
    P[4] = 2 * (x * y + w * z); P[5] = 1 - 2 * (x * x + z * z); P[6] = 2 * (y * z - w * x);
    P[8] = 2 * (x * z - w * y); P[9] = 2 * (y * z + w * x); P[10] = 1 - 2 * (x * x + y * y);
    P[3] = p[0] * s; P[7] = p[1] * s; P[11] = p[2] * s;
    P[12] = 0; P[13] = 0; P[14] = 0; P[15] = 1.f;
    cam_trans[v * 3 + 0] = p[0] * s; cam_trans[v * 3 + 1] = p[1] * s; cam_trans[v * 3 + 2] = p[2] * s;
    cam_quats[v * 4 + 0] = p[3]; cam_quats[v * 4 + 1] = p[4]; cam_quats[v * 4 + 2] = p[5];
    cam_quats[v * 4 + 3] = p[6];
  }
}

// ---- edge mask, stage 1: per-pixel normal + normal validity (points_to_normals, geometry.py:1788-1851)
struct F3 { float x, y, z; };
__device__ __forceinline__ F3 f3sub(F3 a, F3 b) {
  return {__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)};
}
__device__ __forceinline__ F3 f3cross(F3 a, F3 b) {  // np.cross: a1*b2 - a2*b1, ...
  return {__fsub_rn(__fmul_rn(a.y, b.z), __fmul_rn(a.z, b.y)),
          __fsub_rn(__fmul_rn(a.z, b.x), __fmul_rn(a.x, b.z)),
          __fsub_rn(__fmul_rn(a.x, b.y), __fmul_rn(a.y, b.x))};
}
__device__ __forceinline__ float f3norm(F3 a) {
  return sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(a.x, a.x), __fmul_rn(a.y, a.y)), __fmul_rn(a.z, a.z)));
}
__device__ __forceinline__ F3 f3unit(F3 a) {  // n / (|n| + 1e-12)
  const float d = __fadd_rn(f3norm(a), 1e-12f);
  return {a.x / d, a.y / d, a.z / d};
}

__global__ void normals_kernel(const float* __restrict__ pts, const unsigned char* __restrict__ fmask,
                               int views, int H, int W, float* __restrict__ normals,
                               unsigned char* __restrict__ nmask) {
  const long long total = (long long)views * H * W;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int x = i % W, y = (i / W) % H;
    const long long vb = i - (long long)y * W - x;  // first pixel of this view
    auto P = [&](int yy, int xx) -> F3 {  // zero padded (geometry.py:1813-1814)
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) return {0.f, 0.f, 0.f};
      const float* p = pts + (vb + (long long)yy * W + xx) * 3;
      return {p[0], p[1], p[2]};
    };
    auto M = [&](int yy, int xx) -> bool {
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) return false;
      return fmask[vb + (long long)yy * W + xx] != 0;
    };
    const F3 c = P(y, x);
    const F3 up = f3sub(P(y - 1, x), c), left = f3sub(P(y, x - 1), c);
    const F3 down = f3sub(P(y + 1, x), c), right = f3sub(P(y, x + 1), c);
    const F3 n0 = f3unit(f3cross(up, left)), n1 = f3unit(f3cross(left, down));
    const F3 n2 = f3unit(f3cross(down, right)), n3 = f3unit(f3cross(right, up));
    const bool mc = M(y, x), mu = M(y - 1, x), ml = M(y, x - 1), md = M(y + 1, x), mr = M(y, x + 1);
    const float v0 = (mu && ml && mc) ? 1.f : 0.f, v1 = (ml && md && mc) ? 1.f : 0.f;
    const float v2 = (md && mr && mc) ? 1.f : 0.f, v3 = (mr && mu && mc) ? 1.f : 0.f;
    F3 s;
    s.x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(n0.x, v0), __fmul_rn(n1.x, v1)), __fmul_rn(n2.x, v2)), __fmul_rn(n3.x, v3));
    s.y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(n0.y, v0), __fmul_rn(n1.y, v1)), __fmul_rn(n2.y, v2)), __fmul_rn(n3.y, v3));
    s.z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(n0.z, v0), __fmul_rn(n1.z, v1)), __fmul_rn(n2.z, v2)), __fmul_rn(n3.z, v3));
    F3 n = f3unit(s);
    const bool any = (v0 + v1 + v2 + v3) > 0.f;
    if (!any) n = {0.f, 0.f, 0.f};
    // normals_edge re-normalises its input (geometry.py:2216)
    n = f3unit(n);
    normals[i * 3 + 0] = n.x; normals[i * 3 + 1] = n.y; normals[i * 3 + 2] = n.z;
    nmask[i] = any ? 1 : 0;
  }
}

// ---- stage 2: max angle to the 3x3 neighbourhood (edge padded), NaN-propagating like np.max
__global__ void normal_angle_kernel(const float* __restrict__ normals,
                                    const unsigned char* __restrict__ nmask, int views, int H, int W,
                                    float* __restrict__ ang) {
  const long long total = (long long)views * H * W;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int x = i % W, y = (i / W) % H;
    const long long vb = i - (long long)y * W - x;
    const float cx = normals[i * 3], cy = normals[i * 3 + 1], cz = normals[i * 3 + 2];
    float best = 0.f;
    bool nan = false;
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        const int yy = min(max(y + dy, 0), H - 1), xx = min(max(x + dx, 0), W - 1);
        const long long j = vb + (long long)yy * W + xx;
        float a = 0.f;
        if (nmask[j]) {
          const float d = __fadd_rn(__fadd_rn(__fmul_rn(cx, normals[j * 3]), __fmul_rn(cy, normals[j * 3 + 1])),
                                    __fmul_rn(cz, normals[j * 3 + 2]));
          a = acosf(d);
        }
        if (a != a) nan = true;
        best = fmaxf(best, a);
      }
    ang[i] = nan ? __int_as_float(0x7fc00000) : best;
  }
}

// ---- stage 3: pooled normal edge + depth edge -> final mask, applied in place to the dense outputs
__global__ void edge_mask_apply_kernel(const float* __restrict__ ang, const unsigned char* __restrict__ fmask,
                                       int views, int H, int W, int use_edges, double normal_tol_rad,
                                       float depth_rtol, float* __restrict__ pts3d,
                                       float* __restrict__ pts3d_cam, float* __restrict__ depth_ray,
                                       float* __restrict__ depth_z, unsigned char* __restrict__ mask_out) {
  const long long total = (long long)views * H * W;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int x = i % W, y = (i / W) % H;
    const long long vb = i - (long long)y * W - x;
    bool keep = fmask[i] != 0;
    if (use_edges) {
      // max_pool_2d with NaN padding + nanmax (geometry.py:1965-2099)
      float amax = -INFINITY, dmax = -INFINITY, nmax = -INFINITY;
      bool any_a = false;
      for (int dy = -1; dy <= 1; ++dy)
        for (int dx = -1; dx <= 1; ++dx) {
          const int yy = y + dy, xx = x + dx;
          if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
          const long long j = vb + (long long)yy * W + xx;
          const float a = ang[j];
          if (a == a) { amax = fmaxf(amax, a); any_a = true; }
          if (fmask[j]) {
            const float d = pts3d_cam[j * 3 + 2];
            dmax = fmaxf(dmax, d);
            nmax = fmaxf(nmax, -d);
          }
        }
      const bool normal_edge = any_a && ((double)amax > normal_tol_rad);
      const float diff = __fadd_rn(dmax, nmax);
      const bool depth_edge = (diff / pts3d_cam[i * 3 + 2]) > depth_rtol;
      keep = keep && !(depth_edge && normal_edge);
    }
    mask_out[i] = keep ? 1 : 0;
  }
}

// ---- confidence mask (inference.py:440-452): per-view torch.quantile(conf, q) with linear
// interpolation = exact order statistics k_lo, k_hi of the view's H*W confidences, found by an
// 8-bit-digit radix select over the order-preserving integer image of the floats.
// grid (2, views): blockIdx.x selects k_lo / k_hi.  stat[view*2 + which] receives the value.
__global__ void conf_order_stat_kernel(const float* __restrict__ conf, long long npix_view, long long k_lo,
                                       long long k_hi, float* __restrict__ stat) {
  __shared__ unsigned int hist[256];
  __shared__ unsigned int s_prefix;
  __shared__ long long s_k;
  const float* c = conf + (long long)blockIdx.y * npix_view;
  if (threadIdx.x == 0) { s_prefix = 0u; s_k = blockIdx.x == 0 ? k_lo : k_hi; }
  unsigned int mask = 0u;
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int b = threadIdx.x; b < 256; b += blockDim.x) hist[b] = 0u;
    __syncthreads();
    const unsigned int prefix = s_prefix;
    for (long long i = threadIdx.x; i < npix_view; i += blockDim.x) {
      unsigned int u = __float_as_uint(c[i]);
      u ^= (u >> 31) ? 0xffffffffu : 0x80000000u;
      if ((u & mask) == prefix) atomicAdd(&hist[(u >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      long long k = s_k;
      int b = 0;
      for (; b < 255; ++b) {
        if (k < (long long)hist[b]) break;
        k -= hist[b];
      }
      s_k = k;
      s_prefix = prefix | ((unsigned int)b << shift);
    }
    mask |= 0xffu << shift;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    unsigned int u = s_prefix;
    u ^= (u >> 31) ? 0x80000000u : 0xffffffffu;
    stat[blockIdx.y * 2 + blockIdx.x] = __uint_as_float(u);
  }
}

// fm = non_ambiguous & (conf > lerp(stat_lo, stat_hi, w))   (at::lerp's two-sided formula)
__global__ void conf_mask_kernel(const float* __restrict__ conf, const unsigned char* __restrict__ fmask,
                                 const float* __restrict__ stat, float w, long long npix_view, long long total,
                                 unsigned char* __restrict__ out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long v = i / npix_view;
    const float a = stat[v * 2], b = stat[v * 2 + 1];
    const float thr = w < 0.5f ? __fmaf_rn(w, __fsub_rn(b, a), a) : __fsub_rn(b, __fmul_rn(__fsub_rn(b, a), __fsub_rn(1.0f, w)));
    out[i] = (fmask[i] && conf[i] > thr) ? 1 : 0;
  }
}

// depth_z = pts3d_cam[..., 2] and masking happen in a second pass so that stage 3 reads unmasked z
__global__ void apply_mask_kernel(const unsigned char* __restrict__ mask, long long npix, int apply,
                                  float* __restrict__ pts3d, float* __restrict__ pts3d_cam,
                                  float* __restrict__ depth_ray, float* __restrict__ depth_z) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npix;
       i += (long long)gridDim.x * blockDim.x) {
    const float m = (!apply || mask[i]) ? 1.f : 0.f;
    const float z = pts3d_cam[i * 3 + 2];
    depth_z[i] = z * m;
    if (apply) {
      pts3d[i * 3] *= m; pts3d[i * 3 + 1] *= m; pts3d[i * 3 + 2] *= m;
      pts3d_cam[i * 3] *= m; pts3d_cam[i * 3 + 1] *= m; pts3d_cam[i * 3 + 2] *= m;
      depth_ray[i] *= m;
    }
  }
}

}  // namespace mab
