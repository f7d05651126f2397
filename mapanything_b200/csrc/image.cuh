// Image input path on the GPU: what `load_images`
// (mapanything/utils/image.py:134-332) does per image on the CPU after decoding -- PIL LANCZOS /
// BICUBIC resize (mapanything/utils/cropping.py:230-244), centred crop (:441-447), torchvision
// ToTensor + Normalize (image.py:297-300) -- as two byte/integer kernels that reproduce Pillow's
// fixed-point resampling bit for bit (Pillow src/libImaging/Resample.c: 22-bit coefficients,
// accumulation from 2^21, >> 22, clip; horizontal pass first with an 8-bit intermediate).
// Only the rows / columns that survive the crop are computed.  HBM bound: the source image is read
// once (3 B/pixel), the intermediate is rows_needed x out_w x 3 B, the output 12 B/pixel.
#pragma once
#include "common.cuh"

namespace mab {

constexpr int IMG_PRECISION_BITS = 32 - 8 - 2;

__device__ __forceinline__ unsigned char img_clip8(int v) {
  v >>= IMG_PRECISION_BITS;
  return (unsigned char)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

// horizontal pass: src [in_h, in_w, 3] u8 -> tmp [rows, out_w, 3] u8 for source rows y0 .. y0+rows-1
// and the out_w output columns of the crop.  bounds [out_w, 2] (xmin, count), kk [out_w, ksize].
__global__ void image_resample_h_kernel(const unsigned char* __restrict__ src, int in_w, int y0, int rows,
                                        int out_w, const int* __restrict__ bounds, const int* __restrict__ kk,
                                        int ksize, unsigned char* __restrict__ tmp) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (x >= out_w || r >= rows) return;
  const int xmin = bounds[2 * x], n = bounds[2 * x + 1];
  const unsigned char* line = src + ((long long)(y0 + r) * in_w + xmin) * 3;
  const int* k = kk + (long long)x * ksize;
  int s0 = 1 << (IMG_PRECISION_BITS - 1), s1 = s0, s2 = s0;
  for (int i = 0; i < n; ++i) {
    const int c = k[i];
    s0 += line[3 * i + 0] * c; s1 += line[3 * i + 1] * c; s2 += line[3 * i + 2] * c;
  }
  unsigned char* o = tmp + ((long long)r * out_w + x) * 3;
  o[0] = img_clip8(s0); o[1] = img_clip8(s1); o[2] = img_clip8(s2);
}

// vertical pass (or a plain copy when the height is unchanged: ksize == 0) + ToTensor + Normalize:
// tmp [rows, out_w, 3] u8 -> out [3, out_h, out_w] fp32 = (u8 / 255 - mean) / std, optional u8 crop
__global__ void image_resample_v_norm_kernel(const unsigned char* __restrict__ tmp, int y0, int out_w, int out_h,
                                             const int* __restrict__ bounds, const int* __restrict__ kk, int ksize,
                                             float m0, float m1, float m2, float d0, float d1, float d2,
                                             float* __restrict__ out, unsigned char* __restrict__ out_u8) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= out_w || y >= out_h) return;
  unsigned char p0, p1, p2;
  if (ksize > 0) {
    const int ymin = bounds[2 * y], n = bounds[2 * y + 1];
    const int* k = kk + (long long)y * ksize;
    int s0 = 1 << (IMG_PRECISION_BITS - 1), s1 = s0, s2 = s0;
    for (int i = 0; i < n; ++i) {
      const unsigned char* px = tmp + ((long long)(ymin + i - y0) * out_w + x) * 3;
      const int c = k[i];
      s0 += px[0] * c; s1 += px[1] * c; s2 += px[2] * c;
    }
    p0 = img_clip8(s0); p1 = img_clip8(s1); p2 = img_clip8(s2);
  } else {
    const unsigned char* px = tmp + ((long long)y * out_w + x) * 3;
    p0 = px[0]; p1 = px[1]; p2 = px[2];
  }
  const long long plane = (long long)out_h * out_w, o = (long long)y * out_w + x;
  // torchvision: img.to(float32).div(255) ; tensor.sub_(mean).div_(std)
  out[o] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)p0, 255.0f), m0), d0);
  out[plane + o] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)p1, 255.0f), m1), d1);
  out[2 * plane + o] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)p2, 255.0f), m2), d2);
  if (out_u8) { out_u8[o * 3] = p0; out_u8[o * 3 + 1] = p1; out_u8[o * 3 + 2] = p2; }
}

// cv2.resize(..., INTER_NEAREST) of a one-channel fp32 map followed by the crop (cropping.py:246-253,
// 345-349): out[y, x] = src[min(floor((y + top) * ify), in_h - 1), min(floor((x + left) * ifx), in_w - 1)]
// with ifx = 1 / (resize_w / in_w) in double, as OpenCV's resizeNN computes its offset table.
__global__ void resize_nearest_crop_kernel(const float* __restrict__ src, int in_h, int in_w, double ifx,
                                           double ify, int left, int top, int out_w, int out_h,
                                           float* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= out_w || y >= out_h) return;
  int sx = (int)floor((double)(x + left) * ifx), sy = (int)floor((double)(y + top) * ify);
  sx = sx < in_w - 1 ? sx : in_w - 1;
  sy = sy < in_h - 1 ? sy : in_h - 1;
  out[(long long)y * out_w + x] = src[(long long)sy * in_w + sx];
}

}  // namespace mab
