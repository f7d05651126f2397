// Dense export of the predictions: depth map -> world-frame point map
// (mapanything/utils/geometry.py:24-114 `depthmap_to_camera_frame` / `depthmap_to_world_frame`) and
// the mask-filtered, colour-tagged point cloud every demo builds from it on the CPU
// (scripts/demo_local_weight.py:122-140: mask &= depth > 0; mapanything/utils/viz.py:233-239,275-279:
// points[mask], (rgb * 255).astype(uint8)[mask], frame by frame).
//
// Ordered stream compaction in three launches: per-tile counts, one exclusive scan over the tiles,
// then a scatter pass that recomputes the flags.  The output order is numpy's boolean-mask order
// (view-major, row-major), so indices and colours are bit-exact; the points are fp32 arithmetic in
// the reference's operand order.  Bandwidth bound: 4+1 B read twice + 12 B read + 15 B written per
// kept pixel.
#pragma once
#include "common.cuh"

namespace mab {

constexpr int EXP_THREADS = 256;
constexpr int EXP_PPT = 4;                          // consecutive pixels per thread
constexpr int EXP_TILE = EXP_THREADS * EXP_PPT;     // pixels per block; tiles never straddle views

__device__ __forceinline__ bool export_valid(const float* depth, const unsigned char* mask, long long i) {
  return (mask == nullptr || mask[i] != 0) && depth[i] > 0.0f;
}

// pass 1: tile_counts[view * tiles_per_view + tile] = kept pixels of the tile
__global__ void export_count_kernel(const float* __restrict__ depth, const unsigned char* __restrict__ mask,
                                    int npix_view, int tiles_per_view, int* __restrict__ tile_counts) {
  const int view = blockIdx.x / tiles_per_view, tile = blockIdx.x - view * tiles_per_view;
  const int p0 = tile * EXP_TILE + threadIdx.x * EXP_PPT;
  int n = 0;
#pragma unroll
  for (int k = 0; k < EXP_PPT; ++k)
    if (p0 + k < npix_view) n += export_valid(depth, mask, (long long)view * npix_view + p0 + k) ? 1 : 0;
  __shared__ int wsum[EXP_THREADS / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
#pragma unroll
    for (int w = 0; w < EXP_THREADS / 32; ++w) t += wsum[w];
    tile_counts[blockIdx.x] = t;
  }
}

// pass 2 (one block): exclusive scan of the tile counts -> tile_offsets (int64), per-view counts and
// the total: counts[0..views) = kept pixels per view, counts[views] = total
__global__ void export_scan_kernel(const int* __restrict__ tile_counts, int num_tiles, int tiles_per_view,
                                   int views, long long* __restrict__ tile_offsets,
                                   long long* __restrict__ counts) {
  __shared__ long long wtot[32];
  __shared__ long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int base = 0; base < num_tiles; base += blockDim.x) {
    const int i = base + threadIdx.x;
    const long long v = i < num_tiles ? tile_counts[i] : 0;
    long long incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const long long u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    if (lane == 31) wtot[warp] = incl;
    __syncthreads();
    long long wbase = 0;
    for (int w = 0; w < warp; ++w) wbase += wtot[w];
    if (i < num_tiles) tile_offsets[i] = carry + wbase + incl - v;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) {
      long long t = 0;
      for (int w = 0; w < nwarp; ++w) t += wtot[w];
      carry += t;
    }
    __syncthreads();
  }
  // per-view counts from the offsets of each view's first tile
  for (int v = threadIdx.x; v <= views; v += blockDim.x) {
    const long long lo = v < views ? tile_offsets[v * tiles_per_view] : carry;
    if (v > 0) counts[v - 1] = lo - tile_offsets[(v - 1) * tiles_per_view];
    if (v == views) counts[views] = carry;
  }
}

// pass 3: world points (dense, optional) and the compacted cloud
__global__ void export_scatter_kernel(const float* __restrict__ depth, const float* __restrict__ intr,
                                      const float* __restrict__ pose, const unsigned char* __restrict__ mask,
                                      const float* __restrict__ rgb, int W, int npix_view, int tiles_per_view,
                                      const long long* __restrict__ tile_offsets, long long capacity,
                                      float* __restrict__ world_dense, float* __restrict__ out_xyz,
                                      unsigned char* __restrict__ out_rgb) {
  const int view = blockIdx.x / tiles_per_view, tile = blockIdx.x - view * tiles_per_view;
  const int p0 = tile * EXP_TILE + threadIdx.x * EXP_PPT;
  const float* K = intr + view * 9;
  const float fx = K[0], fy = K[4], cx = K[2], cy = K[5];
  float P[12];
  if (pose) {
#pragma unroll
    for (int k = 0; k < 12; ++k) P[k] = pose[view * 16 + k];
  }
  bool keep[EXP_PPT];
  float wx[EXP_PPT], wy[EXP_PPT], wz[EXP_PPT];
  int n = 0;
#pragma unroll
  for (int k = 0; k < EXP_PPT; ++k) {
    keep[k] = false;
    const int pidx = p0 + k;
    if (pidx >= npix_view) continue;
    const long long i = (long long)view * npix_view + pidx;
    const float z = depth[i];
    const int y = pidx / W, x = pidx - y * W;
    // geometry.py:60-62: xx = (x_grid - cx) * depth_z / fx (no FMA contraction, as in ATen)
    const float xx = __fdiv_rn(__fmul_rn(__fsub_rn((float)x, cx), z), fx);
    const float yy = __fdiv_rn(__fmul_rn(__fsub_rn((float)y, cy), z), fy);
    if (pose) {  // geometry.py:98-103: pose @ [x, y, z, 1]
      wx[k] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(P[0], xx), __fmul_rn(P[1], yy)), __fmul_rn(P[2], z)), P[3]);
      wy[k] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(P[4], xx), __fmul_rn(P[5], yy)), __fmul_rn(P[6], z)), P[7]);
      wz[k] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(P[8], xx), __fmul_rn(P[9], yy)), __fmul_rn(P[10], z)), P[11]);
    } else {
      wx[k] = xx; wy[k] = yy; wz[k] = z;
    }
    if (world_dense) {
      world_dense[i * 3 + 0] = wx[k]; world_dense[i * 3 + 1] = wy[k]; world_dense[i * 3 + 2] = wz[k];
    }
    keep[k] = export_valid(depth, mask, i);
    n += keep[k] ? 1 : 0;
  }
  if (!out_xyz && !out_rgb) return;
  // exclusive offset of this thread inside the tile: warp scan + warp totals in shared memory
  __shared__ int wsum[EXP_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += u;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  int wbase = 0;
  for (int w = 0; w < warp; ++w) wbase += wsum[w];
  long long dst = tile_offsets[blockIdx.x] + wbase + incl - n;
#pragma unroll
  for (int k = 0; k < EXP_PPT; ++k) {
    if (!keep[k]) continue;
    if (dst < capacity) {
      if (out_xyz) { out_xyz[dst * 3 + 0] = wx[k]; out_xyz[dst * 3 + 1] = wy[k]; out_xyz[dst * 3 + 2] = wz[k]; }
      if (out_rgb && rgb) {
        const long long i = (long long)view * npix_view + p0 + k;
        // viz.py:239: (colors * 255).astype(np.uint8) -- truncation of the fp32 product
        out_rgb[dst * 3 + 0] = (unsigned char)(int)__fmul_rn(rgb[i * 3 + 0], 255.0f);
        out_rgb[dst * 3 + 1] = (unsigned char)(int)__fmul_rn(rgb[i * 3 + 1], 255.0f);
        out_rgb[dst * 3 + 2] = (unsigned char)(int)__fmul_rn(rgb[i * 3 + 2], 255.0f);
      }
    }
    ++dst;
  }
}

}  // namespace mab
