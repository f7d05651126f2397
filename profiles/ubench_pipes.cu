// Issue-rate micro-benchmark of the instructions the attention softmax is made of (sm_100a).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build_variants/ubench_pipes profiles/ubench_pipes.cu
// One CTA per SM, W warps per CTA, 8 independent dependency chains per thread, clock64 around the loop:
// prints warp-instructions per clock per SM and elements per clock per SM.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define ITERS 4096

template <int OP>
__device__ __forceinline__ void step(uint32_t (&r)[8], uint64_t (&q)[4]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(r[i]));
    if (OP == 1) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(r[i]));
    if (OP == 2) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(r[i]));
    if (OP == 4) asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r[i]) : "f"(__uint_as_float(r[(i + 1) & 7])), "f"(__uint_as_float(r[(i + 2) & 7])));
    if (OP == 5) asm volatile("add.rn.f16x2 %0, %0, %0;" : "+r"(r[i]));
    if (OP == 6) asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(q[i & 3]));
    if (OP == 7) asm volatile("{.reg .f32 a; mov.b32 a, %0; max.f32 a, a, a, a; mov.b32 %0, a;}" : "+r"(r[i]));
    if (OP == 9) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r[i]) : "f"(__uint_as_float(r[(i + 1) & 7])), "f"(__uint_as_float(r[(i + 2) & 7])));
    if (OP == 10) asm volatile("{.reg .f32 a; mov.b32 a, %0; fma.rn.f32 a, a, a, a; mov.b32 %0, a;}" : "+r"(r[i]));
  }
}

template <int OP>
__global__ void bench(uint32_t* out, long long* cycles) {
  uint32_t r[8];
  uint64_t q[4];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = 0x3c003c00u + threadIdx.x + i;
#pragma unroll
  for (int i = 0; i < 4; ++i) q[i] = 0x3f0000003f000000ull + threadIdx.x + i;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) step<OP>(r, q);
  const long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc ^= r[i];
#pragma unroll
  for (int i = 0; i < 4; ++i) acc ^= (uint32_t)q[i] ^ (uint32_t)(q[i] >> 32);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int elems_per_instr, uint32_t* out, long long* cyc) {
  for (int warps : {4, 8, 16}) {
    bench<OP><<<148, warps * 32>>>(out, cyc);
    cudaDeviceSynchronize();
    bench<OP><<<148, warps * 32>>>(out, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double c = 0;
    for (int i = 0; i < 148; ++i) c += h[i];
    c /= 148;
    const double winstr = (double)warps * ITERS * 8;
    printf("%-24s warps/SM %2d: %8.0f clk  %.3f warp-instr/clk/SM  %.1f elem/clk/SM  (%s)\n", name, warps, c,
           winstr / c, winstr * 32 * elems_per_instr / c, cudaGetErrorString(e));
  }
}

int main() {
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 512 * 4);
  cudaMalloc(&cyc, 148 * 8);
  run<0>("ex2.approx.ftz.f32", 1, out, cyc);
  run<1>("ex2.approx.ftz.bf16x2", 2, out, cyc);
  run<2>("ex2.approx.f16x2", 2, out, cyc);
  run<4>("cvt.rn.f16x2.f32", 2, out, cyc);
  run<9>("cvt.rn.bf16x2.f32", 2, out, cyc);
  run<5>("add.rn.f16x2", 2, out, cyc);
  run<6>("fma.rn.f32x2", 2, out, cyc);
  run<10>("fma.rn.f32", 1, out, cyc);
  run<7>("max.f32 (3-input)", 1, out, cyc);
  return 0;
}
