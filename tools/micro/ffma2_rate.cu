// Microbenchmark: issue rate of FFMA (3-register) vs FFMA2 (packed f32x2) per SM sub-partition on sm_100a.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_rate ffma2_rate.cu ; run: ./ffma2_rate
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(float* out, long long* cyc, float a, float b, int iters) {
  float x[16];
  float2 y[8];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = threadIdx.x + i;
#pragma unroll
  for (int i = 0; i < 8; ++i) y[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
  float w = a + threadIdx.x, v = b - threadIdx.x;   // register operands (not immediates)
  const float2 w2 = make_float2(w, v), v2 = make_float2(v, w);
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] = fmaf(x[i], w, v);          // 64 FFMA, 16 independent chains
    } else {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 8; ++i) y[i] = __ffma2_rn(y[i], w2, v2);    // 32 FFMA2 = 64 FMAs, 8 independent chains
    }
  }
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += x[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) s += y[i].x + y[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
  float* out; long long* cyc;
  cudaMalloc(&out, 1 << 22); cudaMalloc(&cyc, 8);
  const int iters = 2000;
  for (int warps = 4; warps <= 32; warps *= 2) {
    for (int mode = 0; mode < 2; ++mode) {
      long long c = 0;
      for (int rep = 0; rep < 2; ++rep) {
        if (mode == 0) k<0><<<148, warps * 32>>>(out, cyc, 1.0001f, 0.5f, iters);
        else k<1><<<148, warps * 32>>>(out, cyc, 1.0001f, 0.5f, iters);
        cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
      }
      const double fma_per_clk_sm = (double)iters * 64 * 32 * warps / (double)c;
      printf("warps/SM %2d  %s  cycles %lld  -> %.1f fp32 FMA lanes/clk/SM (%.2f warp-instr/clk/SMSP)\n", warps, mode ? "FFMA2" : "FFMA ", c,
             fma_per_clk_sm, (double)iters * (mode ? 32 : 64) * warps / 4.0 / (double)c);
    }
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
