// Micro-benchmark: scalar FFMA vs packed FFMA2 (fma.rn.f32x2) on sm_100a, alone and mixed with integer work - decides whether
// the convolution kernels should pack their CG arithmetic (DESIGN.md finding 13).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 ffma2.cu && ./ffma2
#include <cuda_runtime.h>
#include <stdio.h>

constexpr int ITERS = 4096;

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float a, float b) {
  float acc[16];
  float2 acc2[16];
  int iacc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    acc[i] = threadIdx.x * 0.001f + i;
    acc2[i] = make_float2(acc[i], acc[i] + 1.f);
    iacc[i] = threadIdx.x + i;
  }
  const float2 a2 = make_float2(a, a * 1.0001f), b2 = make_float2(b, b * 1.0001f);
  for (int it = 0; it < ITERS; ++it) {
    if (MODE == 0 || MODE == 2) {   // 32 scalar FFMA
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], a, b);
    }
    if (MODE == 1 || MODE == 3) {   // 16 FFMA2 = the same 32 FMAs
#pragma unroll
      for (int i = 0; i < 16; ++i) acc2[i] = __ffma2_rn(acc2[i], a2, b2);
    }
    if (MODE >= 2) {                // + 16 x (xor, add) on the integer pipe
#pragma unroll
      for (int j = 0; j < 16; ++j) iacc[j] = (iacc[j] ^ it) + j;
    }
  }
  float s = 0.f;
  int t = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) { s += acc[i] + acc2[i].x + acc2[i].y; t += iacc[i]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + t;
}

template <int MODE> void run(const char* name, float* out) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int grid = 148 * 8;
  k<MODE><<<grid, 256>>>(out, 0.999f, 0.0f);
  cudaEventRecord(e0);
  k<MODE><<<grid, 256>>>(out, 0.999f, 0.0f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double fma = (double)grid * 256 * ITERS * 32;
  printf("%-28s %8.3f ms  %7.2f TFMA/s\n", name, ms, fma / ms * 1e-9);
}

int main() {
  float* out;
  cudaMalloc(&out, 148 * 8 * 256 * sizeof(float));
  run<0>("32 FFMA", out);
  run<1>("16 FFMA2", out);
  run<2>("32 FFMA + 16 int(2 op)", out);
  run<3>("16 FFMA2 + 16 int(2 op)", out);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
