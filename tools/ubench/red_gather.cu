// Micro-benchmarks that decide the dx-accumulation design of the fused backward kernel:
//  (1) RED.ADD.F32 scatter of 480-float rows (scalar and v4) at random rows of a [N,480] array
//  (2) gather of 480-float rows (LDG.128) at neighbour-like random rows
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_gather red_gather.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

__global__ void red_scalar(const int* __restrict__ rows, int nedge, float* __restrict__ dx, int width) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nw = (gridDim.x * blockDim.x) >> 5;
  for (int e = warp; e < nedge; e += nw) {
    float* r = dx + (size_t)rows[e] * width;
    for (int c = lane; c < width; c += 32) atomicAdd(r + c, 1.0f);
  }
}
__global__ void red_v4(const int* __restrict__ rows, int nedge, float* __restrict__ dx, int width) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nw = (gridDim.x * blockDim.x) >> 5;
  for (int e = warp; e < nedge; e += nw) {
    float* r = dx + (size_t)rows[e] * width;
    for (int c = lane * 4; c < width; c += 128)
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(r + c), "f"(1.f), "f"(1.f), "f"(1.f), "f"(1.f) : "memory");
  }
}
__global__ void gather_v4(const int* __restrict__ rows, int nedge, const float* __restrict__ x, int width, float* out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nw = (gridDim.x * blockDim.x) >> 5;
  float4 acc = make_float4(0, 0, 0, 0);
  for (int e = warp; e < nedge; e += nw) {
    const float* r = x + (size_t)rows[e] * width;
    for (int c = lane * 4; c < width; c += 128) {
      float4 v = __ldg(reinterpret_cast<const float4*>(r + c));
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  if (acc.x + acc.y + acc.z + acc.w == 123.456f) out[0] = acc.x;
}

int main(int argc, char** argv) {
  const int N = 100000, width = 480, deg = 48;
  const int E = N * deg;
  // locality model: atom i's neighbours are i + uniform(-W, W) (sorted atoms) or uniform random (unsorted)
  for (int mode = 0; mode < 2; ++mode) {
    std::vector<int> rows(E);
    srand(1);
    for (int i = 0; i < N; ++i)
      for (int k = 0; k < deg; ++k) {
        int j = mode == 0 ? rand() % N : ((i + rand() % 4000 - 2000) % N + N) % N;
        rows[(size_t)i * deg + k] = j;
      }
    int* d_rows; float *d_x, *d_out;
    cudaMalloc(&d_rows, sizeof(int) * E);
    cudaMalloc(&d_x, sizeof(float) * (size_t)N * width);
    cudaMalloc(&d_out, 16);
    cudaMemcpy(d_rows, rows.data(), sizeof(int) * E, cudaMemcpyHostToDevice);
    cudaMemset(d_x, 0, sizeof(float) * (size_t)N * width);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float ms;
    for (int k = 0; k < 3; ++k) {
      const char* nm[3] = {"red.f32 ", "red.v4  ", "gather  "};
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        if (k == 0) red_scalar<<<148 * 8, 256>>>(d_rows, E, d_x, width);
        if (k == 1) red_v4<<<148 * 8, 256>>>(d_rows, E, d_x, width);
        if (k == 2) gather_v4<<<148 * 8, 256>>>(d_rows, E, d_x, width, d_out);
        cudaEventRecord(b); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
      }
      printf("%s mode=%s  %.3f ms  %.1f GB/s (rows of %d B, %d edges) err=%d\n", nm[k], mode ? "local " : "random", ms,
             (double)E * width * 4 / ms / 1e6, width * 4, E, (int)cudaGetLastError());
    }
    cudaFree(d_rows); cudaFree(d_x); cudaFree(d_out);
  }
  return 0;
}
