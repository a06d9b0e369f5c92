// Per-node irreps linear (block-diagonal in l): y[n, o, m] = scale * sum_k x[n, k, m] W[k, o].
//
// Reference: sevenn/nn/linear.py:46-49 (IrrepsLinear -> e3nn o3.Linear), sevenn/nn/self_connection.py:60-62.
// Frozen-graph arithmetic: per output block, x_l[N, mul_in, 2l+1] contracted over mul_in with
// W_l[mul_in, mul_out] / sqrt(fan_in); flat weight = concatenation of W_l in instruction order.
//
// CUDA-core SGEMM (fp32): rows of the GEMM are (node, m) pairs; 64x64 tiles, 4x4 register micro-tiles.
#include <stdint.h>
#include <stdlib.h>

#include "ops.h"

namespace b200 {

constexpr int LT_M = 64, LT_N = 64, LT_K = 32, LT_T = 256;

template <int D>
__device__ __forceinline__ void linear_block(const LinBlock& b, int in_dim, int out_dim, int n_nodes,
                                             const float* __restrict__ x, const float* __restrict__ W,
                                             float* __restrict__ y, int transpose, int accumulate, float* As,
                                             float* Bs) {
  constexpr int TN = LT_M / D;          // nodes per tile
  const int K = transpose ? b.mul_out : b.mul_in;
  const int Nn = transpose ? b.mul_in : b.mul_out;
  const int xoff = transpose ? b.out_off : b.in_off, yoff = transpose ? b.in_off : b.out_off;
  const int ldx = transpose ? out_dim : in_dim, ldy = transpose ? in_dim : out_dim;
  const int n0 = blockIdx.x * TN, o0 = blockIdx.y * LT_N;
  if (o0 >= Nn) return;
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += LT_K) {
    // A tile: As[k][r], r = nn*D + m
    for (int it = threadIdx.x; it < LT_M * LT_K; it += LT_T) {
      const int nn = it / (LT_K * D), c = it % (LT_K * D);
      const int k = c / D, m = c % D;
      if (nn >= TN) continue;
      float v = 0.f;
      if (n0 + nn < n_nodes && k0 + k < K) v = __ldg(x + (size_t)(n0 + nn) * ldx + xoff + (k0 + k) * D + m);
      As[k * LT_M + nn * D + m] = v;
    }
    if (TN * D < LT_M)
      for (int it = threadIdx.x; it < LT_K * (LT_M - TN * D); it += LT_T) {
        const int k = it / (LT_M - TN * D), r = TN * D + it % (LT_M - TN * D);
        As[k * LT_M + r] = 0.f;
      }
    // B tile: Bs[k][o] = W[k0+k][o0+o] (or its transpose)
    for (int it = threadIdx.x; it < LT_K * LT_N; it += LT_T) {
      int k, o;
      if (!transpose) { k = it / LT_N; o = it % LT_N; } else { o = it / LT_K; k = it % LT_K; }
      float v = 0.f;
      if (k0 + k < K && o0 + o < Nn)
        v = transpose ? __ldg(W + b.w_off + (size_t)(o0 + o) * b.mul_out + k0 + k)
                      : __ldg(W + b.w_off + (size_t)(k0 + k) * b.mul_out + o0 + o);
      Bs[k * (LT_N + 4) + o] = v;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < LT_K; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(As + k * LT_M + 4 * ty);
      const float4 bb = *reinterpret_cast<const float4*>(Bs + k * (LT_N + 4) + 4 * tx);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = 4 * ty + i, nn = r / D, m = r % D;
    if (nn >= TN || n0 + nn >= n_nodes) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int o = o0 + 4 * tx + j;
      if (o >= Nn) continue;
      float* p = y + (size_t)(n0 + nn) * ldy + yoff + o * D + m;
      const float v = acc[i][j] * b.scale;
      *p = accumulate ? (*p + v) : v;
    }
  }
}

__global__ void __launch_bounds__(LT_T)
irreps_linear_kernel(LinDesc d, int n_nodes, const float* __restrict__ x, const float* __restrict__ W,
                     float* __restrict__ y, int transpose, int accumulate) {
  __shared__ __align__(16) float As[LT_K * LT_M];
  __shared__ __align__(16) float Bs[LT_K * (LT_N + 4)];
  const LinBlock& b = d.blk[blockIdx.z];
  // grid.x is sized for D = 5 (fewest nodes per tile); surplus CTAs of wider tiles exit
  if (b.d == 1) {
    if (blockIdx.x * (LT_M / 1) < n_nodes) linear_block<1>(b, d.in_dim, d.out_dim, n_nodes, x, W, y, transpose, accumulate, As, Bs);
  } else if (b.d == 3) {
    if (blockIdx.x * (LT_M / 3) < n_nodes) linear_block<3>(b, d.in_dim, d.out_dim, n_nodes, x, W, y, transpose, accumulate, As, Bs);
  } else {
    if (blockIdx.x * (LT_M / 5) < n_nodes) linear_block<5>(b, d.in_dim, d.out_dim, n_nodes, x, W, y, transpose, accumulate, As, Bs);
  }
}

// dW[k][o] (+)= scale * sum_{n,m} x[n,k,m] dy[n,o,m]: one CTA per (32x32 tile of W, block), loop over nodes
__global__ void __launch_bounds__(256)
irreps_linear_wgrad_kernel(LinDesc d, int n_nodes, const float* __restrict__ x, const float* __restrict__ dy,
                           float* __restrict__ dW, int accumulate) {
  const LinBlock& b = d.blk[blockIdx.z];
  const int k0 = blockIdx.x * 32, o0 = blockIdx.y * 32;
  if (k0 >= b.mul_in || o0 >= b.mul_out) return;
  __shared__ float xs[32][33], ys[32][33];   // [row r in chunk][k or o]
  const int tk = threadIdx.x / 16, to = threadIdx.x % 16;  // each thread: 2x2 outputs
  double acc[2][2] = {};
  const int D = b.d;
  const long long R = (long long)n_nodes * D;
  for (long long r0 = 0; r0 < R; r0 += 32) {
    for (int it = threadIdx.x; it < 32 * 32; it += 256) {
      const int rr = it / 32, c = it % 32;
      const long long r = r0 + rr;
      float xv = 0.f, yv = 0.f;
      if (r < R) {
        const long long n = r / D;
        const int m = (int)(r % D);
        if (k0 + c < b.mul_in) xv = __ldg(x + n * d.in_dim + b.in_off + (k0 + c) * D + m);
        if (o0 + c < b.mul_out) yv = __ldg(dy + n * d.out_dim + b.out_off + (o0 + c) * D + m);
      }
      xs[rr][c] = xv;
      ys[rr][c] = yv;
    }
    __syncthreads();
    float part[2][2] = {};
#pragma unroll 8
    for (int rr = 0; rr < 32; ++rr) {
      const float a0 = xs[rr][tk], a1 = xs[rr][tk + 16], b0 = ys[rr][to], b1 = ys[rr][to + 16];
      part[0][0] = fmaf(a0, b0, part[0][0]); part[0][1] = fmaf(a0, b1, part[0][1]);
      part[1][0] = fmaf(a1, b0, part[1][0]); part[1][1] = fmaf(a1, b1, part[1][1]);
    }
    acc[0][0] += part[0][0]; acc[0][1] += part[0][1]; acc[1][0] += part[1][0]; acc[1][1] += part[1][1];
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int k = k0 + tk + 16 * i, o = o0 + to + 16 * j;
      if (k < b.mul_in && o < b.mul_out) {
        float* p = dW + b.w_off + (size_t)k * b.mul_out + o;
        const float v = (float)(acc[i][j] * (double)b.scale);
        *p = accumulate ? (*p + v) : v;
      }
    }
}

// Split-K weight gradient for many rows (the radial MLP's weight gradient runs over every EDGE of a training batch,
// sevenn/nn/convolution.py:97-106 under loss.backward(), trainer.py:185): dW = X^T G with the reduction over rows.
// 64x64 tile of W per CTA on the tensor cores: warp-level m16n8k8 tf32 MMAs with every operand split hi + lo
// (3xTF32: hi*hi + hi*lo + lo*hi, fp32-grade products), 8 warps x (16 x 32) outputs; the row range is cut into
// `splits` slices (grid.z) whose partial tiles are reduced in fp64 in a fixed order (wgrad_reduce_kernel).
__device__ __forceinline__ void split_tf32(float v, uint32_t& hi, uint32_t& lo) {
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(v));
  const float r = v - __uint_as_float(hi);
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(r));
}
__device__ __forceinline__ void mma_tf32_16x8x8(float* c, const uint32_t* a, const uint32_t* b) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int D>
__global__ void __launch_bounds__(256)
wgrad_splitk_kernel(LinBlock b, int in_dim, int out_dim, int n_nodes, int nodes_per_split, const float* __restrict__ x,
                    const float* __restrict__ dy, float* __restrict__ part) {
  constexpr int KR = 64;        // GEMM rows (node, m) staged per chunk
  constexpr int NCH = KR / D;   // nodes per chunk
  constexpr int LD = 72;        // row stride == 8 (mod 32): the fragment loads below hit 32 distinct banks
  const int k0 = blockIdx.x * 64, o0 = blockIdx.y * 64;
  const int nb = blockIdx.z * nodes_per_split, ne = min(n_nodes, nb + nodes_per_split);
  __shared__ __align__(16) float xs[KR * LD], ys[KR * LD];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int m0 = (warp & 3) * 16, n0 = (warp >> 2) * 32;
  float acc[4][4] = {};
  const int wk = min(64, b.mul_in - k0) * D, wo = min(64, b.mul_out - o0) * D;   // valid contiguous floats per node
  for (int n0n = nb; n0n < ne; n0n += NCH) {
    const int nn = min(NCH, ne - n0n);
    for (int j = threadIdx.x; j < NCH * 64 * D; j += 256) {
      const int node = j / (64 * D), rem = j % (64 * D), c = rem / D, m = rem % D;
      const size_t row = (size_t)(n0n + node);
      const bool in = node < nn;
      xs[(node * D + m) * LD + c] = (in && rem < wk) ? __ldg(x + row * in_dim + b.in_off + k0 * D + rem) : 0.f;
      ys[(node * D + m) * LD + c] = (in && rem < wo) ? __ldg(dy + row * out_dim + b.out_off + o0 * D + rem) : 0.f;
    }
    if constexpr (NCH * D < KR)   // rows the node loop does not cover (D = 3: 63, D = 5: 60 of 64)
      for (int j = threadIdx.x; j < (KR - NCH * D) * 64; j += 256) {
        xs[(NCH * D + j / 64) * LD + j % 64] = 0.f;
        ys[(NCH * D + j / 64) * LD + j % 64] = 0.f;
      }
    __syncthreads();
#pragma unroll 2
    for (int k8 = 0; k8 < KR; k8 += 8) {
      uint32_t ah[4], al[4];
      split_tf32(xs[(k8 + t) * LD + m0 + g], ah[0], al[0]);
      split_tf32(xs[(k8 + t) * LD + m0 + g + 8], ah[1], al[1]);
      split_tf32(xs[(k8 + t + 4) * LD + m0 + g], ah[2], al[2]);
      split_tf32(xs[(k8 + t + 4) * LD + m0 + g + 8], ah[3], al[3]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t bh[2], bl[2];
        split_tf32(ys[(k8 + t) * LD + n0 + j * 8 + g], bh[0], bl[0]);
        split_tf32(ys[(k8 + t + 4) * LD + n0 + j * 8 + g], bh[1], bl[1]);
        mma_tf32_16x8x8(acc[j], al, bh);
        mma_tf32_16x8x8(acc[j], ah, bl);
        mma_tf32_16x8x8(acc[j], ah, bh);
      }
    }
    __syncthreads();
  }
  float* pt = part + (size_t)blockIdx.z * b.mul_in * b.mul_out;
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = k0 + m0 + g + (q >> 1) * 8, o = o0 + n0 + j * 8 + 2 * t + (q & 1);
      if (k < b.mul_in && o < b.mul_out) pt[(size_t)k * b.mul_out + o] = acc[j][q];
    }
}

__global__ void wgrad_reduce_kernel(int numel, int splits, float scale, const float* __restrict__ part, float* __restrict__ dW,
                                    int accumulate) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= numel) return;
  double s = 0.0;
  for (int k = 0; k < splits; ++k) s += (double)part[(size_t)k * numel + i];
  const float v = (float)(s * (double)scale);
  dW[i] = accumulate ? dW[i] + v : v;
}

static bool lin_ok(const LinDesc& d) {
  if (d.nblk < 1 || d.nblk > 4) return false;
  for (int i = 0; i < d.nblk; ++i)
    if (!(d.blk[i].d == 1 || d.blk[i].d == 3 || d.blk[i].d == 5) || d.blk[i].mul_in < 1 || d.blk[i].mul_out < 1) return false;
  return true;
}

int irreps_linear(const LinDesc& d, int n_nodes, const float* x, const float* W, float* y, int transpose, int accumulate,
                  cudaStream_t st, float** pack_cache) {
  B200_REQUIRE(lin_ok(d), "irreps_linear: 1..4 blocks with 2l+1 in {1,3,5}");
  if (n_nodes == 0) return B200_OK;
  // dense channel mixing belongs on the tensor cores (linear_tc.cu); tiny blocks and tiny inputs stay on the FFMA kernel
  static const bool force_ffma = getenv("B200_LINEAR_FFMA") && atoi(getenv("B200_LINEAR_FFMA")) != 0;
  if (!force_ffma && n_nodes >= 64 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && irreps_linear_tc_supported(d, transpose))
    return irreps_linear_tc(d, n_nodes, x, W, y, transpose, accumulate, st, 0, pack_cache);
  int maxn = 0;
  for (int i = 0; i < d.nblk; ++i) maxn = max(maxn, transpose ? d.blk[i].mul_in : d.blk[i].mul_out);
  dim3 grid((n_nodes + (LT_M / 5) - 1) / (LT_M / 5), (maxn + LT_N - 1) / LT_N, d.nblk);
  irreps_linear_kernel<<<grid, LT_T, 0, st>>>(d, n_nodes, x, W, y, transpose, accumulate);
  return check_launch("irreps_linear");
}

int irreps_linear_wgrad(const LinDesc& d, int n_nodes, const float* x, const float* dy, float* dW, int accumulate,
                        cudaStream_t st) {
  B200_REQUIRE(lin_ok(d), "irreps_linear_wgrad: 1..4 blocks with 2l+1 in {1,3,5}");
  if (n_nodes >= 2048) {   // many rows: split the reduction over the rows across CTAs
    keep_async_pool();
    for (int i = 0; i < d.nblk; ++i) {
      const LinBlock& b = d.blk[i];
      const int tiles = ((b.mul_in + 63) / 64) * ((b.mul_out + 63) / 64);
      int splits = max(1, min((4 * 148 + tiles - 1) / tiles, n_nodes / 32));   // >= 32 nodes per slice
      const int per = (n_nodes + splits - 1) / splits;
      splits = (n_nodes + per - 1) / per;
      const int numel = b.mul_in * b.mul_out;
      float* part = nullptr;
      if (cudaMallocAsync((void**)&part, sizeof(float) * (size_t)splits * numel, st) != cudaSuccess) {
        set_error("irreps_linear_wgrad: cudaMallocAsync of the split-K partials failed");
        return B200_ERR_CUDA;
      }
      dim3 grid((b.mul_in + 63) / 64, (b.mul_out + 63) / 64, splits);
      if (b.d == 1) wgrad_splitk_kernel<1><<<grid, 256, 0, st>>>(b, d.in_dim, d.out_dim, n_nodes, per, x, dy, part);
      else if (b.d == 3) wgrad_splitk_kernel<3><<<grid, 256, 0, st>>>(b, d.in_dim, d.out_dim, n_nodes, per, x, dy, part);
      else wgrad_splitk_kernel<5><<<grid, 256, 0, st>>>(b, d.in_dim, d.out_dim, n_nodes, per, x, dy, part);
      wgrad_reduce_kernel<<<(numel + 255) / 256, 256, 0, st>>>(numel, splits, b.scale, part, dW + b.w_off, accumulate);
      cudaFreeAsync(part, st);
    }
    return check_launch("irreps_linear_wgrad", 2 * d.nblk);
  }
  int mk = 0, mo = 0;
  for (int i = 0; i < d.nblk; ++i) { mk = max(mk, d.blk[i].mul_in); mo = max(mo, d.blk[i].mul_out); }
  dim3 grid((mk + 31) / 32, (mo + 31) / 32, d.nblk);
  irreps_linear_wgrad_kernel<<<grid, 256, 0, st>>>(d, n_nodes, x, dy, dW, accumulate);
  return check_launch("irreps_linear_wgrad");
}

}  // namespace b200
