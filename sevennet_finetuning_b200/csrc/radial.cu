// Radial MLP of the convolution: per edge, emb[nb] -> H1 -> H2 -> weight_numel, SiLU with the e3nn
// second-moment constant, no bias, weights scaled by 1/sqrt(fan_in) at run time.
//
// Reference: sevenn/nn/convolution.py:97-106 (weight_nn = e3nn FullyConnectedNet([nb, H1, H2, wn], act));
// frozen-graph arithmetic: h = c*silu(x W / sqrt(fan_in)) for hidden layers, last layer linear.
//
// CUDA-core SGEMM version (fp32 FFMA, 64-edge tiles, 4x4 register micro-tiles).  The hidden
// activations never leave shared memory; only the [E, wn] weights are written.
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

constexpr int RT_M = 64;    // edges per CTA tile
constexpr int RT_H = 64;    // hidden width (padded)
constexpr int RT_T = 256;   // threads
constexpr int LDT = 68;     // padded leading dimension for transposed tiles

// acc[i][j] += sum_k A[k*lda + m0+i] * B[k*ldb + n0+j]
__device__ __forceinline__ void mm_tile_kmajor(const float* A, int lda, const float* B, int ldb, int K, int m0, int n0,
                                               float acc[4][4]) {
#pragma unroll 8
  for (int k = 0; k < K; ++k) {
    const float4 a = *reinterpret_cast<const float4*>(A + k * lda + m0);
    const float4 b = *reinterpret_cast<const float4*>(B + k * ldb + n0);
    const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
  }
}

// stage W[k0:k0+K, n0:n0+64] * scale (row-major [Kdim, ldw]) into Bs[k][n], zero padded
__device__ __forceinline__ void load_B(float* Bs, const float* __restrict__ W, int ldw, int Kdim, int Ndim, int k0, int K,
                                       int n0, float scale) {
  for (int it = threadIdx.x; it < K * RT_H; it += RT_T) {
    const int k = it / RT_H, n = it % RT_H;
    float v = 0.f;
    if (k0 + k < Kdim && n0 + n < Ndim) v = __ldg(W + (size_t)(k0 + k) * ldw + n0 + n) * scale;
    Bs[k * RT_H + n] = v;
  }
}
// stage W^T: Bs[k][n] = W[n0+n, k0+k] * scale  (W row-major [Ndim_rows, ldw]); padded leading dimension LDT
__device__ __forceinline__ void load_BT(float* Bs, const float* __restrict__ W, int ldw, int rows, int cols, int k0,
                                        int n0, float scale) {
  for (int it = threadIdx.x; it < RT_H * RT_H; it += RT_T) {
    const int n = it / RT_H, k = it % RT_H;  // lanes run along k: coalesced global reads
    float v = 0.f;
    if (n0 + n < rows && k0 + k < cols) v = __ldg(W + (size_t)(n0 + n) * ldw + k0 + k) * scale;
    Bs[k * LDT + n] = v;
  }
}

struct MlpSmem {
  float A0[RT_H * RT_M];   // [k][m] activations / gradients (k-major)
  float A1[RT_H * RT_M];
  float B[RT_H * LDT];     // weight tile
};
struct MlpSmemBwd : MlpSmem {
  float D1[RT_H * RT_M];   // c*silu'(z1) as [h][m]
  float D2[RT_H * RT_M];   // c*silu'(z2)
  float AT[RT_M * LDT];    // dw tile as [m][k]
};

// hidden layers for one tile: s.A0 <- emb^T, returns h2 in s.A0 ([k][m]); optionally derivatives in D1/D2
template <bool WITH_DERIV, class SM>
__device__ __forceinline__ void mlp_hidden(SM& s, const MlpDesc& d, long long e0, int ne,
                                           const float* __restrict__ emb, const float* __restrict__ W0,
                                           const float* __restrict__ W1) {
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  // emb tile, transposed to [k][m]
  for (int it = threadIdx.x; it < RT_M * d.nb; it += RT_T) {
    const int m = it / d.nb, k = it % d.nb;
    s.A0[k * RT_M + m] = (m < ne) ? __ldg(emb + (e0 + m) * d.nb + k) : 0.f;
  }
  load_B(s.B, W0, d.h1, d.nb, d.h1, 0, d.nb, 0, (1.0f / sqrtf((float)d.nb)));
  __syncthreads();
  float acc[4][4] = {};
  mm_tile_kmajor(s.A0, RT_M, s.B, RT_H, d.nb, 4 * ty, 4 * tx, acc);
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v, dv;
      silu_vd(acc[i][j], v, dv);
      const bool in = (4 * tx + j) < d.h1;
      s.A1[(4 * tx + j) * RT_M + 4 * ty + i] = in ? v : 0.f;
      if constexpr (WITH_DERIV) s.D1[(4 * tx + j) * RT_M + 4 * ty + i] = in ? dv : 0.f;
    }
  load_B(s.B, W1, d.h2, d.h1, d.h2, 0, d.h1, 0, (1.0f / sqrtf((float)d.h1)));
  __syncthreads();
  float acc2[4][4] = {};
  mm_tile_kmajor(s.A1, RT_M, s.B, RT_H, d.h1, 4 * ty, 4 * tx, acc2);
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v, dv;
      silu_vd(acc2[i][j], v, dv);
      const bool in = (4 * tx + j) < d.h2;
      s.A0[(4 * tx + j) * RT_M + 4 * ty + i] = in ? v : 0.f;   // h2 now lives in A0
      if constexpr (WITH_DERIV) s.D2[(4 * tx + j) * RT_M + 4 * ty + i] = in ? dv : 0.f;
    }
  __syncthreads();
}

__global__ void __launch_bounds__(RT_T)
radial_mlp_fwd_kernel(MlpDesc d, long long E, const float* __restrict__ emb, const float* __restrict__ W0,
                      const float* __restrict__ W1, const float* __restrict__ W2, float* __restrict__ w) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  MlpSmem& s = *reinterpret_cast<MlpSmem*>(smem_raw);
  const long long e0 = (long long)blockIdx.x * RT_M;
  const int ne = (int)min((long long)RT_M, E - e0);
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  mlp_hidden<false>(s, d, e0, ne, emb, W0, W1);
  const float sc = (1.0f / sqrtf((float)d.h2));
  for (int n0 = 0; n0 < d.wn; n0 += RT_H) {
    load_B(s.B, W2, d.wn, d.h2, d.wn, 0, d.h2, n0, sc);
    __syncthreads();
    float acc[4][4] = {};
    mm_tile_kmajor(s.A0, RT_M, s.B, RT_H, d.h2, 4 * ty, 4 * tx, acc);
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = 4 * ty + i;
      if (m >= ne) continue;
      float* row = w + (size_t)(e0 + m) * d.wn + n0 + 4 * tx;
      if (n0 + 4 * tx + 3 < d.wn && (d.wn % 4) == 0) {
        *reinterpret_cast<float4*>(row) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (n0 + 4 * tx + j < d.wn) row[j] = acc[i][j];
      }
    }
  }
}

// demb (+)= d/d emb of sum_e <dw[e], w[e]>   (data gradient only; parameter gradients: radial_mlp_wgrad)
__global__ void __launch_bounds__(RT_T)
radial_mlp_bwd_kernel(MlpDesc d, long long E, const float* __restrict__ emb, const float* __restrict__ W0,
                      const float* __restrict__ W1, const float* __restrict__ W2, const float* __restrict__ dw,
                      float* __restrict__ demb, int accumulate) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  MlpSmemBwd& s = *reinterpret_cast<MlpSmemBwd*>(smem_raw);
  const long long e0 = (long long)blockIdx.x * RT_M;
  const int ne = (int)min((long long)RT_M, E - e0);
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  mlp_hidden<true>(s, d, e0, ne, emb, W0, W1);   // fills D1, D2; A0/A1 are free afterwards
  // dh2[m][h] = sum_k dw[m][k] * W2[h][k] / sqrt(H2)
  float acc[4][4] = {};
  const float sc2 = (1.0f / sqrtf((float)d.h2));
  for (int k0 = 0; k0 < d.wn; k0 += RT_H) {
    for (int it = threadIdx.x; it < RT_M * RT_H; it += RT_T) {
      const int m = it / RT_H, k = it % RT_H;
      float v = 0.f;
      if (m < ne && k0 + k < d.wn) v = __ldg(dw + (size_t)(e0 + m) * d.wn + k0 + k);
      s.AT[m * LDT + k] = v;
    }
    load_BT(s.B, W2, d.wn, d.h2, d.wn, k0, 0, sc2);
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < RT_H; ++k) {
      const float4 b = *reinterpret_cast<const float4*>(s.B + k * LDT + 4 * tx);
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float a = s.AT[(4 * ty + i) * LDT + k];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a, bv[j], acc[i][j]);
      }
    }
    __syncthreads();
  }
  // dz2 = dh2 * act'(z2) -> A0 as [h][m]
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int idx = (4 * tx + j) * RT_M + 4 * ty + i;
      s.A0[idx] = acc[i][j] * s.D2[idx];
    }
  // dh1[m][g] = sum_h dz2[m][h] * W1[g][h] / sqrt(H1)
  load_BT(s.B, W1, d.h2, d.h1, d.h2, 0, 0, (1.0f / sqrtf((float)d.h1)));
  __syncthreads();
  float acc1[4][4] = {};
  mm_tile_kmajor(s.A0, RT_M, s.B, LDT, RT_H, 4 * ty, 4 * tx, acc1);
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int idx = (4 * tx + j) * RT_M + 4 * ty + i;
      s.A1[idx] = acc1[i][j] * s.D1[idx];   // dz1 as [g][m]
    }
  __syncthreads();
  // demb[m][n] = sum_g dz1[m][g] * W0[n][g] / sqrt(nb)
  const float sc0 = (1.0f / sqrtf((float)d.nb));
  for (int it = threadIdx.x; it < RT_M * d.nb; it += RT_T) {
    const int n = it / RT_M, m = it % RT_M;
    if (m >= ne) continue;
    float t = 0.f;
    for (int g = 0; g < d.h1; ++g) t = fmaf(s.A1[g * RT_M + m], __ldg(W0 + n * d.h1 + g), t);
    t *= sc0;
    float* p = demb + (size_t)(e0 + m) * d.nb + n;
    *p = accumulate ? (*p + t) : t;
  }
}

// tensor-core (tcgen05, 3xTF32) versions live in radial_tc.cu; the FFMA kernels above serve the shapes they do
// not cover (hidden widths != 64).  B200_RADIAL_FFMA=1 forces the FFMA kernels (A/B measurements, tests).
bool radial_tc_supported(const MlpDesc& d);
int radial_mlp_fwd_tc(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                      float* w, cudaStream_t st);
int radial_mlp_bwd_tc(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                      const float* dw, float* demb, int accumulate, cudaStream_t st);
static bool use_tc(const MlpDesc& d) {
  const char* e = getenv("B200_RADIAL_FFMA");
  return radial_tc_supported(d) && !(e && e[0] == '1');
}

static bool mlp_ok(const MlpDesc& d) {
  return d.nb >= 1 && d.nb <= RT_H && d.h1 >= 1 && d.h1 <= RT_H && d.h2 >= 1 && d.h2 <= RT_H && d.wn >= 1;
}

int radial_mlp_fwd(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                   float* w, cudaStream_t st) {
  B200_REQUIRE(mlp_ok(d), "radial_mlp_fwd: supported shape is [nb<=64, H1<=64, H2<=64, wn]");
  if (E == 0) return B200_OK;
  if (use_tc(d)) return radial_mlp_fwd_tc(d, E, emb, W0, W1, W2, w, st);
  cudaFuncSetAttribute(radial_mlp_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(MlpSmem));
  radial_mlp_fwd_kernel<<<(unsigned)((E + RT_M - 1) / RT_M), RT_T, sizeof(MlpSmem), st>>>(d, E, emb, W0, W1, W2, w);
  return check_launch("radial_mlp_fwd");
}

int radial_mlp_bwd(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                   const float* dw, float* demb, int accumulate, cudaStream_t st) {
  B200_REQUIRE(mlp_ok(d), "radial_mlp_bwd: supported shape is [nb<=64, H1<=64, H2<=64, wn]");
  if (E == 0) return B200_OK;
  if (use_tc(d)) return radial_mlp_bwd_tc(d, E, emb, W0, W1, W2, dw, demb, accumulate, st);
  cudaFuncSetAttribute(radial_mlp_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(MlpSmemBwd));
  radial_mlp_bwd_kernel<<<(unsigned)((E + RT_M - 1) / RT_M), RT_T, sizeof(MlpSmemBwd), st>>>(d, E, emb, W0, W1, W2, dw,
                                                                                              demb, accumulate);
  return check_launch("radial_mlp_bwd");
}

}  // namespace b200
