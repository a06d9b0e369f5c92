// Radial MLP on the tcgen05 tensor cores (3xTF32, fp32 accumulation in TMEM).
//
// Reference: sevenn/nn/convolution.py:97-106 (weight_nn = e3nn FullyConnectedNet([nb, 64, 64, wn])).
// The two hidden layers (8->64->64, 4.6 kMAC/edge) stay on the CUDA cores; the last layer
// (64 -> weight_numel, 61 kMAC/edge for the 960-wide layers) and its transpose in the backward pass
// are tcgen05.mma GEMMs.
//
// Forward, one CTA per tile of 128 edges:   W^T[wn x 64] . h2^T[64 x 128 edges] -> D[channel, edge]
//   A = 128-channel slab of W2^T (rows = weight channels, K = hidden unit), streamed from L2,
//   B = h2 tile (rows = edges, K = hidden unit), resident in shared memory,
//   D = 128 lanes (channels) x 128 columns (edges) in TMEM, double buffered.
// TMEM lane = weight channel, so the epilogue thread of a lane owns one channel for all edges of the
// tile and its global stores are coalesced across the warp (the same ownership the convolution uses).
#include "common.cuh"
#include "tc.cuh"

namespace b200 {

using namespace tc;

constexpr int TCN = 128;              // edges per tile
constexpr int TCH = 64;               // hidden width (K of the big GEMM)
constexpr int TCT = 256;              // threads
constexpr int CHB = TCN * 16;         // bytes of one 16-byte K-chunk column of a 128-row operand
constexpr int OPB = (TCH / 4) * CHB;  // one [128 x 64] fp32 operand: 32 KB

struct TcFwdSmem {
  uint8_t Bhi[OPB], Blo[OPB];   // h2 tile, hi / lo parts
  uint8_t A[2][2][OPB];         // [stage][hi/lo] slab of W2^T; stage 1 doubles as scratch for the hidden layers
  uint64_t mma_done[2];
  uint32_t tmem_base;
};

// scratch carved out of A[1] while the hidden layers run
constexpr int TC_MAX_NB = 16;
struct HiddenScratch {
  float embT[TC_MAX_NB][TCN];   // [basis function][edge]
  float h1T[TCH][TCN];          // [hidden unit][edge]
};

// 128 x 64 tile product on the CUDA cores: acc[i][j] += sum_k A[k][ty+16i] * B[k*ldb + 4tx+j]
__device__ __forceinline__ void hidden_mm(const float (*A)[TCN], const float* __restrict__ B, int ldb, int K, int ty, int tx,
                                          float acc[8][4]) {
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    const float4 b = *reinterpret_cast<const float4*>(B + k * ldb + 4 * tx);
    const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float a = A[k][ty + 16 * i];
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a, bv[j], acc[i][j]);
    }
  }
}

// stage the 128-channel slab mt of W2^T (scaled) into hi/lo operands: row r = channel mt*128+r, K = hidden unit
__device__ __forceinline__ void load_w2_slab(uint8_t* Ahi, uint8_t* Alo, const float* __restrict__ W2, int wn, int h2dim,
                                             int mt, float scale) {
  const int r = threadIdx.x & 127, half = threadIdx.x >> 7;
  const int ch = mt * TCN + r;
#pragma unroll 2
  for (int c4 = half * 8; c4 < half * 8 + 8; ++c4) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ch < wn) {
      const int k = 4 * c4;
      if (k + 0 < h2dim) v.x = __ldg(W2 + (size_t)(k + 0) * wn + ch) * scale;
      if (k + 1 < h2dim) v.y = __ldg(W2 + (size_t)(k + 1) * wn + ch) * scale;
      if (k + 2 < h2dim) v.z = __ldg(W2 + (size_t)(k + 2) * wn + ch) * scale;
      if (k + 3 < h2dim) v.w = __ldg(W2 + (size_t)(k + 3) * wn + ch) * scale;
    }
    float4 hi, lo;
    split4(v, hi, lo);
    *reinterpret_cast<float4*>(Ahi + c4 * CHB + r * 16) = hi;
    *reinterpret_cast<float4*>(Alo + c4 * CHB + r * 16) = lo;
  }
}

// hidden layers for the CTA's 128 edges; writes h2 (hi/lo) into the B operand and, optionally,
// c*silu'(z2) as D2[h][m].  emb/W0/W1 in the reference layouts ([E,nb], [nb,64], [64,64]).
template <bool WITH_D2, bool WITH_H2>
__device__ __forceinline__ void hidden_layers(HiddenScratch& hs, uint8_t* Bhi, uint8_t* Blo, float (*D2)[TCN], int nb,
                                              long long e0, int ne, const float* __restrict__ emb,
                                              const float* __restrict__ W0, const float* __restrict__ W1) {
  const int ty = threadIdx.x & 15, tx = threadIdx.x >> 4;
  for (int it = threadIdx.x; it < TCN * nb; it += TCT) {
    const int m = it / nb, k = it % nb;
    hs.embT[k][m] = (m < ne) ? __ldg(emb + (e0 + m) * nb + k) : 0.f;
  }
  __syncthreads();
  float acc[8][4] = {};
  hidden_mm(hs.embT, W0, TCH, nb, ty, tx, acc);
  const float s0 = 1.0f / sqrtf((float)nb), s1 = 1.0f / sqrtf((float)TCH);
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) hs.h1T[4 * tx + j][ty + 16 * i] = kSilu2Mom * silu_f(acc[i][j] * s0);
  __syncthreads();
  float acc2[8][4] = {};
  hidden_mm(hs.h1T, W1, TCH, TCH, ty, tx, acc2);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float4 h;
    float v[4], d[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) silu_vd(acc2[i][j] * s1, v[j], d[j]);
    h = make_float4(v[0], v[1], v[2], v[3]);
    const int m = ty + 16 * i;
    if constexpr (WITH_H2) {
      float4 hi, lo;
      split4(h, hi, lo);
      *reinterpret_cast<float4*>(Bhi + tx * CHB + m * 16) = hi;   // chunk tx holds hidden units 4tx..4tx+3
      *reinterpret_cast<float4*>(Blo + tx * CHB + m * 16) = lo;
    }
    if constexpr (WITH_D2) {
#pragma unroll
      for (int j = 0; j < 4; ++j) D2[4 * tx + j][m] = d[j];
    }
  }
}

__global__ void __launch_bounds__(TCT, 1)
radial_mlp_fwd_tc_kernel(int nb, int wn, long long E, const float* __restrict__ emb, const float* __restrict__ W0,
                         const float* __restrict__ W1, const float* __restrict__ W2, float* __restrict__ w) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  TcFwdSmem& s = *reinterpret_cast<TcFwdSmem*>(smem_raw);
  HiddenScratch& hs = *reinterpret_cast<HiddenScratch*>(&s.A[1][0][0]);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * TCN;
  const int ne = (int)min((long long)TCN, E - e0);
  const int NT = (wn + TCN - 1) / TCN;
  constexpr uint32_t idesc = make_idesc_tf32(128, TCN);

  if (warp == 0) tmem_alloc(&s.tmem_base, 256);
  if (threadIdx.x == 32) {
    mbar_init(&s.mma_done[0], 1);
    mbar_init(&s.mma_done[1], 1);
    fence_mbar_init();
  }
  load_w2_slab(s.A[0][0], s.A[0][1], W2, wn, TCH, 0, 0.125f);
  hidden_layers<false, true>(hs, s.Bhi, s.Blo, nullptr, nb, e0, ne, emb, W0, W1);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  for (int mt = 0; mt <= NT; ++mt) {
    const int st = mt & 1;
    if (mt < NT) {
      if (mt > 0) {
        // stage st was last read by MMA(mt-2), whose completion was observed in the epilogue of iteration mt-1
        load_w2_slab(s.A[st][0], s.A[st][1], W2, wn, TCH, mt, 0.125f);
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
      }
      if (threadIdx.x == 0) {
        const uint32_t d = tmem + (uint32_t)st * TCN;
        const uint32_t ahi = smem_u32(s.A[st][0]), alo = smem_u32(s.A[st][1]);
        const uint32_t bhi = smem_u32(s.Bhi), blo = smem_u32(s.Blo);
#pragma unroll
        for (int kk = 0; kk < TCH / 8; ++kk) {
          const uint32_t off = kk * 2 * CHB;
          const uint64_t dah = make_desc(ahi + off, CHB, 128), dal = make_desc(alo + off, CHB, 128);
          const uint64_t dbh = make_desc(bhi + off, CHB, 128), dbl = make_desc(blo + off, CHB, 128);
          mma_tf32(d, dal, dbh, idesc, kk > 0);
          mma_tf32(d, dah, dbl, idesc, 1);
          mma_tf32(d, dah, dbh, idesc, 1);
        }
        mma_commit(&s.mma_done[st]);
      }
    }
    if (mt > 0) {
      // epilogue of tile mt-1: lanes = channels, columns = edges
      const int pt = mt - 1, ps = pt & 1;
      mbar_wait(&s.mma_done[ps], (pt >> 1) & 1);
      tc_fence_after();
      const int q = warp & 3, hh = warp >> 2;
      const int ch = pt * TCN + 32 * q + lane;
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int col = hh * 64 + cc * 32;
        float v[32];
        tmem_ld32(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(ps * TCN + col), v);
        if (ch < wn) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col + j < ne) w[(size_t)(e0 + col + j) * wn + ch] = v[j];
        }
      }
      tc_fence_before();
    }
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// ---- backward (data gradient): demb (+)= d/d emb sum_e <dw[e], w[e]> -------------------------------------
// dh2[e, h] = sum_c dw[e, c] W2[h, c] / 8 is the tensor-core GEMM: A = dw tile (rows = edges, K = channel),
// B = W2 (rows = hidden unit, K = channel), D = 128 lanes (edges) x 64 columns in TMEM, K streamed in
// chunks of 32 channels through a two-stage shared-memory ring.
constexpr int BK = 32;                 // channels per K chunk
constexpr int BCH = TCH * 16;          // bytes of one 16-byte K-chunk column of the 64-row B operand
constexpr int SLD = BK + 4;            // padded row length of the staging tiles (conflict-free float4 columns)

struct TcBwdSmem {
  uint8_t A[2][2][(BK / 4) * CHB];     // [stage][hi/lo] dw chunk      64 KB
  uint8_t B[2][2][(BK / 4) * BCH];     // [stage][hi/lo] W2 chunk      32 KB
  float stage_dw[TCN][SLD];            // 18 KB
  float stage_w2[TCH][SLD];            //  9 KB
  HiddenScratch hs;                    // 40 KB
  float D2[TCH][TCN];                  // 32 KB: c*silu'(z2), later dz1
  uint64_t mma_done[2];
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(TCT, 1)
radial_mlp_bwd_tc_kernel(int nb, int wn, long long E, const float* __restrict__ emb, const float* __restrict__ W0,
                         const float* __restrict__ W1, const float* __restrict__ W2, const float* __restrict__ dw,
                         float* __restrict__ demb, int accumulate) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  TcBwdSmem& s = *reinterpret_cast<TcBwdSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * TCN;
  const int ne = (int)min((long long)TCN, E - e0);
  const int NK = (wn + BK - 1) / BK;
  constexpr uint32_t idesc = make_idesc_tf32(128, TCH);

  if (warp == 0) tmem_alloc(&s.tmem_base, 64);
  if (threadIdx.x == 32) {
    mbar_init(&s.mma_done[0], 1);
    mbar_init(&s.mma_done[1], 1);
    fence_mbar_init();
  }
  hidden_layers<true, false>(s.hs, nullptr, nullptr, s.D2, nb, e0, ne, emb, W0, W1);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  for (int i = 0; i < NK; ++i) {
    const int st = i & 1, k0 = i * BK;
    // global -> staging, coalesced along the channel axis
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int idx = threadIdx.x + TCT * q, row = idx >> 3, c4 = idx & 7;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row < ne && k0 + 4 * c4 < wn) v = __ldg(reinterpret_cast<const float4*>(dw + (size_t)(e0 + row) * wn + k0 + 4 * c4));
      *reinterpret_cast<float4*>(&s.stage_dw[row][4 * c4]) = v;
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int idx = threadIdx.x + TCT * q, row = idx >> 3, c4 = idx & 7;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (k0 + 4 * c4 < wn) v = __ldg(reinterpret_cast<const float4*>(W2 + (size_t)row * wn + k0 + 4 * c4));
      *reinterpret_cast<float4*>(&s.stage_w2[row][4 * c4]) = v;
    }
    // the MMAs that read stage st two iterations ago must have finished before it is overwritten
    if (i >= 2) mbar_wait(&s.mma_done[st], ((i - 2) >> 1) & 1);
    __syncthreads();
    // staging -> hi/lo operands in the canonical K-major layout (lanes run along rows)
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int idx = threadIdx.x + TCT * q, row = idx & 127, c4 = idx >> 7;
      const float4 v = *reinterpret_cast<const float4*>(&s.stage_dw[row][4 * c4]);
      float4 hi, lo;
      split4(v, hi, lo);
      *reinterpret_cast<float4*>(s.A[st][0] + c4 * CHB + row * 16) = hi;
      *reinterpret_cast<float4*>(s.A[st][1] + c4 * CHB + row * 16) = lo;
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int idx = threadIdx.x + TCT * q, row = idx & 63, c4 = idx >> 6;
      float4 v = *reinterpret_cast<const float4*>(&s.stage_w2[row][4 * c4]);
      v.x *= 0.125f; v.y *= 0.125f; v.z *= 0.125f; v.w *= 0.125f;
      float4 hi, lo;
      split4(v, hi, lo);
      *reinterpret_cast<float4*>(s.B[st][0] + c4 * BCH + row * 16) = hi;
      *reinterpret_cast<float4*>(s.B[st][1] + c4 * BCH + row * 16) = lo;
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (threadIdx.x == 0) {
      const uint32_t ahi = smem_u32(s.A[st][0]), alo = smem_u32(s.A[st][1]);
      const uint32_t bhi = smem_u32(s.B[st][0]), blo = smem_u32(s.B[st][1]);
#pragma unroll
      for (int kk = 0; kk < BK / 8; ++kk) {
        const uint64_t dah = make_desc(ahi + kk * 2 * CHB, CHB, 128), dal = make_desc(alo + kk * 2 * CHB, CHB, 128);
        const uint64_t dbh = make_desc(bhi + kk * 2 * BCH, BCH, 128), dbl = make_desc(blo + kk * 2 * BCH, BCH, 128);
        mma_tf32(tmem, dal, dbh, idesc, (i > 0 || kk > 0) ? 1u : 0u);
        mma_tf32(tmem, dah, dbl, idesc, 1);
        mma_tf32(tmem, dah, dbh, idesc, 1);
      }
      mma_commit(&s.mma_done[st]);
    }
  }
  // all MMAs retire in order: the last commit covers every chunk
  mbar_wait(&s.mma_done[(NK - 1) & 1], ((NK - 1) >> 1) & 1);
  tc_fence_after();
  {
    // dz2[h][m] = dh2[m][h] * c*silu'(z2[m][h])  -> h1T (free now)
    const int q = warp & 3, hh = warp >> 2, m = 32 * q + lane;
    float v[32];
    tmem_ld32(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(hh * 32), v);
#pragma unroll
    for (int j = 0; j < 32; ++j) s.hs.h1T[hh * 32 + j][m] = v[j] * s.D2[hh * 32 + j][m];
    tc_fence_before();
  }
  __syncthreads();
  {
    // dh1[m][g] = sum_h dz2[h][m] W1[g][h] / 8 ;  dz1 = dh1 * c*silu'(z1), z1 recomputed from emb (K = nb)
    const int ty = threadIdx.x & 15, tx = threadIdx.x >> 4;
    float acc[8][4] = {};
#pragma unroll 4
    for (int k = 0; k < TCH; ++k) {
      float bv[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = __ldg(W1 + (4 * tx + j) * TCH + k);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float a = s.hs.h1T[k][ty + 16 * i];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a, bv[j], acc[i][j]);
      }
    }
    float z1[8][4] = {};
    hidden_mm(s.hs.embT, W0, TCH, nb, ty, tx, z1);
    const float s0 = 1.0f / sqrtf((float)nb);
    __syncthreads();   // every thread is done reading D2 (epilogue above) before it is reused
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float v, d;
        silu_vd(z1[i][j] * s0, v, d);
        s.D2[4 * tx + j][ty + 16 * i] = acc[i][j] * 0.125f * d;
      }
  }
  __syncthreads();
  {
    const float s0 = 1.0f / sqrtf((float)nb);
    for (int it = threadIdx.x; it < TCN * nb; it += TCT) {
      const int m = it & 127, n = it >> 7;
      if (m >= ne) continue;
      float t = 0.f;
#pragma unroll 8
      for (int g = 0; g < TCH; ++g) t = fmaf(s.D2[g][m], __ldg(W0 + n * TCH + g), t);
      t *= s0;
      float* p = demb + (size_t)(e0 + m) * nb + n;
      *p = accumulate ? (*p + t) : t;
    }
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 64);
}

bool radial_tc_supported(const MlpDesc& d) {
  return d.h1 == TCH && d.h2 == TCH && d.nb >= 1 && d.nb <= TC_MAX_NB && d.wn >= 4 && (d.wn % 4) == 0;
}

int radial_mlp_fwd_tc(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                      float* w, cudaStream_t st) {
  if (E == 0) return B200_OK;
  cudaFuncSetAttribute(radial_mlp_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TcFwdSmem));
  radial_mlp_fwd_tc_kernel<<<(unsigned)((E + TCN - 1) / TCN), TCT, sizeof(TcFwdSmem), st>>>(d.nb, d.wn, E, emb, W0, W1, W2, w);
  return check_launch("radial_mlp_fwd_tc");
}

int radial_mlp_bwd_tc(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                      const float* dw, float* demb, int accumulate, cudaStream_t st) {
  if (E == 0) return B200_OK;
  cudaFuncSetAttribute(radial_mlp_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TcBwdSmem));
  radial_mlp_bwd_tc_kernel<<<(unsigned)((E + TCN - 1) / TCN), TCT, sizeof(TcBwdSmem), st>>>(d.nb, d.wn, E, emb, W0, W1, W2, dw,
                                                                                             demb, accumulate);
  return check_launch("radial_mlp_bwd_tc");
}

}  // namespace b200
