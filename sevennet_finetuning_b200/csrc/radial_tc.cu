// Radial MLP on the tcgen05 tensor cores (3xTF32, fp32 accumulation in TMEM).
//
// Reference: sevenn/nn/convolution.py:97-106 (weight_nn = e3nn FullyConnectedNet([nb, 64, 64, wn])).
// The two hidden layers (8->64->64, 4.6 kMAC/edge) stay on the CUDA cores; the last layer
// (64 -> weight_numel, 61 kMAC/edge for the 960-wide layers) and its transpose in the backward pass
// are tcgen05.mma GEMMs.
//
// Forward, one CTA per tile of 128 edges:   W^T[wn x 64] . h2^T[64 x 128 edges] -> D[channel, edge]
//   A = 128-channel slab of W2^T (rows = weight channels, K = hidden unit): pre-split into tf32 hi/lo in
//       the canonical operand layout by a prepack kernel and streamed by cp.async.bulk (two-stage ring),
//   B = h2 tile (rows = edges, K = hidden unit), resident in shared memory,
//   D = 128 lanes (channels) x 128 columns (edges) in TMEM, double buffered; the two small cross terms
//       (lo*hi, hi*lo) go to a separate accumulator so the main one sees 8 instead of 24 roundings.
// TMEM lane = weight channel, so the epilogue thread of a lane owns one channel for all edges of the
// tile and its global stores are coalesced across the warp (the same ownership the convolution uses).
#include "common.cuh"
#include "tc.cuh"

namespace b200 {

using namespace tc;

constexpr int TCN = 128;              // edges per tile
constexpr int TCH = 64;               // hidden width (K of the big GEMM)
constexpr int TCT = 512;              // threads (16 warps: 4 per scheduler to hide FFMA/LDS latency)
constexpr int CHB = TCN * 16;         // bytes of one 16-byte K-chunk column of a 128-row operand
constexpr int OPB = (TCH / 4) * CHB;  // one [128 x 64] fp32 operand: 32 KB
constexpr int TC_MAX_NB = 16;

__device__ long long g_tc_dbg[128];
#define TCDBG(i) do { if (blockIdx.x == 1000 && threadIdx.x == 0) g_tc_dbg[i] = clock64(); } while (0)

struct HiddenScratch {
  float embT[TC_MAX_NB][TCN];   // [basis function][edge]
  float h1T[TCH][TCN];          // [hidden unit][edge]
};

struct TcFwdSmem {
  uint8_t Bhi[OPB], Blo[OPB];   // h2 tile, hi / lo parts
  uint8_t A[2][2][OPB];         // [stage][hi/lo] slab of W2^T; stage 1 doubles as scratch for the hidden layers
  uint64_t full[2];             // slab landed (bulk copy complete_tx)
  uint64_t mma_done[2];         // MMAs of the tile retired (tcgen05.commit)
  uint32_t tmem_base;
};

// ---- prepack: W2[64][wn] (reference layout) -> operand slabs ---------------------------------------------
// fwd: [NT][hi|lo][16 chunks][128 channel rows][4]  (row r of slab t = channel 128 t + r, K = hidden unit)
// bwd: [NK][hi|lo][ 8 chunks][ 64 hidden rows ][4]  (K = channel 32 i + k)
__global__ void radial_prepack_kernel(int wn, int NT, int NK, const float* __restrict__ W2, float* __restrict__ pf,
                                      float* __restrict__ pb) {
  const int nf = NT * 16 * 128, nbk = NK * 8 * 64;
  for (int it = blockIdx.x * blockDim.x + threadIdx.x; it < nf + nbk; it += gridDim.x * blockDim.x) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 hi, lo;
    if (it < nf) {
      const int r = it & 127, c4 = (it >> 7) & 15, t = it >> 11;
      const int ch = t * 128 + r;
      if (ch < wn) {
        v.x = W2[(size_t)(4 * c4 + 0) * wn + ch]; v.y = W2[(size_t)(4 * c4 + 1) * wn + ch];
        v.z = W2[(size_t)(4 * c4 + 2) * wn + ch]; v.w = W2[(size_t)(4 * c4 + 3) * wn + ch];
      }
      v.x *= 0.125f; v.y *= 0.125f; v.z *= 0.125f; v.w *= 0.125f;
      split4(v, hi, lo);
      float* dst = pf + (size_t)t * (2 * OPB / 4) + (size_t)c4 * (CHB / 4) + r * 4;
      *reinterpret_cast<float4*>(dst) = hi;
      *reinterpret_cast<float4*>(dst + OPB / 4) = lo;
    } else {
      const int j = it - nf;
      const int h = j & 63, c4 = (j >> 6) & 7, i = j >> 9;
      const int c = i * 32 + 4 * c4;
      if (c + 0 < wn) v.x = W2[(size_t)h * wn + c + 0];
      if (c + 1 < wn) v.y = W2[(size_t)h * wn + c + 1];
      if (c + 2 < wn) v.z = W2[(size_t)h * wn + c + 2];
      if (c + 3 < wn) v.w = W2[(size_t)h * wn + c + 3];
      v.x *= 0.125f; v.y *= 0.125f; v.z *= 0.125f; v.w *= 0.125f;
      split4(v, hi, lo);
      float* dst = pb + (size_t)i * (2 * 8 * 64 * 4) + (size_t)c4 * (64 * 4) + h * 4;
      *reinterpret_cast<float4*>(dst) = hi;
      *reinterpret_cast<float4*>(dst + 8 * 64 * 4) = lo;
    }
  }
}

// 128 x 64 tile product on the CUDA cores: acc[i][j] += sum_k A[k][ty+32i] * B[k*ldb + 4tx+j]
// (ty = lane: rows ty, ty+32, ty+64, ty+96; tx = warp: 4 columns; the B row is a warp-uniform broadcast)
constexpr int HR = 4;
__device__ __forceinline__ void hidden_mm(const float (*A)[TCN], const float* __restrict__ B, int ldb, int K, int ty, int tx,
                                          float acc[HR][4]) {
#pragma unroll 8
  for (int k = 0; k < K; ++k) {
    const float4 b = *reinterpret_cast<const float4*>(B + k * ldb + 4 * tx);
    const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < HR; ++i) {
      const float a = A[k][ty + 32 * i];
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a, bv[j], acc[i][j]);
    }
  }
}

// hidden layers for the CTA's 128 edges; writes h2 (hi/lo) into the B operand and/or c*silu'(z2) as D2[h][m].
// emb/W0/W1 in the reference layouts ([E,nb], [nb,64], [64,64]).
template <bool WITH_D2, bool WITH_H2>
__device__ __forceinline__ void hidden_layers(HiddenScratch& hs, uint8_t* Bhi, uint8_t* Blo, float (*D2)[TCN], int nb,
                                              long long e0, int ne, const float* __restrict__ emb,
                                              const float* __restrict__ W0, const float* __restrict__ W1) {
  const int ty = threadIdx.x & 31, tx = threadIdx.x >> 5;
  for (int it = threadIdx.x; it < TCN * nb; it += TCT) {
    const int m = it / nb, k = it % nb;
    hs.embT[k][m] = (m < ne) ? __ldg(emb + (e0 + m) * nb + k) : 0.f;
  }
  __syncthreads();
  float acc[HR][4] = {};
  hidden_mm(hs.embT, W0, TCH, nb, ty, tx, acc);
  const float s0 = 1.0f / sqrtf((float)nb), s1 = 1.0f / sqrtf((float)TCH);
#pragma unroll
  for (int i = 0; i < HR; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) hs.h1T[4 * tx + j][ty + 32 * i] = kSilu2Mom * silu_f(acc[i][j] * s0);
  __syncthreads();
  float acc2[HR][4] = {};
  hidden_mm(hs.h1T, W1, TCH, TCH, ty, tx, acc2);
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    float v[4], d[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) silu_vd(acc2[i][j] * s1, v[j], d[j]);
    const int m = ty + 32 * i;
    if constexpr (WITH_H2) {
      float4 hi, lo;
      split4(make_float4(v[0], v[1], v[2], v[3]), hi, lo);
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
                         const float* __restrict__ W1, const float* __restrict__ pack, float* __restrict__ w) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  TcFwdSmem& s = *reinterpret_cast<TcFwdSmem*>(smem_raw);
  HiddenScratch& hs = *reinterpret_cast<HiddenScratch*>(&s.A[1][0][0]);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * TCN;
  const int ne = (int)min((long long)TCN, E - e0);
  const int NT = (wn + TCN - 1) / TCN;
  constexpr uint32_t idesc = make_idesc_tf32(128, TCN);
  constexpr uint32_t SLAB = 2 * OPB;   // hi + lo, contiguous in the pack and in A[stage]

  TCDBG(0);
  if (warp == 0) tmem_alloc(&s.tmem_base, 512);
  if (threadIdx.x == 32) {
    mbar_init(&s.full[0], 1); mbar_init(&s.full[1], 1);
    mbar_init(&s.mma_done[0], 1); mbar_init(&s.mma_done[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_expect_tx(&s.full[0], SLAB);
    bulk_copy_g2s(s.A[0][0], pack, SLAB, &s.full[0]);
  }
  hidden_layers<false, true>(hs, s.Bhi, s.Blo, nullptr, nb, e0, ne, emb, W0, W1);
  fence_proxy_async();   // h2 operand (and the scratch in stage 1) were written through the generic proxy
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;
  if (threadIdx.x == 0 && NT > 1) {
    mbar_expect_tx(&s.full[1], SLAB);
    bulk_copy_g2s(s.A[1][0], pack + (size_t)(SLAB / 4), SLAB, &s.full[1]);
  }

  TCDBG(1);
  for (int mt = 0; mt <= NT; ++mt) {
    const int st = mt & 1;
    TCDBG(2 + 5 * mt);
    if (mt < NT && threadIdx.x == 0) {
      mbar_wait(&s.full[st], (mt >> 1) & 1);
      TCDBG(3 + 5 * mt);
      tc_fence_after();
      const uint32_t dm = tmem + (uint32_t)st * 256, dc = dm + 128;
      const uint32_t ahi = smem_u32(s.A[st][0]), alo = smem_u32(s.A[st][1]);
      const uint32_t bhi = smem_u32(s.Bhi), blo = smem_u32(s.Blo);
#pragma unroll
      for (int kk = 0; kk < TCH / 8; ++kk) {
        const uint32_t off = kk * 2 * CHB;
        const uint64_t dah = make_desc(ahi + off, CHB, 128), dal = make_desc(alo + off, CHB, 128);
        const uint64_t dbh = make_desc(bhi + off, CHB, 128), dbl = make_desc(blo + off, CHB, 128);
        mma_tf32(dc, dal, dbh, idesc, kk > 0);
        mma_tf32(dc, dah, dbl, idesc, 1);
        mma_tf32(dm, dah, dbh, idesc, kk > 0);
      }
      mma_commit(&s.mma_done[st]);
    }
    TCDBG(4 + 5 * mt);
    if (mt > 0) {
      // epilogue of tile mt-1: lanes = channels, columns = edges
      const int pt = mt - 1, ps = pt & 1;
      mbar_wait(&s.mma_done[ps], (pt >> 1) & 1);
      TCDBG(5 + 5 * mt);
      tc_fence_after();
      if (threadIdx.x == 0 && mt + 1 < NT) {   // stage ps is free again: prefetch slab mt+1 behind the epilogue
        mbar_expect_tx(&s.full[ps], SLAB);
        bulk_copy_g2s(s.A[ps][0], pack + (size_t)(mt + 1) * (SLAB / 4), SLAB, &s.full[ps]);
      }
      const int q = warp & 3, col = (warp >> 2) * 32;
      const int ch = pt * TCN + 32 * q + lane;
      {
        float v[32], c[32];
        const uint32_t ta = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(ps * 256 + col);
        tmem_ld32(ta, v);
        tmem_ld32(ta + 128, c);
        if (ch < wn) {
          float* wb = w + (size_t)e0 * wn;           // CTA-uniform base, 32-bit offsets below
          uint32_t off = (uint32_t)col * (uint32_t)wn + (uint32_t)ch;
          if (col + 32 <= ne) {
#pragma unroll
            for (int j = 0; j < 32; ++j, off += wn) wb[off] = v[j] + c[j];
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j, off += wn)
              if (col + j < ne) wb[off] = v[j] + c[j];
          }
        }
      }
      tc_fence_before();
    }
    TCDBG(6 + 5 * mt);
    __syncthreads();   // TMEM stage of tile mt-1 is drained before MMA(mt+1) overwrites it
  }
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// ---- backward (data gradient): demb (+)= d/d emb sum_e <dw[e], w[e]> -------------------------------------
// dh2[e, h] = sum_c dw[e, c] W2[h, c] / 8 is the tensor-core GEMM: A = dw tile (rows = edges, K = channel),
// B = W2 (rows = hidden unit, K = channel; prepacked, bulk-copied), D = 128 lanes (edges) x 64 columns.
// K is streamed in chunks of 32 channels: cp.async into a padded staging tile one chunk ahead, split into
// tf32 hi/lo operands, MMA.  Six round-robin main accumulators + one for the cross terms keep the number of
// tensor-core roundings per partial sum small; they are added in fp32 (round-to-nearest) in the epilogue.
constexpr int BK = 32;                 // channels per K chunk
constexpr int BCH = TCH * 16;          // bytes of one 16-byte K-chunk column of the 64-row B operand
constexpr int SLD = BK + 4;            // padded row length of the staging tile (conflict-free float4 columns)
constexpr int NACC = 6;                // main accumulators
constexpr int BOPB = (BK / 4) * BCH;   // one [64 x 32] operand part: 8 KB

struct TcBwdSmem {
  uint8_t A[2][2][(BK / 4) * CHB];     // [stage][hi/lo] dw chunk      64 KB
  uint8_t B[3][2][BOPB];               // [stage][hi/lo] W2 chunk      48 KB
  float stage_dw[2][TCN][SLD];         // 36 KB
  HiddenScratch hs;                    // 40 KB
  float D2[TCH][TCN];                  // 32 KB: c*silu'(z2), later dz1
  uint64_t fullB[3];
  uint64_t mma_done[2];
  uint32_t tmem_base;
};

__device__ __forceinline__ void cp_async16(void* dst, const void* src, bool valid) {
  const uint32_t n = valid ? 16u : 0u;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(n) : "memory");
}

__global__ void __launch_bounds__(TCT, 1)
radial_mlp_bwd_tc_kernel(int nb, int wn, long long E, const float* __restrict__ emb, const float* __restrict__ W0,
                         const float* __restrict__ W1, const float* __restrict__ pack, const float* __restrict__ dw,
                         float* __restrict__ demb, int accumulate) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  TcBwdSmem& s = *reinterpret_cast<TcBwdSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * TCN;
  const int ne = (int)min((long long)TCN, E - e0);
  const int NK = (wn + BK - 1) / BK;
  constexpr uint32_t idesc = make_idesc_tf32(128, TCH);
  constexpr uint32_t BSLAB = 2 * BOPB;

  auto stage_chunk = [&](int i) {   // global dw chunk i -> stage_dw[i & 1], coalesced along the channel axis
    const int k0 = i * BK;
#pragma unroll
    for (int q = 0; q < (TCN * 8) / TCT; ++q) {
      const int idx = threadIdx.x + TCT * q, row = idx >> 3, c4 = idx & 7;
      const bool ok = row < ne && k0 + 4 * c4 < wn;
      const float* src = ok ? dw + (size_t)(e0 + row) * wn + k0 + 4 * c4 : dw;
      cp_async16(&s.stage_dw[i & 1][row][4 * c4], src, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  if (warp == 0) tmem_alloc(&s.tmem_base, 512);
  if (threadIdx.x == 32) {
    for (int k = 0; k < 3; ++k) mbar_init(&s.fullB[k], 1);
    mbar_init(&s.mma_done[0], 1); mbar_init(&s.mma_done[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  stage_chunk(0);
  if (threadIdx.x == 0) {
    mbar_expect_tx(&s.fullB[0], BSLAB);
    bulk_copy_g2s(s.B[0][0], pack, BSLAB, &s.fullB[0]);
  }
  hidden_layers<true, false>(s.hs, nullptr, nullptr, s.D2, nb, e0, ne, emb, W0, W1);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  for (int i = 0; i < NK; ++i) {
    const int st = i & 1;
    if (i + 1 < NK) {
      stage_chunk(i + 1);   // stage_dw[(i+1)&1] was consumed by the split pass of iteration i-1
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    // operand stage st (and B stage (i+1)%3) were last read by the MMAs of chunk i-2
    if (i >= 2) mbar_wait(&s.mma_done[st], ((i - 2) >> 1) & 1);
    if (threadIdx.x == 0 && i + 1 < NK) {
      const int bs = (i + 1) % 3;
      mbar_expect_tx(&s.fullB[bs], BSLAB);
      bulk_copy_g2s(s.B[bs][0], pack + (size_t)(i + 1) * (BSLAB / 4), BSLAB, &s.fullB[bs]);
    }
    __syncthreads();   // every thread's cp.async data for chunk i is visible
    // staging -> hi/lo operands in the canonical K-major layout (lanes run along rows)
#pragma unroll
    for (int q = 0; q < (TCN * 8) / TCT; ++q) {
      const int idx = threadIdx.x + TCT * q, row = idx & 127, c4 = idx >> 7;
      const float4 v = *reinterpret_cast<const float4*>(&s.stage_dw[st][row][4 * c4]);
      float4 hi, lo;
      split4(v, hi, lo);
      *reinterpret_cast<float4*>(s.A[st][0] + c4 * CHB + row * 16) = hi;
      *reinterpret_cast<float4*>(s.A[st][1] + c4 * CHB + row * 16) = lo;
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (threadIdx.x == 0) {
      const int bs = i % 3;
      mbar_wait(&s.fullB[bs], (i / 3) & 1);
      tc_fence_after();
      const uint32_t dm = tmem + (uint32_t)(i % NACC) * TCH, dc = tmem + (uint32_t)NACC * TCH;
      const uint32_t ahi = smem_u32(s.A[st][0]), alo = smem_u32(s.A[st][1]);
      const uint32_t bhi = smem_u32(s.B[bs][0]), blo = smem_u32(s.B[bs][1]);
#pragma unroll
      for (int kk = 0; kk < BK / 8; ++kk) {
        const uint64_t dah = make_desc(ahi + kk * 2 * CHB, CHB, 128), dal = make_desc(alo + kk * 2 * CHB, CHB, 128);
        const uint64_t dbh = make_desc(bhi + kk * 2 * BCH, BCH, 128), dbl = make_desc(blo + kk * 2 * BCH, BCH, 128);
        mma_tf32(dc, dal, dbh, idesc, (i > 0 || kk > 0) ? 1u : 0u);
        mma_tf32(dc, dah, dbl, idesc, 1);
        mma_tf32(dm, dah, dbh, idesc, (i >= NACC || kk > 0) ? 1u : 0u);
      }
      mma_commit(&s.mma_done[st]);
    }
    __syncwarp();
  }
  // all MMAs retire in order: the last commit covers every chunk
  mbar_wait(&s.mma_done[(NK - 1) & 1], ((NK - 1) >> 1) & 1);
  tc_fence_after();
  if (warp < 8) {
    // dz2[h][m] = dh2[m][h] * c*silu'(z2[m][h])  -> h1T (free now); dh2 = sum of the partial accumulators
    const int q = warp & 3, hh = warp >> 2, m = 32 * q + lane;
    const uint32_t ta = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(hh * 32);
    float v[32], t[32];
    tmem_ld32(ta + NACC * TCH, v);
    const int np = NK < NACC ? NK : NACC;
    for (int p = 0; p < np; ++p) {
      tmem_ld32(ta + p * TCH, t);
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] += t[j];
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) s.hs.h1T[hh * 32 + j][m] = v[j] * s.D2[hh * 32 + j][m];
    tc_fence_before();
  }
  __syncthreads();
  {
    // dh1[m][g] = sum_h dz2[h][m] W1[g][h] / 8 ;  dz1 = dh1 * c*silu'(z1), z1 recomputed from emb (K = nb)
    const int ty = threadIdx.x & 31, tx = threadIdx.x >> 5;
    float acc[HR][4] = {};
#pragma unroll 4
    for (int k = 0; k < TCH; ++k) {
      float bv[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = __ldg(W1 + (4 * tx + j) * TCH + k);
#pragma unroll
      for (int i = 0; i < HR; ++i) {
        const float a = s.hs.h1T[k][ty + 32 * i];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a, bv[j], acc[i][j]);
      }
    }
    float z1[HR][4] = {};
    hidden_mm(s.hs.embT, W0, TCH, nb, ty, tx, z1);
    const float s0 = 1.0f / sqrtf((float)nb);
    __syncthreads();   // every thread is done reading D2 (epilogue above) before it is reused
#pragma unroll
    for (int i = 0; i < HR; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float v, d;
        silu_vd(z1[i][j] * s0, v, d);
        s.D2[4 * tx + j][ty + 32 * i] = acc[i][j] * 0.125f * d;
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
  if (warp == 0) tmem_dealloc(tmem, 512);
}

extern "C" void b200_debug_tc(long long* out, int n) { cudaMemcpyFromSymbol(out, g_tc_dbg, sizeof(long long) * n); }

bool radial_tc_supported(const MlpDesc& d) {
  return d.h1 == TCH && d.h2 == TCH && d.nb >= 1 && d.nb <= TC_MAX_NB && d.wn >= 4 && (d.wn % 4) == 0;
}

// operand slabs of W2 for one call (stream-ordered scratch: prepack -> kernel -> free)
static int with_pack(const MlpDesc& d, const float* W2, cudaStream_t st, float** pf, float** pb, float** base) {
  const int NT = (d.wn + TCN - 1) / TCN, NK = (d.wn + BK - 1) / BK;
  const size_t nf = (size_t)NT * 2 * OPB, nbk = (size_t)NK * 2 * BOPB;
  if (cudaMallocAsync((void**)base, nf + nbk, st) != cudaSuccess) {
    set_error("radial_mlp: cudaMallocAsync of the operand pack failed");
    return B200_ERR_CUDA;
  }
  *pf = *base;
  *pb = *base + nf / 4;
  radial_prepack_kernel<<<148, 256, 0, st>>>(d.wn, NT, NK, W2, *pf, *pb);
  return B200_OK;
}

int radial_mlp_fwd_tc(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                      float* w, cudaStream_t st) {
  if (E == 0) return B200_OK;
  float *pf, *pb, *base;
  if (int rc = with_pack(d, W2, st, &pf, &pb, &base)) return rc;
  cudaFuncSetAttribute(radial_mlp_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TcFwdSmem));
  radial_mlp_fwd_tc_kernel<<<(unsigned)((E + TCN - 1) / TCN), TCT, sizeof(TcFwdSmem), st>>>(d.nb, d.wn, E, emb, W0, W1, pf, w);
  cudaFreeAsync(base, st);
  return check_launch("radial_mlp_fwd_tc", 2);
}

int radial_mlp_bwd_tc(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                      const float* dw, float* demb, int accumulate, cudaStream_t st) {
  if (E == 0) return B200_OK;
  float *pf, *pb, *base;
  if (int rc = with_pack(d, W2, st, &pf, &pb, &base)) return rc;
  cudaFuncSetAttribute(radial_mlp_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TcBwdSmem));
  radial_mlp_bwd_tc_kernel<<<(unsigned)((E + TCN - 1) / TCN), TCT, sizeof(TcBwdSmem), st>>>(d.nb, d.wn, E, emb, W0, W1, pb, dw,
                                                                                             demb, accumulate);
  cudaFreeAsync(base, st);
  return check_launch("radial_mlp_bwd_tc", 2);
}

}  // namespace b200
