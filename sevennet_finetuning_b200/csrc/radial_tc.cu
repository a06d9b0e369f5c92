// Radial MLP on the tcgen05 tensor cores (3xTF32, fp32 accumulation in TMEM).
//
// Reference: sevenn/nn/convolution.py:97-106 (weight_nn = e3nn FullyConnectedNet([nb, 64, 64, wn])).
// Only the first layer (nb -> 64, K = 8) stays on the CUDA cores.  The 64 -> 64 layer, the last layer
// (64 -> weight_numel, 61 kMAC/edge for the 960-wide layers) and their transposes in the backward pass are
// tcgen05.mma GEMMs.
//
// Both kernels run 16 worker warps and one issuer warp per CTA of 128 edges.  The issuer's single thread
// owns the TMA (cp.async.bulk) loads and every tcgen05.mma; it talks to the workers through mbarriers only,
// so a full tensor-pipe queue never stalls the warps that convert operands and drain TMEM.
//
// Forward:   W2^T[wn x 64] . h2^T[64 x 128 edges] -> D[channel, edge]
//   A = 128-channel slab of W2^T (rows = weight channels, K = hidden unit): pre-split into tf32 hi/lo in
//       the canonical operand layout by a prepack kernel and streamed by cp.async.bulk (two-stage ring),
//   B = h2 tile (rows = edges, K = hidden unit), resident in shared memory,
//   D = 128 lanes (channels) x 128 columns (edges) in TMEM, double buffered; the two small cross terms
//       (lo*hi, hi*lo) go to a separate accumulator so the main one sees 8 instead of 24 roundings.
// TMEM lane = weight channel, so the epilogue thread of a lane owns one channel for all edges of the
// tile and its global stores are coalesced across the warp (the same ownership the convolution uses).
//
// The tensor core accumulates with truncation (a bias of about -2.4e-8 relative per MMA into one
// accumulator), so products that feed the energy are spread over several TMEM accumulators that see at
// most 8 MMAs each and are summed in registers with round-to-nearest adds.
#include "common.cuh"
#include "tc.cuh"

namespace b200 {

using namespace tc;

constexpr int TCN = 128;              // edges per tile
constexpr int TCH = 64;               // hidden width (K of the big GEMM)
constexpr int TCW = 512;              // worker threads (16 warps: 4 per scheduler)
constexpr int TCT = TCW + 32;         // + the issuer warp
constexpr int CHB = TCN * 16;         // bytes of one 16-byte K-chunk column of a 128-row operand
constexpr int OPB = (TCH / 4) * CHB;  // one [128 x 64] fp32 operand part: 32 KB
constexpr int BCH = TCH * 16;         // bytes of one K-chunk column of a 64-row operand
constexpr int W1B = (TCH / 4) * BCH;  // one [64 x 64] operand part: 16 KB
constexpr int TC_MAX_NB = 16;
constexpr int BK = 32;                 // backward: channels per K chunk
constexpr int BOPB = (BK / 4) * BCH;   // one [64 x 32] operand part: 8 KB

__device__ long long g_tc_dbg[128];
__device__ long long g_tc_cta[3 * 8192];   // per CTA: start / end (globaltimer ns), SM id
__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ int smid() { int t; asm volatile("mov.u32 %0, %smid;" : "=r"(t)); return t; }
#define TCCTA(k) do { if (threadIdx.x == 0 && blockIdx.x < 8192) { g_tc_cta[3 * blockIdx.x + (k)] = gtimer(); g_tc_cta[3 * blockIdx.x + 2] = smid(); } } while (0)
#define TCDBG(i, t) do { if (blockIdx.x == 1000 && threadIdx.x == (t)) g_tc_dbg[i] = clock64(); } while (0)

// ---- prepack: W2[64][wn], W1[64][64] (reference layouts) -> operand slabs ---------------------------------
// fwd: [NT][hi|lo][16 chunks][128 channel rows][4]  (row r of slab t = channel 128 t + r, K = hidden unit)
// bwd: [NK][ 8 chunks][128 rows: hi of hidden unit 0..63 | lo of hidden unit 0..63][4]  (K = channel 32 i + k)
// w1t: [16 chunks][128 rows: hi | lo of output unit h][4]  (K = input unit g)    z2 = h1 . W1
// w1 : [16 chunks][128 rows: hi | lo of input unit g ][4]  (K = output unit h)   dh1 = dz2 . W1^T
// The 64-row operands carry their tf32 hi and lo parts as rows 0..63 and 64..127 of ONE 128-row B operand: a single
// N = 128 MMA then yields x.hi in columns 0..63 and x.lo in columns 64..127 at the cost of the N = 64 instruction
// (an M = 128, K = 8 tf32 MMA is bound by the 4 KB A-operand read, about 64 clocks, for every N <= 128).
// every slab carries the 1/sqrt(64) of the layer it belongs to.
__global__ void radial_prepack_kernel(int wn, int NT, int NK, const float* __restrict__ W1, const float* __restrict__ W2,
                                      float* __restrict__ pf, float* __restrict__ pb, float* __restrict__ pw) {
  const int nf = NT * 16 * 128, nbk = NK * 8 * 64, nw = 2 * 16 * 64;
  for (int it = blockIdx.x * blockDim.x + threadIdx.x; it < nf + nbk + nw; it += gridDim.x * blockDim.x) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    float* dst;
    int lo_off;
    if (it < nf) {
      const int r = it & 127, c4 = (it >> 7) & 15, t = it >> 11;
      const int ch = t * 128 + r;
      if (ch < wn) {
        v.x = W2[(size_t)(4 * c4 + 0) * wn + ch]; v.y = W2[(size_t)(4 * c4 + 1) * wn + ch];
        v.z = W2[(size_t)(4 * c4 + 2) * wn + ch]; v.w = W2[(size_t)(4 * c4 + 3) * wn + ch];
      }
      dst = pf + (size_t)t * (2 * OPB / 4) + (size_t)c4 * (CHB / 4) + r * 4;
      lo_off = OPB / 4;
    } else if (it < nf + nbk) {
      const int j = it - nf;
      const int h = j & 63, c4 = (j >> 6) & 7, i = j >> 9;
      const int c = i * 32 + 4 * c4;
      if (c + 0 < wn) v.x = W2[(size_t)h * wn + c + 0];
      if (c + 1 < wn) v.y = W2[(size_t)h * wn + c + 1];
      if (c + 2 < wn) v.z = W2[(size_t)h * wn + c + 2];
      if (c + 3 < wn) v.w = W2[(size_t)h * wn + c + 3];
      dst = pb + (size_t)i * (2 * BOPB / 4) + (size_t)c4 * (CHB / 4) + h * 4;
      lo_off = TCH * 4;
    } else {
      const int j = it - nf - nbk;
      const int row = j & 63, c4 = (j >> 6) & 15, which = j >> 10;
      if (which == 0) {
        v.x = W1[(4 * c4 + 0) * TCH + row]; v.y = W1[(4 * c4 + 1) * TCH + row];
        v.z = W1[(4 * c4 + 2) * TCH + row]; v.w = W1[(4 * c4 + 3) * TCH + row];
      } else {
        v = *reinterpret_cast<const float4*>(W1 + row * TCH + 4 * c4);
      }
      dst = pw + (size_t)which * (2 * W1B / 4) + (size_t)c4 * (CHB / 4) + row * 4;
      lo_off = TCH * 4;
    }
    v.x *= 0.125f; v.y *= 0.125f; v.z *= 0.125f; v.w *= 0.125f;
    float4 hi, lo;
    split4(v, hi, lo);
    *reinterpret_cast<float4*>(dst) = hi;
    *reinterpret_cast<float4*>(dst + lo_off) = lo;
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

// workers: emb tile -> embT, z1 = emb . W0 (K = nb, CUDA cores), h1 = c silu(z1 / sqrt(nb)) written as the
// hi/lo operand [128 edges x 64] at h1 (hi) and h1 + OPB (lo).  emb / W0 in the reference layouts.
// The emb values and this thread's W0 columns are requested first thing, so both round trips overlap.
__device__ __forceinline__ void hidden1_operand(float (*embT)[TCN], uint8_t* h1, int nb, long long e0, int ne,
                                                const float* __restrict__ emb, const float* __restrict__ W0) {
  const int ty = threadIdx.x & 31, tx = threadIdx.x >> 5;
  // element it = tid + 512 j of the [128 x nb] tile (nb <= 16: at most 4 per thread), coalesced
  float ev[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int it = threadIdx.x + TCW * j;
    ev[j] = (it < TCN * nb && it / nb < ne) ? __ldg(emb + e0 * nb + it) : 0.f;
  }
  constexpr int WPRE = 8;   // basis functions whose W0 rows are preloaded (SevenNet-0: all eight)
  float4 wv[WPRE];
#pragma unroll
  for (int k = 0; k < WPRE; ++k)
    wv[k] = k < nb ? __ldg(reinterpret_cast<const float4*>(W0 + k * TCH + 4 * tx)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int it = threadIdx.x + TCW * j;
    if (it < TCN * nb) embT[it % nb][it / nb] = ev[j];
  }
  bar_sync(1, TCW);
  float acc[HR][4] = {};
#pragma unroll
  for (int k = 0; k < WPRE; ++k) {
    if (k < nb) {
      const float bv[4] = {wv[k].x, wv[k].y, wv[k].z, wv[k].w};
#pragma unroll
      for (int i = 0; i < HR; ++i) {
        const float a = embT[k][ty + 32 * i];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a, bv[j], acc[i][j]);
      }
    }
  }
  if (nb > WPRE) hidden_mm(embT + WPRE, W0 + WPRE * TCH, TCH, nb - WPRE, ty, tx, acc);
  const float s0 = 1.0f / sqrtf((float)nb);
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    float4 v, hi, lo;
    v.x = kSilu2Mom * silu_f(acc[i][0] * s0); v.y = kSilu2Mom * silu_f(acc[i][1] * s0);
    v.z = kSilu2Mom * silu_f(acc[i][2] * s0); v.w = kSilu2Mom * silu_f(acc[i][3] * s0);
    split4(v, hi, lo);
    uint8_t* p = h1 + tx * CHB + (ty + 32 * i) * 16;   // chunk tx holds hidden units 4tx..4tx+3
    *reinterpret_cast<float4*>(p) = hi;
    *reinterpret_cast<float4*>(p + OPB) = lo;
  }
}

// issuer: D[128 rows x 64] = A[128 x 64] . B[64 x 64]^T with A (hi at a, lo at a + OPB) and the stacked hi|lo B
// operand at b in shared memory.  A.hi . [B.hi | B.lo] alternates between the accumulators at TMEM columns 0 and 128
// (four K-steps each), A.lo . [B.hi | B.lo] goes to column 256.
__device__ __forceinline__ void issue_64x64(uint32_t tmem, uint32_t a, uint32_t b) {
  constexpr uint32_t idesc = make_idesc_tf32(128, 128);
#pragma unroll
  for (int kk = 0; kk < TCH / 8; ++kk) {
    const uint64_t dah = make_desc(a + kk * 2 * CHB, CHB, 128), dal = make_desc(a + OPB + kk * 2 * CHB, CHB, 128);
    const uint64_t db = make_desc(b + kk * 2 * CHB, CHB, 128);
    mma_tf32(tmem + 256, dal, db, idesc, kk > 0);
    mma_tf32(tmem + (uint32_t)(kk & 1) * 128, dah, db, idesc, kk >= 2);
  }
}
// worker warp (lane quarter q, column group cg): its 16 columns of the product above, partial sums added in fp32
// (small terms first)
__device__ __forceinline__ void read_64x64(uint32_t tmem, int q, int cg, float (&z)[16]) {
  const uint32_t ta = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(cg * 16);
  float t[16];
  tmem_ld16(ta + 256 + TCH, z);
  const uint32_t rest[5] = {256, TCH, 128 + TCH, 0, 128};
#pragma unroll
  for (int a = 0; a < 5; ++a) {
    tmem_ld16(ta + rest[a], t);
#pragma unroll
    for (int j = 0; j < 16; ++j) z[j] += t[j];
  }
}

// ---- forward --------------------------------------------------------------------------------------------
struct TcFwdSmem {
  uint8_t B[2][OPB];            // h2 tile, hi / lo parts; before that: W1^T operand in B[0], embT in B[1]
  uint8_t A[2][2][OPB];         // [stage][hi/lo] slab of W2^T; stage 1 first holds the h1 operand
  uint64_t full[2];             // slab landed (bulk copy complete_tx)
  uint64_t mma_done[2];         // MMAs of the slab retired (tcgen05.commit)
  uint64_t tmem_empty[2];       // the 16 worker warps drained the accumulator stage
  uint64_t w1_full, h1_ready, z2_done, h2_ready;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(TCT, 1)
radial_mlp_fwd_tc_kernel(int nb, int wn, long long E, const float* __restrict__ emb, const float* __restrict__ W0,
                         const float* __restrict__ pack, const float* __restrict__ pw, float* __restrict__ w) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  TcFwdSmem& s = *reinterpret_cast<TcFwdSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * TCN;
  const int ne = (int)min((long long)TCN, E - e0);
  const int NT = (wn + TCN - 1) / TCN;
  constexpr uint32_t SLAB = 2 * OPB;   // hi + lo, contiguous in the pack and in A[stage]
  // every CTA walks the slabs in its own rotation, so that the CTAs running at the same time do not all ask
  // the same L2 slices for the same slab
  const int rot = blockIdx.x % NT;
  auto slab_of = [&](int mt) { const int t = mt + rot; return t >= NT ? t - NT : t; };

  TCCTA(0);
  if (warp == 16) tmem_alloc(&s.tmem_base, 512);
  if (threadIdx.x == 0) {
    for (int k = 0; k < 2; ++k) { mbar_init(&s.full[k], 1); mbar_init(&s.mma_done[k], 1); mbar_init(&s.tmem_empty[k], 16); }
    mbar_init(&s.w1_full, 1); mbar_init(&s.z2_done, 1);
    mbar_init(&s.h1_ready, 16); mbar_init(&s.h2_ready, 16);
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  if (warp == 16) {
    if (elect_one()) {
      mbar_expect_tx(&s.w1_full, 2 * W1B);
      bulk_copy_g2s(s.B[0], pw, 2 * W1B, &s.w1_full);
      mbar_expect_tx(&s.full[0], SLAB);
      bulk_copy_g2s(s.A[0][0], pack + (size_t)slab_of(0) * (SLAB / 4), SLAB, &s.full[0]);
      TCDBG(40, 512);
      mbar_wait(&s.w1_full, 0);
      mbar_wait(&s.h1_ready, 0);
      TCDBG(41, 512);
      tc_fence_after();
      issue_64x64(tmem, smem_u32(s.A[1][0]), smem_u32(s.B[0]));
      mma_commit(&s.z2_done);
      mbar_wait(&s.z2_done, 0);
      TCDBG(42, 512);            // the h1 operand is consumed: its stage can take slab 1
      if (NT > 1) {
        mbar_expect_tx(&s.full[1], SLAB);
        bulk_copy_g2s(s.A[1][0], pack + (size_t)slab_of(1) * (SLAB / 4), SLAB, &s.full[1]);
      }
      mbar_wait(&s.h2_ready, 0);
      TCDBG(43, 512);
      constexpr uint32_t idesc = make_idesc_tf32(128, TCN);
      const uint32_t bhi = smem_u32(s.B[0]), blo = smem_u32(s.B[1]);
      for (int mt = 0; mt < NT; ++mt) {
        const int st = mt & 1;
        TCDBG(44 + 4 * mt, 512);
        mbar_wait(&s.full[st], (mt >> 1) & 1);
        TCDBG(45 + 4 * mt, 512);
        if (mt >= 2) mbar_wait(&s.tmem_empty[st], ((mt >> 1) - 1) & 1);
        TCDBG(46 + 4 * mt, 512);
        tc_fence_after();
        const uint32_t dm = tmem + (uint32_t)st * 256, dc = dm + 128;
        const uint32_t ahi = smem_u32(s.A[st][0]), alo = smem_u32(s.A[st][1]);
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
        TCDBG(47 + 4 * mt, 512);
        if (mt >= 1 && mt + 1 < NT) {      // the other stage is free once slab mt-1 has retired
          mbar_wait(&s.mma_done[st ^ 1], ((mt - 1) >> 1) & 1);
          mbar_expect_tx(&s.full[st ^ 1], SLAB);
          bulk_copy_g2s(s.A[st ^ 1][0], pack + (size_t)slab_of(mt + 1) * (SLAB / 4), SLAB, &s.full[st ^ 1]);
        }
      }
    }
  } else {
    const int q = warp & 3, cg = warp >> 2;
    TCDBG(0, 0);
    hidden1_operand(reinterpret_cast<float(*)[TCN]>(s.B[1]), s.A[1][0], nb, e0, ne, emb, W0);
    fence_proxy_async();
    warp_arrive(&s.h1_ready);
    TCDBG(1, 0);
    mbar_wait(&s.z2_done, 0);
    TCDBG(2, 0);
    tc_fence_after();
    {
      // h2 = c silu(z2) -> B operand (rows = edges, K = hidden unit); the W1^T operand in B[0] is consumed
      float z[16];
      read_64x64(tmem, q, cg, z);
      const int m = 32 * q + lane;
#pragma unroll
      for (int j4 = 0; j4 < 4; ++j4) {
        float4 v, hi, lo;
        v.x = kSilu2Mom * silu_f(z[4 * j4 + 0]); v.y = kSilu2Mom * silu_f(z[4 * j4 + 1]);
        v.z = kSilu2Mom * silu_f(z[4 * j4 + 2]); v.w = kSilu2Mom * silu_f(z[4 * j4 + 3]);
        split4(v, hi, lo);
        *reinterpret_cast<float4*>(s.B[0] + (cg * 4 + j4) * CHB + m * 16) = hi;
        *reinterpret_cast<float4*>(s.B[1] + (cg * 4 + j4) * CHB + m * 16) = lo;
      }
    }
    fence_proxy_async();
    tc_fence_before();
    warp_arrive(&s.h2_ready);
    TCDBG(3, 0);
    // epilogue: lanes = channels, columns = edges; warp (q, cg) drains channels 32q.. of edges 32cg..
    const int col = cg * 32;
    float* wb = w + (size_t)e0 * wn;           // CTA-uniform base, 32-bit offsets below
    for (int mt = 0; mt < NT; ++mt) {
      const int st = mt & 1;
      const int ch = slab_of(mt) * TCN + 32 * q + lane;
      TCDBG(4 + 3 * mt, 0);
      mbar_wait(&s.mma_done[st], (mt >> 1) & 1);
      TCDBG(5 + 3 * mt, 0);
      tc_fence_after();
      float v[32], c[32];
      const uint32_t ta = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(st * 256 + col);
      tmem_ld32(ta, v);
      tmem_ld32(ta + 128, c);
      tc_fence_before();
      warp_arrive(&s.tmem_empty[st]);
      TCDBG(6 + 3 * mt, 0);
      if (ch < wn) {
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
  }
  TCDBG(30, 0);
  TCCTA(1);
  tc_fence_before();
  __syncthreads();
  if (warp == 16) tmem_dealloc(tmem, 512);
}

// ---- backward (data gradient): demb (+)= d/d emb sum_e <dw[e], w[e]> -------------------------------------
// dh2[e, h] = sum_c dw[e, c] W2[h, c] / 8 is the big GEMM: A = dw tile (rows = edges, K = channel),
// B = W2 (rows = hidden unit, K = channel; prepacked, bulk-copied), D = 128 lanes (edges) x 64 columns.
// K is streamed in chunks of 32 channels.  Every worker keeps its two float4 of the next PD chunks in registers
// (the only HBM stream of the kernel: PD x 16 KB in flight per SM), splits them into tf32 hi/lo and stores them
// straight into the operand ring in the K-major SWIZZLE_128B layout (a row's 32 channels are its 128-byte swizzle
// atom), so that the eight 16-byte units a quarter-warp writes for one row fall into distinct banks and the tensor
// core reads the operand at full rate (a padded no-swizzle layout costs 2.7x in MMA time).  W2 is the stacked hi|lo operand (N = 128), so a chunk
// costs two MMAs per K-step: dw.hi goes to three round-robin accumulators, dw.lo to a fourth; the workers add the
// partial sums in fp32 (round-to-nearest).  These sums feed forces only (1e-5 relative is the gate), so 40 MMAs per
// accumulator are acceptable here, unlike in the forward kernel.  z2 (prologue) and dh1 = dz2 . W1^T (epilogue) are 64 x 64 tensor-core
// products as in the forward kernel.
constexpr int NACC = 3;                // main accumulators (128 TMEM columns each: dw.hi . [W2.hi | W2.lo])
constexpr int APART = TCN * BK * 4;    // one part (hi or lo) of an A stage: 128 rows x 128 B, SWIZZLE_128B
constexpr int BST = 3;                 // operand ring depth
constexpr int PD = 6;                  // chunks of dw held in registers ahead of the split

struct TcBwdSmem {
  uint8_t A[BST][2][APART];            // [stage][hi/lo] dw chunk (97 KB); h1 / dz2 operands outside the main loop
  uint8_t B[BST][2][BOPB];             // [stage][hi/lo] W2 chunk (48 KB); W1^T / W1 operands outside the main loop
  float embT[TC_MAX_NB][TCN];          // 8 KB
  float D2[TCH][TCN];                  // 32 KB: c*silu'(z2), later the partial demb sums
  uint64_t fullB[BST], a_ready[BST], mma_done[BST];
  uint64_t w1_full, h1_ready, z2_done, dz2_ready, dh1_done;
  uint32_t tmem_base;
};
static_assert(BST * 2 * APART >= 2 * OPB && BST * 2 * BOPB >= 2 * W1B, "operand rings double as hidden-layer operands");

__global__ void __launch_bounds__(TCT, 1)
radial_mlp_bwd_tc_kernel(int nb, int wn, long long E, const float* __restrict__ emb, const float* __restrict__ W0,
                         const float* __restrict__ pack, const float* __restrict__ pw, const float* __restrict__ dw,
                         float* __restrict__ demb, int accumulate) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  TcBwdSmem& s = *reinterpret_cast<TcBwdSmem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * TCN;
  const int ne = (int)min((long long)TCN, E - e0);
  const int NK = (wn + BK - 1) / BK;
  constexpr uint32_t BSLAB = 2 * BOPB;
  uint8_t* const opA = &s.A[0][0][0];   // [128 x 64] hi | lo operand (h1, later dz2)
  uint8_t* const opB = &s.B[0][0][0];   // [ 64 x 64] hi | lo operand (W1^T, later W1)

  TCCTA(0);
  if (warp == 16) tmem_alloc(&s.tmem_base, 512);
  if (threadIdx.x == 0) {
    for (int k = 0; k < BST; ++k) { mbar_init(&s.fullB[k], 1); mbar_init(&s.a_ready[k], 16); mbar_init(&s.mma_done[k], 1); }
    mbar_init(&s.w1_full, 1); mbar_init(&s.z2_done, 1); mbar_init(&s.dh1_done, 1);
    mbar_init(&s.h1_ready, 16); mbar_init(&s.dz2_ready, 16);
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;
  const int last_st = (NK - 1) % BST, last_par = ((NK - 1) / BST) & 1;
  // every CTA walks the K chunks in its own rotation (see the forward kernel)
  const int rot = blockIdx.x % NK;
  auto chunk_of = [&](int i) { const int c = i + rot; return c >= NK ? c - NK : c; };

  if (warp == 16) {
    if (elect_one()) {
      mbar_expect_tx(&s.w1_full, 2 * W1B);
      bulk_copy_g2s(opB, pw, 2 * W1B, &s.w1_full);
      mbar_wait(&s.w1_full, 0);
      mbar_wait(&s.h1_ready, 0);
      tc_fence_after();
      issue_64x64(tmem, smem_u32(opA), smem_u32(opB));
      mma_commit(&s.z2_done);
      mbar_wait(&s.z2_done, 0);            // both rings are free now
      TCDBG(40, 512);
      for (int j = 0; j < 2 && j < NK; ++j) {
        mbar_expect_tx(&s.fullB[j], BSLAB);
        bulk_copy_g2s(s.B[j][0], pack + (size_t)chunk_of(j) * (BSLAB / 4), BSLAB, &s.fullB[j]);
      }
      constexpr uint32_t idesc = make_idesc_tf32(128, 128), idesc_lo = make_idesc_tf32(128, TCH);
      for (int i = 0; i < NK; ++i) {
        const int st = i % BST;
        if (i < 8) TCDBG(44 + 4 * i, 512);
        mbar_wait(&s.fullB[st], (i / BST) & 1);
        if (i < 8) TCDBG(45 + 4 * i, 512);
        mbar_wait(&s.a_ready[st], (i / BST) & 1);
        if (i < 8) TCDBG(46 + 4 * i, 512);
        if (blockIdx.x == 1000 && i >= 8 && i < 16) g_tc_dbg[64 + i] = clock64();
        tc_fence_after();
        const uint32_t dm = tmem + (uint32_t)(i % NACC) * 128, dc = tmem + (uint32_t)NACC * 128;
        const uint32_t ahi = smem_u32(s.A[st][0]), alo = ahi + APART;
        const uint32_t bb = smem_u32(s.B[st][0]);
#pragma unroll
        for (int kk = 0; kk < BK / 8; ++kk) {
          const uint64_t dah = make_desc_sw128(ahi + kk * 32), dal = make_desc_sw128(alo + kk * 32);
          const uint64_t db = make_desc(bb + kk * 2 * CHB, CHB, 128);
          mma_tf32(dc, dal, db, idesc_lo, (i > 0 || kk > 0) ? 1u : 0u);   // dw.lo . W2.hi only (rows 0..63 of the stack)
          mma_tf32(dm, dah, db, idesc, (i >= NACC || kk > 0) ? 1u : 0u);
        }
        mma_commit(&s.mma_done[st]);
        if (i < 8) TCDBG(47 + 4 * i, 512);
        if (blockIdx.x == 1000 && i >= 8 && i < 16) { mbar_wait(&s.mma_done[st], (i / BST) & 1); TCDBG(80 + i, 512); g_tc_dbg[96 + i] = clock64(); }
        if (blockIdx.x == 1000 && i >= 8 && i < 16) TCDBG(104 + i, 512);
        if (i + 2 < NK) {                  // W2 chunk i+2 goes where chunk i-1 was
          const int bs = (i + 2) % BST;
          if (i >= 1) mbar_wait(&s.mma_done[bs], ((i - 1) / BST) & 1);
          mbar_expect_tx(&s.fullB[bs], BSLAB);
          bulk_copy_g2s(s.B[bs][0], pack + (size_t)chunk_of(i + 2) * (BSLAB / 4), BSLAB, &s.fullB[bs]);
        }
      }
      TCDBG(41, 512);
      mbar_wait(&s.mma_done[last_st], last_par);
      TCDBG(42, 512);
      mbar_expect_tx(&s.w1_full, 2 * W1B);
      bulk_copy_g2s(opB, pw + 2 * W1B / 4, 2 * W1B, &s.w1_full);
      mbar_wait(&s.w1_full, 1);
      mbar_wait(&s.dz2_ready, 0);
      tc_fence_after();
      issue_64x64(tmem, smem_u32(opA), smem_u32(opB));
      mma_commit(&s.dh1_done);
    }
  } else {
    const int q = warp & 3, cg = warp >> 2, m = 32 * q + lane;
    // this thread's share of a chunk: rows (tid>>3) and (tid>>3)+64, 16-byte column tid&7 (coalesced 128-byte rows)
    const int lrow = threadIdx.x >> 3, lc4 = threadIdx.x & 7;
    const float* dwp = dw + (size_t)e0 * wn + 4 * lc4;
    auto gload = [&](int i, float4 (&buf)[2]) {
      const int c = i < NK ? chunk_of(i) : 0;
      const int k = c * BK + 4 * lc4;
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int row = lrow + 64 * r;
        buf[r] = (i < NK && row < ne && k < wn) ? __ldcs(reinterpret_cast<const float4*>(dwp + (size_t)row * wn + c * BK))
                                                : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    };
    TCDBG(0, 0);
    hidden1_operand(s.embT, opA, nb, e0, ne, emb, W0);
    fence_proxy_async();
    warp_arrive(&s.h1_ready);
    // the first chunks of dw travel while the issuer runs the z2 product (about 4 k clocks until they are needed)
    float4 pf0[2], pf1[2], pf2[2], pf3[2], pf4[2], pf5[2];
    gload(0, pf0); gload(1, pf1); gload(2, pf2); gload(3, pf3); gload(4, pf4); gload(5, pf5);
    TCDBG(1, 0);
    mbar_wait(&s.z2_done, 0);
    TCDBG(2, 0);
    tc_fence_after();
    {
      float z[16];
      read_64x64(tmem, q, cg, z);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float v, d;
        silu_vd(z[j], v, d);
        s.D2[cg * 16 + j][m] = d;
      }
      tc_fence_before();
    }
    TCDBG(3, 0);
    auto step = [&](int i, float4 (&buf)[2]) {
      const int st = i % BST;
      if (i < 8) TCDBG(4 + 3 * i, 0);
      const float4 v0 = buf[0], v1 = buf[1];
      if (i < 8) TCDBG(5 + 3 * i, 0);
      if (i >= BST) mbar_wait(&s.mma_done[st], (i / BST - 1) & 1);   // stage last read by the MMAs of chunk i-3
      if (i < 8) TCDBG(6 + 3 * i, 0);
      float4 hi, lo;
      uint8_t* a = s.A[st][0] + lrow * 128 + ((lc4 ^ (lrow & 7)) << 4);
      split4(v0, hi, lo);
      *reinterpret_cast<float4*>(a) = hi;
      *reinterpret_cast<float4*>(a + APART) = lo;
      split4(v1, hi, lo);
      *reinterpret_cast<float4*>(a + 64 * 128) = hi;
      *reinterpret_cast<float4*>(a + 64 * 128 + APART) = lo;
      fence_proxy_async();
      warp_arrive(&s.a_ready[st]);
    };
    // the loads of three consecutive chunks (384 contiguous bytes of every row) are issued together: requests to one
    // DRAM page arrive close in time instead of one chunk period apart
    for (int i = 0; i < NK; i += PD) {
      if (i > 0) { gload(i + 3, pf3); gload(i + 4, pf4); gload(i + 5, pf5); }
      step(i, pf0);
      if (i + 1 < NK) step(i + 1, pf1);
      if (i + 2 < NK) step(i + 2, pf2);
      gload(i + 6, pf0); gload(i + 7, pf1); gload(i + 8, pf2);
      if (i + 3 < NK) step(i + 3, pf3);
      if (i + 4 < NK) step(i + 4, pf4);
      if (i + 5 < NK) step(i + 5, pf5);
    }
    // all MMAs retire in order: the last commit covers every chunk
    TCDBG(31, 0);
    mbar_wait(&s.mma_done[last_st], last_par);
    TCDBG(32, 0);
    tc_fence_after();
    {
      // dz2[m][h] = dh2[m][h] * c*silu'(z2[m][h]) as the A operand of dh1 = dz2 . W1^T
      const uint32_t ta = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(cg * 16);
      float v[16], t[16];
      tmem_ld16(ta + NACC * 128, v);                   // dw.lo . W2.hi
      const int np = NK < NACC ? NK : NACC;
      for (int p = 0; p < 2 * np; ++p) {               // dw.hi . W2.lo, then the main terms
        const uint32_t col = p < np ? p * 128 + TCH : (p - np) * 128;
        tmem_ld16(ta + col, t);
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] += t[j];
      }
#pragma unroll
      for (int j4 = 0; j4 < 4; ++j4) {
        float4 x, hi, lo;
        x.x = v[4 * j4 + 0] * s.D2[cg * 16 + 4 * j4 + 0][m]; x.y = v[4 * j4 + 1] * s.D2[cg * 16 + 4 * j4 + 1][m];
        x.z = v[4 * j4 + 2] * s.D2[cg * 16 + 4 * j4 + 2][m]; x.w = v[4 * j4 + 3] * s.D2[cg * 16 + 4 * j4 + 3][m];
        split4(x, hi, lo);
        *reinterpret_cast<float4*>(opA + (cg * 4 + j4) * CHB + m * 16) = hi;
        *reinterpret_cast<float4*>(opA + OPB + (cg * 4 + j4) * CHB + m * 16) = lo;
      }
    }
    fence_proxy_async();
    tc_fence_before();
    warp_arrive(&s.dz2_ready);
    TCDBG(33, 0);
    // z1 of this thread's 16 hidden units (K = nb), overlapped with the dh1 MMAs
    const float s0 = 1.0f / sqrtf((float)nb);
    float z1[16] = {};
    for (int k = 0; k < nb; ++k) {
      const float a = s.embT[k][m];
#pragma unroll
      for (int j4 = 0; j4 < 4; ++j4) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(W0 + k * TCH + cg * 16 + 4 * j4));
        z1[4 * j4 + 0] = fmaf(a, b.x, z1[4 * j4 + 0]); z1[4 * j4 + 1] = fmaf(a, b.y, z1[4 * j4 + 1]);
        z1[4 * j4 + 2] = fmaf(a, b.z, z1[4 * j4 + 2]); z1[4 * j4 + 3] = fmaf(a, b.w, z1[4 * j4 + 3]);
      }
    }
    TCDBG(34, 0);
    mbar_wait(&s.dh1_done, 0);
    TCDBG(35, 0);
    tc_fence_after();
    {
      // dz1 = dh1 * c*silu'(z1);  partial demb[m][n] = sum over this warp's 16 hidden units of dz1 W0[n][.] / sqrt(nb)
      float dh1[16];
      read_64x64(tmem, q, cg, dh1);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float v, d;
        silu_vd(z1[j] * s0, v, d);
        dh1[j] *= d;
      }
      // D2 is consumed: every warp read it before dz2_ready, which dh1_done follows
      float (*part)[TC_MAX_NB][TCN] = reinterpret_cast<float(*)[TC_MAX_NB][TCN]>(&s.D2[0][0]);
      for (int n = 0; n < nb; ++n) {
        float t = 0.f;
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(W0 + n * TCH + cg * 16 + 4 * j4));
          t = fmaf(dh1[4 * j4 + 0], b.x, t); t = fmaf(dh1[4 * j4 + 1], b.y, t);
          t = fmaf(dh1[4 * j4 + 2], b.z, t); t = fmaf(dh1[4 * j4 + 3], b.w, t);
        }
        part[cg][n][m] = t * s0;
      }
      bar_sync(1, TCW);
      for (int it = threadIdx.x; it < TCN * nb; it += TCW) {
        const int mm = it & 127, n = it >> 7;
        if (mm >= ne) continue;
        const float t = (part[0][n][mm] + part[1][n][mm]) + (part[2][n][mm] + part[3][n][mm]);
        float* p = demb + (size_t)(e0 + mm) * nb + n;
        *p = accumulate ? (*p + t) : t;
      }
    }
  }
  TCDBG(30, 0);
  TCCTA(1);
  tc_fence_before();
  __syncthreads();
  if (warp == 16) tmem_dealloc(tmem, 512);
}

extern "C" void b200_debug_tc_cta(long long* out, int n) { cudaMemcpyFromSymbol(out, g_tc_cta, sizeof(long long) * n); }
extern "C" void b200_debug_tc(long long* out, int n) { cudaMemcpyFromSymbol(out, g_tc_dbg, sizeof(long long) * n); }

bool radial_tc_supported(const MlpDesc& d) {
  return d.h1 == TCH && d.h2 == TCH && d.nb >= 1 && d.nb <= TC_MAX_NB && d.wn >= 4 && (d.wn % 4) == 0;
}

// operand slabs of W1 / W2 for one call (stream-ordered scratch: prepack -> kernel -> free)
struct Pack {
  float *base, *fwd, *bwd, *w1;
};
static int with_pack(const MlpDesc& d, const float* W1, const float* W2, cudaStream_t st, Pack& p) {
  const int NT = (d.wn + TCN - 1) / TCN, NK = (d.wn + BK - 1) / BK;
  const size_t nf = (size_t)NT * 2 * OPB, nbk = (size_t)NK * 2 * BOPB, nw = (size_t)4 * W1B;
  keep_async_pool();
  if (cudaMallocAsync((void**)&p.base, nf + nbk + nw, st) != cudaSuccess) {
    set_error("radial_mlp: cudaMallocAsync of the operand pack failed");
    return B200_ERR_CUDA;
  }
  p.fwd = p.base;
  p.bwd = p.base + nf / 4;
  p.w1 = p.bwd + nbk / 4;
  radial_prepack_kernel<<<148, 256, 0, st>>>(d.wn, NT, NK, W1, W2, p.fwd, p.bwd, p.w1);
  return B200_OK;
}

int radial_mlp_fwd_tc(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                      float* w, cudaStream_t st) {
  if (E == 0) return B200_OK;
  Pack p;
  if (int rc = with_pack(d, W1, W2, st, p)) return rc;
  cudaFuncSetAttribute(radial_mlp_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TcFwdSmem));
  radial_mlp_fwd_tc_kernel<<<(unsigned)((E + TCN - 1) / TCN), TCT, sizeof(TcFwdSmem), st>>>(d.nb, d.wn, E, emb, W0, p.fwd, p.w1, w);
  cudaFreeAsync(p.base, st);
  return check_launch("radial_mlp_fwd_tc", 2);
}

int radial_mlp_bwd_tc(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                      const float* dw, float* demb, int accumulate, cudaStream_t st) {
  if (E == 0) return B200_OK;
  Pack p;
  if (int rc = with_pack(d, W1, W2, st, p)) return rc;
  cudaFuncSetAttribute(radial_mlp_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TcBwdSmem) + 1024);
  radial_mlp_bwd_tc_kernel<<<(unsigned)((E + TCN - 1) / TCN), TCT, sizeof(TcBwdSmem) + 1024, st>>>(d.nb, d.wn, E, emb, W0, p.bwd, p.w1,
                                                                                             dw, demb, accumulate);
  cudaFreeAsync(p.base, st);
  return check_launch("radial_mlp_bwd_tc", 2);
}

}  // namespace b200
