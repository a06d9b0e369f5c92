// Fused convolution kernels: radial MLP output layer on tcgen05 -> per-edge weights in TMEM -> channel-wise
// (uvu) Clebsch-Gordan product -> segment-reduced scatter, forward and backward, without the per-edge
// weight tensor [E, weight_numel] ever reaching HBM.
//
// Reference semantics: sevenn/nn/convolution.py:105-121
//   weight  = weight_nn(edge_embedding)                      (e3nn FullyConnectedNet [nb, 64, 64, weight_numel])
//   message = TensorProduct_uvu(x[edge_src], edge_attr, weight);  x = scatter_sum(message, edge_dst) / denominator
// and, for the reverse pass (ForceStressOutputFromEdge, nn/force_output.py:158-215), dE/d edge_attr, dE/d r and
// dE/d x[edge_src].
//
// Work split (one CTA per tile of edges, edges sorted by destination):
//   * `fused_radial_kernel` (CUDA cores): the two hidden layers of the radial MLP for every edge, value h2 and
//     tangent h2' = d h2 / d r (forward mode: the MLP input depends on the edge only through r), written as
//     ready-to-use tcgen05 operand tiles (tf32 hi/lo split, K-major canonical layout, 32 edges per tile).
//   * `fused_conv_fwd_kernel`, 64 edges per CTA: W2^T is streamed in slabs of 128 rows (cp.async.bulk, hi/lo
//     prepacked, two shared-memory stages) and D[row, edge] = W2^T_slab . h2^T (3xTF32) is accumulated for ALL slabs
//     into TMEM (8 slabs x 64 columns = the 512 columns of an SM).  The slab rows are permuted so that a TMEM lane
//     (= thread of an epilogue warp) holds, slab after slab, every path weight of ONE channel u of one input irrep:
//     the thread gathers x[src, u] once per edge (coalesced over the warp), reads its weights with tcgen05.ld,
//     applies the sparse CG contractions of all its paths and keeps the partial sums of the current destination row
//     in registers; rows are flushed with coalesced RED.ADD (out is zeroed first).
//   * `fused_conv_bwd_kernel`, 32 edges per CTA, D = [w | w'] (w' = d w / d r from the tangent tile).  Per edge and
//     channel the thread forms d x[src, u] (complete over its paths: one RED per component) and its share of
//     d r = sum (d w) w' and d Y = sum w T, which are summed over the warp and accumulated per edge in shared memory.
#include "common.cuh"
#include "tc.cuh"

namespace b200 {

using namespace tc;
#include "cg_direct.cuh"

constexpr int FT = 32;                  // edges per operand tile
constexpr int FH = 64;                  // hidden width = K of the GEMM
constexpr int F_ACH = 128 * 16;         // bytes of one 16-byte K-chunk column of a 128-row A slab
constexpr int F_BCH = FT * 16;          // ... of a 32-row B tile
constexpr int F_APART = (FH / 4) * F_ACH;   // 32 KB: one part (hi or lo) of an A slab
constexpr int F_BPART = (FH / 4) * F_BCH;   // 8 KB: one part of a B tile
constexpr int F_TILE_FLOATS = 4 * F_BPART / 4;   // h2 hi | h2 lo | h2' hi | h2' lo
constexpr int F_MAXSLAB = 8;            // 8 slabs x 64 TMEM columns
constexpr int F_NF = 64;                // edges per forward CTA
constexpr int F_NB = 32;                // edges per backward CTA
constexpr int F_SLABCOLS = 64;          // TMEM columns per slab (fwd: 64 edges; bwd: 32 x {w, w'})
constexpr int F_WARPS = 16;
constexpr int F_THREADS = F_WARPS * 32;
constexpr int F_MAXJOB = 6;

// A job: one thread-quarter (32 TMEM lanes = 32 channels ubase..ubase+31 of input irrep `role`) and the subset `mask`
// of that irrep's paths whose weights sit in consecutive slabs starting at first_slab (bit i = i-th path of the role).
struct FusedJob {
  int first_slab, role, ubase, mask;
};
struct FusedPlan {
  int nslab;
  int njob[4];
  FusedJob job[4][F_MAXJOB];
  signed char row_path[F_MAXSLAB][4];   // path id held by the 32 rows (slab, quarter), -1: zero rows
  short row_ubase[F_MAXSLAB][4];
};

static bool build_plan(const ConvDesc& d, FusedPlan& pl) {
  if (d.m0 % 32 || d.m1 % 32 || d.m2 % 32 || d.m0 > 128 || d.m1 > 64 || d.m2 > 32) return false;
  static const int L1[kNumPaths] = {0, 0, 0, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2};
  // rough relative cost of a path per (channel, edge): balances the four thread-quarters
  static const int COST[kNumPaths] = {1, 1, 1, 2, 2, 3, 5, 5, 7, 3, 5, 7, 3, 7, 10};
  struct Group { int role, ubase, np, pid[6], bit[6]; };
  Group groups[16];
  int ng = 0, total_cost = 0, total_units = 0;
  const int mul[3] = {d.m0, d.m1, d.m2};
  for (int role = 2; role >= 0; --role)          // heavy groups first
    for (int ub = 0; ub < mul[role]; ub += 32) {
      Group& g = groups[ng];
      g.role = role; g.ubase = ub; g.np = 0;
      int bit = 0;
      for (int p = 0; p < kNumPaths; ++p) {
        if (L1[p] != role || d.woff[p] < 0) continue;
        g.pid[g.np] = p; g.bit[g.np] = bit++; ++g.np;
        total_cost += COST[p]; ++total_units;
      }
      if (g.np) ++ng;
    }
  if (total_units == 0 || total_units > 4 * F_MAXSLAB) return false;
  const int cap = F_MAXSLAB;
  const int target = (total_cost + 3) / 4 + 1;
  int qcost[4] = {0, 0, 0, 0}, qslab[4] = {0, 0, 0, 0};
  for (int q = 0; q < 4; ++q) pl.njob[q] = 0;
  for (int s = 0; s < F_MAXSLAB; ++s)
    for (int q = 0; q < 4; ++q) { pl.row_path[s][q] = -1; pl.row_ubase[s][q] = 0; }
  for (int gi = 0; gi < ng; ++gi) {
    Group& g = groups[gi];
    int left = g.np;       // paths are handed out from the last (most expensive) to the first
    while (left > 0) {
      int q = -1;
      for (int k = 0; k < 4; ++k)
        if (qslab[k] < cap && (q < 0 || qcost[k] < qcost[q])) q = k;
      if (q < 0) return false;
      int take = 0, cost = 0;
      while (take < left && qslab[q] + take < cap) {
        const int c = COST[g.pid[left - 1 - take]];
        if (take > 0 && qcost[q] + cost + c > target) break;
        cost += c; ++take;
      }
      if (pl.njob[q] >= F_MAXJOB) return false;
      FusedJob& job = pl.job[q][pl.njob[q]++];
      job.first_slab = qslab[q]; job.role = g.role; job.ubase = g.ubase; job.mask = 0;
      for (int k = left - take; k < left; ++k) {   // ascending path order inside the job
        job.mask |= 1 << g.bit[k];
        pl.row_path[qslab[q]][q] = (signed char)g.pid[k];
        pl.row_ubase[qslab[q]][q] = (short)g.ubase;
        ++qslab[q];
      }
      qcost[q] += cost;
      left -= take;
    }
  }
  pl.nslab = 0;
  for (int q = 0; q < 4; ++q) pl.nslab = qslab[q] > pl.nslab ? qslab[q] : pl.nslab;
  return true;
}

// ---- operand pack of W2 for a plan: [slab][hi|lo][16 chunks][128 rows][4], scaled by 1/sqrt(64) ------------
__global__ void fused_prepack_kernel(FusedPlan pl, ConvDesc d, const float* __restrict__ W2, float* __restrict__ pack) {
  const int n = pl.nslab * 16 * 128;
  const float sc = 1.0f / sqrtf((float)FH);
  for (int it = blockIdx.x * blockDim.x + threadIdx.x; it < n; it += gridDim.x * blockDim.x) {
    const int r = it & 127, c4 = (it >> 7) & 15, s = it >> 11;
    const int p = pl.row_path[s][r >> 5];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p >= 0) {
      const int ch = d.woff[p] + pl.row_ubase[s][r >> 5] + (r & 31);
      v.x = W2[(size_t)(4 * c4 + 0) * d.wn + ch] * sc; v.y = W2[(size_t)(4 * c4 + 1) * d.wn + ch] * sc;
      v.z = W2[(size_t)(4 * c4 + 2) * d.wn + ch] * sc; v.w = W2[(size_t)(4 * c4 + 3) * d.wn + ch] * sc;
    }
    float4 hi, lo;
    split4(v, hi, lo);
    float* dst = pack + (size_t)s * (2 * F_APART / 4) + (size_t)c4 * (F_ACH / 4) + r * 4;
    *reinterpret_cast<float4*>(dst) = hi;
    *reinterpret_cast<float4*>(dst + F_APART / 4) = lo;
  }
}

// ---- hidden layers of the radial MLP, value and d/dr tangent, as operand tiles -------------------------------
// radial basis and its r-derivative (sevenn/nn/edge_embedding.py:114-116 BesselBasis times the cutoff envelope)
__device__ __forceinline__ void cutoff_vd_f(const EdgeParams& p, float r, float& f, float& df) {
  if (p.cutoff_kind == B200_CUTOFF_XPLOR) {
    const float r2 = r * r, ron2 = p.r_on * p.r_on, rc2 = p.r_cut * p.r_cut;
    if (r < p.r_on) {
      f = 1.0f;
      df = 0.0f;
    } else {
      const float den = (rc2 - ron2) * (rc2 - ron2) * (rc2 - ron2);
      const float a = rc2 - r2;
      f = a * a * (rc2 + 2.0f * r2 - 3.0f * ron2) / den;
      df = 12.0f * r * a * (ron2 - r2) / den;
    }
  } else {
    const float x = r / p.r_cut, pp = p.poly_p;
    const float c0 = (pp + 1.0f) * (pp + 2.0f) * 0.5f, c1 = pp * (pp + 2.0f), c2 = pp * (pp + 1.0f) * 0.5f;
    const float xp1 = powf(x, pp - 1.0f), xp = xp1 * x, xq = xp * x, xr = xq * x;
    f = 1.0f - c0 * xp + c1 * xq - c2 * xr;
    df = (-c0 * pp * xp1 + c1 * (pp + 1.0f) * xp - c2 * (pp + 2.0f) * xq) / p.r_cut;
  }
}

constexpr int FR_E = 64;    // edges per CTA of the radial kernel (two operand tiles)
constexpr int FR_T = 256;   // threads: 64 edges x 4 groups of 16 hidden units
constexpr int FR_MAXNB = 16;

__global__ void __launch_bounds__(FR_T)
fused_radial_kernel(EdgeParams ep, int nb, long long E, const float* __restrict__ vec, const float* __restrict__ coeffs,
                    const float* __restrict__ W0, const float* __restrict__ W1, float* __restrict__ tiles) {
  extern __shared__ __align__(16) float fr_smem[];
  float* sW1 = fr_smem;                                                         // [FH][FH]
  float* sW0 = sW1 + FH * FH;                                                   // [nb][FH]
  float (*sE)[FR_MAXNB][FR_E] = reinterpret_cast<float (*)[FR_MAXNB][FR_E]>(sW0 + FR_MAXNB * FH);   // basis value / d basis/dr
  float (*sH)[FH][FR_E] = reinterpret_cast<float (*)[FH][FR_E]>(sW0 + FR_MAXNB * FH + 2 * FR_MAXNB * FR_E);   // h1 / h1'
  const int tid = threadIdx.x, e = tid & 63, hg = tid >> 6;
  const long long e0 = (long long)blockIdx.x * FR_E;
  for (int i = tid; i < nb * FH; i += FR_T) sW0[i] = W0[i];
  for (int i = tid; i < FH * FH; i += FR_T) sW1[i] = W1[i];
  if (tid < FR_E) {
    float r = 1.0f;
    const bool ok = e0 + tid < E;
    if (ok) {
      const float vx = vec[(e0 + tid) * 3], vy = vec[(e0 + tid) * 3 + 1], vz = vec[(e0 + tid) * 3 + 2];
      r = sqrtf(vx * vx + vy * vy + vz * vz);
    }
    float f, df;
    cutoff_vd_f(ep, r, f, df);
    const float pre = 2.0f / ep.r_cut;
    for (int n = 0; n < nb; ++n) {
      const float k = coeffs[n];
      float s, c;
      sincosf(k * r, &s, &c);
      const float b = pre * s / r, db = pre * (k * c * r - s) / (r * r);
      sE[0][n][tid] = ok ? b * f : 0.f;
      sE[1][n][tid] = ok ? db * f + b * df : 0.f;
    }
  }
  __syncthreads();
  const float s0 = 1.0f / sqrtf((float)nb), s1 = 1.0f / sqrtf((float)FH);
  {
    float z[16], dz[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { z[j] = 0.f; dz[j] = 0.f; }
    for (int k = 0; k < nb; ++k) {
      const float a = sE[0][k][e], da = sE[1][k][e];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float wv = sW0[k * FH + hg * 16 + j];
        z[j] = fmaf(a, wv, z[j]);
        dz[j] = fmaf(da, wv, dz[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float v, dv;
      silu_vd(z[j] * s0, v, dv);
      sH[0][hg * 16 + j][e] = v;
      sH[1][hg * 16 + j][e] = dv * dz[j] * s0;
    }
  }
  __syncthreads();
  float z[16], dz[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) { z[j] = 0.f; dz[j] = 0.f; }
#pragma unroll 4
  for (int k = 0; k < FH; ++k) {
    const float a = sH[0][k][e], da = sH[1][k][e];
    const float4* wr = reinterpret_cast<const float4*>(&sW1[k * FH + hg * 16]);
#pragma unroll
    for (int j4 = 0; j4 < 4; ++j4) {
      const float4 wv = wr[j4];
      z[4 * j4 + 0] = fmaf(a, wv.x, z[4 * j4 + 0]); dz[4 * j4 + 0] = fmaf(da, wv.x, dz[4 * j4 + 0]);
      z[4 * j4 + 1] = fmaf(a, wv.y, z[4 * j4 + 1]); dz[4 * j4 + 1] = fmaf(da, wv.y, dz[4 * j4 + 1]);
      z[4 * j4 + 2] = fmaf(a, wv.z, z[4 * j4 + 2]); dz[4 * j4 + 2] = fmaf(da, wv.z, dz[4 * j4 + 2]);
      z[4 * j4 + 3] = fmaf(a, wv.w, z[4 * j4 + 3]); dz[4 * j4 + 3] = fmaf(da, wv.w, dz[4 * j4 + 3]);
    }
  }
  float* tile = tiles + (size_t)(2 * blockIdx.x + (e >> 5)) * F_TILE_FLOATS;   // two operand tiles of 32 edges per CTA
  const int row = e & 31;
  const bool ok = e0 + e < E;
#pragma unroll
  for (int j4 = 0; j4 < 4; ++j4) {
    float v[4], t[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float dv;
      silu_vd(z[4 * j4 + j] * s1, v[j], dv);
      t[j] = dv * dz[4 * j4 + j] * s1;
      if (!ok) { v[j] = 0.f; t[j] = 0.f; }
    }
    float4 hi, lo;
    const int chunk = hg * 4 + j4;
    float* base = tile + chunk * (F_BCH / 4) + row * 4;
    split4(make_float4(v[0], v[1], v[2], v[3]), hi, lo);
    *reinterpret_cast<float4*>(base) = hi;
    *reinterpret_cast<float4*>(base + F_BPART / 4) = lo;
    split4(make_float4(t[0], t[1], t[2], t[3]), hi, lo);
    *reinterpret_cast<float4*>(base + 2 * (F_BPART / 4)) = hi;
    *reinterpret_cast<float4*>(base + 3 * (F_BPART / 4)) = lo;
  }
}

// ---- shared pieces of the two convolution kernels ---------------------------------------------------------------
// whole warp: 8 consecutive fp32 columns of this warp's 32 TMEM lanes
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
  uint32_t r[8];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

// D[128 rows, 64 edges] (+)= A_slab . B_tile^T as 3xTF32 (lo*hi + hi*lo + hi*hi into one fp32 accumulator)
__device__ __forceinline__ void issue_slab_mma(uint32_t d_tmem, const uint8_t* a_hi, const uint8_t* b_hi) {
  constexpr uint32_t idesc = make_idesc_tf32(128, FT);
  const uint32_t ah = smem_u32(a_hi), al = ah + F_APART, bh = smem_u32(b_hi), bl = bh + F_BPART;
#pragma unroll
  for (int kk = 0; kk < FH / 8; ++kk) {
    const uint64_t dah = make_desc(ah + kk * 2 * F_ACH, F_ACH, 128), dal = make_desc(al + kk * 2 * F_ACH, F_ACH, 128);
    const uint64_t dbh = make_desc(bh + kk * 2 * F_BCH, F_BCH, 128), dbl = make_desc(bl + kk * 2 * F_BCH, F_BCH, 128);
    mma_tf32(d_tmem, dal, dbh, idesc, kk > 0);
    mma_tf32(d_tmem, dah, dbl, idesc, 1);
    mma_tf32(d_tmem, dah, dbh, idesc, 1);
  }
}

__device__ __forceinline__ int path_l1(int p) { return p < 3 ? 0 : (p < 9 ? 1 : 2); }
__device__ __forceinline__ int conv_xoff(const ConvDesc& d, int l1) { return l1 == 0 ? 0 : (l1 == 1 ? d.m0 : d.m0 + 3 * d.m1); }


template <int... PIDS> struct PL;
// ---- TMEM loads: issue a batch, then one wait -------------------------------------------------------------------
template <int N> __device__ __forceinline__ void tmem_ld_nowait(uint32_t taddr, float* v);
template <> __device__ __forceinline__ void tmem_ld_nowait<4>(uint32_t taddr, float* v) {
  uint32_t r[4];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(r[i]);
}
template <> __device__ __forceinline__ void tmem_ld_nowait<2>(uint32_t taddr, float* v) {
  uint32_t r[2];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(taddr) : "memory");
  v[0] = __uint_as_float(r[0]);
  v[1] = __uint_as_float(r[1]);
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// bit j: edge e_lo + j starts a new destination row within this warp's range (bit 0 always set)
__device__ __forceinline__ unsigned break_mask(const int* dst, int e_lo, int e_hi) {
  const int lane = threadIdx.x & 31, e = e_lo + lane;
  const bool b = e < e_hi && (lane == 0 || dst[e] != dst[e - 1]);
  return __ballot_sync(0xffffffffu, b);
}
__device__ __forceinline__ void load_Y9(const float* rec, float* Y) {
  const float4 a = *reinterpret_cast<const float4*>(rec), b = *reinterpret_cast<const float4*>(rec + 4);
  Y[0] = a.x; Y[1] = a.y; Y[2] = a.z; Y[3] = a.w; Y[4] = b.x; Y[5] = b.y; Y[6] = b.z; Y[7] = b.w; Y[8] = rec[8];
}

// sums over the warp; the lanes named in the comments end with the component, the others hold garbage
__device__ __forceinline__ float warp_sum1(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// 8 values: lanes with (lane & 3) == 0 hold component ((lane>>4)&1)*4 + ((lane>>3)&1)*2 + ((lane>>2)&1)
__device__ __forceinline__ float warp_sum8(float* v, int lane) {
  const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float send = h16 ? v[i] : v[i + 4], keep = h16 ? v[i + 4] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float send = h8 ? v[i] : v[i + 2], keep = h8 ? v[i + 2] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  {
    const float send = h4 ? v[0] : v[1], keep = h4 ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
  return v[0];
}

// ---- path lists: everything a thread does for the paths of its channel ------------------------------------------------
// `mask` selects the paths of the list that belong to the job (bit BIT = this path); their weights sit in consecutive
// slabs.  w[i][j] = weight of list entry i for edge j of the current block (entries outside the mask stay unset).
template <> struct PL<> {
  static constexpr int NP = 0, NACC = 0;
  template <int BIT, int NB> __device__ __forceinline__ static void ld_w(int, int&, uint32_t, float (*)[NB]) {}
  template <int BIT, int NB> __device__ __forceinline__ static void fwd(int, const float*, const float*, const float (*)[NB], int, float*) {}
  template <int BIT> __device__ __forceinline__ static void flush(int, const ConvDesc&, float*, int, float, float*) {}
  template <int BIT> __device__ __forceinline__ static void load_g(int, const ConvDesc&, const float*, int, float, float*) {}
  template <int BIT, int NB>
  __device__ __forceinline__ static void bwd(int, bool, const float*, const float*, const float (*)[NB], const float (*)[NB], int,
                                             const float*, float&, float*, float*) {}
};
template <int PID, int... REST> struct PL<PID, REST...> {
  using C = CgD<PID>;
  using R = PL<REST...>;
  static constexpr int NP = 1 + R::NP, NACC = C::D3 + R::NACC;
  template <int BIT, int NB>
  __device__ __forceinline__ static void ld_w(int mask, int& slab, uint32_t tcol, float (*w)[NB]) {
    if ((mask >> BIT) & 1) {
      tmem_ld_nowait<NB>(tcol + (uint32_t)(slab * F_SLABCOLS), w[0]);
      ++slab;
    }
    R::template ld_w<BIT + 1, NB>(mask, slab, tcol, w + 1);
  }
  template <int BIT, int NB>
  __device__ __forceinline__ static void fwd(int mask, const float* x, const float* Y, const float (*w)[NB], int jj, float* acc) {
    if ((mask >> BIT) & 1) {
      float X[C::D1];
#pragma unroll
      for (int a = 0; a < C::D1; ++a) X[a] = w[0][jj] * x[a];
      C::fwd(X, Y, acc);
    }
    R::template fwd<BIT + 1, NB>(mask, x, Y, w + 1, jj, acc + C::D3);
  }
  template <int BIT>
  __device__ __forceinline__ static void flush(int mask, const ConvDesc& d, float* orow, int u, float inv, float* acc) {
    if ((mask >> BIT) & 1) {
      float* o = orow + d.ooff[PID] + u * C::D3;
#pragma unroll
      for (int c = 0; c < C::D3; ++c) { atomicAdd(o + c, acc[c] * inv); acc[c] = 0.f; }
    }
    R::template flush<BIT + 1>(mask, d, orow, u, inv, acc + C::D3);
  }
  template <int BIT>
  __device__ __forceinline__ static void load_g(int mask, const ConvDesc& d, const float* grow, int u, float inv, float* g) {
    if ((mask >> BIT) & 1) {
      const float* p = grow + d.ooff[PID] + u * C::D3;
#pragma unroll
      for (int c = 0; c < C::D3; ++c) g[c] = __ldg(p + c) * inv;
    }
    R::template load_g<BIT + 1>(mask, d, grow, u, inv, g + C::D3);
  }
  // one edge: dr += (d w) w', dY += w T, dx += w (C Y g)
  template <int BIT, int NB>
  __device__ __forceinline__ static void bwd(int mask, bool want_dx, const float* x, const float* Y, const float (*w)[NB],
                                             const float (*wp)[NB], int jj, const float* g, float& dr, float* dY, float* dx) {
    if ((mask >> BIT) & 1) {
      const float wv = w[0][jj];
      float T[C::D2];
      C::tvec(x, g, T);
      float dwv = 0.f;
      if constexpr (C::L2 == 0) {
        dwv = T[0];
      } else {
#pragma unroll
        for (int b = 0; b < C::D2; ++b) {
          dwv = fmaf(Y[C::YO + b], T[b], dwv);
          dY[C::YO + b] = fmaf(wv, T[b], dY[C::YO + b]);
        }
      }
      dr = fmaf(dwv, wp[0][jj], dr);
      if (want_dx) {
        float t[C::D1];
        C::dxv(Y, g, t);
#pragma unroll
        for (int a = 0; a < C::D1; ++a) dx[a] = fmaf(wv, t[a], dx[a]);
      }
    }
    R::template bwd<BIT + 1, NB>(mask, want_dx, x, Y, w + 1, wp + 1, jj, g + C::D3, dr, dY, dx);
  }
};


// paths of each input irrep (role) in the reference instruction order, for the two output truncations
template <int ROLE, int LOUT> struct RolePaths;
template <> struct RolePaths<0, 2> { using L = PL<P000, P011, P022>; };
template <> struct RolePaths<1, 2> { using L = PL<P101, P110, P111, P112, P121, P122>; };
template <> struct RolePaths<2, 2> { using L = PL<P202, P211, P212, P220, P221, P222>; };
template <> struct RolePaths<0, 0> { using L = PL<P000>; };
template <> struct RolePaths<1, 0> { using L = PL<P110>; };
template <> struct RolePaths<2, 0> { using L = PL<P220>; };

// ================================================ forward ======================================================
struct FusedFwdSmem {
  uint8_t A[2][2 * F_APART];     // two stages of W2^T slabs (hi | lo)
  uint8_t B[2][2 * F_BPART];     // two 32-edge operand tiles (h2 hi | h2 lo)
  float Y[F_NF][12];
  const float* xrow[F_NF];       // x + src * xdim
  int dst[F_NF];
  uint64_t full_a[2], full_b, slab_done[F_MAXSLAB];
  uint32_t tmem_base;
};

template <int ROLE, int LOUT>
__device__ __forceinline__ void fwd_job(const ConvDesc& d, const FusedFwdSmem& s, const FusedJob& job, uint32_t tq, int e_lo,
                                        int e_hi, unsigned brk, float inv_denom, float* __restrict__ out) {
  using L = typename RolePaths<ROLE, LOUT>::L;
  constexpr int D1 = 2 * ROLE + 1, NB = 4;
  if (e_lo >= e_hi) return;
  const int lane = threadIdx.x & 31;
  const int u = job.ubase + lane, mask = job.mask;
  const int xo = conv_xoff(d, ROLE) + u * D1;
  float acc[L::NACC];
#pragma unroll
  for (int c = 0; c < L::NACC; ++c) acc[c] = 0.f;
  int cur = -1;
  float xq[NB][D1];   // gathers run NB edges ahead of the arithmetic: a slot is refilled as soon as it is consumed
#pragma unroll
  for (int jj = 0; jj < NB; ++jj) {
    const float* xr = s.xrow[min(e_lo + jj, e_hi - 1)] + xo;
#pragma unroll
    for (int a = 0; a < D1; ++a) xq[jj][a] = __ldg(xr + a);
  }
#pragma unroll 1
  for (int eb = e_lo; eb < e_hi; eb += NB) {
    float w[L::NP][NB];
    int slab = job.first_slab;
    L::template ld_w<0, NB>(mask, slab, tq + (uint32_t)eb, w);
    tmem_wait_ld();
    const unsigned bm = brk >> (eb - e_lo);
#pragma unroll
    for (int jj = 0; jj < NB; ++jj) {
      const int e = eb + jj;
      if (e < e_hi) {
        if ((bm >> jj) & 1u) {
          if (cur >= 0) L::template flush<0>(mask, d, out + (size_t)cur * d.odim, u, inv_denom, acc);
          cur = s.dst[e];
        }
        float Yl[9], xv[D1];
        load_Y9(s.Y[e], Yl);
#pragma unroll
        for (int a = 0; a < D1; ++a) xv[a] = xq[jj][a];
        {
          const float* xr = s.xrow[min(e + NB, e_hi - 1)] + xo;
#pragma unroll
          for (int a = 0; a < D1; ++a) xq[jj][a] = __ldg(xr + a);
        }
        L::template fwd<0, NB>(mask, xv, Yl, w, jj, acc);
      }
    }
  }
  L::template flush<0>(mask, d, out + (size_t)cur * d.odim, u, inv_denom, acc);
}

__global__ void __launch_bounds__(F_THREADS, 1)
fused_conv_fwd_kernel(FusedPlan pl, ConvDesc d, long long E, const int* __restrict__ dstv, const int* __restrict__ srcv,
                      const float* __restrict__ x, const float* __restrict__ Yg, const float* __restrict__ tiles,
                      const float* __restrict__ pack, float inv_denom, float* __restrict__ out) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  FusedFwdSmem& s = *reinterpret_cast<FusedFwdSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * F_NF;
  const int ne = (int)min((long long)F_NF, E - e0);
  const int ntile = (ne + FT - 1) / FT;
  const int NS = pl.nslab;
  constexpr uint32_t SLAB = 2 * F_APART;

  if (warp == 0) {
    tmem_alloc(&s.tmem_base, 512);
    if (lane == 0) {
      mbar_init(&s.full_a[0], 1); mbar_init(&s.full_a[1], 1); mbar_init(&s.full_b, 1);
      for (int k = 0; k < F_MAXSLAB; ++k) mbar_init(&s.slab_done[k], 1);
      fence_mbar_init();
    }
  }
  for (int i = threadIdx.x; i < ne; i += F_THREADS) {
    s.xrow[i] = x + (size_t)srcv[e0 + i] * d.xdim;
    s.dst[i] = dstv[e0 + i];
  }
  for (int i = threadIdx.x; i < ne * 9; i += F_THREADS) s.Y[i / 9][i % 9] = Yg[e0 * 9 + i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  if (threadIdx.x == 0) {
    // all slabs of the tile are accumulated up front (8 slabs x 64 columns fill the TMEM of the SM)
    auto issue = [&](int sl) {
      const int st = sl & 1;
      mbar_wait(&s.full_a[st], (sl >> 1) & 1);
      tc_fence_after();
      for (int j = 0; j < ntile; ++j) issue_slab_mma(tmem + (uint32_t)(sl * F_SLABCOLS + j * FT), s.A[st], s.B[j]);
      mma_commit(&s.slab_done[sl]);
    };
    mbar_expect_tx(&s.full_b, (uint32_t)ntile * 2 * F_BPART);
    for (int j = 0; j < ntile; ++j)
      bulk_copy_g2s(s.B[j], tiles + (size_t)(2 * blockIdx.x + j) * F_TILE_FLOATS, 2 * F_BPART, &s.full_b);
    for (int k = 0; k < 2 && k < NS; ++k) {
      mbar_expect_tx(&s.full_a[k], SLAB);
      bulk_copy_g2s(s.A[k], pack + (size_t)k * (SLAB / 4), SLAB, &s.full_a[k]);
    }
    mbar_wait(&s.full_b, 0);
    issue(0);
    for (int sl = 0; sl < NS; ++sl) {
      if (sl + 1 < NS) issue(sl + 1);
      if (sl + 2 < NS) {   // MMA(sl) has consumed A[sl & 1]: refill it with slab sl + 2
        mbar_wait(&s.slab_done[sl], 0);
        mbar_expect_tx(&s.full_a[sl & 1], SLAB);
        bulk_copy_g2s(s.A[sl & 1], pack + (size_t)(sl + 2) * (SLAB / 4), SLAB, &s.full_a[sl & 1]);
      }
    }
  }
  __syncwarp();
  const int q = warp & 3, cg = warp >> 2;
  const int e_lo = 16 * cg, e_hi = min(16 * cg + 16, ne);
  const unsigned brk = break_mask(s.dst, e_lo, e_hi);
  const uint32_t tq = tmem + ((uint32_t)(32 * q) << 16);
  for (int j = 0; j < pl.njob[q]; ++j) {
    const FusedJob job = pl.job[q][j];
    mbar_wait(&s.slab_done[job.first_slab + __popc(job.mask) - 1], 0);
    tc_fence_after();
    if (d.lout == 2) {
      if (job.role == 0) fwd_job<0, 2>(d, s, job, tq, e_lo, e_hi, brk, inv_denom, out);
      else if (job.role == 1) fwd_job<1, 2>(d, s, job, tq, e_lo, e_hi, brk, inv_denom, out);
      else fwd_job<2, 2>(d, s, job, tq, e_lo, e_hi, brk, inv_denom, out);
    } else {
      if (job.role == 0) fwd_job<0, 0>(d, s, job, tq, e_lo, e_hi, brk, inv_denom, out);
      else if (job.role == 1) fwd_job<1, 0>(d, s, job, tq, e_lo, e_hi, brk, inv_denom, out);
      else fwd_job<2, 0>(d, s, job, tq, e_lo, e_hi, brk, inv_denom, out);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// ================================================ backward =====================================================
struct FusedBwdSmem {
  uint8_t A[2][2 * F_APART];     // two stages of W2^T slabs (hi | lo)
  uint8_t B[4 * F_BPART];        // one 32-edge operand tile: h2 hi | h2 lo | h2' hi | h2' lo
  float Y[F_NB][12];
  float acc[F_NB][12];           // per edge: [0] d r, [1..8] d Y
  const float* xrow[F_NB];       // x + src * xdim
  const float* xbase;
  int dst[F_NB];
  uint64_t full_a[2], full_b, slab_done[F_MAXSLAB];
  uint32_t tmem_base;
};

template <int ROLE, int LOUT>
__device__ __forceinline__ void bwd_job(const ConvDesc& d, FusedBwdSmem& s, const FusedJob& job, uint32_t tq, int e_lo, int e_hi,
                                        unsigned brk, const float* __restrict__ gout, float inv_denom,
                                        float* __restrict__ dxo) {
  using L = typename RolePaths<ROLE, LOUT>::L;
  constexpr int D1 = 2 * ROLE + 1, NB = 2;
  if (e_lo >= e_hi) return;
  const int lane = threadIdx.x & 31;
  const int u = job.ubase + lane, mask = job.mask;
  const int xo = conv_xoff(d, ROLE) + u * D1;
  const bool want_dx = dxo != nullptr;
  float g[L::NACC];
  float xq[NB][D1];
#pragma unroll
  for (int jj = 0; jj < NB; ++jj) {
    const float* xr = s.xrow[min(e_lo + jj, e_hi - 1)] + xo;
#pragma unroll
    for (int a = 0; a < D1; ++a) xq[jj][a] = __ldg(xr + a);
  }
#pragma unroll 1
  for (int eb = e_lo; eb < e_hi; eb += NB) {
    float w[L::NP][NB], wp[L::NP][NB];
    int slab = job.first_slab;
    L::template ld_w<0, NB>(mask, slab, tq + (uint32_t)eb, w);
    slab = job.first_slab;
    L::template ld_w<0, NB>(mask, slab, tq + (uint32_t)(F_NB + eb), wp);
    tmem_wait_ld();
    const unsigned bm = brk >> (eb - e_lo);
#pragma unroll
    for (int jj = 0; jj < NB; ++jj) {
      const int e = eb + jj;
      if (e < e_hi) {
        if ((bm >> jj) & 1u) L::template load_g<0>(mask, d, gout + (size_t)s.dst[e] * d.odim, u, inv_denom, g);
        float Yl[9], xv[D1];
        load_Y9(s.Y[e], Yl);
#pragma unroll
        for (int a = 0; a < D1; ++a) xv[a] = xq[jj][a];
        {
          const float* xr = s.xrow[min(e + NB, e_hi - 1)] + xo;
#pragma unroll
          for (int a = 0; a < D1; ++a) xq[jj][a] = __ldg(xr + a);
        }
        float dr = 0.f, dY[9], dx[D1];
#pragma unroll
        for (int b = 0; b < 9; ++b) dY[b] = 0.f;
#pragma unroll
        for (int a = 0; a < D1; ++a) dx[a] = 0.f;
        L::template bwd<0, NB>(mask, want_dx, xv, Yl, w, wp, jj, g, dr, dY, dx);
        if (want_dx) {   // complete over this thread's paths of channel u: one RED per component
          float* p = dxo + (s.xrow[e] - s.xbase) + xo;
#pragma unroll
          for (int a = 0; a < D1; ++a) atomicAdd(p + a, dx[a]);
        }
        // sums over the 32 channels of the warp, then per-edge accumulation in shared memory
        const float rs = warp_sum1(dr);
        if (lane == 0) atomicAdd(&s.acc[e][0], rs);
        const float ys = warp_sum8(dY + 1, lane);
        if ((lane & 3) == 0) atomicAdd(&s.acc[e][1 + ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1)], ys);
      }
    }
  }
}

__global__ void __launch_bounds__(F_THREADS, 1)
fused_conv_bwd_kernel(FusedPlan pl, ConvDesc d, long long E, const int* __restrict__ dstv, const int* __restrict__ srcv,
                      const float* __restrict__ x, const float* __restrict__ Yg, const float* __restrict__ tiles,
                      const float* __restrict__ pack, const float* __restrict__ gout, float inv_denom,
                      float* __restrict__ dxo, float* __restrict__ dYo, float* __restrict__ dro, int accumulate) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  FusedBwdSmem& s = *reinterpret_cast<FusedBwdSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * F_NB;
  const int ne = (int)min((long long)F_NB, E - e0);
  const int NS = pl.nslab;
  constexpr uint32_t SLAB = 2 * F_APART;

  if (warp == 0) {
    tmem_alloc(&s.tmem_base, 512);
    if (lane == 0) {
      mbar_init(&s.full_a[0], 1); mbar_init(&s.full_a[1], 1); mbar_init(&s.full_b, 1);
      for (int k = 0; k < F_MAXSLAB; ++k) mbar_init(&s.slab_done[k], 1);
      fence_mbar_init();
    }
  }
  for (int i = threadIdx.x; i < ne; i += F_THREADS) {
    s.xrow[i] = x + (size_t)srcv[e0 + i] * d.xdim;
    s.dst[i] = dstv[e0 + i];
  }
  if (threadIdx.x == 0) s.xbase = x;
  for (int i = threadIdx.x; i < ne * 9; i += F_THREADS) s.Y[i / 9][i % 9] = Yg[e0 * 9 + i];
  for (int i = threadIdx.x; i < F_NB * 12; i += F_THREADS) (&s.acc[0][0])[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  if (threadIdx.x == 0) {
    auto issue = [&](int sl) {
      const int st = sl & 1;
      mbar_wait(&s.full_a[st], (sl >> 1) & 1);
      tc_fence_after();
      issue_slab_mma(tmem + (uint32_t)(sl * F_SLABCOLS), s.A[st], s.B);                       // w  = W2^T . h2
      issue_slab_mma(tmem + (uint32_t)(sl * F_SLABCOLS + F_NB), s.A[st], s.B + 2 * F_BPART);   // w' = W2^T . h2'
      mma_commit(&s.slab_done[sl]);
    };
    mbar_expect_tx(&s.full_b, 4 * F_BPART);
    bulk_copy_g2s(s.B, tiles + (size_t)blockIdx.x * F_TILE_FLOATS, 4 * F_BPART, &s.full_b);
    for (int k = 0; k < 2 && k < NS; ++k) {
      mbar_expect_tx(&s.full_a[k], SLAB);
      bulk_copy_g2s(s.A[k], pack + (size_t)k * (SLAB / 4), SLAB, &s.full_a[k]);
    }
    mbar_wait(&s.full_b, 0);
    issue(0);
    for (int sl = 0; sl < NS; ++sl) {
      if (sl + 1 < NS) issue(sl + 1);
      if (sl + 2 < NS) {
        mbar_wait(&s.slab_done[sl], 0);
        mbar_expect_tx(&s.full_a[sl & 1], SLAB);
        bulk_copy_g2s(s.A[sl & 1], pack + (size_t)(sl + 2) * (SLAB / 4), SLAB, &s.full_a[sl & 1]);
      }
    }
  }
  __syncwarp();
  const int q = warp & 3, cg = warp >> 2;
  const int e_lo = 8 * cg, e_hi = min(8 * cg + 8, ne);
  const unsigned brk = break_mask(s.dst, e_lo, e_hi);
  const uint32_t tq = tmem + ((uint32_t)(32 * q) << 16);
  for (int j = 0; j < pl.njob[q]; ++j) {
    const FusedJob job = pl.job[q][j];
    mbar_wait(&s.slab_done[job.first_slab + __popc(job.mask) - 1], 0);
    tc_fence_after();
    if (d.lout == 2) {
      if (job.role == 0) bwd_job<0, 2>(d, s, job, tq, e_lo, e_hi, brk, gout, inv_denom, dxo);
      else if (job.role == 1) bwd_job<1, 2>(d, s, job, tq, e_lo, e_hi, brk, gout, inv_denom, dxo);
      else bwd_job<2, 2>(d, s, job, tq, e_lo, e_hi, brk, gout, inv_denom, dxo);
    } else {
      if (job.role == 0) bwd_job<0, 0>(d, s, job, tq, e_lo, e_hi, brk, gout, inv_denom, dxo);
      else if (job.role == 1) bwd_job<1, 0>(d, s, job, tq, e_lo, e_hi, brk, gout, inv_denom, dxo);
      else bwd_job<2, 0>(d, s, job, tq, e_lo, e_hi, brk, gout, inv_denom, dxo);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
  // per-edge results: d r, d Y (each edge belongs to exactly one CTA: plain read-modify-write)
  for (int i = threadIdx.x; i < ne * 9; i += F_THREADS) {
    const int e = i / 9, b = i % 9;
    if (b == 0) {
      float* p = dro + e0 + e;
      *p = accumulate ? *p + s.acc[e][0] : s.acc[e][0];
    } else {
      float* p = dYo + (e0 + e) * 9 + b;
      *p = accumulate ? *p + s.acc[e][b] : s.acc[e][b];
    }
  }
}

// ================================================ launchers ====================================================
constexpr size_t FR_SMEM = sizeof(float) * (FH * FH + FR_MAXNB * FH + 2 * FR_MAXNB * FR_E + 2 * FH * FR_E);

bool fused_supported(const ConvDesc& d, const MlpDesc& m) {
  FusedPlan pl;
  return m.h1 == FH && m.h2 == FH && m.nb >= 1 && m.nb <= FR_MAXNB && m.wn == d.wn && build_plan(d, pl);
}

size_t fused_radial_bytes(long long E) { return (size_t)((E + FR_E - 1) / FR_E) * 2 * F_TILE_FLOATS * sizeof(float); }

int fused_radial(const EdgeParams& ep, const MlpDesc& m, long long E, const float* vec, const float* coeffs, const float* W0,
                 const float* W1, float* tiles, cudaStream_t st) {
  B200_REQUIRE(m.h1 == FH && m.h2 == FH && m.nb >= 1 && m.nb <= FR_MAXNB, "fused_radial: needs a [nb<=16, 64, 64, wn] radial MLP");
  if (E == 0) return B200_OK;
  cudaFuncSetAttribute(fused_radial_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FR_SMEM);
  fused_radial_kernel<<<(unsigned)((E + FR_E - 1) / FR_E), FR_T, FR_SMEM, st>>>(ep, m.nb, E, vec, coeffs, W0, W1, tiles);
  return check_launch("fused_radial");
}

static int make_pack(const FusedPlan& pl, const ConvDesc& d, const float* W2, cudaStream_t st, float** pack) {
  keep_async_pool();
  if (cudaMallocAsync((void**)pack, (size_t)pl.nslab * 2 * F_APART, st) != cudaSuccess) {
    set_error("fused_conv: cudaMallocAsync of the operand pack failed");
    return B200_ERR_CUDA;
  }
  fused_prepack_kernel<<<64, 256, 0, st>>>(pl, d, W2, *pack);
  return B200_OK;
}

int fused_conv_fwd(const ConvDesc& d, int n_dst, long long E, const int* dst, const int* src, const float* x, const float* Y,
                   const float* tiles, const float* W2, float denom, float* out, cudaStream_t st) {
  FusedPlan pl;
  B200_REQUIRE(build_plan(d, pl), "fused_conv_fwd: unsupported irreps (multiplicities must be multiples of 32, <= 128/64/32, at most 32 path units)");
  cudaMemsetAsync(out, 0, sizeof(float) * (size_t)n_dst * d.odim, st);
  if (E == 0 || n_dst == 0) return B200_OK;
  float* pack;
  if (int rc = make_pack(pl, d, W2, st, &pack)) return rc;
  cudaFuncSetAttribute(fused_conv_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FusedFwdSmem));
  fused_conv_fwd_kernel<<<(unsigned)((E + F_NF - 1) / F_NF), F_THREADS, sizeof(FusedFwdSmem), st>>>(
      pl, d, E, dst, src, x, Y, tiles, pack, 1.0f / denom, out);
  cudaFreeAsync(pack, st);
  return check_launch("fused_conv_fwd", 2);
}

int fused_conv_bwd(const ConvDesc& d, int n_src, long long E, const int* dst, const int* src, const float* x, const float* Y,
                   const float* tiles, const float* W2, const float* gout, float denom, float* dx, float* dY, float* dr,
                   int accumulate, cudaStream_t st) {
  FusedPlan pl;
  B200_REQUIRE(build_plan(d, pl), "fused_conv_bwd: unsupported irreps (multiplicities must be multiples of 32, <= 128/64/32, at most 32 path units)");
  if (dx) cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)n_src * d.xdim, st);
  if (E == 0) return B200_OK;
  float* pack;
  if (int rc = make_pack(pl, d, W2, st, &pack)) return rc;
  const size_t smem = sizeof(FusedBwdSmem);
  cudaFuncSetAttribute(fused_conv_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  fused_conv_bwd_kernel<<<(unsigned)((E + F_NB - 1) / F_NB), F_THREADS, smem, st>>>(pl, d, E, dst, src, x, Y, tiles, pack, gout,
                                                                                   1.0f / denom, dx, dY, dr, accumulate);
  cudaFreeAsync(pack, st);
  return check_launch("fused_conv_bwd", 2);
}

}  // namespace b200
