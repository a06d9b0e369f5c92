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
//     ready-to-use tcgen05 operand tiles (tf32 hi/lo split, K-major canonical layout, 64 edges per tile).
//   * `fused_conv_fwd_kernel`: 128 edges per CTA.  A control warp streams slabs of 128 rows of W2^T
//     (cp.async.bulk, hi/lo prepacked) and issues D[row, edge] = W2^T_slab . h2^T into TMEM (3xTF32).
//     16 epilogue warps: TMEM lane = one (path, channel) pair, so a thread owns a weight channel for 32 edges,
//     gathers x[src] (coalesced over the warp), applies the sparse CG contraction and keeps the partial sum of the
//     current destination row in registers; rows are flushed with coalesced RED.ADD (out is zeroed first).
//   * `fused_conv_bwd_kernel`: 64 edges per CTA, D = [w | w'] (w' = d w / d r from the tangent tile).  Per edge:
//     d r += sum_ch (d w) w', d Y += w T, and d x[src]; the per-edge d x row is combined over the paths of a channel
//     in thread-private shared-memory staging and written with one coalesced RED.ADD row per edge.
#include "common.cuh"
#include "tc.cuh"

namespace b200 {

using namespace tc;
#include "cg_direct.cuh"

constexpr int FT = 64;                  // edges per operand tile
constexpr int FH = 64;                  // hidden width = K of the GEMM
constexpr int F_ACH = 128 * 16;         // bytes of one 16-byte K-chunk column of a 128-row A slab
constexpr int F_BCH = FT * 16;          // ... of a 64-row B tile
constexpr int F_APART = (FH / 4) * F_ACH;   // 32 KB: one part (hi or lo) of an A slab
constexpr int F_BPART = (FH / 4) * F_BCH;   // 16 KB: one part of a B tile
constexpr int F_TILE_FLOATS = 4 * F_BPART / 4;   // h2 hi | h2 lo | h2' hi | h2' lo
constexpr int F_MAXSLAB = 12;
constexpr int F_EPI_WARPS = 16;
constexpr int F_THREADS = F_EPI_WARPS * 32;

struct FusedPlan {
  int nslab;
  int stage_width;                 // floats per edge of the backward d x staging (3 m1 + 5 m2)
  signed char path[F_MAXSLAB][4];  // path id of the 32-row unit (slab, quarter), -1: empty
  short ubase[F_MAXSLAB][4];       // first channel of the unit
};

static bool build_plan(const ConvDesc& d, FusedPlan& pl) {
  if (d.m0 % 32 || d.m1 % 32 || d.m2 % 32 || d.m0 > 128 || d.m1 > 64 || d.m2 > 32) return false;
  static const int L1[kNumPaths] = {0, 0, 0, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2};
  // quarter 0: role-1 channels 0..31, quarter 1: role-1 channels 32..63, quarter 2: role-2 channels 0..31 (each
  // (role, channel) keeps one owner thread through the tile: the backward staging needs no synchronisation);
  // role-0 units fill quarter 3 and whatever is left.
  int q_list[3][kNumPaths], q_n[3] = {0, 0, 0}, r0[64][2], r0_n = 0;
  for (int p = 0; p < kNumPaths; ++p) {
    if (d.woff[p] < 0) continue;
    if (L1[p] == 0) {
      for (int u = 0; u < d.m0; u += 32) { r0[r0_n][0] = p; r0[r0_n][1] = u; ++r0_n; }
    } else if (L1[p] == 1) {
      q_list[0][q_n[0]++] = p;
      if (d.m1 > 32) q_list[1][q_n[1]++] = p;
    } else {
      q_list[2][q_n[2]++] = p;
    }
  }
  int taken[3] = {0, 0, 0}, r0_taken = 0, s = 0;
  for (;; ++s) {
    const bool more = taken[0] < q_n[0] || taken[1] < q_n[1] || taken[2] < q_n[2] || r0_taken < r0_n;
    if (!more) break;
    if (s >= F_MAXSLAB) return false;
    for (int q = 0; q < 4; ++q) {
      pl.path[s][q] = -1;
      pl.ubase[s][q] = 0;
      if (q < 3 && taken[q] < q_n[q]) {
        pl.path[s][q] = (signed char)q_list[q][taken[q]++];
        pl.ubase[s][q] = (short)(q == 1 ? 32 : 0);
      } else if (r0_taken < r0_n) {
        pl.path[s][q] = (signed char)r0[r0_taken][0];
        pl.ubase[s][q] = (short)r0[r0_taken][1];
        ++r0_taken;
      }
    }
  }
  pl.nslab = s;
  pl.stage_width = 3 * d.m1 + 5 * d.m2;
  return s > 0;
}

// ---- operand pack of W2 for a plan: [slab][hi|lo][16 chunks][128 rows][4], scaled by 1/sqrt(64) ------------
__global__ void fused_prepack_kernel(FusedPlan pl, ConvDesc d, const float* __restrict__ W2, float* __restrict__ pack) {
  const int n = pl.nslab * 16 * 128;
  const float sc = 1.0f / sqrtf((float)FH);
  for (int it = blockIdx.x * blockDim.x + threadIdx.x; it < n; it += gridDim.x * blockDim.x) {
    const int r = it & 127, c4 = (it >> 7) & 15, s = it >> 11;
    const int p = pl.path[s][r >> 5];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p >= 0) {
      const int ch = d.woff[p] + pl.ubase[s][r >> 5] + (r & 31);
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

constexpr int FR_T = 256;   // threads: 64 edges x 4 groups of 16 hidden units
constexpr int FR_MAXNB = 16;

__global__ void __launch_bounds__(FR_T)
fused_radial_kernel(EdgeParams ep, int nb, long long E, const float* __restrict__ vec, const float* __restrict__ coeffs,
                    const float* __restrict__ W0, const float* __restrict__ W1, float* __restrict__ tiles) {
  extern __shared__ __align__(16) float fr_smem[];
  float* sW1 = fr_smem;                                                         // [FH][FH]
  float* sW0 = sW1 + FH * FH;                                                   // [nb][FH]
  float (*sE)[FR_MAXNB][FT] = reinterpret_cast<float (*)[FR_MAXNB][FT]>(sW0 + FR_MAXNB * FH);   // basis value / d basis/dr
  float (*sH)[FH][FT] = reinterpret_cast<float (*)[FH][FT]>(sW0 + FR_MAXNB * FH + 2 * FR_MAXNB * FT);   // h1 / h1'
  const int tid = threadIdx.x, e = tid & 63, hg = tid >> 6;
  const long long e0 = (long long)blockIdx.x * FT;
  for (int i = tid; i < nb * FH; i += FR_T) sW0[i] = W0[i];
  for (int i = tid; i < FH * FH; i += FR_T) sW1[i] = W1[i];
  if (tid < FT) {
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
  float* tile = tiles + (size_t)blockIdx.x * F_TILE_FLOATS;
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
    float* base = tile + chunk * (F_BCH / 4) + e * 4;
    split4(make_float4(v[0], v[1], v[2], v[3]), hi, lo);
    *reinterpret_cast<float4*>(base) = hi;
    *reinterpret_cast<float4*>(base + F_BPART / 4) = lo;
    split4(make_float4(t[0], t[1], t[2], t[3]), hi, lo);
    *reinterpret_cast<float4*>(base + 2 * (F_BPART / 4)) = hi;
    *reinterpret_cast<float4*>(base + 3 * (F_BPART / 4)) = lo;
  }
}

// ---- shared pieces of the two convolution kernels ---------------------------------------------------------------
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
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

#define F_DISPATCH_PATH(p, CALL)                                  \
  switch (p) {                                                    \
    case P000: { CALL(P000); } break;                             \
    case P011: { CALL(P011); } break;                             \
    case P022: { CALL(P022); } break;                             \
    case P101: { CALL(P101); } break;                             \
    case P110: { CALL(P110); } break;                             \
    case P111: { CALL(P111); } break;                             \
    case P112: { CALL(P112); } break;                             \
    case P121: { CALL(P121); } break;                             \
    case P122: { CALL(P122); } break;                             \
    case P202: { CALL(P202); } break;                             \
    case P211: { CALL(P211); } break;                             \
    case P212: { CALL(P212); } break;                             \
    case P220: { CALL(P220); } break;                             \
    case P221: { CALL(P221); } break;                             \
    case P222: { CALL(P222); } break;                             \
    default: break;                                               \
  }

// ================================================ forward ======================================================
// whole warp: 4 consecutive fp32 columns of this warp's 32 TMEM lanes
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float* v) {
  uint32_t r[4];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(r[i]);
}

// the harmonics a path needs, from the padded shared-memory record (compile-time indices only)
template <int L2> __device__ __forceinline__ void load_Y(const float* rec, float* Y) {
  if constexpr (L2 == 1) {
    const float4 v = *reinterpret_cast<const float4*>(rec);
    Y[1] = v.y; Y[2] = v.z; Y[3] = v.w;
  } else if constexpr (L2 == 2) {
    const float4 v = *reinterpret_cast<const float4*>(rec + 4);
    Y[4] = v.x; Y[5] = v.y; Y[6] = v.z; Y[7] = v.w;
    Y[8] = rec[8];
  }
}

// bit j: edge e_lo + j starts a new destination row within this warp's range (bit 0 always set)
__device__ __forceinline__ unsigned break_mask(const int* dst, int e_lo, int e_hi) {
  const int lane = threadIdx.x & 31, e = e_lo + lane;
  const bool b = e < e_hi && (lane == 0 || dst[e] != dst[e - 1]);
  return __ballot_sync(0xffffffffu, b);
}

struct FusedFwdSmem {
  uint8_t A[2][2 * F_APART];     // two stages of W2^T slabs (hi | lo)
  uint8_t B[2][2 * F_BPART];     // two 64-edge operand tiles (h2 hi | h2 lo)
  float Y[2 * FT][12];
  const float* xrow[2 * FT];     // x + src * xdim
  int dst[2 * FT];
  uint64_t full_a[2], full_b, mma_done[2], tmem_free[2];
  uint32_t tmem_base;
};

template <int PID>
__device__ __forceinline__ void fwd_unit(const ConvDesc& d, const FusedFwdSmem& s, uint32_t taddr, int e_lo, int e_hi,
                                         unsigned brk, int ubase, float inv_denom, float* __restrict__ out) {
  using C = CgD<PID>;
  if (e_lo >= e_hi) return;
  const int lane = threadIdx.x & 31;
  const int u = ubase + lane;
  const int xo = conv_xoff(d, C::L1) + u * C::D1;
  const size_t oo = (size_t)d.ooff[PID] + (size_t)u * C::D3;
  float acc[C::D3];
#pragma unroll
  for (int c = 0; c < C::D3; ++c) acc[c] = 0.f;
  int cur = -1;
  float xq[4][C::D1];   // gathers run 4 edges ahead of the arithmetic: slot jj is refilled as soon as it is consumed
  auto load_x = [&](float (&xq)[4][C::D1], int eb) {
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const float* xr = s.xrow[min(eb + jj, e_hi - 1)] + xo;
#pragma unroll
      for (int a = 0; a < C::D1; ++a) xq[jj][a] = __ldg(xr + a);
    }
  };
  auto flush = [&]() {
    float* o = out + (size_t)cur * d.odim + oo;
#pragma unroll
    for (int c = 0; c < C::D3; ++c) { atomicAdd(o + c, acc[c] * inv_denom); acc[c] = 0.f; }
  };
  auto compute = [&](int eb) {
    float w4[4];
    tmem_ld4(taddr + (uint32_t)(eb - e_lo), w4);
    const unsigned bm = brk >> (eb - e_lo);
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const int e = eb + jj;
      if (e < e_hi) {
        if ((bm >> jj) & 1u) {
          if (cur >= 0) flush();
          cur = s.dst[e];
        }
        float Yl[12], X[C::D1];
        load_Y<C::L2>(s.Y[e], Yl);
#pragma unroll
        for (int a = 0; a < C::D1; ++a) X[a] = w4[jj] * xq[jj][a];
        {
          const float* xr = s.xrow[min(e + 4, e_hi - 1)] + xo;
#pragma unroll
          for (int a = 0; a < C::D1; ++a) xq[jj][a] = __ldg(xr + a);
        }
        C::fwd(X, Yl, acc);
      }
    }
  };
  load_x(xq, e_lo);
#pragma unroll 1
  for (int eb = e_lo; eb < e_hi; eb += 4) compute(eb);
  flush();
}

__global__ void __launch_bounds__(F_THREADS, 1)
fused_conv_fwd_kernel(FusedPlan pl, ConvDesc d, long long E, const int* __restrict__ dstv, const int* __restrict__ srcv,
                      const float* __restrict__ x, const float* __restrict__ Yg, const float* __restrict__ tiles,
                      const float* __restrict__ pack, float inv_denom, float* __restrict__ out) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  FusedFwdSmem& s = *reinterpret_cast<FusedFwdSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * (2 * FT);
  const int ne = (int)min((long long)(2 * FT), E - e0);
  const int nsub = (ne + FT - 1) / FT;
  const int NS = pl.nslab;
  constexpr uint32_t SLAB = 2 * F_APART;

  if (warp == 0) {
    tmem_alloc(&s.tmem_base, 256);
    if (lane == 0) {
      mbar_init(&s.full_a[0], 1); mbar_init(&s.full_a[1], 1); mbar_init(&s.full_b, 1);
      mbar_init(&s.mma_done[0], 1); mbar_init(&s.mma_done[1], 1);
      mbar_init(&s.tmem_free[0], F_EPI_WARPS); mbar_init(&s.tmem_free[1], F_EPI_WARPS);
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

  // slab sl + 1 is launched (operand refill + MMA issue) by warp (sl + 1) % 16 at the start of its iteration sl
  auto launch_slab = [&](int sl) {   // one thread; sl < NS
    const int st = sl & 1;
    mbar_wait(&s.full_a[st], (sl >> 1) & 1);
    if (sl >= 2) mbar_wait(&s.tmem_free[st], ((sl - 2) >> 1) & 1);
    tc_fence_after();
    for (int j = 0; j < nsub; ++j) issue_slab_mma(tmem + (uint32_t)(st * 128 + j * FT), s.A[st], s.B[j]);
    mma_commit(&s.mma_done[st]);
  };
  if (threadIdx.x == 0) {
    mbar_expect_tx(&s.full_b, (uint32_t)nsub * 2 * F_BPART);
    for (int j = 0; j < nsub; ++j)
      bulk_copy_g2s(s.B[j], tiles + (size_t)(2 * blockIdx.x + j) * F_TILE_FLOATS, 2 * F_BPART, &s.full_b);
    for (int k = 0; k < 2 && k < NS; ++k) {
      mbar_expect_tx(&s.full_a[k], SLAB);
      bulk_copy_g2s(s.A[k], pack + (size_t)k * (SLAB / 4), SLAB, &s.full_a[k]);
    }
    mbar_wait(&s.full_b, 0);
    launch_slab(0);
  }
  __syncwarp();
  const int q = warp & 3, cg = warp >> 2;
  const int e_lo = 32 * cg, e_hi = min(32 * cg + 32, ne);
  const unsigned brk = break_mask(s.dst, e_lo, e_hi);
  for (int sl = 0; sl < NS; ++sl) {
    const int st = sl & 1;
    mbar_wait(&s.mma_done[st], (sl >> 1) & 1);
    if (warp == ((sl + 1) & (F_EPI_WARPS - 1)) && sl + 1 < NS) {
      if (lane == 0) {
        if (sl + 2 < NS) {   // MMA(sl) has consumed A[st]: refill it with slab sl + 2
          mbar_expect_tx(&s.full_a[st], SLAB);
          bulk_copy_g2s(s.A[st], pack + (size_t)(sl + 2) * (SLAB / 4), SLAB, &s.full_a[st]);
        }
        launch_slab(sl + 1);
      }
      __syncwarp();
    }
    tc_fence_after();
    const int p = pl.path[sl][q];
    const int ub = pl.ubase[sl][q];
    const uint32_t taddr = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(st * 128 + e_lo);
#define F_CALL(PID) fwd_unit<PID>(d, s, taddr, e_lo, e_hi, brk, ub, inv_denom, out)
    F_DISPATCH_PATH(p, F_CALL)
#undef F_CALL
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&s.tmem_free[st]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// ================================================ backward =====================================================
struct FusedBwdSmem {
  uint8_t A[2 * F_APART];        // one stage of W2^T slabs (hi | lo)
  uint8_t B[4 * F_BPART];        // one 64-edge operand tile: h2 hi | h2 lo | h2' hi | h2' lo
  float Y[FT][12];
  float acc[FT][12];             // per edge: [0] d r, [1..8] d Y
  const float* xrow[FT + 1];     // x + src * xdim; [FT] = x
  int src[FT], dst[FT];
  uint64_t full_a, full_b, mma_done[2], tmem_free[2];
  uint32_t tmem_base;
  float stage[1];                // [FT][stage_width] thread-private d x staging (dynamic tail)
};

// sums over the warp; the lane named in the comment ends with the component, the others hold garbage
__device__ __forceinline__ float warp_sum1(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// 4 values: lanes with (lane & 7) == 0 hold component ((lane>>4)&1)*2 + ((lane>>3)&1)
__device__ __forceinline__ float warp_sum4(float* v, int lane) {
  const bool h16 = lane & 16, h8 = lane & 8;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float send = h16 ? v[i] : v[i + 2], keep = h16 ? v[i + 2] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
  {
    const float send = h8 ? v[0] : v[1], keep = h8 ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 4);
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
  return v[0];
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

template <int PID>
__device__ __forceinline__ void bwd_unit(const ConvDesc& d, FusedBwdSmem& s, int stage_width, uint32_t taddr, int e_lo, int e_hi,
                                         unsigned brk, int ubase, const float* __restrict__ gout, float inv_denom,
                                         float* __restrict__ dxo) {
  using C = CgD<PID>;
  if (e_lo >= e_hi) return;
  const int lane = threadIdx.x & 31;
  const int u = ubase + lane;
  const int xo = conv_xoff(d, C::L1) + u * C::D1;
  const size_t oo = (size_t)d.ooff[PID] + (size_t)u * C::D3;
  const int so = (C::L1 == 1 ? 0 : 3 * d.m1) + u * C::D1;   // staging column (roles 1, 2)
  float g[C::D3];
  float xq[4][C::D1];
  auto load_x = [&](float (&xq)[4][C::D1], int eb) {
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const float* xr = s.xrow[min(eb + jj, e_hi - 1)] + xo;
#pragma unroll
      for (int a = 0; a < C::D1; ++a) xq[jj][a] = __ldg(xr + a);
    }
  };
  auto compute = [&](int eb) {
    float w4[4], wp4[4];
    tmem_ld4(taddr + (uint32_t)(eb - e_lo), w4);
    tmem_ld4(taddr + (uint32_t)(FT + eb - e_lo), wp4);
    const unsigned bm = brk >> (eb - e_lo);
    float r0[4];   // d r of the 4 edges (reduced together when the path has no harmonics to differentiate)
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const int e = eb + jj;
      r0[jj] = 0.f;
      if (e < e_hi) {
        if ((bm >> jj) & 1u) {
          const float* gr = gout + (size_t)s.dst[e] * d.odim + oo;
#pragma unroll
          for (int c = 0; c < C::D3; ++c) g[c] = __ldg(gr + c) * inv_denom;
        }
        const float w = w4[jj], wp = wp4[jj];
        float Yl[12], T[C::D2];
        load_Y<C::L2>(s.Y[e], Yl);
        C::tvec(xq[jj], g, T);
        {
          const float* xr = s.xrow[min(e + 4, e_hi - 1)] + xo;
#pragma unroll
          for (int a = 0; a < C::D1; ++a) xq[jj][a] = __ldg(xr + a);
        }
        float dwv = 0.f;
        if constexpr (C::L2 == 0) {
          dwv = T[0];
        } else {
#pragma unroll
          for (int b = 0; b < C::D2; ++b) dwv = fmaf(Yl[C::YO + b], T[b], dwv);
        }
        if (dxo) {   // not needed for the first layer: its input depends on the species only
          float dx[C::D1];
          C::dxv(Yl, g, dx);
          if constexpr (C::L1 == 0) {
            atomicAdd(dxo + (s.xrow[e] - s.xrow[FT]) + xo, w * dx[0]);
          } else {
            float* sp = s.stage + e * stage_width + so;
#pragma unroll
            for (int a = 0; a < C::D1; ++a) sp[a] = fmaf(w, dx[a], sp[a]);
          }
        }
        // reduction over the 32 channels of the unit: d r and d Y of this edge
        if constexpr (C::L2 == 0) {
          r0[jj] = dwv * wp;
        } else if constexpr (C::L2 == 1) {
          float v[4] = {dwv * wp, w * T[0], w * T[1], w * T[2]};
          const float r = warp_sum4(v, lane);
          if ((lane & 7) == 0) atomicAdd(&s.acc[e][((lane >> 4) & 1) * 2 + ((lane >> 3) & 1)], r);
        } else {
          float v[8] = {dwv * wp, w * T[0], w * T[1], w * T[2], w * T[3], w * T[4], 0.f, 0.f};
          const float r = warp_sum8(v, lane);
          const int m = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
          if ((lane & 3) == 0 && m < 6) atomicAdd(&s.acc[e][m == 0 ? 0 : 3 + m], r);
        }
      }
    }
    if constexpr (C::L2 == 0) {
      const float r = warp_sum4(r0, lane);
      const int jj = ((lane >> 4) & 1) * 2 + ((lane >> 3) & 1);
      if ((lane & 7) == 0 && eb + jj < e_hi) atomicAdd(&s.acc[eb + jj][0], r);
    }
  };
  load_x(xq, e_lo);
#pragma unroll 1
  for (int eb = e_lo; eb < e_hi; eb += 4) compute(eb);
}

__global__ void __launch_bounds__(F_THREADS, 1)
fused_conv_bwd_kernel(FusedPlan pl, ConvDesc d, long long E, const int* __restrict__ dstv, const int* __restrict__ srcv,
                      const float* __restrict__ x, const float* __restrict__ Yg, const float* __restrict__ tiles,
                      const float* __restrict__ pack, const float* __restrict__ gout, float inv_denom,
                      float* __restrict__ dxo, float* __restrict__ dYo, float* __restrict__ dro, int accumulate) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  FusedBwdSmem& s = *reinterpret_cast<FusedBwdSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long e0 = (long long)blockIdx.x * FT;
  const int ne = (int)min((long long)FT, E - e0);
  const int NS = pl.nslab, SW = pl.stage_width;
  constexpr uint32_t SLAB = 2 * F_APART;

  if (warp == 0) {
    tmem_alloc(&s.tmem_base, 256);
    if (lane == 0) {
      mbar_init(&s.full_a, 1); mbar_init(&s.full_b, 1);
      mbar_init(&s.mma_done[0], 1); mbar_init(&s.mma_done[1], 1);
      mbar_init(&s.tmem_free[0], F_EPI_WARPS); mbar_init(&s.tmem_free[1], F_EPI_WARPS);
      fence_mbar_init();
    }
  }
  for (int i = threadIdx.x; i < ne; i += F_THREADS) {
    const int sj = srcv[e0 + i];
    s.src[i] = sj;
    s.xrow[i] = x + (size_t)sj * d.xdim;
    s.dst[i] = dstv[e0 + i];
  }
  if (threadIdx.x == 0) s.xrow[FT] = x;   // base, to turn a row pointer back into an offset
  for (int i = threadIdx.x; i < ne * 9; i += F_THREADS) s.Y[i / 9][i % 9] = Yg[e0 * 9 + i];
  for (int i = threadIdx.x; i < FT * 12; i += F_THREADS) (&s.acc[0][0])[i] = 0.f;
  for (int i = threadIdx.x; i < FT * SW; i += F_THREADS) s.stage[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  auto launch_slab = [&](int sl) {   // one thread: A(sl) must have been requested
    const int st = sl & 1;
    mbar_wait(&s.full_a, sl & 1);
    if (sl >= 2) mbar_wait(&s.tmem_free[st], ((sl - 2) >> 1) & 1);
    tc_fence_after();
    issue_slab_mma(tmem + (uint32_t)(st * 128), s.A, s.B);                      // w  = W2^T . h2
    issue_slab_mma(tmem + (uint32_t)(st * 128 + FT), s.A, s.B + 2 * F_BPART);   // w' = W2^T . h2'
    mma_commit(&s.mma_done[st]);
  };
  if (threadIdx.x == 0) {
    mbar_expect_tx(&s.full_b, 4 * F_BPART);
    bulk_copy_g2s(s.B, tiles + (size_t)blockIdx.x * F_TILE_FLOATS, 4 * F_BPART, &s.full_b);
    mbar_expect_tx(&s.full_a, SLAB);
    bulk_copy_g2s(s.A, pack, SLAB, &s.full_a);
    mbar_wait(&s.full_b, 0);
    launch_slab(0);
  }
  __syncwarp();
  const int q = warp & 3, cg = warp >> 2;
  const int e_lo = 16 * cg, e_hi = min(16 * cg + 16, ne);
  const unsigned brk = break_mask(s.dst, e_lo, e_hi);
  for (int sl = 0; sl < NS; ++sl) {
    const int st = sl & 1;
    mbar_wait(&s.mma_done[st], (sl >> 1) & 1);
    if (warp == ((sl + 1) & (F_EPI_WARPS - 1)) && sl + 1 < NS) {
      if (lane == 0) {   // MMA(sl) has consumed the single A stage: stream slab sl + 1 and issue it
        mbar_expect_tx(&s.full_a, SLAB);
        bulk_copy_g2s(s.A, pack + (size_t)(sl + 1) * (SLAB / 4), SLAB, &s.full_a);
        launch_slab(sl + 1);
      }
      __syncwarp();
    }
    tc_fence_after();
    const int p = pl.path[sl][q];
    const int ub = pl.ubase[sl][q];
    const uint32_t taddr = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(st * 128 + e_lo);
#define F_CALL(PID) bwd_unit<PID>(d, s, SW, taddr, e_lo, e_hi, brk, ub, gout, inv_denom, dxo)
    F_DISPATCH_PATH(p, F_CALL)
#undef F_CALL
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&s.tmem_free[st]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
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
  // staged d x rows (roles 1 and 2): one coalesced RED row per edge
  if (SW > 0 && dxo) {
    for (int e = warp; e < ne; e += F_EPI_WARPS) {
      float* row = dxo + (size_t)s.src[e] * d.xdim + d.m0;
      for (int c = lane; c < SW; c += 32) atomicAdd(row + c, s.stage[e * SW + c]);
    }
  }
}

// ================================================ launchers ====================================================
constexpr size_t FR_SMEM = sizeof(float) * (FH * FH + FR_MAXNB * FH + 2 * FR_MAXNB * FT + 2 * FH * FT);

bool fused_supported(const ConvDesc& d, const MlpDesc& m) {
  FusedPlan pl;
  return m.h1 == FH && m.h2 == FH && m.nb >= 1 && m.nb <= FR_MAXNB && m.wn == d.wn && build_plan(d, pl);
}

size_t fused_radial_bytes(long long E) { return (size_t)((E + FT - 1) / FT) * F_TILE_FLOATS * sizeof(float); }

int fused_radial(const EdgeParams& ep, const MlpDesc& m, long long E, const float* vec, const float* coeffs, const float* W0,
                 const float* W1, float* tiles, cudaStream_t st) {
  B200_REQUIRE(m.h1 == FH && m.h2 == FH && m.nb >= 1 && m.nb <= FR_MAXNB, "fused_radial: needs a [nb<=16, 64, 64, wn] radial MLP");
  if (E == 0) return B200_OK;
  cudaFuncSetAttribute(fused_radial_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FR_SMEM);
  fused_radial_kernel<<<(unsigned)((E + FT - 1) / FT), FR_T, FR_SMEM, st>>>(ep, m.nb, E, vec, coeffs, W0, W1, tiles);
  return check_launch("fused_radial");
}

static int make_pack(const FusedPlan& pl, const ConvDesc& d, const float* W2, cudaStream_t st, float** pack) {
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
  B200_REQUIRE(build_plan(d, pl), "fused_conv_fwd: unsupported irreps (multiplicities must be multiples of 32, <= 128/64/32)");
  cudaMemsetAsync(out, 0, sizeof(float) * (size_t)n_dst * d.odim, st);
  if (E == 0 || n_dst == 0) return B200_OK;
  float* pack;
  if (int rc = make_pack(pl, d, W2, st, &pack)) return rc;
  cudaFuncSetAttribute(fused_conv_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FusedFwdSmem));
  fused_conv_fwd_kernel<<<(unsigned)((E + 2 * FT - 1) / (2 * FT)), F_THREADS, sizeof(FusedFwdSmem), st>>>(
      pl, d, E, dst, src, x, Y, tiles, pack, 1.0f / denom, out);
  cudaFreeAsync(pack, st);
  return check_launch("fused_conv_fwd", 2);
}

int fused_conv_bwd(const ConvDesc& d, int n_src, long long E, const int* dst, const int* src, const float* x, const float* Y,
                   const float* tiles, const float* W2, const float* gout, float denom, float* dx, float* dY, float* dr,
                   int accumulate, cudaStream_t st) {
  FusedPlan pl;
  B200_REQUIRE(build_plan(d, pl), "fused_conv_bwd: unsupported irreps (multiplicities must be multiples of 32, <= 128/64/32)");
  if (dx) cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)n_src * d.xdim, st);
  if (E == 0) return B200_OK;
  float* pack;
  if (int rc = make_pack(pl, d, W2, st, &pack)) return rc;
  const size_t smem = offsetof(FusedBwdSmem, stage) + sizeof(float) * (size_t)FT * pl.stage_width + 16;
  B200_REQUIRE(smem <= 232448, "fused_conv_bwd: staging does not fit in shared memory");
  cudaFuncSetAttribute(fused_conv_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  fused_conv_bwd_kernel<<<(unsigned)((E + FT - 1) / FT), F_THREADS, smem, st>>>(pl, d, E, dst, src, x, Y, tiles, pack, gout,
                                                                                   1.0f / denom, dx, dY, dr, accumulate);
  cudaFreeAsync(pack, st);
  return check_launch("fused_conv_bwd", 2);
}

}  // namespace b200
