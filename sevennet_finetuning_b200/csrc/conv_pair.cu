// Channel-pair convolution kernels of the MD engine (tabulated radial weights, radial_table.cu).
//
// Same semantics as conv.cu (sevenn/nn/convolution.py:19-32,105-121: uvu tensor product, weight multiply, sorted
// segment sum to the destination atom, / denominator) with a different work split, chosen from what ncu showed for the
// per-role kernels (forward bound by 16-byte-per-weight table reads from L2, reverse pass by instruction issue, one-warp
// CTAs, three passes over every edge):
//   * ONE kernel per layer and direction.  A CTA owns one destination row and has two self-sufficient warps that never
//     meet at a CTA barrier: warp A = the 1e channels plus the first half of the 0e channels, warp B = the 2e channels
//     (its two half-warps take half of the paths each) plus the second half of the 0e channels.  Each warp stages what
//     it alone needs (Y, its four coupling matrices, the table interval) once per 8-edge chunk.
//   * A thread owns a PAIR of adjacent channels and computes with packed fp32 (FFMA2, fma.rn.f32x2): coupling-matrix
//     entries are scalars shared by the pair, x, weights and accumulators are pairs.  Half the issue slots per channel
//     for the CG arithmetic, the table polynomial, the shared-memory reads, the shuffle reductions and the RED.ADDs
//     (red.global.add.v2.f32).
//   * The packed table (radial_table_build_packed) costs 12 bytes per weight: one 16-byte load {a0,a1,b0,b1} and one
//     8-byte load (4 x fp16 {c0,c1,d0,d1}) per channel pair and path.
//   * The source rows x[src] of a chunk are brought into shared memory by cp.async.bulk + mbarrier, one chunk ahead,
//     off the LSU issue path (two contiguous segments per edge and warp).
#include <cuda_fp16.h>

#include "cg_generated.cuh"
#include "common.cuh"
#include "ops.h"
#include "tc.cuh"

namespace b200 {
namespace pairk {

#ifndef B200_PAIR_TE
#define B200_PAIR_TE 8
#endif
#ifndef B200_PAIR_FWD_UNROLL
#define B200_PAIR_FWD_UNROLL 2
#endif
#ifndef B200_PAIR_BWD_UNROLL
#define B200_PAIR_BWD_UNROLL 2
#endif
#ifndef B200_PAIR_FWD_MINB
#define B200_PAIR_FWD_MINB 3
#endif
#ifndef B200_PAIR_BWD_MINB
#define B200_PAIR_BWD_MINB 3
#endif
constexpr int kFwdUnroll = B200_PAIR_FWD_UNROLL, kBwdUnroll = B200_PAIR_BWD_UNROLL;   // the two edges of a pair: unrolled or looped
constexpr int TE = B200_PAIR_TE;          // edges per chunk (even, <= 8)
constexpr float kInvSqrt3 = 0.5773502691896258f;
constexpr float kInvSqrt5 = 0.4472135954999579f;

// per-edge shared-memory record (floats): Y[9] padded to 12, then the warp's four coupling matrices
constexpr int REC_A = 68, REC_B = 100;

// A channel pair lives in ONE 64-bit register for its whole life (packed once, unpacked at the end): with float2 values
// the compiler re-assembles the register pair with MOVs in front of every FFMA2.
using f2 = unsigned long long;
__device__ __forceinline__ f2 mk(float a, float b) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ float lo(f2 v) {
  float a, b;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
  return a;
}
__device__ __forceinline__ float hi(f2 v) {
  float a, b;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
  return b;
}
__device__ __forceinline__ float hsum(f2 v) { return lo(v) + hi(v); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
  f2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ f2 fma2(float s, f2 b, f2 c) { return fma2(mk(s, s), b, c); }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
  f2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f2 mul2(float s, f2 b) { return mul2(mk(s, s), b); }
#define B200_ZERO2 0ull

// OFF: offset of the coupling matrix in the warp's record (-1: the path needs only Y)
template <int PID> struct PT;
#define B200_PT(PID, L1_, L2_, L3_, OFF_)                                                  \
  template <> struct PT<PID> {                                                              \
    static constexpr int L1 = L1_, L2 = L2_, L3 = L3_, D1 = 2 * L1_ + 1, D3 = 2 * L3_ + 1; \
    static constexpr int OFF = OFF_;                                                        \
    static constexpr int YO = L2_ == 1 ? 1 : 4;                                             \
  };
B200_PT(P000, 0, 0, 0, -1)
B200_PT(P011, 0, 1, 1, -1)
B200_PT(P022, 0, 2, 2, -1)
B200_PT(P101, 1, 0, 1, -1)
B200_PT(P110, 1, 1, 0, -1)
B200_PT(P111, 1, 1, 1, 12)
B200_PT(P112, 1, 1, 2, 24)
B200_PT(P121, 1, 2, 1, 40)
B200_PT(P122, 1, 2, 2, 52)
B200_PT(P202, 2, 0, 2, -1)
B200_PT(P211, 2, 1, 1, 12)
B200_PT(P212, 2, 1, 2, 28)
B200_PT(P220, 2, 2, 0, -1)
B200_PT(P221, 2, 2, 1, 56)
B200_PT(P222, 2, 2, 2, 72)
#undef B200_PT

template <int PID> __device__ __forceinline__ void load_M(const float* rec, float* M) {
  constexpr int N4 = (PT<PID>::D1 * PT<PID>::D3 + 3) / 4;
  const float4* q = reinterpret_cast<const float4*>(rec + PT<PID>::OFF);
#pragma unroll
  for (int i = 0; i < N4; ++i) {
    const float4 v = q[i];
    M[4 * i] = v.x; M[4 * i + 1] = v.y; M[4 * i + 2] = v.z; M[4 * i + 3] = v.w;
  }
}
template <int PID> __device__ __forceinline__ void cg_M(const float* Y, float* M) {
  if constexpr (PID == P111) cg_M_111(Y, M);
  else if constexpr (PID == P112) cg_M_112(Y, M);
  else if constexpr (PID == P121) cg_M_121(Y, M);
  else if constexpr (PID == P122) cg_M_122(Y, M);
  else if constexpr (PID == P211) cg_M_211(Y, M);
  else if constexpr (PID == P212) cg_M_212(Y, M);
  else if constexpr (PID == P221) cg_M_221(Y, M);
  else if constexpr (PID == P222) cg_M_222(Y, M);
}
template <int PID> __device__ __forceinline__ void cg_dY(const f2* G, f2* dY) {
  if constexpr (PID == P111) cg_dY_111(G, dY);
  else if constexpr (PID == P112) cg_dY_112(G, dY);
  else if constexpr (PID == P121) cg_dY_121(G, dY);
  else if constexpr (PID == P122) cg_dY_122(G, dY);
  else if constexpr (PID == P211) cg_dY_211(G, dY);
  else if constexpr (PID == P212) cg_dY_212(G, dY);
  else if constexpr (PID == P221) cg_dY_221(G, dY);
  else if constexpr (PID == P222) cg_dY_222(G, dY);
}

// ---- per-path arithmetic on a channel pair ---------------------------------------------------------
// forward: acc[c] += w * sum_a M[a][c] x[a]
template <int PID> __device__ __forceinline__ void p_fwd(const float* rec, const f2* x, f2 w, f2* acc) {
  using P = PT<PID>;
  if constexpr (PID == P000) {
    acc[0] = fma2(w, x[0], acc[0]);
  } else if constexpr (PID == P011 || PID == P022) {
    const f2 t = mul2(w, x[0]);
#pragma unroll
    for (int c = 0; c < P::D3; ++c) acc[c] = fma2(rec[P::YO + c], t, acc[c]);
  } else if constexpr (PID == P101 || PID == P202) {
#pragma unroll
    for (int c = 0; c < P::D3; ++c) acc[c] = fma2(w, x[c], acc[c]);
  } else if constexpr (PID == P110 || PID == P220) {
    f2 s = mul2(rec[P::YO], x[0]);
#pragma unroll
    for (int a = 1; a < P::D1; ++a) s = fma2(rec[P::YO + a], x[a], s);
    acc[0] = fma2(mul2((PID == P110) ? kInvSqrt3 : kInvSqrt5, w), s, acc[0]);
  } else {
    float M[(P::D1 * P::D3 + 3) / 4 * 4];
    load_M<PID>(rec, M);
#pragma unroll
    for (int c = 0; c < P::D3; ++c) {
      f2 s = mul2(M[c], x[0]);
#pragma unroll
      for (int a = 1; a < P::D1; ++a) s = fma2(M[a * P::D3 + c], x[a], s);
      acc[c] = fma2(w, s, acc[c]);
    }
  }
}

// reverse: dw = sum_{a,c} M[a][c] x[a] g[c];  dY[b] += w sum_{a,c} C[a,b,c] x[a] g[c];  dx[a] += w sum_c M[a][c] g[c]
template <int PID, bool DX>
__device__ __forceinline__ void p_bwd(const float* rec, const f2* x, const f2* g, f2 w, f2& dw, f2* dY, f2* dx) {
  using P = PT<PID>;
  if constexpr (PID == P000) {
    dw = mul2(x[0], g[0]);
    if constexpr (DX) dx[0] = fma2(w, g[0], dx[0]);
  } else if constexpr (PID == P011 || PID == P022) {
    f2 s = B200_ZERO2;
    const f2 t = mul2(w, x[0]);
#pragma unroll
    for (int c = 0; c < P::D3; ++c) {
      s = fma2(rec[P::YO + c], g[c], s);
      dY[P::YO + c] = fma2(t, g[c], dY[P::YO + c]);
    }
    dw = mul2(x[0], s);
    if constexpr (DX) dx[0] = fma2(w, s, dx[0]);
  } else if constexpr (PID == P101 || PID == P202) {
    f2 s = B200_ZERO2;
#pragma unroll
    for (int c = 0; c < P::D3; ++c) {
      s = fma2(x[c], g[c], s);
      if constexpr (DX) dx[c] = fma2(w, g[c], dx[c]);
    }
    dw = s;
  } else if constexpr (PID == P110 || PID == P220) {
    constexpr float k = (PID == P110) ? kInvSqrt3 : kInvSqrt5;
    const f2 gk = mul2(k, g[0]);
    const f2 t = mul2(w, gk);
    f2 s = B200_ZERO2;
#pragma unroll
    for (int a = 0; a < P::D1; ++a) {
      s = fma2(rec[P::YO + a], x[a], s);
      dY[P::YO + a] = fma2(t, x[a], dY[P::YO + a]);
      if constexpr (DX) dx[a] = fma2(rec[P::YO + a], t, dx[a]);
    }
    dw = mul2(gk, s);
  } else {
    float M[(P::D1 * P::D3 + 3) / 4 * 4];
    load_M<PID>(rec, M);
    f2 G[P::D1 * P::D3];
    f2 s = B200_ZERO2;
#pragma unroll
    for (int a = 0; a < P::D1; ++a) {
      f2 ta = B200_ZERO2;   // sum_c M[a][c] g[c]: shared by dw and dx
#pragma unroll
      for (int c = 0; c < P::D3; ++c) {
        G[a * P::D3 + c] = mul2(x[a], g[c]);
        ta = fma2(M[a * P::D3 + c], g[c], ta);
      }
      s = fma2(x[a], ta, s);
      if constexpr (DX) dx[a] = fma2(w, ta, dx[a]);
    }
    dw = s;
    f2 t[9];
#pragma unroll
    for (int b = 0; b < 9; ++b) t[b] = B200_ZERO2;
    cg_dY<PID>(G, t);
#pragma unroll
    for (int b = 0; b < 2 * P::L2 + 1; ++b) dY[P::YO + b] = fma2(w, t[P::YO + b], dY[P::YO + b]);
  }
}

// ---- packed table: one channel pair of one path -----------------------------------------------------
struct TabEntry {
  float4 ab;    // {a0, a1, b0, b1}
  unsigned c;   // 2 x fp16 {c0, c1} * kTabCDScale
  unsigned dd;  // 2 x fp16 {d0, d1} * kTabCDScale (reverse pass only)
};
__device__ __forceinline__ f2 half2_as_f2(unsigned u) {
  const float2 v = __half22float2(*reinterpret_cast<const __half2*>(&u));
  return mk(v.x, v.y);
}
// forward: w(t) = a + t (b + st c),  st = t / kTabCDScale; the cubic term (4e-8 of w) is left out of the VALUE
__device__ __forceinline__ f2 tab_w(const TabEntry& e, float t, float st) {
  return fma2(t, fma2(st, half2_as_f2(e.c), mk(e.ab.z, e.ab.w)), mk(e.ab.x, e.ab.y));
}
// reverse: w with the cubic term and dw/dt = b + 2 st (c + 1.5 t d)
__device__ __forceinline__ void tab_wd(const TabEntry& e, float t, float st, f2& w, f2& wd) {
  const f2 c = half2_as_f2(e.c), dd = half2_as_f2(e.dd);
  const f2 b = mk(e.ab.z, e.ab.w);
  w = fma2(t, fma2(st, fma2(t, dd, c), b), mk(e.ab.x, e.ab.y));
  wd = fma2(2.f * st, fma2(1.5f * t, dd, c), b);
}

struct Tab {
  const float4* AB;
  const unsigned* C;
  const unsigned* D;
  int np;        // channel pairs per table row (wn / 2)
  float inv_h;
  int n_grid;
};

// ---- path lists ------------------------------------------------------------------------------------------
template <int... PIDS> struct PL;
template <> struct PL<> {
  static constexpr int NACC = 0, NP = 0;
  template <bool WITH_D> __device__ __forceinline__ static void load_tab(const ConvDesc&, const Tab&, int, int, TabEntry*) {}
  __device__ __forceinline__ static void fwd(const float*, const f2*, const TabEntry*, float, float, f2*) {}
  template <bool DX>
  __device__ __forceinline__ static void bwd(const float*, const f2*, const f2*, const TabEntry*, float, float, f2*, f2*, f2&) {}
  __device__ __forceinline__ static void load_g(const ConvDesc&, const float*, int, float, f2*) {}
  __device__ __forceinline__ static void store_o(const ConvDesc&, float*, int, float, const f2*) {}
};
template <int PID, int... REST> struct PL<PID, REST...> {
  using P = PT<PID>;
  using R = PL<REST...>;
  static constexpr int NACC = P::D3 + R::NACC;
  static constexpr int NP = 1 + sizeof...(REST);
  // e[0..NP) <- table entries of channel pair q in table row `row`
  template <bool WITH_D>
  __device__ __forceinline__ static void load_tab(const ConvDesc& d, const Tab& tb, int row, int q, TabEntry* e) {
    const size_t idx = (size_t)row * tb.np + (d.woff[PID] >> 1) + q;
    e[0].ab = __ldg(tb.AB + idx);
    e[0].c = __ldg(tb.C + idx);
    if constexpr (WITH_D) e[0].dd = __ldg(tb.D + idx);
    R::template load_tab<WITH_D>(d, tb, row, q, e + 1);
  }
  __device__ __forceinline__ static void fwd(const float* rec, const f2* x, const TabEntry* e, float t, float st, f2* acc) {
    p_fwd<PID>(rec, x, tab_w(e[0], t, st), acc);
    R::fwd(rec, x, e + 1, t, st, acc + P::D3);
  }
  // drp += dw * dw/dt of every path: the weight gradient never leaves the registers
  template <bool DX>
  __device__ __forceinline__ static void bwd(const float* rec, const f2* x, const f2* g, const TabEntry* e, float t, float st,
                                             f2* dY, f2* dx, f2& drp) {
    f2 w, wd, dw;
    tab_wd(e[0], t, st, w, wd);
    p_bwd<PID, DX>(rec, x, g, w, dw, dY, dx);
    drp = fma2(dw, wd, drp);
    R::template bwd<DX>(rec, x, g + P::D3, e + 1, t, st, dY, dx, drp);
  }
  // g[c] <- scale * row[slot of channels (2q, 2q+1)]
  __device__ __forceinline__ static void load_g(const ConvDesc& d, const float* __restrict__ row, int q, float scale, f2* g) {
    const float* r = row + d.ooff[PID] + 2 * q * P::D3;
#pragma unroll
    for (int c = 0; c < P::D3; ++c) g[c] = mk(scale * __ldg(r + c), scale * __ldg(r + P::D3 + c));
    R::load_g(d, row, q, scale, g + P::D3);
  }
  __device__ __forceinline__ static void store_o(const ConvDesc& d, float* __restrict__ row, int q, float scale, const f2* o) {
    float* r = row + d.ooff[PID] + 2 * q * P::D3;
#pragma unroll
    for (int c = 0; c < P::D3; ++c) {
      r[c] = lo(o[c]) * scale;
      r[P::D3 + c] = hi(o[c]) * scale;
    }
    R::store_o(d, row, q, scale, o + P::D3);
  }
};

template <int ROLE, int LOUT> struct RolePaths;
template <> struct RolePaths<0, 2> { using L = PL<P000, P011, P022>; };
template <> struct RolePaths<1, 2> { using L = PL<P101, P110, P111, P121, P112, P122>; };
template <> struct RolePaths<2, 2> { using L = PL<P212, P222, P211, P202, P221, P220>; };
template <> struct RolePaths<0, 0> { using L = PL<P000>; };
template <> struct RolePaths<1, 0> { using L = PL<P110>; };
template <> struct RolePaths<2, 0> { using L = PL<P220>; };

// Source rows reach these kernels COMPONENT-MAJOR: inside every l > 0 block the layout is [m][channel] instead of the
// reference's [channel][m] (a transpose inside the block, done in place by block_transpose_kernel once per layer and
// direction).  X[m] = (channel 2q, channel 2q+1) is then one aligned 8-byte shared-memory load, conflict-free across the
// lanes, and the gradient of a component goes out as ONE red.global.add.v2.f32 per lane to 256 contiguous bytes per warp.
// (With [channel][m] rows the 32 lanes of a RED touched 24-40 sectors for 8 useful ones, and an L2 slice retires RED
// sectors at half the rate of reads: ncu, 40 % of the reverse pass's L2 time went into them.)
template <int D1> __device__ __forceinline__ void load_x_pair(const float* blk, int mul, int q, f2* X) {
#pragma unroll
  for (int a = 0; a < D1; ++a) X[a] = *reinterpret_cast<const f2*>(blk + a * mul + 2 * q);
}
template <int D1> __device__ __forceinline__ void red_x_pair(float* blk, int mul, int q, const f2* dx) {
#pragma unroll
  for (int a = 0; a < D1; ++a)
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(blk + a * mul + 2 * q), "f"(lo(dx[a])), "f"(hi(dx[a]))
                 : "memory");
}

// in-place [channel][m] <-> [m][channel] transpose of the 1e and 2e blocks of every row; one CTA per row
__global__ void block_transpose_kernel(float* __restrict__ x, int xdim, int m0, int m1, int m2, int inverse) {
  extern __shared__ float row_s[];
  float* r = x + (size_t)blockIdx.x * xdim;
  const int n = 3 * m1 + 5 * m2;
  for (int i = threadIdx.x; i < n; i += blockDim.x) row_s[i] = r[m0 + i];
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    // i indexes the OUTPUT; find the input element
    const bool first = i < 3 * m1;
    const int D = first ? 3 : 5, mul = first ? m1 : m2, off = first ? 0 : 3 * m1, j = i - off;
    int src;
    if (!inverse) {   // output [m][channel] <- input [channel][m]
      const int a = j / mul, u = j - a * mul;
      src = u * D + a;
    } else {          // output [channel][m] <- input [m][channel]
      const int u = j / D, a = j - u * D;
      src = a * mul + u;
    }
    r[m0 + i] = row_s[off + src];
  }
}

// ---- per-warp machinery ------------------------------------------------------------------------------------
struct Args {
  const int* rowptr;
  const int* src;
  const float* x;     // [n_src, xdim], component-major blocks (see above)
  const float* Y;
  const float* len;
  Tab tab;
  float scale;        // forward: 1 / denominator applied to the output; reverse: 1 / denominator applied to gout
  float* out;         // forward output [n_dst, odim]
  const float* gout;  // reverse: d E / d out
  float* dY;          // reverse: [E, 9], added into
  float* dr;          // reverse: [E], added into
  float* dx;          // reverse: [n_src, xdim] component-major blocks, added into (zeroed by the launcher); nullptr: not needed
};

// which parts of a source row a warp needs: its share of the 0e block and its main block (1e for warp A, 2e for warp B)
template <int WARP> struct Seg {
  int n0, g0, nM, gM, xw;
  // main / r0: the warp works on its main block / on its share of the 0e block (a launch may do only one of them)
  __device__ __host__ Seg(const ConvDesc& d, bool main, bool r0) {
    n0 = !r0 ? 0 : (WARP == 0 ? (d.m0 < 64 ? d.m0 : 64) : (d.m0 > 64 ? d.m0 - 64 : 0));
    g0 = WARP * 64;
    nM = !main ? 0 : (WARP == 0 ? 3 * d.m1 : 5 * d.m2);
    gM = WARP == 0 ? d.m0 : d.m0 + 3 * d.m1;
    xw = n0 + nM;
  }
};

// shared memory of one warp: 2 barriers, interval / coordinate per edge, source atoms of the two chunks in flight,
// [TE][REC] records, [2][TE][xw] source-row segments
template <int WARP> struct WarpSmem {
  static constexpr int REC = WARP == 0 ? REC_A : REC_B;
  float* xbuf;
  float* rec;
  int* ti;
  float* tt;
  int* sidx;
  uint64_t* bar;
  __device__ explicit WarpSmem(unsigned char* base) {
    bar = reinterpret_cast<uint64_t*>(base);
    ti = reinterpret_cast<int*>(base + 16);
    tt = reinterpret_cast<float*>(base + 16 + 4 * TE);
    sidx = reinterpret_cast<int*>(base + 16 + 8 * TE);
    rec = reinterpret_cast<float*>(base + 16 + 16 * TE);
    xbuf = rec + TE * REC;
  }
  __host__ __device__ static size_t bytes(int xw) { return 16 + 16 * TE + sizeof(float) * (TE * REC + 2 * TE * xw); }
};

// bulk copies of the chunk's source-row segments (lane k = edge k of the chunk), completion on `bar`.  An odd chunk gets
// one zero row behind its last edge: the edge loop runs over pairs of edges without a bounds test.
template <int WARP>
__device__ __forceinline__ void issue_x(const Seg<WARP>& sg, const float* __restrict__ x, int xdim, int s, int ne, int lane,
                                        float* dstbuf, uint64_t* bar, int* sidx) {
  if (lane == 0) tc::mbar_expect_tx(bar, (uint32_t)(ne * sg.xw * sizeof(float)));
  __syncwarp();
  if (lane < ne) {
    sidx[lane] = s;
    const float* row = x + (size_t)s * xdim;
    float* dst = dstbuf + lane * sg.xw;
    if (sg.n0 > 0) tc::bulk_copy_g2s(dst, row + sg.g0, sg.n0 * sizeof(float), bar);
    if (sg.nM > 0) tc::bulk_copy_g2s(dst + sg.n0, row + sg.gM, sg.nM * sizeof(float), bar);
  }
  if (ne & 1) {
    float* z = dstbuf + ne * sg.xw;
    for (int i = lane; i < sg.xw; i += 32) z[i] = 0.f;
  }
}

// staging of a chunk's records: lane = (edge k = lane >> 2, matrix j = lane & 3); split in a load half (issued before the
// compute of the previous chunk) and a compute + store half (after it).  Edges behind the end of the row get zero records.
struct StageIn {
  float Y[9];
  float r;
};
__device__ __forceinline__ void stage_load(const float* __restrict__ Y, const float* __restrict__ len, int c0, int ne, int lane,
                                           StageIn& in) {
  const int k = lane >> 2;
#pragma unroll
  for (int b = 0; b < 9; ++b) in.Y[b] = 0.f;
  in.r = 0.f;
  if (k < ne) {
    const float* Yg = Y + (size_t)(c0 + k) * 9;
#pragma unroll
    for (int b = 0; b < 9; ++b) in.Y[b] = __ldg(Yg + b);
    in.r = __ldg(len + c0 + k);
  }
}
template <int WARP, int LOUT, bool MAIN>
__device__ __forceinline__ void stage_store(const WarpSmem<WARP>& sm, const Tab& tb, int lane, const StageIn& in) {
  constexpr int REC = WarpSmem<WARP>::REC;
  const int k = lane >> 2, j = lane & 3;
  if (k >= TE) return;
  float* r = sm.rec + k * REC;
  if (j == 0) {
#pragma unroll
    for (int b = 0; b < 9; ++b) r[b] = in.Y[b];
    const int ti = min((int)(in.r * tb.inv_h), tb.n_grid - 1);
    sm.ti[k] = ti;
    sm.tt[k] = fmaf(in.r, tb.inv_h, -(float)ti);
  }
  if constexpr (LOUT == 2 && MAIN) {
    if constexpr (WARP == 0) {
      if (j == 0) cg_M<P111>(in.Y, r + PT<P111>::OFF);
      else if (j == 1) cg_M<P112>(in.Y, r + PT<P112>::OFF);
      else if (j == 2) cg_M<P121>(in.Y, r + PT<P121>::OFF);
      else cg_M<P122>(in.Y, r + PT<P122>::OFF);
    } else {
      if (j == 0) cg_M<P211>(in.Y, r + PT<P211>::OFF);
      else if (j == 1) cg_M<P212>(in.Y, r + PT<P212>::OFF);
      else if (j == 2) cg_M<P221>(in.Y, r + PT<P221>::OFF);
      else cg_M<P222>(in.Y, r + PT<P222>::OFF);
    }
  }
}

// what every lane of a warp knows about its share of the row
template <int WARP> struct LaneMap {
  int qM, half, q0, l0;      // main-role pair (clamped), half-warp (warp B), 0e pair in the row, 0e pair in the warp's segment
  bool actM, act0;
  __device__ LaneMap(const ConvDesc& d, int lane) {
    const int mulM = WARP == 0 ? d.m1 : d.m2;
    const int q = WARP == 0 ? lane : (lane & 15);
    half = WARP == 0 ? 0 : (lane >> 4);
    actM = 2 * q < mulM;
    qM = actM ? q : 0;          // idle lanes repeat pair 0 and never store
    act0 = 2 * (WARP * 32 + lane) < d.m0;
    l0 = act0 ? lane : 0;
    q0 = WARP * 32 + l0;
  }
};

// ---- forward ---------------------------------------------------------------------------------------------------
// MAIN: the warp works on its 1e / 2e block; R0: on its share of the 0e block.  Separate launches for the two keep the
// register count (accumulators of one role only) and the loop body small: twice the resident warps.
template <int WARP, int LOUT, bool MAIN, bool R0>
__device__ __forceinline__ void fwd_warp(const ConvDesc& d, const Args& a, unsigned char* smem_base, int row, int lane) {
  constexpr int MR = WARP == 0 ? 1 : 2;   // main role of the warp
  constexpr int D1 = 2 * MR + 1;
  constexpr int REC = WarpSmem<WARP>::REC;
  using LM = typename RolePaths<MR, LOUT>::L;
  using L0 = typename RolePaths<0, LOUT>::L;
  const Seg<WARP> sg(d, MAIN, R0);
  const WarpSmem<WARP> sm(smem_base);
  const LaneMap<WARP> lm(d, lane);
  const int e0 = __ldg(a.rowptr + row), e1 = __ldg(a.rowptr + row + 1);
  f2 accM[MAIN ? LM::NACC : 1], acc0[R0 ? L0::NACC : 1];
#pragma unroll
  for (int c = 0; c < (MAIN ? LM::NACC : 1); ++c) accM[c] = B200_ZERO2;
#pragma unroll
  for (int c = 0; c < (R0 ? L0::NACC : 1); ++c) acc0[c] = B200_ZERO2;

  if (lane == 0) {
    tc::mbar_init(sm.bar, 1);
    tc::mbar_init(sm.bar + 1, 1);
    tc::fence_mbar_init();
  }
  __syncwarp();
  const int nchunk = (e1 - e0 + TE - 1) / TE;
  uint32_t phase = 0;   // bit b: parity to wait for on barrier b
  if (nchunk > 0) {
    const int ne = min(TE, e1 - e0);
    const int s = lane < ne ? __ldg(a.src + e0 + lane) : 0;
    issue_x<WARP>(sg, a.x, d.xdim, s, ne, lane, sm.xbuf, sm.bar, sm.sidx);
    StageIn in;
    stage_load(a.Y, a.len, e0, ne, lane, in);
    stage_store<WARP, LOUT, MAIN>(sm, a.tab, lane, in);
    __syncwarp();
  }
  int s_next = (nchunk > 1 && lane < TE && e0 + TE + lane < e1) ? __ldg(a.src + e0 + TE + lane) : 0;
  for (int n = 0; n < nchunk; ++n) {
    const int c0 = e0 + n * TE, ne = min(TE, e1 - c0), cur = n & 1;
    const int ne_next = min(TE, e1 - c0 - TE);   // <= 0: no next chunk
    StageIn in;
    if (ne_next > 0) {
      issue_x<WARP>(sg, a.x, d.xdim, s_next, ne_next, lane, sm.xbuf + (cur ^ 1) * TE * sg.xw, sm.bar + (cur ^ 1),
                    sm.sidx + (cur ^ 1) * TE);
      stage_load(a.Y, a.len, c0 + TE, ne_next, lane, in);
      s_next = (lane < TE && c0 + 2 * TE + lane < e1) ? __ldg(a.src + c0 + 2 * TE + lane) : 0;
    }
    tc::mbar_wait(sm.bar + cur, (phase >> cur) & 1u);
    phase ^= 1u << cur;
    const float* xb = sm.xbuf + cur * TE * sg.xw;
    // (Loading an edge's table entries one edge ahead into a second register set was tried: slower, 18.2 -> 20.4 ms forward
    // and 38.3 -> 42.6 ms reverse per step - the kernels are bound by L2 throughput of the table reads, not by their latency.)
#pragma unroll 1
    for (int kk = 0; kk < ne; kk += 2) {   // pairs of edges; a zero record / zero row stands behind an odd end
      if constexpr (MAIN) {
        // warp A takes both edges of the pair, the half-warps of warp B one each
#pragma unroll kFwdUnroll
        for (int s2 = 0; s2 < (WARP == 0 ? 2 : 1); ++s2) {
          const int k = kk + (WARP == 0 ? s2 : lm.half);
          const float t = sm.tt[k], st = t * (1.0f / kTabCDScale);
          TabEntry e[LM::NP];
          LM::template load_tab<false>(d, a.tab, sm.ti[k], lm.qM, e);
          f2 X[D1];
          load_x_pair<D1>(xb + k * sg.xw + sg.n0, WARP == 0 ? d.m1 : d.m2, lm.qM, X);
          LM::fwd(sm.rec + k * REC, X, e, t, st, accM);
        }
      }
      if constexpr (R0) {
#pragma unroll kFwdUnroll
        for (int s2 = 0; s2 < 2; ++s2) {
          const int k = kk + s2;
          const float t = sm.tt[k], st = t * (1.0f / kTabCDScale);
          TabEntry e[L0::NP];
          L0::template load_tab<false>(d, a.tab, sm.ti[k], lm.q0, e);
          f2 X[1];
          load_x_pair<1>(xb + k * sg.xw, 0, lm.l0, X);
          L0::fwd(sm.rec + k * REC, X, e, t, st, acc0);
        }
      }
    }
    __syncwarp();
    if (ne_next > 0) stage_store<WARP, LOUT, MAIN>(sm, a.tab, lane, in);
    __syncwarp();
  }
  float* orow = a.out + (size_t)row * d.odim;
  if constexpr (MAIN) {
    if constexpr (WARP == 1) {   // the two half-warps hold partial sums over alternate edges
#pragma unroll
      for (int c = 0; c < LM::NACC; ++c)
        accM[c] = mk(lo(accM[c]) + __shfl_xor_sync(0xffffffffu, lo(accM[c]), 16),
                     hi(accM[c]) + __shfl_xor_sync(0xffffffffu, hi(accM[c]), 16));
    }
    if (lm.actM && lm.half == 0) LM::store_o(d, orow, lm.qM, a.scale, accM);
  }
  if constexpr (R0) {
    if (lm.act0) L0::store_o(d, orow, lm.q0, a.scale, acc0);
  }
}

// One kernel per warp type: all warps of a launch run the same loop, which then fits the instruction cache (with both
// types in one kernel 30 % of the issue slots were lost to instruction fetch, ncu stall_no_inst).  Warp w of a CTA owns
// destination row blockIdx.x * WPC + w and its own slice of the dynamic shared memory.
constexpr int WPC = 4;
template <int WT, int LOUT, bool MAIN, bool R0>
__global__ void __launch_bounds__(32 * WPC, (MAIN ? B200_PAIR_FWD_MINB : 6)) conv_pair_fwd_kernel(ConvDesc d, Args a, int n_dst) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * WPC + warp;
  if (row >= n_dst) return;
  fwd_warp<WT, LOUT, MAIN, R0>(d, a, smem + warp * WarpSmem<WT>::bytes(Seg<WT>(d, MAIN, R0).xw), row, lane);
}

// ---- reverse pass ------------------------------------------------------------------------------------------------
// Sums over the 32 lanes of v[0..15] (two edges x 8 components of d Y) and of the two d E / d t values; afterwards lane l
// holds in v[0] the total of value l >> 1 and in r0 the total of r0 (lanes 0-15) or r1 (lanes 16-31).
__device__ __forceinline__ void warp_reduce16(float* v, float& r0, float r1, int lane) {
  const bool b16 = lane & 16, b8 = lane & 8, b4 = lane & 4, b2 = lane & 2;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float send = b16 ? v[i] : v[i + 8], keep = b16 ? v[i + 8] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
  {
    const float send = b16 ? r0 : r1, keep = b16 ? r1 : r0;
    r0 = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float send = b8 ? v[i] : v[i + 4], keep = b8 ? v[i + 4] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float send = b4 ? v[i] : v[i + 2], keep = b4 ? v[i + 2] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  {
    const float send = b2 ? v[0] : v[1], keep = b2 ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  }
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) r0 += __shfl_xor_sync(0xffffffffu, r0, o);
}

template <int WARP, int LOUT, bool MAIN, bool R0, bool DX>
__device__ __forceinline__ void bwd_warp(const ConvDesc& d, const Args& a, unsigned char* smem_base, int row, int lane) {
  constexpr int MR = WARP == 0 ? 1 : 2;
  constexpr int D1 = 2 * MR + 1;
  constexpr int REC = WarpSmem<WARP>::REC;
  using LM = typename RolePaths<MR, LOUT>::L;
  using L0 = typename RolePaths<0, LOUT>::L;
  const Seg<WARP> sg(d, MAIN, R0);
  const WarpSmem<WARP> sm(smem_base);
  const LaneMap<WARP> lm(d, lane);
  const int e0 = __ldg(a.rowptr + row), e1 = __ldg(a.rowptr + row + 1);
  // output gradient of the row; idle lanes hold zeros, so everything they compute (linear in g) vanishes
  f2 gM[MAIN ? LM::NACC : 1], g0[R0 ? L0::NACC : 1];
  {
    const float* grow = a.gout + (size_t)row * d.odim;
    if constexpr (MAIN) LM::load_g(d, grow, lm.qM, lm.actM ? a.scale : 0.f, gM);
    if constexpr (R0) L0::load_g(d, grow, lm.q0, lm.act0 ? a.scale : 0.f, g0);
  }
  if (lane == 0) {
    tc::mbar_init(sm.bar, 1);
    tc::mbar_init(sm.bar + 1, 1);
    tc::fence_mbar_init();
  }
  __syncwarp();
  const int nchunk = (e1 - e0 + TE - 1) / TE;
  uint32_t phase = 0;
  if (nchunk > 0) {
    const int ne = min(TE, e1 - e0);
    const int s = lane < ne ? __ldg(a.src + e0 + lane) : 0;
    issue_x<WARP>(sg, a.x, d.xdim, s, ne, lane, sm.xbuf, sm.bar, sm.sidx);
    StageIn in;
    stage_load(a.Y, a.len, e0, ne, lane, in);
    stage_store<WARP, LOUT, MAIN>(sm, a.tab, lane, in);
    __syncwarp();
  }
  int s_next = (nchunk > 1 && lane < TE && e0 + TE + lane < e1) ? __ldg(a.src + e0 + TE + lane) : 0;
  for (int n = 0; n < nchunk; ++n) {
    const int c0 = e0 + n * TE, ne = min(TE, e1 - c0), cur = n & 1;
    const int ne_next = min(TE, e1 - c0 - TE);
    StageIn in;
    if (ne_next > 0) {
      issue_x<WARP>(sg, a.x, d.xdim, s_next, ne_next, lane, sm.xbuf + (cur ^ 1) * TE * sg.xw, sm.bar + (cur ^ 1),
                    sm.sidx + (cur ^ 1) * TE);
      stage_load(a.Y, a.len, c0 + TE, ne_next, lane, in);
      s_next = (lane < TE && c0 + 2 * TE + lane < e1) ? __ldg(a.src + c0 + 2 * TE + lane) : 0;
    }
    tc::mbar_wait(sm.bar + cur, (phase >> cur) & 1u);
    phase ^= 1u << cur;
    const float* xb = sm.xbuf + cur * TE * sg.xw;
    const int* sidx = sm.sidx + cur * TE;
#pragma unroll 1
    for (int kk = 0; kk < ne; kk += 2) {
      float v[16], r0 = 0.f, r1 = 0.f;   // d Y[1..8] of edges kk, kk + 1 and their d E / d t: this lane's share
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = 0.f;
      if constexpr (MAIN) {
#pragma unroll kBwdUnroll
        for (int s2 = 0; s2 < (WARP == 0 ? 2 : 1); ++s2) {
          const int k = kk + (WARP == 0 ? s2 : lm.half);
          const float t = sm.tt[k], st = t * (1.0f / kTabCDScale);
          TabEntry e[LM::NP];
          LM::template load_tab<true>(d, a.tab, sm.ti[k], lm.qM, e);
          f2 X[D1], dxv[D1], dY[9], drp = B200_ZERO2;
          load_x_pair<D1>(xb + k * sg.xw + sg.n0, WARP == 0 ? d.m1 : d.m2, lm.qM, X);
#pragma unroll
          for (int b = 0; b < 9; ++b) dY[b] = B200_ZERO2;
#pragma unroll
          for (int b = 0; b < D1; ++b) dxv[b] = B200_ZERO2;
          LM::template bwd<DX>(sm.rec + k * REC, X, gM, e, t, st, dY, dxv, drp);
          if constexpr (DX) {
            if (k < ne && lm.actM) red_x_pair<D1>(a.dx + (size_t)sidx[k] * d.xdim + sg.gM, WARP == 0 ? d.m1 : d.m2, lm.qM, dxv);
          }
          if constexpr (WARP == 0) {
#pragma unroll
            for (int m = 0; m < 8; ++m) {   // (no v[8 * s2 + m]: the loop over s2 may stay rolled)
              const float s = hsum(dY[1 + m]);
              if (s2 == 0) v[m] = s;
              else v[8 + m] = s;
            }
            if (s2 == 0) r0 = hsum(drp);
            else r1 = hsum(drp);
          } else {   // half-warp h holds edge kk + h
#pragma unroll
            for (int m = 0; m < 8; ++m) {
              const float s = hsum(dY[1 + m]);
              v[m] = lm.half ? 0.f : s;
              v[8 + m] = lm.half ? s : 0.f;
            }
            const float s = hsum(drp);
            r0 = lm.half ? 0.f : s;
            r1 = lm.half ? s : 0.f;
          }
        }
      }
      if constexpr (R0) {
#pragma unroll kBwdUnroll
      for (int s2 = 0; s2 < 2; ++s2) {
        const int k = kk + s2;
        const float t = sm.tt[k], st = t * (1.0f / kTabCDScale);
        TabEntry e[L0::NP];
        L0::template load_tab<true>(d, a.tab, sm.ti[k], lm.q0, e);
        f2 X[1], dxv[1], dY[9], drp = B200_ZERO2;
        load_x_pair<1>(xb + k * sg.xw, 0, lm.l0, X);
#pragma unroll
        for (int b = 0; b < 9; ++b) dY[b] = B200_ZERO2;
        dxv[0] = B200_ZERO2;
        L0::template bwd<DX>(sm.rec + k * REC, X, g0, e, t, st, dY, dxv, drp);
        if constexpr (DX) {
          if (k < ne && lm.act0) red_x_pair<1>(a.dx + (size_t)sidx[k] * d.xdim + sg.g0, 0, lm.l0, dxv);
        }
        if constexpr (LOUT == 2) {
#pragma unroll
          for (int m = 0; m < 8; ++m) {
            const float s = hsum(dY[1 + m]);
            if (s2 == 0) v[m] += s;
            else v[8 + m] += s;
          }
        }
        if (s2 == 0) r0 += hsum(drp);
        else r1 += hsum(drp);
      }
      }
      warp_reduce16(v, r0, r1, lane);
      const float tot_r = r0 * a.tab.inv_h;   // d t / d r = 1 / h
      if ((MAIN || LOUT == 2) && (lane & 1) == 0) {
        const int idx = lane >> 1, ke = kk + (idx >> 3), m = idx & 7;
        if (ke < ne) atomicAdd(a.dY + (size_t)(c0 + ke) * 9 + 1 + m, v[0]);
      }
      if ((lane & 15) == 0) {
        const int ke = kk + (lane >> 4);
        if (ke < ne) atomicAdd(a.dr + c0 + ke, tot_r);
      }
    }
    __syncwarp();
    if (ne_next > 0) stage_store<WARP, LOUT, MAIN>(sm, a.tab, lane, in);
    __syncwarp();
  }
}

template <int WT, int LOUT, bool MAIN, bool R0, bool DX>
__global__ void __launch_bounds__(32 * WPC, (MAIN ? B200_PAIR_BWD_MINB : 4)) conv_pair_bwd_kernel(ConvDesc d, Args a, int n_dst) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * WPC + warp;
  if (row >= n_dst) return;
  bwd_warp<WT, LOUT, MAIN, R0, DX>(d, a, smem + warp * WarpSmem<WT>::bytes(Seg<WT>(d, MAIN, R0).xw), row, lane);
}

}  // namespace pairk

// Shapes the channel-pair kernels take: multiplicities that fill at most the two warps of a row's CTA and keep every
// bulk-copied segment a multiple of 16 bytes (SevenNet-0: 128x0e [+ 64x1e + 32x2e]).  Others use the kernels of conv.cu.
bool conv_pair_supported(const ConvDesc& d) {
  const bool main_ok = (d.m1 == 0 && d.m2 == 0) || (d.m1 > 0 && d.m2 > 0);
  return (d.lout == 0 || d.lout == 2) && d.m0 > 0 && d.m0 <= 128 && d.m0 % 4 == 0 && d.m1 >= 0 && d.m1 <= 64 &&
         d.m1 % 4 == 0 && d.m2 >= 0 && d.m2 <= 32 && d.m2 % 4 == 0 && main_ok && d.wn % 2 == 0;
}

// in place: standard irreps-blocked rows <-> the component-major rows the kernels read (x) and write (d x)
int conv_pair_interleave(const ConvDesc& d, long long n_rows, float* x, int inverse, cudaStream_t st) {
  const int n = 3 * d.m1 + 5 * d.m2;
  if (n_rows == 0 || n == 0) return B200_OK;
  pairk::block_transpose_kernel<<<(unsigned)n_rows, 128, n * sizeof(float), st>>>(x, d.xdim, d.m0, d.m1, d.m2, inverse);
  return check_launch("conv_pair_interleave");
}

template <int WT> static size_t pair_smem_bytes(const ConvDesc& d, bool main, bool r0) {
  return pairk::WPC * pairk::WarpSmem<WT>::bytes(pairk::Seg<WT>(d, main, r0).xw);
}

static pairk::Tab make_tab(const ConvDesc& d, const void* packed, float inv_h, int n_grid) {
  pairk::Tab t;
  t.AB = reinterpret_cast<const float4*>(packed);
  t.np = d.wn / 2;
  t.C = reinterpret_cast<const unsigned*>(t.AB + (size_t)n_grid * t.np);
  t.D = t.C + (size_t)n_grid * t.np;
  t.inv_h = inv_h;
  t.n_grid = n_grid;
  return t;
}

template <int WT, class K>
static int launch_pair(K kernel, const ConvDesc& d, const pairk::Args& a, int n_dst, bool main, bool r0, cudaStream_t st) {
  const size_t sm = pair_smem_bytes<WT>(d, main, r0);
  if (sm > 48 * 1024 && cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm) != cudaSuccess) {
    set_error("conv_pair: cannot reserve %zu bytes of shared memory", sm);
    return B200_ERR_CUDA;
  }
  kernel<<<(n_dst + pairk::WPC - 1) / pairk::WPC, 32 * pairk::WPC, sm, st>>>(d, a, n_dst);
  return B200_OK;
}
// one launch per warp type and role set: the 0e share of both warp types, then (if the row has 1e / 2e channels) the main
// blocks.  KERNEL<WT, LOUT, MAIN, R0, ...>
#define B200_PAIR_ROLES(KERNEL, LOUT, ...)                                                                              \
  do {                                                                                                                  \
    rc = launch_pair<0>(pairk::KERNEL<0, LOUT, false, true __VA_ARGS__>, d, a, n_dst, false, true, st);                 \
    if (rc == B200_OK) rc = launch_pair<1>(pairk::KERNEL<1, LOUT, false, true __VA_ARGS__>, d, a, n_dst, false, true, st); \
    if (rc == B200_OK && main)                                                                                          \
      rc = launch_pair<0>(pairk::KERNEL<0, LOUT, true, false __VA_ARGS__>, d, a, n_dst, true, false, st);               \
    if (rc == B200_OK && main)                                                                                          \
      rc = launch_pair<1>(pairk::KERNEL<1, LOUT, true, false __VA_ARGS__>, d, a, n_dst, true, false, st);               \
  } while (0)

int conv_pair_fwd(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
                  const float* len, const void* packed_table, float inv_h, int n_grid, float denom, float* out,
                  cudaStream_t st) {
  B200_REQUIRE(conv_pair_supported(d), "conv_pair_fwd: unsupported irreps");
  if (n_dst == 0) return B200_OK;
  pairk::Args a{};
  a.rowptr = rowptr; a.src = src; a.x = x; a.Y = Y; a.len = len;
  a.tab = make_tab(d, packed_table, inv_h, n_grid);
  a.scale = 1.0f / denom;
  a.out = out;
  const bool main = d.m1 > 0;
  int rc;
  if (d.lout == 2) B200_PAIR_ROLES(conv_pair_fwd_kernel, 2);
  else B200_PAIR_ROLES(conv_pair_fwd_kernel, 0);
  if (rc != B200_OK) return rc;
  return check_launch("conv_pair_fwd", main ? 4 : 2);
}

int conv_pair_bwd(const ConvDesc& d, int n_dst, int n_src, const int* rowptr, const int* src, const float* x, const float* Y,
                  const float* len, const void* packed_table, float inv_h, int n_grid, const float* gout, float denom,
                  float* dY, float* dr, float* dx, cudaStream_t st) {
  B200_REQUIRE(conv_pair_supported(d), "conv_pair_bwd: unsupported irreps");
  if (dx) cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)n_src * d.xdim, st);
  if (n_dst == 0) return B200_OK;
  pairk::Args a{};
  a.rowptr = rowptr; a.src = src; a.x = x; a.Y = Y; a.len = len;
  a.tab = make_tab(d, packed_table, inv_h, n_grid);
  a.scale = 1.0f / denom;
  a.gout = gout; a.dY = dY; a.dr = dr; a.dx = dx;
  const bool main = d.m1 > 0;
  int rc;
#define B200_COMMA ,
  if (d.lout == 2) {
    if (dx) B200_PAIR_ROLES(conv_pair_bwd_kernel, 2, B200_COMMA true);
    else B200_PAIR_ROLES(conv_pair_bwd_kernel, 2, B200_COMMA false);
  } else {
    if (dx) B200_PAIR_ROLES(conv_pair_bwd_kernel, 0, B200_COMMA true);
    else B200_PAIR_ROLES(conv_pair_bwd_kernel, 0, B200_COMMA false);
  }
#undef B200_COMMA
  if (rc != B200_OK) return rc;
  return check_launch("conv_pair_bwd", main ? 4 : 2);
}

}  // namespace b200
