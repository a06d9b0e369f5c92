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

namespace b200 {
namespace pairk {

constexpr int TE = 8;          // edges per chunk
constexpr float kInvSqrt3 = 0.5773502691896258f;
constexpr float kInvSqrt5 = 0.4472135954999579f;

// per-edge shared-memory record (floats): Y[9] padded to 12, then the warp's four coupling matrices
constexpr int REC_A = 68, REC_B = 100;

using f2 = float2;
__device__ __forceinline__ f2 mk(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ f2 fma2(float s, f2 b, f2 c) { return __ffma2_rn(make_float2(s, s), b, c); }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ f2 mul2(float s, f2 b) { return __fmul2_rn(make_float2(s, s), b); }
#define B200_ZERO2 make_float2(0.f, 0.f)

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
