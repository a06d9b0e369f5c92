// Channel-wise (uvu) Clebsch-Gordan tensor product fused with the weight multiply and the
// sorted-segment sum to destination atoms, plus its two backward kernels.
//
// Reference semantics: sevenn/nn/convolution.py:105-121 (IrrepsConvolution.forward):
//   message = TensorProduct_uvu(x[edge_src], edge_attr, weight);  x = scatter_sum(message, edge_dst) / denominator
// No per-edge message or outer product ever reaches HBM: one CTA owns one destination atom (one CSR
// row), a thread owns one channel u of one input irrep block and a sub-list of its paths, walks the
// row's edges sequentially and keeps the (2 l3+1) partial sums of each path in registers.  The
// per-edge coupling matrices M_p[m1][m3] = sum_m2 C_p[m1,m2,m3] Y[m2] are shared by all channels and
// are staged once per edge in shared memory.
//
// Thread layout: blockDim = NG * mulpad, group g = tid / mulpad (warp-uniform), channel u = tid % mulpad.
#include "cg_generated.cuh"
#include "common.cuh"

namespace b200 {

// edges staged per chunk: the widest records (role 2, 400 B per edge, one warp per CTA) are staged 16 at a time so
// that shared memory does not cap the resident CTAs below the 32-per-SM limit
template <int ROLE> __host__ __device__ constexpr int te_of() { return ROLE == 2 ? 16 : 32; }
constexpr float kInvSqrt3 = 0.5773502691896258f;
constexpr float kInvSqrt5 = 0.4472135954999579f;

// ---- compile-time path table ------------------------------------------------------------------
template <int PID> struct PathT;
#define B200_PATH(PID, L1_, L2_, L3_, TAB_)                                                \
  template <> struct PathT<PID> {                                                          \
    static constexpr int L1 = L1_, L2 = L2_, L3 = L3_, D1 = 2 * L1_ + 1, D3 = 2 * L3_ + 1; \
    static constexpr int TAB = TAB_; /* offset of M in the per-edge record, -1: trivial */  \
  };
B200_PATH(P000, 0, 0, 0, -1)
B200_PATH(P011, 0, 1, 1, -1)
B200_PATH(P022, 0, 2, 2, -1)
B200_PATH(P101, 1, 0, 1, -1)
B200_PATH(P110, 1, 1, 0, -1)
B200_PATH(P111, 1, 1, 1, 12)
B200_PATH(P112, 1, 1, 2, 24)
B200_PATH(P121, 1, 2, 1, 40)
B200_PATH(P122, 1, 2, 2, 52)
B200_PATH(P202, 2, 0, 2, -1)
B200_PATH(P211, 2, 1, 1, 12)
B200_PATH(P212, 2, 1, 2, 28)
B200_PATH(P220, 2, 2, 0, -1)
B200_PATH(P221, 2, 2, 1, 56)
B200_PATH(P222, 2, 2, 2, 72)
#undef B200_PATH

// per-edge shared-memory record: Y[9] (padded to 12) followed by the role's coupling matrices
template <int ROLE, int LOUT> struct RoleT {
  static constexpr int REC = (LOUT == 0 || ROLE == 0) ? 12 : (ROLE == 1 ? 68 : 100);
  static constexpr int NTAB = (LOUT == 0 || ROLE == 0) ? 0 : 4;
  static constexpr int D1 = 2 * ROLE + 1;
};

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
template <int PID> __device__ __forceinline__ void cg_dY(const float* G, float* dY) {
  if constexpr (PID == P111) cg_dY_111(G, dY);
  else if constexpr (PID == P112) cg_dY_112(G, dY);
  else if constexpr (PID == P121) cg_dY_121(G, dY);
  else if constexpr (PID == P122) cg_dY_122(G, dY);
  else if constexpr (PID == P211) cg_dY_211(G, dY);
  else if constexpr (PID == P212) cg_dY_212(G, dY);
  else if constexpr (PID == P221) cg_dY_221(G, dY);
  else if constexpr (PID == P222) cg_dY_222(G, dY);
}

template <int PID> __device__ __forceinline__ void load_M(const float* rec, float* M) {
  constexpr int N = PathT<PID>::D1 * PathT<PID>::D3;
  constexpr int N4 = (N + 3) / 4;
  const float4* q = reinterpret_cast<const float4*>(rec + PathT<PID>::TAB);
#pragma unroll
  for (int i = 0; i < N4; ++i) {
    float4 v = q[i];
    M[4 * i] = v.x; M[4 * i + 1] = v.y; M[4 * i + 2] = v.z; M[4 * i + 3] = v.w;
  }
}

// ---- per-path arithmetic ------------------------------------------------------------------------
// forward: acc[c] += w * sum_a M[a][c] x[a]
template <int PID> __device__ __forceinline__ void p_fwd(const float* rec, const float* x, float w, float* acc) {
  using P = PathT<PID>;
  if constexpr (PID == P000) {
    acc[0] = fmaf(w, x[0], acc[0]);
  } else if constexpr (PID == P011 || PID == P022) {
    const float t = w * x[0];
    constexpr int yo = (PID == P011) ? 1 : 4;
#pragma unroll
    for (int c = 0; c < P::D3; ++c) acc[c] = fmaf(t, rec[yo + c], acc[c]);
  } else if constexpr (PID == P101 || PID == P202) {
#pragma unroll
    for (int c = 0; c < P::D3; ++c) acc[c] = fmaf(w, x[c], acc[c]);
  } else if constexpr (PID == P110 || PID == P220) {
    constexpr int yo = (PID == P110) ? 1 : 4;
    float s = 0.f;
#pragma unroll
    for (int a = 0; a < P::D1; ++a) s = fmaf(x[a], rec[yo + a], s);
    acc[0] = fmaf(w * ((PID == P110) ? kInvSqrt3 : kInvSqrt5), s, acc[0]);
  } else {
    float M[(P::D1 * P::D3 + 3) / 4 * 4];
    load_M<PID>(rec, M);
#pragma unroll
    for (int c = 0; c < P::D3; ++c) {
      float s = 0.f;
#pragma unroll
      for (int a = 0; a < P::D1; ++a) s = fmaf(M[a * P::D3 + c], x[a], s);
      acc[c] = fmaf(w, s, acc[c]);
    }
  }
}

// backward wrt x: dx[a] += w * sum_c M[a][c] g[c]
template <int PID> __device__ __forceinline__ void p_bwdx(const float* rec, const float* g, float w, float* dx) {
  using P = PathT<PID>;
  if constexpr (PID == P000) {
    dx[0] = fmaf(w, g[0], dx[0]);
  } else if constexpr (PID == P011 || PID == P022) {
    constexpr int yo = (PID == P011) ? 1 : 4;
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < P::D3; ++c) s = fmaf(rec[yo + c], g[c], s);
    dx[0] = fmaf(w, s, dx[0]);
  } else if constexpr (PID == P101 || PID == P202) {
#pragma unroll
    for (int c = 0; c < P::D3; ++c) dx[c] = fmaf(w, g[c], dx[c]);
  } else if constexpr (PID == P110 || PID == P220) {
    constexpr int yo = (PID == P110) ? 1 : 4;
    const float t = w * g[0] * ((PID == P110) ? kInvSqrt3 : kInvSqrt5);
#pragma unroll
    for (int a = 0; a < P::D1; ++a) dx[a] = fmaf(t, rec[yo + a], dx[a]);
  } else {
    float M[(P::D1 * P::D3 + 3) / 4 * 4];
    load_M<PID>(rec, M);
#pragma unroll
    for (int a = 0; a < P::D1; ++a) {
      float s = 0.f;
#pragma unroll
      for (int c = 0; c < P::D3; ++c) s = fmaf(M[a * P::D3 + c], g[c], s);
      dx[a] = fmaf(w, s, dx[a]);
    }
  }
}

// backward wrt the per-edge weight and the spherical harmonics:
//   dw = sum_{a,c} M[a][c] x[a] g[c];   dY[b] += w * sum_{a,c} C[a,b,c] x[a] g[c]
// and, with DX, the gradient wrt the gathered source row from the same intermediates:
//   dx[a] += w * sum_c M[a][c] g[c]
template <int PID, bool DX>
__device__ __forceinline__ void p_bwde(const float* rec, const float* x, const float* g, float w, float& dw, float* dY,
                                       float* dx) {
  using P = PathT<PID>;
  if constexpr (PID == P000) {
    dw = x[0] * g[0];
    if constexpr (DX) dx[0] = fmaf(w, g[0], dx[0]);
  } else if constexpr (PID == P011 || PID == P022) {
    constexpr int yo = (PID == P011) ? 1 : 4;
    float s = 0.f;
    const float t = w * x[0];
#pragma unroll
    for (int c = 0; c < P::D3; ++c) {
      s = fmaf(rec[yo + c], g[c], s);
      dY[yo + c] = fmaf(t, g[c], dY[yo + c]);
    }
    dw = x[0] * s;
    if constexpr (DX) dx[0] = fmaf(w, s, dx[0]);
  } else if constexpr (PID == P101 || PID == P202) {
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < P::D3; ++c) {
      s = fmaf(x[c], g[c], s);
      if constexpr (DX) dx[c] = fmaf(w, g[c], dx[c]);
    }
    dw = s;
  } else if constexpr (PID == P110 || PID == P220) {
    constexpr int yo = (PID == P110) ? 1 : 4;
    constexpr float k = (PID == P110) ? kInvSqrt3 : kInvSqrt5;
    float s = 0.f;
    const float t = w * g[0] * k;
#pragma unroll
    for (int a = 0; a < P::D1; ++a) {
      s = fmaf(x[a], rec[yo + a], s);
      dY[yo + a] = fmaf(t, x[a], dY[yo + a]);
      if constexpr (DX) dx[a] = fmaf(t, rec[yo + a], dx[a]);
    }
    dw = g[0] * k * s;
  } else {
    float M[(P::D1 * P::D3 + 3) / 4 * 4];
    load_M<PID>(rec, M);
    float G[P::D1 * P::D3];
    float s = 0.f;
#pragma unroll
    for (int a = 0; a < P::D1; ++a) {
      float ta = 0.f;   // sum_c M[a][c] g[c]: shared by dw and dx
#pragma unroll
      for (int c = 0; c < P::D3; ++c) {
        G[a * P::D3 + c] = x[a] * g[c];
        ta = fmaf(M[a * P::D3 + c], g[c], ta);
      }
      s = fmaf(x[a], ta, s);
      if constexpr (DX) dx[a] = fmaf(w, ta, dx[a]);
    }
    dw = s;
    float t[9];
#pragma unroll
    for (int b = 0; b < 9; ++b) t[b] = 0.f;
    cg_dY<PID>(G, t);
    constexpr int yo = (P::L2 == 1) ? 1 : 4;
#pragma unroll
    for (int b = 0; b < 2 * P::L2 + 1; ++b) dY[yo + b] = fmaf(w, t[yo + b], dY[yo + b]);
  }
}

// ---- path lists (one per thread group) ----------------------------------------------------------
// Variants of the passes over per-edge weights for the second-order pass of the training path (higher_order.py,
// ConvB.backward): the weights of a path are scaled on the way in and its weight gradient on the way out (the variation of
// the filter acts on the l2 > 0 paths only: scale 0 on the others), and the weight gradient can be added into its array or
// left out.
struct ConvMode {
  float wscale[kNumPaths];
  int dw_mode;   // 0: d weight overwritten, 1: added into, 2: not computed
};
static ConvMode plain_mode() {
  ConvMode m;
  for (int p = 0; p < kNumPaths; ++p) m.wscale[p] = 1.f;
  m.dw_mode = 0;
  return m;
}

template <int... PIDS> struct PL;
template <> struct PL<> {
  static constexpr int NACC = 0;
  static constexpr int NP = 0;
  template <bool TAB> __device__ __forceinline__ static void load_w(const ConvDesc&, const float*, int, float, float*) {}
  __device__ __forceinline__ static void load_wd(const ConvDesc&, const float*, int, float, float*, float*) {}
  __device__ __forceinline__ static void fwd(const float*, const float*, const float*, float*) {}
  __device__ __forceinline__ static void bwdx(const float*, const float*, const float*, float*) {}
  template <bool DX>
  __device__ __forceinline__ static void bwde(const float*, const float*, const float*, const float*, float*, float*, float*) {}
  __device__ __forceinline__ static void load_o(const ConvDesc&, const float*, int, float, float*) {}
  __device__ __forceinline__ static void store_o(const ConvDesc&, float*, int, float, const float*) {}
  __device__ __forceinline__ static void store_w(const ConvDesc&, float*, int, const float*) {}
  __device__ __forceinline__ static void load_ws(const ConvDesc&, const ConvMode&, const float*, int, float*) {}
  __device__ __forceinline__ static void store_ws(const ConvDesc&, const ConvMode&, float*, int, const float*) {}
};
template <int PID, int... REST> struct PL<PID, REST...> {
  using P = PathT<PID>;
  using R = PL<REST...>;
  static constexpr int NACC = P::D3 + R::NACC;
  static constexpr int NP = 1 + sizeof...(REST);
  // wv[0..NP) <- weights of channel u for this edge.  TAB: wrow is the edge's row of the tabulated radial function
  // (float4 {a, b, c, d} per weight, radial_table.cu) and the weight is the cubic a + b t + c t^2 + d t^3 at the edge's t
  template <bool TAB>
  __device__ __forceinline__ static void load_w(const ConvDesc& d, const float* __restrict__ wrow, int u, float t, float* wv) {
    const int k = d.woff[PID] + u;
    if constexpr (TAB) {
      const float4 cf = __ldg(reinterpret_cast<const float4*>(wrow) + k);
      wv[0] = fmaf(fmaf(fmaf(cf.w, t, cf.z), t, cf.y), t, cf.x);
    } else {
      wv[0] = __ldg(wrow + k);
    }
    R::template load_w<TAB>(d, wrow, u, t, wv + 1);
  }
  // tabulated weights and their derivative with respect to t
  __device__ __forceinline__ static void load_wd(const ConvDesc& d, const float* __restrict__ wrow, int u, float t, float* wv,
                                                 float* wdv) {
    const int k = d.woff[PID] + u;
    const float4 cf = __ldg(reinterpret_cast<const float4*>(wrow) + k);
    wv[0] = fmaf(fmaf(fmaf(cf.w, t, cf.z), t, cf.y), t, cf.x);
    wdv[0] = fmaf(fmaf(3.f * cf.w, t, 2.f * cf.z), t, cf.y);
    R::load_wd(d, wrow, u, t, wv + 1, wdv + 1);
  }
  __device__ __forceinline__ static void store_w(const ConvDesc& d, float* __restrict__ wrow, int u, const float* wv) {
    wrow[d.woff[PID] + u] = wv[0];
    R::store_w(d, wrow, u, wv + 1);
  }
  // per-edge weights from memory, scaled per path; their gradient scaled the same way, stored or added
  __device__ __forceinline__ static void load_ws(const ConvDesc& d, const ConvMode& md, const float* __restrict__ wrow, int u,
                                                 float* wv) {
    wv[0] = md.wscale[PID] * __ldg(wrow + d.woff[PID] + u);
    R::load_ws(d, md, wrow, u, wv + 1);
  }
  __device__ __forceinline__ static void store_ws(const ConvDesc& d, const ConvMode& md, float* __restrict__ wrow, int u,
                                                  const float* wv) {
    float* p = wrow + d.woff[PID] + u;
    const float v = md.wscale[PID] * wv[0];
    *p = md.dw_mode == 1 ? *p + v : v;
    R::store_ws(d, md, wrow, u, wv + 1);
  }
  __device__ __forceinline__ static void fwd(const float* rec, const float* x, const float* wv, float* acc) {
    p_fwd<PID>(rec, x, wv[0], acc);
    R::fwd(rec, x, wv + 1, acc + P::D3);
  }
  __device__ __forceinline__ static void bwdx(const float* rec, const float* g, const float* wv, float* dx) {
    p_bwdx<PID>(rec, g, wv[0], dx);
    R::bwdx(rec, g + P::D3, wv + 1, dx);
  }
  template <bool DX>
  __device__ __forceinline__ static void bwde(const float* rec, const float* x, const float* g, const float* wv, float* dwv,
                                              float* dY, float* dx) {
    p_bwde<PID, DX>(rec, x, g, wv[0], dwv[0], dY, dx);
    R::template bwde<DX>(rec, x, g + P::D3, wv + 1, dwv + 1, dY, dx);
  }
  // o[0..NACC) <- scale * row[slot of channel u]
  __device__ __forceinline__ static void load_o(const ConvDesc& d, const float* __restrict__ row, int u, float scale, float* o) {
#pragma unroll
    for (int c = 0; c < P::D3; ++c) o[c] = scale * __ldg(row + d.ooff[PID] + u * P::D3 + c);
    R::load_o(d, row, u, scale, o + P::D3);
  }
  __device__ __forceinline__ static void store_o(const ConvDesc& d, float* __restrict__ row, int u, float denom, const float* o) {
#pragma unroll
    for (int c = 0; c < P::D3; ++c) row[d.ooff[PID] + u * P::D3 + c] = o[c] / denom;
    R::store_o(d, row, u, denom, o + P::D3);
  }
};

// group -> path list, per (ROLE, LOUT)
template <int ROLE, int LOUT, int G> struct GroupT { using L = PL<>; };
template <> struct GroupT<0, 2, 0> { using L = PL<P000, P011, P022>; };
// One group per role: a thread owns ALL paths of its channel, so x[src], the dY warp reduction and the RED.ADD of
// d x[src] happen once per (edge, channel).  (Splitting a role's paths over 2-4 thread groups was measured slower:
// 2.4x more RED traffic and one shuffle reduction per group.)  The group index is kept for other splits.
template <> struct GroupT<1, 2, 0> { using L = PL<P101, P110, P111, P121, P112, P122>; };
template <> struct GroupT<2, 2, 0> { using L = PL<P212, P222, P211, P202, P221, P220>; };
template <> struct GroupT<0, 0, 0> { using L = PL<P000>; };
template <> struct GroupT<1, 0, 0> { using L = PL<P110>; };
template <> struct GroupT<2, 0, 0> { using L = PL<P220>; };
template <int ROLE, int LOUT> __host__ __device__ constexpr int num_groups() { return 1; }
constexpr int kMaxAcc = 22;   // output components a thread accumulates: 3+1+3+3+5+5 (role 1), 5+5+3+5+3+1 (role 2)

// ---- staging of one chunk of edges ----------------------------------------------------------------
// eid: optional indirection (edge ids sorted by source, for the backward-x pass); gidx: row to gather per edge
struct TabArg {          // tabulated radial weights (radial_table.cu); len == nullptr: per-edge weights come from memory
  const float* len;      // edge lengths [E]
  float inv_h;
  int n_grid;
  float* dr;             // reverse pass: d E / d r per edge
  int acc_dr;            // add into dr instead of overwriting
};

template <int ROLE, int LOUT>
__device__ __forceinline__ void stage_chunk(float* rec, int* s_e, int* s_g, const float* __restrict__ Y,
                                            const int* __restrict__ eid, const int* __restrict__ gidx, int c0, int ne,
                                            const TabArg* targ = nullptr, int* s_ti = nullptr, float* s_tt = nullptr) {
  using RT = RoleT<ROLE, LOUT>;
  constexpr int NI = RT::NTAB + 1;
  for (int it = threadIdx.x; it < ne * NI; it += blockDim.x) {
    const int k = it / NI, tab = it % NI;
    const int e = eid ? __ldg(eid + c0 + k) : c0 + k;
    const float* Yg = Y + (size_t)e * 9;
    float* r = rec + k * RT::REC;
    if (tab == RT::NTAB) {
#pragma unroll
      for (int b = 0; b < 9; ++b) r[b] = __ldg(Yg + b);
      s_e[k] = e;
      s_g[k] = __ldg(gidx + e);
      if (targ) {   // interval and local coordinate of the edge in the radial table
        const float r = __ldg(targ->len + e);
        const int ti = min((int)(r * targ->inv_h), targ->n_grid - 1);
        s_ti[k] = ti;
        s_tt[k] = fmaf(r, targ->inv_h, -(float)ti);
      }
    } else if constexpr (RT::NTAB > 0) {
      float Yl[9];
#pragma unroll
      for (int b = 0; b < 9; ++b) Yl[b] = __ldg(Yg + b);
      if constexpr (ROLE == 1) {
        if (tab == 0) cg_M<P111>(Yl, r + PathT<P111>::TAB);
        else if (tab == 1) cg_M<P112>(Yl, r + PathT<P112>::TAB);
        else if (tab == 2) cg_M<P121>(Yl, r + PathT<P121>::TAB);
        else cg_M<P122>(Yl, r + PathT<P122>::TAB);
      } else {
        if (tab == 0) cg_M<P211>(Yl, r + PathT<P211>::TAB);
        else if (tab == 1) cg_M<P212>(Yl, r + PathT<P212>::TAB);
        else if (tab == 2) cg_M<P221>(Yl, r + PathT<P221>::TAB);
        else cg_M<P222>(Yl, r + PathT<P222>::TAB);
      }
    }
  }
}

template <int ROLE> __device__ __forceinline__ int role_mul(const ConvDesc& d) { return ROLE == 0 ? d.m0 : (ROLE == 1 ? d.m1 : d.m2); }
template <int ROLE> __device__ __forceinline__ int role_xoff(const ConvDesc& d) {
  return ROLE == 0 ? 0 : (ROLE == 1 ? d.m0 : d.m0 + 3 * d.m1);
}

// ---- forward ------------------------------------------------------------------------------------
template <int ROLE, int LOUT, int G, bool TAB>
__device__ __forceinline__ void fwd_group(const ConvDesc& d, const float* rec, const int* s_e, const int* s_g, int ne,
                                          const float* __restrict__ x, const float* __restrict__ w, int u, float* acc,
                                          const int* s_ti, const float* s_tt, const ConvMode& md) {
  using L = typename GroupT<ROLE, LOUT, G>::L;
  constexpr int D1 = RoleT<ROLE, LOUT>::D1;
  const int xo = role_xoff<ROLE>(d) + u * D1;
#pragma unroll 2
  for (int k = 0; k < ne; ++k) {
    float xv[D1], wv[L::NP > 0 ? L::NP : 1];
    const float* xr = x + (size_t)s_g[k] * d.xdim + xo;
#pragma unroll
    for (int a = 0; a < D1; ++a) xv[a] = __ldg(xr + a);
    if constexpr (TAB) L::template load_w<true>(d, w + (size_t)s_ti[k] * 4 * d.wn, u, s_tt[k], wv);
    else L::load_ws(d, md, w + (size_t)s_e[k] * d.wn, u, wv);
    L::fwd(rec + k * RoleT<ROLE, LOUT>::REC, xv, wv, acc);
  }
}

template <int ROLE, int LOUT, bool TAB>
__global__ void __launch_bounds__(256)
conv_fwd_kernel(ConvDesc d, int mulpad, const int* __restrict__ rowptr, const int* __restrict__ src,
                const float* __restrict__ x, const float* __restrict__ Y, const float* __restrict__ w, float denom,
                float* __restrict__ out, TabArg tab, ConvMode md) {
  using RT = RoleT<ROLE, LOUT>;
  constexpr int TE = te_of<ROLE>();
  __shared__ __align__(16) float rec[TE * RT::REC];
  __shared__ int s_e[TE], s_g[TE];
  __shared__ int s_ti[TAB ? TE : 1];
  __shared__ float s_tt[TAB ? TE : 1];
  const int i = blockIdx.x;
  const int e0 = rowptr[i], e1 = rowptr[i + 1];
  const int grp = threadIdx.x / mulpad, u = threadIdx.x % mulpad;
  const bool active = u < role_mul<ROLE>(d);
  float acc[kMaxAcc];
#pragma unroll
  for (int c = 0; c < kMaxAcc; ++c) acc[c] = 0.f;
  for (int c0 = e0; c0 < e1; c0 += TE) {
    const int ne = min(TE, e1 - c0);
    __syncthreads();
    stage_chunk<ROLE, LOUT>(rec, s_e, s_g, Y, nullptr, src, c0, ne, TAB ? &tab : nullptr, s_ti, s_tt);
    __syncthreads();
    if (active) {
      if (grp == 0) fwd_group<ROLE, LOUT, 0, TAB>(d, rec, s_e, s_g, ne, x, w, u, acc, s_ti, s_tt, md);
      else if (grp == 1) fwd_group<ROLE, LOUT, 1, TAB>(d, rec, s_e, s_g, ne, x, w, u, acc, s_ti, s_tt, md);
      else if (grp == 2) fwd_group<ROLE, LOUT, 2, TAB>(d, rec, s_e, s_g, ne, x, w, u, acc, s_ti, s_tt, md);
      else fwd_group<ROLE, LOUT, 3, TAB>(d, rec, s_e, s_g, ne, x, w, u, acc, s_ti, s_tt, md);
    }
  }
  if (active) {
    float* orow = out + (size_t)i * d.odim;
    if (grp == 0) GroupT<ROLE, LOUT, 0>::L::store_o(d, orow, u, denom, acc);
    else if (grp == 1) GroupT<ROLE, LOUT, 1>::L::store_o(d, orow, u, denom, acc);
    else if (grp == 2) GroupT<ROLE, LOUT, 2>::L::store_o(d, orow, u, denom, acc);
    else GroupT<ROLE, LOUT, 3>::L::store_o(d, orow, u, denom, acc);
  }
}

// ---- backward wrt x[src] --------------------------------------------------------------------------
// One CTA per SOURCE atom j: walks the edges whose source is j (eid = edge ids sorted by source,
// rrowptr = their CSR), gathers the output-gradient row of each edge's destination.
template <int ROLE, int LOUT, int G>
__device__ __forceinline__ void bwdx_group(const ConvDesc& d, const float* rec, const int* s_e, const int* s_g, int ne,
                                           const float* __restrict__ gout, const float* __restrict__ w, float inv_denom,
                                           int u, float* dx) {
  using L = typename GroupT<ROLE, LOUT, G>::L;
#pragma unroll 2
  for (int k = 0; k < ne; ++k) {
    float gv[L::NACC > 0 ? L::NACC : 1], wv[L::NP > 0 ? L::NP : 1];
    L::load_o(d, gout + (size_t)s_g[k] * d.odim, u, inv_denom, gv);
    L::template load_w<false>(d, w + (size_t)s_e[k] * d.wn, u, 0.f, wv);
    L::bwdx(rec + k * RoleT<ROLE, LOUT>::REC, gv, wv, dx);
  }
}

template <int ROLE, int LOUT>
__global__ void __launch_bounds__(256)
conv_bwdx_kernel(ConvDesc d, int mulpad, const int* __restrict__ rrowptr, const int* __restrict__ eid,
                 const int* __restrict__ dst, const float* __restrict__ gout, const float* __restrict__ Y,
                 const float* __restrict__ w, float inv_denom, float* __restrict__ dxo) {
  using RT = RoleT<ROLE, LOUT>;
  constexpr int D1 = RT::D1;
  constexpr int NG = num_groups<ROLE, LOUT>();
  constexpr int TE = te_of<ROLE>();
  __shared__ __align__(16) float rec[TE * RT::REC];
  __shared__ int s_e[TE], s_g[TE];
  extern __shared__ float s_part[];  // [(NG-1)][mulpad][D1] partial sums of groups 1..NG-1
  const int j = blockIdx.x;
  const int e0 = rrowptr[j], e1 = rrowptr[j + 1];
  const int grp = threadIdx.x / mulpad, u = threadIdx.x % mulpad;
  const bool active = u < role_mul<ROLE>(d);
  float dx[D1];
#pragma unroll
  for (int a = 0; a < D1; ++a) dx[a] = 0.f;
  for (int c0 = e0; c0 < e1; c0 += TE) {
    const int ne = min(TE, e1 - c0);
    __syncthreads();
    stage_chunk<ROLE, LOUT>(rec, s_e, s_g, Y, eid, dst, c0, ne);
    __syncthreads();
    if (active) {
      if (grp == 0) bwdx_group<ROLE, LOUT, 0>(d, rec, s_e, s_g, ne, gout, w, inv_denom, u, dx);
      else if (grp == 1) bwdx_group<ROLE, LOUT, 1>(d, rec, s_e, s_g, ne, gout, w, inv_denom, u, dx);
      else if (grp == 2) bwdx_group<ROLE, LOUT, 2>(d, rec, s_e, s_g, ne, gout, w, inv_denom, u, dx);
      else bwdx_group<ROLE, LOUT, 3>(d, rec, s_e, s_g, ne, gout, w, inv_denom, u, dx);
    }
  }
  if constexpr (NG > 1) {
    if (active && grp > 0) {
#pragma unroll
      for (int a = 0; a < D1; ++a) s_part[((grp - 1) * mulpad + u) * D1 + a] = dx[a];
    }
    __syncthreads();
    if (active && grp == 0) {
      for (int g = 1; g < NG; ++g)
#pragma unroll
        for (int a = 0; a < D1; ++a) dx[a] += s_part[((g - 1) * mulpad + u) * D1 + a];
    }
  }
  if (active && grp == 0) {
    float* r = dxo + (size_t)j * d.xdim + role_xoff<ROLE>(d) + u * D1;
#pragma unroll
    for (int a = 0; a < D1; ++a) r[a] = dx[a];
  }
}

// ---- backward wrt per-edge weights and spherical harmonics ---------------------------------------------
// lanes with (lane & 3) == 0 end up with the warp sum of component m = ((lane>>4)&1)*4 + ((lane>>3)&1)*2 + ((lane>>2)&1)
__device__ __forceinline__ float warp_reduce8(float* v, int lane) {
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

template <int ROLE, int LOUT, int G, bool DX, bool TAB>
__device__ __forceinline__ void bwde_group(const ConvDesc& d, const float* rec, const int* s_e, const int* s_g, int ne,
                                           const float* __restrict__ x, const float* __restrict__ w,
                                           float* __restrict__ dw, const float* gv, int u, bool active, float* s_dY,
                                           float* __restrict__ dxo, const int* s_ti, const float* s_tt, float* s_dr,
                                           const ConvMode& md) {
  using L = typename GroupT<ROLE, LOUT, G>::L;
  constexpr int D1 = RoleT<ROLE, LOUT>::D1;
  const int xo = role_xoff<ROLE>(d) + u * D1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int k = 0; k < ne; ++k) {
    float dY[9];
#pragma unroll
    for (int b = 0; b < 9; ++b) dY[b] = 0.f;
    float drp = 0.f;   // TAB: this channel's share of d E / d t of the edge
    if (active) {
      constexpr int NPW = L::NP > 0 ? L::NP : 1;
      float xv[D1], wv[NPW], dwv[NPW], wdv[TAB ? NPW : 1];
      const float* xr = x + (size_t)s_g[k] * d.xdim + xo;
#pragma unroll
      for (int a = 0; a < D1; ++a) xv[a] = __ldg(xr + a);
      if constexpr (TAB) L::load_wd(d, w + (size_t)s_ti[k] * 4 * d.wn, u, s_tt[k], wv, wdv);
      else L::load_ws(d, md, w + (size_t)s_e[k] * d.wn, u, wv);
      float dxv[D1];
#pragma unroll
      for (int a = 0; a < D1; ++a) dxv[a] = 0.f;
      L::template bwde<DX>(rec + k * RoleT<ROLE, LOUT>::REC, xv, gv, wv, dwv, dY, dxv);
      if constexpr (TAB) {   // the weight gradient never leaves the registers: contract it with d w / d t
#pragma unroll
        for (int q = 0; q < L::NP; ++q) drp = fmaf(dwv[q], wdv[q], drp);
      } else {
        if (md.dw_mode != 2) L::store_ws(d, md, dw + (size_t)s_e[k] * d.wn, u, dwv);
      }
      if constexpr (DX) {   // scatter to the source atom's row (RED.ADD; the row was zeroed by the launcher)
        float* xr = dxo + (size_t)s_g[k] * d.xdim + xo;
#pragma unroll
        for (int a = 0; a < D1; ++a) atomicAdd(xr + a, dxv[a]);
      }
    }
    if constexpr (LOUT == 0 && ROLE == 0) {
      // only path 000: no dependence on Y
    } else {
      const float r = warp_reduce8(dY + 1, lane);
      if ((lane & 3) == 0) s_dY[(warp * te_of<ROLE>() + k) * 8 + (lane >> 2)] = r;  // slot lane>>2 == component m
    }
    if constexpr (TAB) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) drp += __shfl_xor_sync(0xffffffffu, drp, o);
      if (lane == 0) s_dr[warp * te_of<ROLE>() + k] = drp;
    }
  }
}

template <int ROLE, int LOUT, bool DX, bool TAB>
__global__ void __launch_bounds__(256)
conv_bwde_kernel(ConvDesc d, int mulpad, const int* __restrict__ rowptr, const int* __restrict__ src,
                 const float* __restrict__ x, const float* __restrict__ Y, const float* __restrict__ w,
                 const float* __restrict__ gout, float inv_denom, float* __restrict__ dw, float* __restrict__ dYo,
                 int accumulate, float* __restrict__ dxo, TabArg tab, ConvMode md) {
  using RT = RoleT<ROLE, LOUT>;
  constexpr int NG = num_groups<ROLE, LOUT>();
  constexpr int TE = te_of<ROLE>();
  __shared__ __align__(16) float rec[TE * RT::REC];
  __shared__ int s_e[TE], s_g[TE];
  __shared__ int s_ti[TAB ? TE : 1];
  __shared__ float s_tt[TAB ? TE : 1];
  extern __shared__ float s_dY[];  // [nwarps][TE][8], then (TAB) s_dr[nwarps][TE]
  const int i = blockIdx.x;
  const int e0 = rowptr[i], e1 = rowptr[i + 1];
  const int grp = threadIdx.x / mulpad, u = threadIdx.x % mulpad;
  const bool active = u < role_mul<ROLE>(d) && grp < NG;
  const int nwarps = blockDim.x >> 5;
  float gv[kMaxAcc];
#pragma unroll
  for (int c = 0; c < kMaxAcc; ++c) gv[c] = 0.f;
  if (active) {
    const float* grow = gout + (size_t)i * d.odim;
    if (grp == 0) GroupT<ROLE, LOUT, 0>::L::load_o(d, grow, u, inv_denom, gv);
    else if (grp == 1) GroupT<ROLE, LOUT, 1>::L::load_o(d, grow, u, inv_denom, gv);
    else if (grp == 2) GroupT<ROLE, LOUT, 2>::L::load_o(d, grow, u, inv_denom, gv);
    else GroupT<ROLE, LOUT, 3>::L::load_o(d, grow, u, inv_denom, gv);
  }
  for (int c0 = e0; c0 < e1; c0 += TE) {
    const int ne = min(TE, e1 - c0);
    __syncthreads();
    stage_chunk<ROLE, LOUT>(rec, s_e, s_g, Y, nullptr, src, c0, ne, TAB ? &tab : nullptr, s_ti, s_tt);
    __syncthreads();
    float* s_dr = s_dY + nwarps * TE * 8;
    if (grp == 0) bwde_group<ROLE, LOUT, 0, DX, TAB>(d, rec, s_e, s_g, ne, x, w, dw, gv, u, active, s_dY, dxo, s_ti, s_tt, s_dr, md);
    else if (grp == 1) bwde_group<ROLE, LOUT, 1, DX, TAB>(d, rec, s_e, s_g, ne, x, w, dw, gv, u, active, s_dY, dxo, s_ti, s_tt, s_dr, md);
    else if (grp == 2) bwde_group<ROLE, LOUT, 2, DX, TAB>(d, rec, s_e, s_g, ne, x, w, dw, gv, u, active, s_dY, dxo, s_ti, s_tt, s_dr, md);
    else bwde_group<ROLE, LOUT, 3, DX, TAB>(d, rec, s_e, s_g, ne, x, w, dw, gv, u, active, s_dY, dxo, s_ti, s_tt, s_dr, md);
    if constexpr (TAB || !(LOUT == 0 && ROLE == 0)) __syncthreads();
    if constexpr (TAB) {   // d E / d r of the chunk's edges: sum over the warps (channels), times d t / d r = 1 / h
      for (int k = threadIdx.x; k < ne; k += blockDim.x) {
        float tot = 0.f;
        for (int wq = 0; wq < nwarps; ++wq) tot += s_dr[wq * TE + k];
        float* p = tab.dr + (size_t)(c0 + k);
        *p = tab.acc_dr ? fmaf(tot, tab.inv_h, *p) : tot * tab.inv_h;
      }
    }
    if constexpr (!(LOUT == 0 && ROLE == 0)) {
      for (int it = threadIdx.x; it < ne * 8; it += blockDim.x) {
        const int k = it >> 3, s = it & 7, m = s;
        float tot = 0.f;
        for (int wq = 0; wq < nwarps; ++wq) tot += s_dY[(wq * TE + k) * 8 + s];
        float* p = dYo + (size_t)(c0 + k) * 9 + 1 + m;
        *p = accumulate ? (*p + tot) : tot;
      }
    }
  }
}

// ---- launchers ------------------------------------------------------------------------------------
template <int ROLE, int LOUT>
static void launch_fwd(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
                       const float* w, float denom, float* out, cudaStream_t st, const TabArg* tab = nullptr,
                       const ConvMode* mode = nullptr) {
  const int mul = ROLE == 0 ? d.m0 : (ROLE == 1 ? d.m1 : d.m2);
  if (mul == 0 || n_dst == 0) return;
  const int mulpad = round_up(mul, 32);
  const int threads = mulpad * num_groups<ROLE, LOUT>();
  const ConvMode md = mode ? *mode : plain_mode();
  if (tab) conv_fwd_kernel<ROLE, LOUT, true><<<n_dst, threads, 0, st>>>(d, mulpad, rowptr, src, x, Y, w, denom, out, *tab, md);
  else conv_fwd_kernel<ROLE, LOUT, false><<<n_dst, threads, 0, st>>>(d, mulpad, rowptr, src, x, Y, w, denom, out, TabArg{}, md);
}
template <int ROLE, int LOUT>
static void launch_bwdx(const ConvDesc& d, int n_src, const int* rrowptr, const int* eid, const int* dst,
                        const float* gout, const float* Y, const float* w, float inv_denom, float* dx, cudaStream_t st) {
  const int mul = ROLE == 0 ? d.m0 : (ROLE == 1 ? d.m1 : d.m2);
  if (mul == 0 || n_src == 0) return;
  const int mulpad = round_up(mul, 32);
  constexpr int NG = num_groups<ROLE, LOUT>();
  const int threads = mulpad * NG;
  const size_t sm = (size_t)(NG - 1) * mulpad * (2 * ROLE + 1) * sizeof(float);
  conv_bwdx_kernel<ROLE, LOUT><<<n_src, threads, sm, st>>>(d, mulpad, rrowptr, eid, dst, gout, Y, w, inv_denom, dx);
}
template <int ROLE, int LOUT>
static void launch_bwde(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
                        const float* w, const float* gout, float inv_denom, float* dw, float* dY, int accumulate,
                        float* dx, cudaStream_t st, const TabArg* tab = nullptr, const ConvMode* mode = nullptr) {
  const int mul = ROLE == 0 ? d.m0 : (ROLE == 1 ? d.m1 : d.m2);
  if (mul == 0 || n_dst == 0) return;
  const ConvMode md = mode ? *mode : plain_mode();
  const int mulpad = round_up(mul, 32);
  const int threads = mulpad * num_groups<ROLE, LOUT>();
  const size_t sm = (size_t)(threads / 32) * te_of<ROLE>() * (tab ? 9 : 8) * sizeof(float);
  if (tab) {
    if (dx)
      conv_bwde_kernel<ROLE, LOUT, true, true><<<n_dst, threads, sm, st>>>(d, mulpad, rowptr, src, x, Y, w, gout, inv_denom,
                                                                            nullptr, dY, accumulate, dx, *tab, md);
    else
      conv_bwde_kernel<ROLE, LOUT, false, true><<<n_dst, threads, sm, st>>>(d, mulpad, rowptr, src, x, Y, w, gout, inv_denom,
                                                                             nullptr, dY, accumulate, nullptr, *tab, md);
  } else if (dx) {
    conv_bwde_kernel<ROLE, LOUT, true, false><<<n_dst, threads, sm, st>>>(d, mulpad, rowptr, src, x, Y, w, gout, inv_denom, dw,
                                                                           dY, accumulate, dx, TabArg{}, md);
  } else {
    conv_bwde_kernel<ROLE, LOUT, false, false><<<n_dst, threads, sm, st>>>(d, mulpad, rowptr, src, x, Y, w, gout, inv_denom, dw,
                                                                            dY, accumulate, nullptr, TabArg{}, md);
  }
}

static bool desc_ok(const ConvDesc& d) {
  return (d.lout == 0 || d.lout == 2) && d.m0 >= 0 && d.m1 >= 0 && d.m2 >= 0 && d.m0 <= 256 && d.m1 <= 128 &&
         d.m2 <= 64;
}

static int conv_fwd_mode(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
                         const float* w, float denom, float* out, const ConvMode* md, cudaStream_t st) {
  B200_REQUIRE(desc_ok(d), "conv_fwd: unsupported irreps (m0<=256, m1<=128, m2<=64, lout in {0,2})");
  if (d.lout == 2) {
    launch_fwd<0, 2>(d, n_dst, rowptr, src, x, Y, w, denom, out, st, nullptr, md);
    launch_fwd<1, 2>(d, n_dst, rowptr, src, x, Y, w, denom, out, st, nullptr, md);
    launch_fwd<2, 2>(d, n_dst, rowptr, src, x, Y, w, denom, out, st, nullptr, md);
  } else {
    launch_fwd<0, 0>(d, n_dst, rowptr, src, x, Y, w, denom, out, st, nullptr, md);
    launch_fwd<1, 0>(d, n_dst, rowptr, src, x, Y, w, denom, out, st, nullptr, md);
    launch_fwd<2, 0>(d, n_dst, rowptr, src, x, Y, w, denom, out, st, nullptr, md);
  }
  return check_launch("conv_fwd", (d.m0 > 0) + (d.m1 > 0) + (d.m2 > 0));
}
int conv_fwd(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
             const float* w, float denom, float* out, cudaStream_t st) {
  return conv_fwd_mode(d, n_dst, rowptr, src, x, Y, w, denom, out, nullptr, st);
}
static bool fill_mode(const float* path_scale, int dw_mode, ConvMode& md) {
  md = plain_mode();
  if (path_scale)
    for (int p = 0; p < kNumPaths; ++p) md.wscale[p] = path_scale[p];
  md.dw_mode = dw_mode;
  return dw_mode >= 0 && dw_mode <= 2;
}
int conv_fwd_scaled(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
                    const float* w, float denom, float* out, const float* path_scale, cudaStream_t st) {
  ConvMode md;
  fill_mode(path_scale, 0, md);
  return conv_fwd_mode(d, n_dst, rowptr, src, x, Y, w, denom, out, &md, st);
}

int conv_bwd_x(const ConvDesc& d, int n_src, const int* rrowptr, const int* eid, const int* dst, const float* gout,
               const float* Y, const float* w, float denom, float* dx, cudaStream_t st) {
  B200_REQUIRE(desc_ok(d), "conv_bwd_x: unsupported irreps");
  const float inv = 1.0f / denom;
  if (d.lout == 2) {
    launch_bwdx<0, 2>(d, n_src, rrowptr, eid, dst, gout, Y, w, inv, dx, st);
    launch_bwdx<1, 2>(d, n_src, rrowptr, eid, dst, gout, Y, w, inv, dx, st);
    launch_bwdx<2, 2>(d, n_src, rrowptr, eid, dst, gout, Y, w, inv, dx, st);
  } else {
    launch_bwdx<0, 0>(d, n_src, rrowptr, eid, dst, gout, Y, w, inv, dx, st);
    launch_bwdx<1, 0>(d, n_src, rrowptr, eid, dst, gout, Y, w, inv, dx, st);
    launch_bwdx<2, 0>(d, n_src, rrowptr, eid, dst, gout, Y, w, inv, dx, st);
  }
  return check_launch("conv_bwd_x", (d.m0 > 0) + (d.m1 > 0) + (d.m2 > 0));
}

// d weight, d edge_attr and (dx != nullptr) d x[src] in one pass over the destination-sorted edges: the gradient
// rows of the source atoms are accumulated with RED.ADD, so the per-edge weights are read once for all three.
static int conv_bwd_mode(const ConvDesc& d, int n_dst, int n_src, int n_edges, const int* rowptr, const int* src, const float* x,
                         const float* Y, const float* w, const float* gout, float denom, float* dw, float* dY, int accumulate,
                         float* dx, const ConvMode* md, cudaStream_t st);
int conv_bwd_edge_x(const ConvDesc& d, int n_dst, int n_src, int n_edges, const int* rowptr, const int* src, const float* x,
                    const float* Y, const float* w, const float* gout, float denom, float* dw, float* dY, int accumulate,
                    float* dx, cudaStream_t st) {
  return conv_bwd_mode(d, n_dst, n_src, n_edges, rowptr, src, x, Y, w, gout, denom, dw, dY, accumulate, dx, nullptr, st);
}
int conv_bwd_scaled(const ConvDesc& d, int n_dst, int n_src, int n_edges, const int* rowptr, const int* src, const float* x,
                    const float* Y, const float* w, const float* gout, float denom, float* dw, float* dY, int accumulate,
                    float* dx, const float* path_scale, int dw_mode, cudaStream_t st) {
  ConvMode md;
  B200_REQUIRE(fill_mode(path_scale, dw_mode, md), "conv_bwd_scaled: dw_mode must be 0, 1 or 2");
  B200_REQUIRE(dw != nullptr || dw_mode == 2, "conv_bwd_scaled: d_weight is NULL");
  return conv_bwd_mode(d, n_dst, n_src, n_edges, rowptr, src, x, Y, w, gout, denom, dw, dY, accumulate, dx, &md, st);
}
static int conv_bwd_mode(const ConvDesc& d, int n_dst, int n_src, int n_edges, const int* rowptr, const int* src, const float* x,
                         const float* Y, const float* w, const float* gout, float denom, float* dw, float* dY, int accumulate,
                         float* dx, const ConvMode* md, cudaStream_t st) {
  B200_REQUIRE(desc_ok(d), "conv_bwd_edge: unsupported irreps");
  const float inv = 1.0f / denom;
  if (dx) cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)n_src * d.xdim, st);
  // the three role kernels run back to back on one stream: the first may overwrite dY, the rest add
  int acc = accumulate;
  if (d.lout == 2) {
    if (d.m0) { launch_bwde<0, 2>(d, n_dst, rowptr, src, x, Y, w, gout, inv, dw, dY, acc, dx, st, nullptr, md); acc = 1; }
    if (d.m1) { launch_bwde<1, 2>(d, n_dst, rowptr, src, x, Y, w, gout, inv, dw, dY, acc, dx, st, nullptr, md); acc = 1; }
    if (d.m2) { launch_bwde<2, 2>(d, n_dst, rowptr, src, x, Y, w, gout, inv, dw, dY, acc, dx, st, nullptr, md); acc = 1; }
  } else {
    // role 0 (path 000) does not touch dY
    if (!acc && !d.m1 && !d.m2) cudaMemsetAsync(dY, 0, sizeof(float) * 9 * (size_t)n_edges, st);
    launch_bwde<0, 0>(d, n_dst, rowptr, src, x, Y, w, gout, inv, dw, dY, acc, dx, st, nullptr, md);
    if (d.m1) { launch_bwde<1, 0>(d, n_dst, rowptr, src, x, Y, w, gout, inv, dw, dY, acc, dx, st, nullptr, md); acc = 1; }
    if (d.m2) { launch_bwde<2, 0>(d, n_dst, rowptr, src, x, Y, w, gout, inv, dw, dY, acc, dx, st, nullptr, md); acc = 1; }
  }
  return check_launch("conv_bwd_edge", (d.m0 > 0) + (d.m1 > 0) + (d.m2 > 0));
}

// ---- the same two passes with the per-edge weights interpolated from the tabulated radial function (radial_table.cu):
// neither the weights [E, wn] nor their gradient exist in memory; the reverse pass emits d E / d r per edge.
int conv_fwd_tab(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
                 const float* len, const float* table, float inv_h, int n_grid, float denom, float* out, cudaStream_t st) {
  B200_REQUIRE(desc_ok(d), "conv_fwd_tab: unsupported irreps (m0<=256, m1<=128, m2<=64, lout in {0,2})");
  const TabArg tab{len, inv_h, n_grid, nullptr, 0};
  if (d.lout == 2) {
    launch_fwd<0, 2>(d, n_dst, rowptr, src, x, Y, table, denom, out, st, &tab);
    launch_fwd<1, 2>(d, n_dst, rowptr, src, x, Y, table, denom, out, st, &tab);
    launch_fwd<2, 2>(d, n_dst, rowptr, src, x, Y, table, denom, out, st, &tab);
  } else {
    launch_fwd<0, 0>(d, n_dst, rowptr, src, x, Y, table, denom, out, st, &tab);
    launch_fwd<1, 0>(d, n_dst, rowptr, src, x, Y, table, denom, out, st, &tab);
    launch_fwd<2, 0>(d, n_dst, rowptr, src, x, Y, table, denom, out, st, &tab);
  }
  return check_launch("conv_fwd_tab", (d.m0 > 0) + (d.m1 > 0) + (d.m2 > 0));
}

int conv_bwd_tab(const ConvDesc& d, int n_dst, int n_src, int n_edges, const int* rowptr, const int* src, const float* x,
                 const float* Y, const float* len, const float* table, float inv_h, int n_grid, const float* gout, float denom,
                 float* dY, int accumulate_dY, float* dr, int accumulate_dr, float* dx, cudaStream_t st) {
  B200_REQUIRE(desc_ok(d), "conv_bwd_tab: unsupported irreps");
  const float inv = 1.0f / denom;
  if (dx) cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)n_src * d.xdim, st);
  int acc = accumulate_dY;
  TabArg tab{len, inv_h, n_grid, dr, accumulate_dr};
  if (d.lout == 2) {
    if (d.m0) { launch_bwde<0, 2>(d, n_dst, rowptr, src, x, Y, table, gout, inv, nullptr, dY, acc, dx, st, &tab); acc = 1; tab.acc_dr = 1; }
    if (d.m1) { launch_bwde<1, 2>(d, n_dst, rowptr, src, x, Y, table, gout, inv, nullptr, dY, acc, dx, st, &tab); acc = 1; tab.acc_dr = 1; }
    if (d.m2) { launch_bwde<2, 2>(d, n_dst, rowptr, src, x, Y, table, gout, inv, nullptr, dY, acc, dx, st, &tab); acc = 1; tab.acc_dr = 1; }
  } else {
    if (!acc && !d.m1 && !d.m2) cudaMemsetAsync(dY, 0, sizeof(float) * 9 * (size_t)n_edges, st);
    launch_bwde<0, 0>(d, n_dst, rowptr, src, x, Y, table, gout, inv, nullptr, dY, acc, dx, st, &tab);
    tab.acc_dr = 1;
    if (d.m1) { launch_bwde<1, 0>(d, n_dst, rowptr, src, x, Y, table, gout, inv, nullptr, dY, acc, dx, st, &tab); acc = 1; }
    if (d.m2) { launch_bwde<2, 0>(d, n_dst, rowptr, src, x, Y, table, gout, inv, nullptr, dY, acc, dx, st, &tab); acc = 1; }
  }
  return check_launch("conv_bwd_tab", (d.m0 > 0) + (d.m1 > 0) + (d.m2 > 0));
}

int conv_bwd_edge(const ConvDesc& d, int n_dst, int n_edges, const int* rowptr, const int* src, const float* x, const float* Y,
                  const float* w, const float* gout, float denom, float* dw, float* dY, int accumulate,
                  cudaStream_t st) {
  return conv_bwd_edge_x(d, n_dst, 0, n_edges, rowptr, src, x, Y, w, gout, denom, dw, dY, accumulate, nullptr, st);
}

}  // namespace b200
