// Train-mode radial weights from a table: w[e, :] = sum_j q_j(t_e) P[i_e, j, :].
//
// Reference: sevenn/nn/convolution.py:97-106 (weight = weight_nn(edge_embedding), e3nn FullyConnectedNet) under
// sevenn/train/trainer.py:173-186 (loss.backward() through create_graph forces, force_output.py:165-170).
// weight_nn is a function of ONE scalar, the edge length.  In train() mode its parameters change every step, so the
// table is rebuilt every step - by the caller, with differentiable tensor operations over the 4,001 grid points (the
// network itself runs on 4,001 rows instead of one row per edge; sevennet_finetuning_b200/nn.py) - and handed over as per-interval
// polynomial coefficients P[n_grid, 4, wn] = (a, b, c, d):  w(t) = a + b t + c t^2 + d t^3 on interval i, t = r * inv_h - i.
// The edge side is linear in P.  Its derivatives of every order are the three maps below with m = order of the
// derivative in r (q^(0) = {1, t, t^2, t^3}, q^(1) = {0, 1, 2t, 3t^2} / h, q^(2) = {0, 0, 2, 6t} / h^2, q^(3) = {0, 0, 0, 6} / h^3):
//   interp_m (P, r)      -> [E, wn]       out[e, k]  = coef[e] * sum_j q_j^(m)(t_e) P[i_e, j, k]
//   dot_m    (P, r, g)   -> [E]           out[e]     = sum_k g[e, k] interp_m(P, r)[e, k]
//   scatter_m(g, coef, r)-> [n, 4, wn]    G[i, j, k] = sum_{e in i} coef[e] q_j^(m)(t_e) g[e, k]
// d interp_m = (scatter_m, dot_{m+1}), d dot_m = (scatter_m, dot_{m+1}, interp_m), d scatter_m = (interp_m, dot_m, dot_{m+1}):
// a closed family (sevennet_finetuning_b200/higher_order.py).
//
// All three visit the edges in the order `order` (edge ids sorted by interval, nullptr = natural order): consecutive edges
// share table rows (L1 hits), and the scatter keeps the sums of an interval in registers and adds them to G once.
#include "common.cuh"
#include "ops.h"

namespace b200 {

namespace {

__device__ __forceinline__ void locate_t(float r, float inv_h, int n_grid, int& i, float& t) {
  i = min(max((int)(r * inv_h), 0), n_grid - 1);
  t = fmaf(r, inv_h, -(float)i);
}

// q_j^(m)(t) / h^m
__device__ __forceinline__ void basis_m(int m, float t, float inv_h, float q[4]) {
  switch (m) {
    case 0: q[0] = 1.f; q[1] = t; q[2] = t * t; q[3] = t * t * t; break;
    case 1: q[0] = 0.f; q[1] = inv_h; q[2] = 2.f * t * inv_h; q[3] = 3.f * t * t * inv_h; break;
    case 2: q[0] = 0.f; q[1] = 0.f; q[2] = 2.f * inv_h * inv_h; q[3] = 6.f * t * inv_h * inv_h; break;
    case 3: q[0] = 0.f; q[1] = 0.f; q[2] = 0.f; q[3] = 6.f * inv_h * inv_h * inv_h; break;
    default: q[0] = q[1] = q[2] = q[3] = 0.f; break;
  }
}

constexpr int TT_WARPS = 8;
constexpr int TT_EPW = 8;   // edges per warp

// warp = TT_EPW consecutive positions of `order`, one edge at a time, lanes over the weights in float4 steps.  The first
// TT_EPW lanes fetch edge id, interval and local coordinate of the warp's edges up front (order -> r -> table row would
// otherwise be a chain of three dependent global loads in front of every edge).
template <bool DOT>
__global__ void __launch_bounds__(TT_WARPS * 32)
table_interp_kernel(int m, int wn, long long E, float inv_h, int n_grid, const float* __restrict__ P,
                    const float* __restrict__ r, const int* __restrict__ order, const float* __restrict__ coef,
                    const float* __restrict__ g, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const long long w = (long long)blockIdx.x * TT_WARPS + (threadIdx.x >> 5);
  const long long j0 = w * TT_EPW;
  if (j0 >= E) return;
  const int ne = (int)min((long long)TT_EPW, E - j0);
  const int nq = wn >> 2;
  long long my_e = 0;
  int my_i = 0;
  float my_t = 0.f, my_c = 1.f;
  if (lane < ne) {
    my_e = order ? order[j0 + lane] : j0 + lane;
    locate_t(__ldg(r + my_e), inv_h, n_grid, my_i, my_t);
    if (!DOT && coef) my_c = __ldg(coef + my_e);
  }
  for (int j = 0; j < ne; ++j) {
    const long long e = __shfl_sync(0xffffffffu, my_e, j);
    const int i = __shfl_sync(0xffffffffu, my_i, j);
    const float t = __shfl_sync(0xffffffffu, my_t, j);
    float q[4];
    basis_m(m, t, inv_h, q);
    const float4* Pi = reinterpret_cast<const float4*>(P + (size_t)i * 4 * wn);
    if constexpr (DOT) {
      const float4* ge = reinterpret_cast<const float4*>(g + (size_t)e * wn);
      float s = 0.f;
      for (int k = lane; k < nq; k += 32) {
        const float4 a = __ldg(Pi + k), b = __ldg(Pi + nq + k), c = __ldg(Pi + 2 * nq + k), d = __ldg(Pi + 3 * nq + k);
        const float4 gv = __ldcs(ge + k);
        s += gv.x * (q[0] * a.x + q[1] * b.x + q[2] * c.x + q[3] * d.x) + gv.y * (q[0] * a.y + q[1] * b.y + q[2] * c.y + q[3] * d.y) +
             gv.z * (q[0] * a.z + q[1] * b.z + q[2] * c.z + q[3] * d.z) + gv.w * (q[0] * a.w + q[1] * b.w + q[2] * c.w + q[3] * d.w);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) out[e] = s;
    } else {
      const float ce = __shfl_sync(0xffffffffu, my_c, j);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) q[jj] *= ce;
      float4* oe = reinterpret_cast<float4*>(out + (size_t)e * wn);
      for (int k = lane; k < nq; k += 32) {
        const float4 a = __ldg(Pi + k), b = __ldg(Pi + nq + k), c = __ldg(Pi + 2 * nq + k), d = __ldg(Pi + 3 * nq + k);
        float4 v;
        v.x = q[0] * a.x + q[1] * b.x + q[2] * c.x + q[3] * d.x;
        v.y = q[0] * a.y + q[1] * b.y + q[2] * c.y + q[3] * d.y;
        v.z = q[0] * a.z + q[1] * b.z + q[2] * c.z + q[3] * d.z;
        v.w = q[0] * a.w + q[1] * b.w + q[2] * c.w + q[3] * d.w;
        __stcs(oe + k, v);
      }
    }
  }
}

constexpr int TS_T = 256;
constexpr int TS_MAXQ = 2;   // float4 columns per thread: wn <= 4 * 256 * 2
constexpr int TS_EPC = 64;   // edges per CTA

// CTA = TS_EPC consecutive positions of `order`; thread = up to TS_MAXQ float4 columns of the weights.  Edge id, interval
// and the four coefficients coef * q_j(t) of the CTA's edges are staged in shared memory first; the sums of the current
// interval stay in registers and go to G with one atomic add per element when the interval changes.
// DOT: the same pass also emits dot[e] = sum_k g[e, k] d^md w_k / d r^md (r[e]) (the two reverse-pass results that read g).
template <bool DOT>
__global__ void __launch_bounds__(TS_T)
table_scatter_kernel(int m, int wn, long long E, float inv_h, int n_grid, const float* __restrict__ g,
                     const float* __restrict__ coef, const float* __restrict__ r, const int* __restrict__ order,
                     float* __restrict__ G, int md, const float* __restrict__ P, float* __restrict__ dot) {
  __shared__ long long s_e[TS_EPC];
  __shared__ int s_i[TS_EPC];
  __shared__ float s_q[TS_EPC][4];
  __shared__ float s_qd[DOT ? TS_EPC : 1][4];
  __shared__ float s_dot[DOT ? TS_EPC : 1];
  const long long j0 = (long long)blockIdx.x * TS_EPC;
  const int ne = (int)min((long long)TS_EPC, E - j0);
  const int nq = wn >> 2;
  if (threadIdx.x < ne) {
    const long long e = order ? order[j0 + threadIdx.x] : j0 + threadIdx.x;
    int i;
    float t, q[4];
    locate_t(__ldg(r + e), inv_h, n_grid, i, t);
    basis_m(m, t, inv_h, q);
    const float ce = coef ? __ldg(coef + e) : 1.f;
    s_e[threadIdx.x] = e;
    s_i[threadIdx.x] = i;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) s_q[threadIdx.x][jj] = q[jj] * ce;
    if constexpr (DOT) {
      basis_m(md, t, inv_h, q);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) s_qd[threadIdx.x][jj] = q[jj];
      s_dot[threadIdx.x] = 0.f;
    }
  }
  __syncthreads();
  float4 acc[4][TS_MAXQ];
  int cur = -1;
  auto flush = [&]() {
    if (cur < 0) return;
    float* Gi = G + (size_t)cur * 4 * wn;
#pragma unroll
    for (int u = 0; u < TS_MAXQ; ++u) {
      const int k = threadIdx.x + u * TS_T;
      if (k < nq) {
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          if (jj < m) continue;   // q_j^(m) = 0 for j < m
          float* p = Gi + (size_t)jj * wn + 4 * k;
          atomicAdd(p, acc[jj][u].x); atomicAdd(p + 1, acc[jj][u].y); atomicAdd(p + 2, acc[jj][u].z); atomicAdd(p + 3, acc[jj][u].w);
        }
      }
    }
  };
  for (int j = 0; j < ne; j += 4) {
    // the rows of four edges in flight, then the (uniform) interval test per edge
    float4 gv[4][TS_MAXQ];
#pragma unroll
    for (int d = 0; d < 4; ++d)
#pragma unroll
      for (int u = 0; u < TS_MAXQ; ++u) {
        const int k = threadIdx.x + u * TS_T;
        gv[d][u] = (j + d < ne && k < nq) ? __ldcs(reinterpret_cast<const float4*>(g + (size_t)s_e[j + d] * wn) + k)
                                          : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    if constexpr (DOT) {   // the four edges' dot products first: independent chains, one round of shuffles for all
      float sdot[4];
#pragma unroll
      for (int d = 0; d < 4; ++d) {
        sdot[d] = 0.f;
        if (j + d < ne) {
          const float4* Pi = reinterpret_cast<const float4*>(P + (size_t)s_i[j + d] * 4 * wn);
          const float q0 = s_qd[j + d][0], q1 = s_qd[j + d][1], q2 = s_qd[j + d][2], q3 = s_qd[j + d][3];
#pragma unroll
          for (int u = 0; u < TS_MAXQ; ++u) {
            const int k = threadIdx.x + u * TS_T;
            if (k < nq) {
              const float4 a = __ldg(Pi + k), b = __ldg(Pi + nq + k), c = __ldg(Pi + 2 * nq + k), dd = __ldg(Pi + 3 * nq + k);
              const float4 v = gv[d][u];
              sdot[d] += v.x * (q0 * a.x + q1 * b.x + q2 * c.x + q3 * dd.x) + v.y * (q0 * a.y + q1 * b.y + q2 * c.y + q3 * dd.y) +
                         v.z * (q0 * a.z + q1 * b.z + q2 * c.z + q3 * dd.z) + v.w * (q0 * a.w + q1 * b.w + q2 * c.w + q3 * dd.w);
            }
          }
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
#pragma unroll
        for (int d = 0; d < 4; ++d) sdot[d] += __shfl_xor_sync(0xffffffffu, sdot[d], o);
      if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int d = 0; d < 4; ++d)
          if (j + d < ne) atomicAdd(&s_dot[j + d], sdot[d]);
      }
    }
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      if (j + d >= ne) break;
      const int i = s_i[j + d];
      if (i != cur) {
        flush();
        cur = i;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj)
#pragma unroll
          for (int u = 0; u < TS_MAXQ; ++u) acc[jj][u] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const float c = s_q[j + d][jj];
#pragma unroll
        for (int u = 0; u < TS_MAXQ; ++u) {
          acc[jj][u].x = fmaf(c, gv[d][u].x, acc[jj][u].x);
          acc[jj][u].y = fmaf(c, gv[d][u].y, acc[jj][u].y);
          acc[jj][u].z = fmaf(c, gv[d][u].z, acc[jj][u].z);
          acc[jj][u].w = fmaf(c, gv[d][u].w, acc[jj][u].w);
        }
      }
    }
  }
  flush();
  if constexpr (DOT) {
    __syncthreads();
    if (threadIdx.x < ne) dot[s_e[threadIdx.x]] = s_dot[threadIdx.x];
  }
}

// nodal values and slopes (float64, [n_grid + 1, wn]) -> per-interval cubic coefficients (float32, [n_grid, 4, wn]):
//   a = Z_i, b = h Z'_i, c = 3 (Z_{i+1} - Z_i) - 2 h Z'_i - h Z'_{i+1}, d = -2 (Z_{i+1} - Z_i) + h Z'_i + h Z'_{i+1}
// The difference of neighbouring nodal values is formed in float64: in float32 it loses the digits the slope needs.
__global__ void hermite_kernel(int n_grid, int wn, double h, const double* __restrict__ Z, const double* __restrict__ dZ,
                               float* __restrict__ P) {
  const long long id = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (id >= (long long)n_grid * wn) return;
  const int i = (int)(id / wn), k = (int)(id % wn);
  const double v0 = Z[(size_t)i * wn + k], v1 = Z[(size_t)(i + 1) * wn + k];
  const double s0 = h * dZ[(size_t)i * wn + k], s1 = h * dZ[(size_t)(i + 1) * wn + k], dv = v1 - v0;
  float* p = P + (size_t)i * 4 * wn + k;
  p[0] = (float)v0;
  p[wn] = (float)s0;
  p[2 * wn] = (float)(3.0 * dv - 2.0 * s0 - s1);
  p[3 * wn] = (float)(s0 + s1 - 2.0 * dv);
}
// the adjoint map: gradient with respect to P -> gradients with respect to Z and Z'
__global__ void hermite_adj_kernel(int n_grid, int wn, double h, const float* __restrict__ G, double* __restrict__ gZ,
                                   double* __restrict__ gdZ) {
  const long long id = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (id >= (long long)(n_grid + 1) * wn) return;
  const int i = (int)(id / wn), k = (int)(id % wn);
  double z = 0.0, dz = 0.0;
  if (i < n_grid) {       // left node of interval i
    const float* g = G + (size_t)i * 4 * wn + k;
    const double ga = g[0], gb = g[wn], gc = g[2 * wn], gd = g[3 * wn];
    z += ga - 3.0 * gc + 2.0 * gd;
    dz += gb - 2.0 * gc + gd;
  }
  if (i > 0) {            // right node of interval i - 1
    const float* g = G + (size_t)(i - 1) * 4 * wn + k;
    const double gc = g[2 * wn], gd = g[3 * wn];
    z += 3.0 * gc - 2.0 * gd;
    dz += gd - gc;
  }
  gZ[id] = z;
  gdZ[id] = h * dz;
}

int check_shape(const char* what, int m, int wn, int n_grid, float inv_h) {
  B200_REQUIRE(m >= 0 && m <= 4, "%s: derivative order %d", what, m);
  B200_REQUIRE(wn >= 4 && wn % 4 == 0 && n_grid >= 1 && inv_h > 0.f, "%s: wn %% 4 == 0, n_grid >= 1, 1/h > 0", what);
  return B200_OK;
}

}  // namespace

int table_interp(int m, int wn, long long E, float inv_h, int n_grid, const float* P, const float* r, const int* order,
                 const float* coef, float* out, cudaStream_t st) {
  if (int rc = check_shape("table_interp", m, wn, n_grid, inv_h)) return rc;
  if (E == 0) return B200_OK;
  const unsigned blocks = (unsigned)((E + (long long)TT_EPW * TT_WARPS - 1) / ((long long)TT_EPW * TT_WARPS));
  table_interp_kernel<false><<<blocks, TT_WARPS * 32, 0, st>>>(m, wn, E, inv_h, n_grid, P, r, order, coef, nullptr, out);
  return check_launch("table_interp");
}

int table_dot(int m, int wn, long long E, float inv_h, int n_grid, const float* P, const float* r, const int* order,
              const float* g, float* out, cudaStream_t st) {
  if (int rc = check_shape("table_dot", m, wn, n_grid, inv_h)) return rc;
  if (E == 0) return B200_OK;
  const unsigned blocks = (unsigned)((E + (long long)TT_EPW * TT_WARPS - 1) / ((long long)TT_EPW * TT_WARPS));
  table_interp_kernel<true><<<blocks, TT_WARPS * 32, 0, st>>>(m, wn, E, inv_h, n_grid, P, r, order, nullptr, g, out);
  return check_launch("table_dot");
}

int table_scatter(int m, int wn, long long E, float inv_h, int n_grid, const float* g, const float* coef, const float* r,
                  const int* order, float* G, cudaStream_t st) {
  if (int rc = check_shape("table_scatter", m, wn, n_grid, inv_h)) return rc;
  B200_REQUIRE(wn <= 4 * TS_T * TS_MAXQ, "table_scatter: wn <= %d", 4 * TS_T * TS_MAXQ);
  if (cudaMemsetAsync(G, 0, sizeof(float) * (size_t)n_grid * 4 * wn, st) != cudaSuccess) {
    set_error("table_scatter: memset failed");
    return B200_ERR_CUDA;
  }
  if (E == 0) return B200_OK;
  const unsigned blocks = (unsigned)((E + TS_EPC - 1) / TS_EPC);
  table_scatter_kernel<false><<<blocks, TS_T, 0, st>>>(m, wn, E, inv_h, n_grid, g, coef, r, order, G, 0, nullptr, nullptr);
  return check_launch("table_scatter");
}

int table_scatter_dot(int m, int md, int wn, long long E, float inv_h, int n_grid, const float* P, const float* g,
                      const float* coef, const float* r, const int* order, float* G, float* dot, cudaStream_t st) {
  if (int rc = check_shape("table_scatter_dot", m, wn, n_grid, inv_h)) return rc;
  B200_REQUIRE(md >= 0 && md <= 4, "table_scatter_dot: derivative order %d", md);
  B200_REQUIRE(wn <= 4 * TS_T * TS_MAXQ, "table_scatter_dot: wn <= %d", 4 * TS_T * TS_MAXQ);
  if (cudaMemsetAsync(G, 0, sizeof(float) * (size_t)n_grid * 4 * wn, st) != cudaSuccess) {
    set_error("table_scatter_dot: memset failed");
    return B200_ERR_CUDA;
  }
  if (E == 0) return B200_OK;
  const unsigned blocks = (unsigned)((E + TS_EPC - 1) / TS_EPC);
  table_scatter_kernel<true><<<blocks, TS_T, 0, st>>>(m, wn, E, inv_h, n_grid, g, coef, r, order, G, md, P, dot);
  return check_launch("table_scatter_dot");
}

int table_hermite(int n_grid, int wn, double h, const double* Z, const double* dZ, float* P, cudaStream_t st) {
  B200_REQUIRE(n_grid >= 1 && wn >= 1 && h > 0.0, "table_hermite: n_grid >= 1, wn >= 1, h > 0");
  const long long n = (long long)n_grid * wn;
  hermite_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n_grid, wn, h, Z, dZ, P);
  return check_launch("table_hermite");
}

int table_hermite_adj(int n_grid, int wn, double h, const float* G, double* gZ, double* gdZ, cudaStream_t st) {
  B200_REQUIRE(n_grid >= 1 && wn >= 1 && h > 0.0, "table_hermite_adj: n_grid >= 1, wn >= 1, h > 0");
  const long long n = (long long)(n_grid + 1) * wn;
  hermite_adj_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n_grid, wn, h, G, gZ, gdZ);
  return check_launch("table_hermite_adj");
}

}  // namespace b200
