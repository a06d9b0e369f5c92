// Tabulated radial weights for the MD path: the per-edge weights of the convolution are a function of ONE scalar,
//   w_k(r) = weight_nn( BesselBasis(r) * cutoff(r) )_k       (sevenn/nn/convolution.py:105-106, edge_embedding.py:220-230)
// so for fixed parameters (an MD run, a calculator) the 8 -> 64 -> 64 -> weight_numel MLP need not be evaluated per
// edge: it is evaluated once, in float64, on a uniform grid of r with its exact derivative, and every edge
// interpolates with a cubic Hermite spline (value and slope at both ends of its interval):
//   w_k(r) = h00(t) v_i + h10(t) m_i + h01(t) v_{i+1} + h11(t) m_{i+1},   t = r/h - i,   m = h * dw/dr.
// Interpolation error ~ h^4 w''''/384: with h = 1.25e-3 A it is < 1e-9 relative, below fp32 rounding of the MLP
// itself; the XPLOR envelope's curvature jump at r_on sits on a grid node.  The reverse pass needs only
//   dE/dr_e = sum_k dE/dw_k * dw_k/dr(r_e)   (derivative of the same interpolant: energy and force stay consistent).
// A mid-layer table is 4000 x 4 x 960 fp32 = 61 MB; the edges are visited sorted by interval (one counting sort per
// force call, shared by all layers and both passes), so a table row is fetched from L2 once per CTA and then served by
// L1: the per-edge work is one streaming write (forward) or read (reverse) of the [E, wn] tensor.
// Training keeps the explicit MLP (its parameters change every step and need gradients).
#include <algorithm>

#include <cuda_fp16.h>
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"
#include "ops.h"

namespace b200 {

constexpr int kTabMaxHidden = 128;

__device__ __forceinline__ void cutoff_vd_f64(const EdgeParams& p, double r, double& f, double& df) {
  if (p.cutoff_kind == B200_CUTOFF_XPLOR) {
    const double r2 = r * r, ron2 = (double)p.r_on * p.r_on, rc2 = (double)p.r_cut * p.r_cut;
    if (r < (double)p.r_on) {
      f = 1.0;
      df = 0.0;
    } else {
      const double den = (rc2 - ron2) * (rc2 - ron2) * (rc2 - ron2), a = rc2 - r2;
      f = a * a * (rc2 + 2.0 * r2 - 3.0 * ron2) / den;
      df = 12.0 * r * a * (ron2 - r2) / den;
    }
  } else {
    const double x = r / (double)p.r_cut, pp = (double)p.poly_p;
    const double c0 = (pp + 1.0) * (pp + 2.0) * 0.5, c1 = pp * (pp + 2.0), c2 = pp * (pp + 1.0) * 0.5;
    const double xp1 = r > 0.0 ? pow(x, pp - 1.0) : 0.0, xp = xp1 * x, xq = xp * x, xr = xq * x;
    f = 1.0 - c0 * xp + c1 * xq - c2 * xr;
    df = (-c0 * pp * xp1 + c1 * (pp + 1.0) * xp - c2 * (pp + 2.0) * xq) / (double)p.r_cut;
  }
}

// weight_nn and its d/dr at one radius, float64 (frozen-graph arithmetic: h = c silu(x W / sqrt(fan_in))); whole CTA
__device__ void mlp_hidden_f64(const EdgeParams& p, const MlpDesc& d, double r, const float* __restrict__ coeffs,
                               const float* __restrict__ W0, const float* __restrict__ W1, double* e, double* de, double* a1,
                               double* da1, double* a2, double* da2) {
  const double c = 1.6791767923989418;   // e3nn normalize2mom(silu)
  if (threadIdx.x < d.nb) {
    const double k = (double)coeffs[threadIdx.x], pre = 2.0 / (double)p.r_cut;
    double b, db, f, df;
    if (r > 0.0) {
      const double s = sin(k * r), cs = cos(k * r);
      b = pre * s / r;
      db = pre * (k * cs * r - s) / (r * r);
    } else {   // sin(k r) / r -> k, an even function of r
      b = pre * k;
      db = 0.0;
    }
    cutoff_vd_f64(p, r, f, df);
    e[threadIdx.x] = b * f;
    de[threadIdx.x] = db * f + b * df;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < d.h1; j += blockDim.x) {
    double z = 0.0, dz = 0.0;
    for (int n = 0; n < d.nb; ++n) {
      const double w = (double)W0[n * d.h1 + j];
      z += e[n] * w;
      dz += de[n] * w;
    }
    const double sc = 1.0 / sqrt((double)d.nb);
    z *= sc; dz *= sc;
    const double s = 1.0 / (1.0 + exp(-z));
    a1[j] = c * z * s;
    da1[j] = c * s * (1.0 + z * (1.0 - s)) * dz;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < d.h2; j += blockDim.x) {
    double z = 0.0, dz = 0.0;
    for (int n = 0; n < d.h1; ++n) {
      const double w = (double)W1[n * d.h2 + j];
      z += a1[n] * w;
      dz += da1[n] * w;
    }
    const double sc = 1.0 / sqrt((double)d.h1);
    z *= sc; dz *= sc;
    const double s = 1.0 / (1.0 + exp(-z));
    a2[j] = c * z * s;
    da2[j] = c * s * (1.0 + z * (1.0 - s)) * dz;
  }
  __syncthreads();
}

// One CTA per interval [r_i, r_i + h]: the MLP at both ends in float64, then the cubic of the interval in the local
// coordinate t = (r - r_i) / h:  w(t) = a + b t + c t^2 + d t^3 with a = v0, b = m0, c = 3 D - 2 m0 - m1, d = m0 + m1 - 2 D
// (v = value, m = h dw/dr, D = v1 - v0 formed in float64: the fp32 table never differences two nearly equal values, so
// the slope (b + 2 c t + 3 d t^2) / h is as accurate as the value).  Row layout: T[i][k] = float4 {a, b, c, d} of weight k:
// one 16-byte load per weight, consecutive channels read consecutive float4s.
__global__ void __launch_bounds__(128)
radial_table_build_kernel(EdgeParams p, MlpDesc d, double h, const float* __restrict__ coeffs, const float* __restrict__ W0,
                          const float* __restrict__ W1, const float* __restrict__ W2, float* __restrict__ T) {
  __shared__ double e[32], de[32], a1[kTabMaxHidden], da1[kTabMaxHidden], a2[2][kTabMaxHidden], da2[2][kTabMaxHidden];
  const int i = blockIdx.x;
  for (int end = 0; end < 2; ++end)
    mlp_hidden_f64(p, d, (double)(i + end) * h, coeffs, W0, W1, e, de, a1, da1, a2[end], da2[end]);
  float* row = T + (size_t)i * 4 * d.wn;
  const double sc = 1.0 / sqrt((double)d.h2);
  for (int k = threadIdx.x; k < d.wn; k += blockDim.x) {
    double v[2] = {0.0, 0.0}, m[2] = {0.0, 0.0};
    for (int n = 0; n < d.h2; ++n) {
      const double w2 = (double)W2[(size_t)n * d.wn + k];
      v[0] += a2[0][n] * w2; m[0] += da2[0][n] * w2;
      v[1] += a2[1][n] * w2; m[1] += da2[1][n] * w2;
    }
    const double v0 = v[0] * sc, v1 = v[1] * sc, m0 = m[0] * sc * h, m1 = m[1] * sc * h, D = v1 - v0;
    reinterpret_cast<float4*>(row)[k] =
        make_float4((float)v0, (float)m0, (float)(3.0 * D - 2.0 * m0 - m1), (float)(m0 + m1 - 2.0 * D));
  }
}

// Packed layout for the channel-pair convolution kernels (conv_pair.cu): the same interval cubics, 12 bytes per weight
// instead of 16.  Per interval i and weight PAIR q = (2q, 2q+1):  AB[i][q] = float4 {a0, a1, b0, b1} (fp32) and
// C[i][q] = 2 x fp16 {c0, c1}, D[i][q] = 2 x fp16 {d0, d1}, both scaled by kTabCDScale (three planes: the forward pass
// does not read D, whose term is 4e-8 of the weight; the reverse pass needs it for dw/dr).  a and b carry the value and the first-order term and stay
// fp32; c (second order in h) and d (third order) are about 1e-5 and 1e-8 of the weight, so 11 mantissa bits bound their
// contribution to w at 1e-8 relative and to dw/dr at 1e-6 relative (the scale keeps the c and d of small weights out of
// the fp16 subnormals).  The convolution kernels are bound by these reads from L2 (960 weights per edge per pass).
__global__ void __launch_bounds__(128)
radial_table_build_packed_kernel(EdgeParams p, MlpDesc d, double h, const float* __restrict__ coeffs,
                                 const float* __restrict__ W0, const float* __restrict__ W1, const float* __restrict__ W2,
                                 float4* __restrict__ AB, unsigned* __restrict__ Cp, unsigned* __restrict__ Dp) {
  __shared__ double e[32], de[32], a1[kTabMaxHidden], da1[kTabMaxHidden], a2[2][kTabMaxHidden], da2[2][kTabMaxHidden];
  const int i = blockIdx.x;
  for (int end = 0; end < 2; ++end)
    mlp_hidden_f64(p, d, (double)(i + end) * h, coeffs, W0, W1, e, de, a1, da1, a2[end], da2[end]);
  const double sc = 1.0 / sqrt((double)d.h2);
  const int np = d.wn / 2;
  for (int q = threadIdx.x; q < np; q += blockDim.x) {
    float a[2], b[2], c[2], dd[2];
    for (int s = 0; s < 2; ++s) {
      const int k = 2 * q + s;
      double v[2] = {0.0, 0.0}, m[2] = {0.0, 0.0};
      for (int n = 0; n < d.h2; ++n) {
        const double w2 = (double)W2[(size_t)n * d.wn + k];
        v[0] += a2[0][n] * w2; m[0] += da2[0][n] * w2;
        v[1] += a2[1][n] * w2; m[1] += da2[1][n] * w2;
      }
      const double v0 = v[0] * sc, v1 = v[1] * sc, m0 = m[0] * sc * h, m1 = m[1] * sc * h, D = v1 - v0;
      a[s] = (float)v0;
      b[s] = (float)m0;
      c[s] = (float)((3.0 * D - 2.0 * m0 - m1) * (double)kTabCDScale);
      dd[s] = (float)((m0 + m1 - 2.0 * D) * (double)kTabCDScale);
    }
    AB[(size_t)i * np + q] = make_float4(a[0], a[1], b[0], b[1]);
    const __half2 ch = __floats2half2_rn(c[0], c[1]), dh = __floats2half2_rn(dd[0], dd[1]);
    Cp[(size_t)i * np + q] = *reinterpret_cast<const unsigned*>(&ch);
    Dp[(size_t)i * np + q] = *reinterpret_cast<const unsigned*>(&dh);
  }
}

// interval index and local coordinate of an edge length
__device__ __forceinline__ void locate(float r, float inv_h, int n_grid, int& i, float& t) {
  i = min((int)(r * inv_h), n_grid - 1);
  t = fmaf(r, inv_h, -(float)i);   // one rounding; exact when 1/h is representable (h = 1.25e-3 for SevenNet-0)
}

__global__ void radial_bin_kernel(long long E, float inv_h, int n_grid, const float* __restrict__ len, int* __restrict__ bin) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= E) return;
  int i;
  float t;
  locate(len[e], inv_h, n_grid, i, t);
  bin[e] = i;
}

constexpr int kTabEdgesPerCta = 128;   // consecutive (interval-sorted) edges per CTA: they share one or two table rows

// w[e, :] for every edge, visited in the order `order` (edges sorted by interval).  A CTA walks kTabEdgesPerCta
// consecutive positions, 256 / (wn / 4) edges at a time, thread = (edge, 4 consecutive weights): the table row is
// fetched from L2 by the first edge of an interval and served by L1 afterwards; w is written with streaming stores.
__global__ void __launch_bounds__(256)
radial_table_fwd_kernel(int wn, long long E, float inv_h, int n_grid, const float* __restrict__ len, const int* __restrict__ order,
                        const float* __restrict__ T, float* __restrict__ w) {
  const int q = wn >> 2;
  const int epi = max(1, 256 / q);                   // edges per iteration
  const int sub = threadIdx.x / q, k0 = threadIdx.x % q;
  const long long j0 = (long long)blockIdx.x * kTabEdgesPerCta, j1 = min(E, j0 + kTabEdgesPerCta);
  // edge id, interval and local coordinate of the CTA's edges, fetched once (two dependent global loads per edge
  // would otherwise sit in front of every iteration)
  __shared__ int s_e[kTabEdgesPerCta], s_i[kTabEdgesPerCta];
  __shared__ float s_t[kTabEdgesPerCta];
  if (threadIdx.x < kTabEdgesPerCta && j0 + threadIdx.x < j1) {
    const int e = order ? __ldg(order + j0 + threadIdx.x) : (int)(j0 + threadIdx.x);
    s_e[threadIdx.x] = e;
    locate(__ldg(len + e), inv_h, n_grid, s_i[threadIdx.x], s_t[threadIdx.x]);
  }
  __syncthreads();
  if (sub >= epi) return;
  for (long long j = j0 + sub; j < j1; j += epi) {
    const long long e = s_e[j - j0];
    const int i = s_i[j - j0];
    const float t = s_t[j - j0];
    const float4* row = reinterpret_cast<const float4*>(T + (size_t)i * 4 * wn);
    float4* out = reinterpret_cast<float4*>(w + (size_t)e * wn);
    for (int k4 = k0; k4 < q; k4 += 256) {           // one pass unless wn > 1024
      const float4 c0 = __ldg(row + 4 * k4), c1 = __ldg(row + 4 * k4 + 1), c2 = __ldg(row + 4 * k4 + 2), c3 = __ldg(row + 4 * k4 + 3);
      float4 o;
      o.x = fmaf(fmaf(fmaf(c0.w, t, c0.z), t, c0.y), t, c0.x);
      o.y = fmaf(fmaf(fmaf(c1.w, t, c1.z), t, c1.y), t, c1.x);
      o.z = fmaf(fmaf(fmaf(c2.w, t, c2.z), t, c2.y), t, c2.x);
      o.w = fmaf(fmaf(fmaf(c3.w, t, c3.z), t, c3.y), t, c3.x);
      __stcs(out + k4, o);
    }
  }
}

// dE/dr[e] (+)= sum_k dw[e, k] * d w_k / d r: one warp per edge, a CTA walks kTabEdgesPerCta consecutive positions
__global__ void __launch_bounds__(256)
radial_table_bwd_kernel(int wn, long long E, float inv_h, int n_grid, const float* __restrict__ len, const int* __restrict__ order,
                        const float* __restrict__ T, const float* __restrict__ dw, float* __restrict__ dr, int accumulate) {
  const int q = wn >> 2, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long j0 = (long long)blockIdx.x * kTabEdgesPerCta, j1 = min(E, j0 + kTabEdgesPerCta);
  __shared__ int s_e[kTabEdgesPerCta], s_i[kTabEdgesPerCta];
  __shared__ float s_t[kTabEdgesPerCta];
  if (threadIdx.x < kTabEdgesPerCta && j0 + threadIdx.x < j1) {
    const int e = order ? __ldg(order + j0 + threadIdx.x) : (int)(j0 + threadIdx.x);
    s_e[threadIdx.x] = e;
    locate(__ldg(len + e), inv_h, n_grid, s_i[threadIdx.x], s_t[threadIdx.x]);
  }
  __syncthreads();
  for (long long j = j0 + warp; j < j1; j += 8) {
    const long long e = s_e[j - j0];
    const int i = s_i[j - j0];
    const float t = s_t[j - j0];
    const float t2 = 2.f * t, t3 = 3.f * t * t;
    const float4* row = reinterpret_cast<const float4*>(T + (size_t)i * 4 * wn);
    const float4* g = reinterpret_cast<const float4*>(dw + (size_t)e * wn);
    float acc = 0.f;
    for (int k4 = lane; k4 < q; k4 += 32) {
      const float4 c0 = __ldg(row + 4 * k4), c1 = __ldg(row + 4 * k4 + 1), c2 = __ldg(row + 4 * k4 + 2), c3 = __ldg(row + 4 * k4 + 3);
      const float4 gg = __ldcs(g + k4);
      acc = fmaf(gg.x, fmaf(c0.w, t3, fmaf(c0.z, t2, c0.y)), acc);
      acc = fmaf(gg.y, fmaf(c1.w, t3, fmaf(c1.z, t2, c1.y)), acc);
      acc = fmaf(gg.z, fmaf(c2.w, t3, fmaf(c2.z, t2, c2.y)), acc);
      acc = fmaf(gg.w, fmaf(c3.w, t3, fmaf(c3.z, t2, c3.y)), acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) {
      const float v = acc * inv_h;
      dr[e] = accumulate ? dr[e] + v : v;
    }
  }
}

// grid spacing: about r_cut / 4000, adjusted so that the XPLOR switching radius is a node
RadialTableGrid radial_table_grid(const EdgeParams& p) {
  RadialTableGrid g;
  double h = (double)p.r_cut / 4000.0;
  if (p.cutoff_kind == B200_CUTOFF_XPLOR && p.r_on > 0.f && p.r_on < p.r_cut) {
    const double m = floor((double)p.r_on / h + 0.5);
    if (m >= 1.0) h = (double)p.r_on / m;
  }
  g.h = h;
  g.n = (int)ceil((double)p.r_cut / h) + 1;   // intervals 0..n-1 cover r_cut with one interval of slack
  return g;
}

size_t radial_table_floats(const EdgeParams& p, const MlpDesc& d) {
  return (size_t)radial_table_grid(p).n * 4 * d.wn;
}

bool radial_table_supported(const MlpDesc& d) {
  return d.nb >= 1 && d.nb <= 32 && d.h1 >= 1 && d.h1 <= kTabMaxHidden && d.h2 >= 1 && d.h2 <= kTabMaxHidden && d.wn >= 4 &&
         d.wn % 4 == 0;
}

size_t radial_table_packed_bytes(const EdgeParams& p, const MlpDesc& d) {   // AB plane, C plane, D plane
  return (size_t)radial_table_grid(p).n * (d.wn / 2) * (sizeof(float4) + 2 * sizeof(unsigned));
}
int radial_table_build_packed(const EdgeParams& p, const MlpDesc& d, const float* coeffs, const float* W0, const float* W1,
                              const float* W2, void* T, cudaStream_t st) {
  B200_REQUIRE(radial_table_supported(d) && d.wn % 2 == 0, "radial_table_build_packed: unsupported weight_nn shape");
  const RadialTableGrid g = radial_table_grid(p);
  float4* AB = reinterpret_cast<float4*>(T);
  unsigned* Cp = reinterpret_cast<unsigned*>(AB + (size_t)g.n * (d.wn / 2));
  unsigned* Dp = Cp + (size_t)g.n * (d.wn / 2);
  radial_table_build_packed_kernel<<<g.n, 128, 0, st>>>(p, d, g.h, coeffs, W0, W1, W2, AB, Cp, Dp);
  return check_launch("radial_table_build_packed");
}

int radial_table_build(const EdgeParams& p, const MlpDesc& d, const float* coeffs, const float* W0, const float* W1,
                       const float* W2, float* T, cudaStream_t st) {
  B200_REQUIRE(radial_table_supported(d), "radial_table_build: unsupported weight_nn shape");
  const RadialTableGrid g = radial_table_grid(p);
  radial_table_build_kernel<<<g.n, 128, 0, st>>>(p, d, g.h, coeffs, W0, W1, W2, T);
  return check_launch("radial_table_build");
}

__global__ void iota_kernel(long long E, int* __restrict__ v) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e < E) v[e] = (int)e;
}

// Edge ids sorted by table interval.  The order only has to bring edges of the same interval together (L1 reuse of
// the table rows), it carries no arithmetic; one radix-sort pass over the interval bits (CUB, library utility).
// ws: radial_table_order_bytes(E) bytes of scratch.
size_t radial_table_order_bytes(long long E) {
  size_t tmp = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const int*)nullptr, (int*)nullptr, (const int*)nullptr, (int*)nullptr, (int)E, 0, 16);
  return ((tmp + 255) / 256 * 256) + 3 * (((size_t)E * sizeof(int) + 255) / 256 * 256);
}

int radial_table_order(const EdgeParams& p, long long E, const float* len, void* ws, int* order, cudaStream_t st) {
  if (E == 0) return B200_OK;
  const RadialTableGrid g = radial_table_grid(p);
  B200_REQUIRE(g.n < (1 << 16), "radial_table_order: more than 65535 intervals");
  const size_t slab = ((size_t)E * sizeof(int) + 255) / 256 * 256;
  int* bin = (int*)ws;
  int* bin_sorted = (int*)((char*)ws + slab);
  int* ids = (int*)((char*)ws + 2 * slab);
  void* tmp = (char*)ws + 3 * slab;
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const int*)nullptr, (int*)nullptr, (const int*)nullptr, (int*)nullptr, (int)E, 0, 16);
  radial_bin_kernel<<<(unsigned)((E + 255) / 256), 256, 0, st>>>(E, (float)(1.0 / g.h), g.n, len, bin);
  iota_kernel<<<(unsigned)((E + 255) / 256), 256, 0, st>>>(E, ids);
  int bits = 1;
  while ((1 << bits) < g.n) ++bits;
  if (cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, bin, bin_sorted, ids, order, (int)E, 0, bits, st) != cudaSuccess) {
    set_error("radial_table_order: radix sort failed");
    return B200_ERR_CUDA;
  }
  return check_launch("radial_table_order", 3);
}

int radial_table_fwd(const EdgeParams& p, const MlpDesc& d, long long E, const float* len, const int* order, const float* T,
                     float* w, cudaStream_t st) {
  if (E == 0) return B200_OK;
  const RadialTableGrid g = radial_table_grid(p);
  const unsigned blocks = (unsigned)((E + kTabEdgesPerCta - 1) / kTabEdgesPerCta);
  radial_table_fwd_kernel<<<blocks, 256, 0, st>>>(d.wn, E, (float)(1.0 / g.h), g.n, len, order, T, w);
  return check_launch("radial_table_fwd");
}

int radial_table_bwd(const EdgeParams& p, const MlpDesc& d, long long E, const float* len, const int* order, const float* T,
                     const float* dw, float* dr, int accumulate, cudaStream_t st) {
  if (E == 0) return B200_OK;
  const RadialTableGrid g = radial_table_grid(p);
  const unsigned blocks = (unsigned)((E + kTabEdgesPerCta - 1) / kTabEdgesPerCta);
  radial_table_bwd_kernel<<<blocks, 256, 0, st>>>(d.wn, E, (float)(1.0 / g.h), g.n, len, order, T, dw, dr, accumulate);
  return check_launch("radial_table_bwd");
}

}  // namespace b200
