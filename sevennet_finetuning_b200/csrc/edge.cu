// Edge embedding: spherical harmonics to lmax=2 (component normalisation), Bessel radial basis and
// cutoff envelope, forward and backward (d/d edge_vec).
//
// Reference: sevenn/nn/edge_embedding.py:220-230 (EdgeEmbedding.forward), :114-116 (BesselBasis),
// :138-145 (PolynomialCutoff), :163-173 (XPLORCutoff), :190-198 (SphericalEncoding -> e3nn
// SphericalHarmonics(normalize, 'component'); polynomial forms as in the frozen graph).
#include "common.cuh"

namespace b200 {

__device__ __forceinline__ void cutoff_vd(const EdgeParams& p, float r, float& f, float& df) {
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

__global__ void edge_embed_fwd_kernel(EdgeParams p, long long E, const float* __restrict__ vec,
                                      const float* __restrict__ coeffs, float* __restrict__ len,
                                      float* __restrict__ Y, float* __restrict__ emb) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= E) return;
  const float vx = vec[e * 3], vy = vec[e * 3 + 1], vz = vec[e * 3 + 2];
  const float r = sqrtf(vx * vx + vy * vy + vz * vz);
  if (len) len[e] = r;
  float ux = vx, uy = vy, uz = vz;
  if (p.normalize_sph) {
    const float n = fmaxf(r, 1e-12f);
    ux = vx / n; uy = vy / n; uz = vz / n;
  }
  const float s3 = 1.7320508075688772f, s5 = 2.23606797749979f;
  float* y = Y + e * 9;
  y[0] = 1.0f;
  y[1] = s3 * ux; y[2] = s3 * uy; y[3] = s3 * uz;
  const float x2z2 = ux * ux + uz * uz;
  y[4] = s5 * (s3 * ux * uz);
  y[5] = s5 * (s3 * ux * uy);
  y[6] = s5 * (uy * uy - 0.5f * x2z2);
  y[7] = s5 * (s3 * uy * uz);
  y[8] = s5 * (0.5f * s3 * (uz * uz - ux * ux));
  float f, df;
  cutoff_vd(p, r, f, df);
  const float pre = 2.0f / p.r_cut;
  for (int n = 0; n < p.num_basis; ++n) emb[e * p.num_basis + n] = pre * sinf(coeffs[n] * r) / r * f;
}

// dvec = dE/d edge_vec from dE/dY (components 1..8; component 0 is the constant 1) and dE/d emb
__global__ void edge_embed_bwd_kernel(EdgeParams p, long long E, const float* __restrict__ vec,
                                      const float* __restrict__ coeffs, const float* __restrict__ dY,
                                      const float* __restrict__ demb, const float* __restrict__ dr_in,
                                      float* __restrict__ dvec) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= E) return;
  const float vx = vec[e * 3], vy = vec[e * 3 + 1], vz = vec[e * 3 + 2];
  const float r = sqrtf(vx * vx + vy * vy + vz * vz);
  float gx = 0.f, gy = 0.f, gz = 0.f;
  if (dr_in) {   // dE/dr given directly (fused convolution backward)
    const float t = dr_in[e] / r;
    gx = t * vx; gy = t * vy; gz = t * vz;
  }
  if (demb) {
    float f, df;
    cutoff_vd(p, r, f, df);
    const float pre = 2.0f / p.r_cut;
    float dr = 0.f;
    for (int n = 0; n < p.num_basis; ++n) {
      const float k = coeffs[n];
      float s, c;
      sincosf(k * r, &s, &c);
      const float b = pre * s / r;                       // basis value
      const float db = pre * (k * c * r - s) / (r * r);  // its derivative
      dr += demb[e * p.num_basis + n] * (db * f + b * df);
    }
    const float ir = 1.0f / r;
    gx += dr * vx * ir; gy += dr * vy * ir; gz += dr * vz * ir;
  }
  if (dY) {
    const float* g = dY + e * 9;
    const float s3 = 1.7320508075688772f, s5 = 2.23606797749979f, a = 3.872983346207417f;  // sqrt(15)
    float n = 1.0f, ux = vx, uy = vy, uz = vz;
    if (p.normalize_sph) {
      n = fmaxf(r, 1e-12f);
      ux = vx / n; uy = vy / n; uz = vz / n;
    }
    const float g1 = g[1], g2 = g[2], g3 = g[3], g4 = g[4], g5 = g[5], g6 = g[6], g7 = g[7], g8 = g[8];
    float hx = s3 * g1 + a * uz * g4 + a * uy * g5 - s5 * ux * g6 - a * ux * g8;
    float hy = s3 * g2 + a * ux * g5 + 2.0f * s5 * uy * g6 + a * uz * g7;
    float hz = s3 * g3 + a * ux * g4 - s5 * uz * g6 + a * uy * g7 + a * uz * g8;
    if (p.normalize_sph) {
      const float dot = ux * hx + uy * hy + uz * hz;
      hx = (hx - ux * dot) / n; hy = (hy - uy * dot) / n; hz = (hz - uz * dot) / n;
    }
    gx += hx; gy += hy; gz += hz;
  }
  dvec[e * 3] = gx; dvec[e * 3 + 1] = gy; dvec[e * 3 + 2] = gz;
}

// Forces and virial from the per-edge gradient (ForceStressOutputFromEdge, sevenn/nn/force_output.py:158-215):
//   F[i] = sum_{e: edge_index[0]=i} f_e - sum_{e: edge_index[1]=i} f_e ;  virial6 = sum_e [rx fx, ry fy, rz fz, rx fy, ry fz, rz fx]
// One warp per atom, deterministic: row i of the destination CSR and row i of the source CSR.
__global__ void force_virial_kernel(int n_atoms, const int* __restrict__ rowptr, const int* __restrict__ rrowptr,
                                    const int* __restrict__ eid_by_src, const float* __restrict__ vec,
                                    const float* __restrict__ dvec, float* __restrict__ force,
                                    float* __restrict__ virial_atom) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n_atoms) return;
  float fx = 0.f, fy = 0.f, fz = 0.f;
  float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int e = rowptr[warp] + lane; e < rowptr[warp + 1]; e += 32) {
    fx += dvec[(size_t)e * 3]; fy += dvec[(size_t)e * 3 + 1]; fz += dvec[(size_t)e * 3 + 2];
  }
  // the reference scatters the per-edge virial to edge_index[1] (the source atom) before the total sum
  for (int q = rrowptr[warp] + lane; q < rrowptr[warp + 1]; q += 32) {
    const size_t e = eid_by_src[q];
    const float dx = dvec[e * 3], dy = dvec[e * 3 + 1], dz = dvec[e * 3 + 2];
    const float rx = vec[e * 3], ry = vec[e * 3 + 1], rz = vec[e * 3 + 2];
    fx -= dx; fy -= dy; fz -= dz;
    v[0] += rx * dx; v[1] += ry * dy; v[2] += rz * dz; v[3] += rx * dy; v[4] += ry * dz; v[5] += rz * dx;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    fx += __shfl_xor_sync(0xffffffffu, fx, o);
    fy += __shfl_xor_sync(0xffffffffu, fy, o);
    fz += __shfl_xor_sync(0xffffffffu, fz, o);
#pragma unroll
    for (int k = 0; k < 6; ++k) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
  }
  if (lane == 0) {
    force[(size_t)warp * 3] = fx; force[(size_t)warp * 3 + 1] = fy; force[(size_t)warp * 3 + 2] = fz;
    if (virial_atom)
#pragma unroll
      for (int k = 0; k < 6; ++k) virial_atom[(size_t)warp * 6 + k] = v[k];
  }
}

// ---- second-order pass (fine-tuning: the loss differentiates the forces, sevenn/nn/force_output.py:165-170 with
// create_graph=self.training, sevenn/train/trainer.py:173-186) ------------------------------------------------------------
// The first reverse pass is dvec = B(vec, k; gY, gemb) = sum_c gY_c grad Y_c + sum_n gemb_n grad emb_n.  With
// a = dL/d(dvec), S = <a, dvec>:
//   dS/d gY_c   = D_a Y_c,  dS/d gemb_n = D_a emb_n          (directional derivative of the forward map along a)
//   dS/d vec    = D_a B(vec)                                  (Hessian-vector product)
//   dS/d k_n    = sum_e gemb_n D_a (d emb_n / d k_n)          (Bessel frequencies are trainable, edge_embedding.py:95-101)
// All four are the tangent parts of the same forward / reverse formulas evaluated on dual numbers (value, derivative
// along a): the arithmetic below is the float code of the kernels above, templated on the scalar type.
struct Dual {
  float v, d;
};
__device__ __forceinline__ Dual mk(float v, float d = 0.f) { Dual r; r.v = v; r.d = d; return r; }
__device__ __forceinline__ Dual operator+(Dual a, Dual b) { return mk(a.v + b.v, a.d + b.d); }
__device__ __forceinline__ Dual operator-(Dual a, Dual b) { return mk(a.v - b.v, a.d - b.d); }
__device__ __forceinline__ Dual operator*(Dual a, Dual b) { return mk(a.v * b.v, a.d * b.v + a.v * b.d); }
__device__ __forceinline__ Dual operator/(Dual a, Dual b) {
  const float q = a.v / b.v;
  return mk(q, (a.d - q * b.d) / b.v);
}
__device__ __forceinline__ Dual operator*(float a, Dual b) { return mk(a * b.v, a * b.d); }
__device__ __forceinline__ Dual operator*(Dual b, float a) { return mk(a * b.v, a * b.d); }
__device__ __forceinline__ Dual operator+(float a, Dual b) { return mk(a + b.v, b.d); }
__device__ __forceinline__ Dual operator-(float a, Dual b) { return mk(a - b.v, -b.d); }
__device__ __forceinline__ Dual operator-(Dual b, float a) { return mk(b.v - a, b.d); }
__device__ __forceinline__ Dual dsqrt(Dual a) {
  const float s = sqrtf(a.v);
  return mk(s, 0.5f * a.d / s);
}
__device__ __forceinline__ void dsincos(Dual a, Dual& s, Dual& c) {
  float sv, cv;
  sincosf(a.v, &sv, &cv);
  s = mk(sv, cv * a.d);
  c = mk(cv, -sv * a.d);
}
__device__ __forceinline__ Dual dpow(Dual a, float e) { return mk(powf(a.v, e), e * powf(a.v, e - 1.0f) * a.d); }

// cutoff envelope f(r) and f'(r) on duals (same branches and formulas as cutoff_vd)
__device__ __forceinline__ void cutoff_vd_dual(const EdgeParams& p, Dual r, Dual& f, Dual& df) {
  if (p.cutoff_kind == B200_CUTOFF_XPLOR) {
    const float ron2 = p.r_on * p.r_on, rc2 = p.r_cut * p.r_cut;
    if (r.v < p.r_on) {
      f = mk(1.0f);
      df = mk(0.0f);
    } else {
      const float den = (rc2 - ron2) * (rc2 - ron2) * (rc2 - ron2);
      const Dual r2 = r * r, a = rc2 - r2;
      f = a * a * (rc2 + 2.0f * r2 - 3.0f * ron2) * (1.0f / den);
      df = 12.0f * r * a * (ron2 - r2) * (1.0f / den);
    }
  } else {
    const float pp = p.poly_p;
    const Dual x = r * (1.0f / p.r_cut);
    const float c0 = (pp + 1.0f) * (pp + 2.0f) * 0.5f, c1 = pp * (pp + 2.0f), c2 = pp * (pp + 1.0f) * 0.5f;
    const Dual xp1 = dpow(x, pp - 1.0f), xp = xp1 * x, xq = xp * x, xr = xq * x;
    f = 1.0f - c0 * xp + c1 * xq - c2 * xr;
    df = (c1 * (pp + 1.0f) * xp - c0 * pp * xp1 - c2 * (pp + 2.0f) * xq) * (1.0f / p.r_cut);
  }
}

constexpr int kEdge2Threads = 128;

// one thread per edge; dk partial sums per block (fp64) -> partial[blockIdx][num_basis]
__global__ void __launch_bounds__(kEdge2Threads)
edge_embed_bwd2_kernel(EdgeParams p, long long E, const float* __restrict__ vec, const float* __restrict__ coeffs,
                       const float* __restrict__ gY, const float* __restrict__ gemb, const float* __restrict__ a_,
                       float* __restrict__ dvec2, float* __restrict__ dgY, float* __restrict__ dgemb,
                       double* __restrict__ dk_partial) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  __shared__ double red[kEdge2Threads / 32][32];
  float dk_local[32];
#pragma unroll 1
  for (int n = 0; n < p.num_basis; ++n) dk_local[n] = 0.f;
  if (e < E) {
    const Dual vx = mk(vec[e * 3], a_[e * 3]), vy = mk(vec[e * 3 + 1], a_[e * 3 + 1]), vz = mk(vec[e * 3 + 2], a_[e * 3 + 2]);
    const Dual r = dsqrt(vx * vx + vy * vy + vz * vz);
    const float s3 = 1.7320508075688772f, s5 = 2.23606797749979f, sa = 3.872983346207417f;
    Dual n = mk(1.0f), ux = vx, uy = vy, uz = vz;
    if (p.normalize_sph) {
      n = r.v > 1e-12f ? r : mk(1e-12f);
      ux = vx / n; uy = vy / n; uz = vz / n;
    }
    // ---- forward tangents: dS/d gY, dS/d gemb
    if (dgY) {
      float* o = dgY + e * 9;
      const Dual x2z2 = ux * ux + uz * uz;
      o[0] = 0.f;
      o[1] = s3 * ux.d; o[2] = s3 * uy.d; o[3] = s3 * uz.d;
      o[4] = (s5 * s3 * (ux * uz)).d;
      o[5] = (s5 * s3 * (ux * uy)).d;
      o[6] = (s5 * (uy * uy - 0.5f * x2z2)).d;
      o[7] = (s5 * s3 * (uy * uz)).d;
      o[8] = (s5 * 0.5f * s3 * (uz * uz - ux * ux)).d;
    }
    Dual f, df;
    cutoff_vd_dual(p, r, f, df);
    const float pre = 2.0f / p.r_cut;
    Dual gx = mk(0.f), gy = mk(0.f), gz = mk(0.f);
    // ---- radial part: reverse formulas of edge_embed_bwd_kernel on duals
    Dual dr = mk(0.f);
    for (int q = 0; q < p.num_basis; ++q) {
      const float k = coeffs[q];
      Dual s, c;
      dsincos(k * r, s, c);
      const Dual b = pre * s / r;
      const Dual db = pre * (k * c * r - s) / (r * r);
      if (dgemb) dgemb[e * p.num_basis + q] = (b * f).d;
      if (gemb) {
        const float ge = gemb[e * p.num_basis + q];
        dr = dr + ge * (db * f + b * df);
        dk_local[q] = ge * (pre * c * f).d;    // d emb_q / d k_q = pre cos(k r) f(r)
      }
    }
    if (gemb) {
      const Dual ir = mk(1.0f) / r;
      gx = gx + dr * vx * ir; gy = gy + dr * vy * ir; gz = gz + dr * vz * ir;
    }
    if (gY) {
      const float* g = gY + e * 9;
      const float g1 = g[1], g2 = g[2], g3 = g[3], g4 = g[4], g5 = g[5], g6 = g[6], g7 = g[7], g8 = g[8];
      Dual hx = s3 * g1 + (sa * g4) * uz + (sa * g5) * uy - (s5 * g6) * ux - (sa * g8) * ux;
      Dual hy = s3 * g2 + (sa * g5) * ux + (2.0f * s5 * g6) * uy + (sa * g7) * uz;
      Dual hz = s3 * g3 + (sa * g4) * ux - (s5 * g6) * uz + (sa * g7) * uy + (sa * g8) * uz;
      if (p.normalize_sph) {
        const Dual dot = ux * hx + uy * hy + uz * hz;
        hx = (hx - ux * dot) / n; hy = (hy - uy * dot) / n; hz = (hz - uz * dot) / n;
      }
      gx = gx + hx; gy = gy + hy; gz = gz + hz;
    }
    dvec2[e * 3] = gx.d; dvec2[e * 3 + 1] = gy.d; dvec2[e * 3 + 2] = gz.d;
  }
  if (dk_partial) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int q = 0; q < p.num_basis; ++q) {
      double v = (double)dk_local[q];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) red[warp][q] = v;
    }
    __syncthreads();
    if (threadIdx.x < p.num_basis) {
      double t = 0.0;
      for (int w = 0; w < kEdge2Threads / 32; ++w) t += red[w][threadIdx.x];
      dk_partial[(size_t)blockIdx.x * p.num_basis + threadIdx.x] = t;
    }
  }
}

// first-order coefficient gradient d k_n = sum_e gemb[e, n] (2 / r_c) cos(k_n r) f(r), same two-stage reduction
__global__ void __launch_bounds__(kEdge2Threads)
edge_embed_dk_kernel(EdgeParams p, long long E, const float* __restrict__ vec, const float* __restrict__ coeffs,
                     const float* __restrict__ gemb, double* __restrict__ dk_partial) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  __shared__ double red[kEdge2Threads / 32][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float r = 1.f, f = 0.f, df;
  if (e < E) {
    const float vx = vec[e * 3], vy = vec[e * 3 + 1], vz = vec[e * 3 + 2];
    r = sqrtf(vx * vx + vy * vy + vz * vz);
    cutoff_vd(p, r, f, df);
  }
  for (int q = 0; q < p.num_basis; ++q) {
    double v = e < E ? (double)(gemb[e * p.num_basis + q] * (2.0f / p.r_cut) * cosf(coeffs[q] * r) * f) : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][q] = v;
  }
  __syncthreads();
  if (threadIdx.x < p.num_basis) {
    double t = 0.0;
    for (int w = 0; w < kEdge2Threads / 32; ++w) t += red[w][threadIdx.x];
    dk_partial[(size_t)blockIdx.x * p.num_basis + threadIdx.x] = t;
  }
}

__global__ void edge_dk_reduce_kernel(int nblocks, int nb, const double* __restrict__ partial, float* __restrict__ dk) {
  __shared__ double sh[256];
  const int q = blockIdx.x;
  double s = 0.0;
  for (int b = threadIdx.x; b < nblocks; b += blockDim.x) s += partial[(size_t)b * nb + q];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) dk[q] = (float)sh[0];
}

int edge_embed_bwd2(const EdgeParams& p, long long E, const float* vec, const float* coeffs, const float* gY,
                    const float* gemb, const float* a, float* dvec2, float* dgY, float* dgemb, float* dk, cudaStream_t st) {
  B200_REQUIRE(p.num_basis >= 1 && p.num_basis <= 32, "edge_embed_bwd2: num_basis out of range");
  if (E == 0) {
    if (dk) cudaMemsetAsync(dk, 0, sizeof(float) * p.num_basis, st);
    return B200_OK;
  }
  const unsigned nblocks = (unsigned)((E + kEdge2Threads - 1) / kEdge2Threads);
  double* partial = nullptr;
  if (dk) {
    keep_async_pool();
    if (cudaMallocAsync((void**)&partial, sizeof(double) * (size_t)nblocks * p.num_basis, st) != cudaSuccess) {
      set_error("edge_embed_bwd2: cudaMallocAsync failed");
      return B200_ERR_CUDA;
    }
  }
  edge_embed_bwd2_kernel<<<nblocks, kEdge2Threads, 0, st>>>(p, E, vec, coeffs, gY, gemb, a, dvec2, dgY, dgemb, partial);
  if (dk) {
    edge_dk_reduce_kernel<<<p.num_basis, 256, 0, st>>>((int)nblocks, p.num_basis, partial, dk);
    cudaFreeAsync(partial, st);
  }
  return check_launch("edge_embed_bwd2", dk ? 2 : 1);
}

int edge_embed_dk(const EdgeParams& p, long long E, const float* vec, const float* coeffs, const float* gemb, float* dk,
                  cudaStream_t st) {
  B200_REQUIRE(p.num_basis >= 1 && p.num_basis <= 32, "edge_embed_dk: num_basis out of range");
  if (E == 0) {
    cudaMemsetAsync(dk, 0, sizeof(float) * p.num_basis, st);
    return B200_OK;
  }
  const unsigned nblocks = (unsigned)((E + kEdge2Threads - 1) / kEdge2Threads);
  double* partial = nullptr;
  keep_async_pool();
  if (cudaMallocAsync((void**)&partial, sizeof(double) * (size_t)nblocks * p.num_basis, st) != cudaSuccess) {
    set_error("edge_embed_dk: cudaMallocAsync failed");
    return B200_ERR_CUDA;
  }
  edge_embed_dk_kernel<<<nblocks, kEdge2Threads, 0, st>>>(p, E, vec, coeffs, gemb, partial);
  edge_dk_reduce_kernel<<<p.num_basis, 256, 0, st>>>((int)nblocks, p.num_basis, partial, dk);
  cudaFreeAsync(partial, st);
  return check_launch("edge_embed_dk", 2);
}

int edge_embed_fwd(const EdgeParams& p, long long E, const float* vec, const float* coeffs, float* len, float* Y,
                   float* emb, cudaStream_t st) {
  B200_REQUIRE(p.num_basis >= 1 && p.num_basis <= 32, "edge_embed_fwd: num_basis out of range");
  if (E == 0) return B200_OK;
  edge_embed_fwd_kernel<<<(unsigned)((E + 255) / 256), 256, 0, st>>>(p, E, vec, coeffs, len, Y, emb);
  return check_launch("edge_embed_fwd");
}

int edge_embed_bwd(const EdgeParams& p, long long E, const float* vec, const float* coeffs, const float* dY,
                   const float* demb, float* dvec, cudaStream_t st) {
  if (E == 0) return B200_OK;
  edge_embed_bwd_kernel<<<(unsigned)((E + 255) / 256), 256, 0, st>>>(p, E, vec, coeffs, dY, demb, nullptr, dvec);
  return check_launch("edge_embed_bwd");
}

int edge_embed_bwd_full(const EdgeParams& p, long long E, const float* vec, const float* coeffs, const float* dY,
                        const float* demb, const float* dr, float* dvec, cudaStream_t st) {
  if (E == 0) return B200_OK;
  edge_embed_bwd_kernel<<<(unsigned)((E + 255) / 256), 256, 0, st>>>(p, E, vec, coeffs, dY, demb, dr, dvec);
  return check_launch("edge_embed_bwd");
}

int edge_embed_bwd_r(const EdgeParams& p, long long E, const float* vec, const float* dY, const float* dr, float* dvec,
                     cudaStream_t st) {
  if (E == 0) return B200_OK;
  edge_embed_bwd_kernel<<<(unsigned)((E + 255) / 256), 256, 0, st>>>(p, E, vec, nullptr, dY, nullptr, dr, dvec);
  return check_launch("edge_embed_bwd_r");
}

int force_virial(int n_atoms, const int* rowptr, const int* rrowptr, const int* eid_by_src, const float* vec,
                 const float* dvec, float* force, float* virial_atom, cudaStream_t st) {
  if (n_atoms == 0) return B200_OK;
  const long long threads = (long long)n_atoms * 32;
  force_virial_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(n_atoms, rowptr, rrowptr, eid_by_src, vec, dvec,
                                                                         force, virial_atom);
  return check_launch("force_virial");
}

}  // namespace b200
