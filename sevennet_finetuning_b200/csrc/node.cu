// Node-side elementwise pieces: gate nonlinearity, species embedding gather, rescale, column sums.
//
// Reference: sevenn/nn/equivariant_gate.py:59-61 (e3nn Gate; frozen graph: c*silu on scalars and on
// gates, gated[u,:] *= gate[u]); sevenn/nn/node_embedding.py:39-48 + nn/linear.py:37-44;
// sevenn/nn/scale.py:35-38,67-73; sevenn/nn/linear.py:76-90 (AtomReduce).
#include "common.cuh"

namespace b200 {

// in  row: [ns scalars | ng = m1+m2 gates | m1 x 3 | m2 x 5];  out row: [ns | m1 x 3 | m2 x 5]
__global__ void gate_fwd_kernel(int ns, int m1, int m2, long long n_nodes, const float* __restrict__ in,
                                float* __restrict__ out) {
  const int din = ns + m1 + m2 + 3 * m1 + 5 * m2, dout = ns + 3 * m1 + 5 * m2;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= n_nodes * dout) return;
  const long long n = idx / dout;
  const int c = (int)(idx % dout);
  const float* r = in + n * din;
  float v;
  if (c < ns) {
    v = kSilu2Mom * silu_f(r[c]);
  } else {
    int u, off;
    if (c < ns + 3 * m1) { u = (c - ns) / 3; off = c - ns; }
    else { u = m1 + (c - ns - 3 * m1) / 5; off = c - ns; }
    v = kSilu2Mom * silu_f(r[ns + u]) * r[ns + m1 + m2 + off];
  }
  out[idx] = v;
}

__global__ void gate_bwd_kernel(int ns, int m1, int m2, long long n_nodes, const float* __restrict__ in,
                                const float* __restrict__ dout_, float* __restrict__ din_) {
  const int ng = m1 + m2, din = ns + ng + 3 * m1 + 5 * m2, dout = ns + 3 * m1 + 5 * m2;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= n_nodes * din) return;
  const long long n = idx / din;
  const int c = (int)(idx % din);
  const float* r = in + n * din;
  const float* g = dout_ + n * dout;
  float v, d;
  if (c < ns) {
    silu_vd(r[c], v, d);
    din_[idx] = g[c] * d;
  } else if (c < ns + ng) {
    const int u = c - ns;
    const int D = u < m1 ? 3 : 5;
    const int off = u < m1 ? u * 3 : 3 * m1 + (u - m1) * 5;
    float s = 0.f;
    for (int m = 0; m < D; ++m) s = fmaf(g[ns + off + m], r[ns + ng + off + m], s);
    silu_vd(r[c], v, d);
    din_[idx] = s * d;
  } else {
    const int off = c - ns - ng;
    const int u = off < 3 * m1 ? off / 3 : m1 + (off - 3 * m1) / 5;
    din_[idx] = g[ns + off] * kSilu2Mom * silu_f(r[ns + u]);
  }
}

// Second-order pass of the gate (reverse of gate_bwd; force-matching fine-tuning differentiates the forces, which are
// themselves a reverse pass: sevenn/nn/force_output.py:165-170 create_graph=self.training).  With phi = c*silu:
//   gate_bwd:  din_s = phi'(s) g_s,  din_t[u] = phi'(t_u) sum_m v[u,m] g_v[u,m],  din_v[u,m] = phi(t_u) g_v[u,m]
// and a = d L / d din, this kernel returns d L / d in (dx) and d L / d g (dg).
__device__ __forceinline__ void silu_vdd(float z, float& v, float& d, float& dd) {
  const float s = 1.0f / (1.0f + expf(-z));
  v = kSilu2Mom * z * s;
  d = kSilu2Mom * s * (1.0f + z * (1.0f - s));
  dd = kSilu2Mom * s * (1.0f - s) * (2.0f + z * (1.0f - 2.0f * s));
}

__global__ void gate_bwd2_dx_kernel(int ns, int m1, int m2, long long n_nodes, const float* __restrict__ in,
                                    const float* __restrict__ g_, const float* __restrict__ a_, float* __restrict__ dx) {
  const int ng = m1 + m2, din = ns + ng + 3 * m1 + 5 * m2, dout = ns + 3 * m1 + 5 * m2;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= n_nodes * din) return;
  const long long n = idx / din;
  const int c = (int)(idx % din);
  const float* r = in + n * din;
  const float* a = a_ + n * din;
  const float* g = g_ + n * dout;
  float v, d, dd;
  if (c < ns) {
    silu_vdd(r[c], v, d, dd);
    dx[idx] = a[c] * dd * g[c];
  } else if (c < ns + ng) {
    const int u = c - ns;
    const int D = u < m1 ? 3 : 5;
    const int off = u < m1 ? u * 3 : 3 * m1 + (u - m1) * 5;
    float svg = 0.f, sag = 0.f;
    for (int m = 0; m < D; ++m) {
      svg = fmaf(g[ns + off + m], r[ns + ng + off + m], svg);
      sag = fmaf(g[ns + off + m], a[ns + ng + off + m], sag);
    }
    silu_vdd(r[c], v, d, dd);
    dx[idx] = a[c] * dd * svg + d * sag;
  } else {
    const int off = c - ns - ng;
    const int u = off < 3 * m1 ? off / 3 : m1 + (off - 3 * m1) / 5;
    silu_vdd(r[ns + u], v, d, dd);
    dx[idx] = a[ns + u] * d * g[ns + off];
  }
}

__global__ void gate_bwd2_dg_kernel(int ns, int m1, int m2, long long n_nodes, const float* __restrict__ in,
                                    const float* __restrict__ a_, float* __restrict__ dg) {
  const int ng = m1 + m2, din = ns + ng + 3 * m1 + 5 * m2, dout = ns + 3 * m1 + 5 * m2;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= n_nodes * dout) return;
  const long long n = idx / dout;
  const int c = (int)(idx % dout);
  const float* r = in + n * din;
  const float* a = a_ + n * din;
  float v, d, dd;
  if (c < ns) {
    silu_vdd(r[c], v, d, dd);
    dg[idx] = a[c] * d;
  } else {
    const int off = c - ns;
    const int u = off < 3 * m1 ? off / 3 : m1 + (off - 3 * m1) / 5;
    silu_vdd(r[ns + u], v, d, dd);
    dg[idx] = a[ns + u] * d * r[ns + ng + off] + a[ns + ng + off] * v;
  }
}

__global__ void embed_fwd_kernel(long long n_nodes, int n_species, int dim, const int* __restrict__ types,
                                 const float* __restrict__ W, float* __restrict__ x) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= n_nodes * dim) return;
  const long long n = idx / dim;
  const int c = (int)(idx % dim);
  x[idx] = W[(size_t)types[n] * dim + c] * (1.0f / sqrtf((float)n_species));
}

__global__ void rescale_fwd_kernel(int n_nodes, const int* __restrict__ types, const float* __restrict__ scaled,
                                   const float* __restrict__ scale, const float* __restrict__ shift, int species_wise,
                                   float* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_nodes) return;
  const int t = species_wise ? types[n] : 0;
  out[n] = scaled[n] * scale[t] + shift[t];
}

// one CTA per column, fp64 accumulation in a fixed order
__global__ void colsum_kernel(int n_rows, int ncol, const float* __restrict__ in, float* __restrict__ out) {
  __shared__ double sh[1024];
  const int c = blockIdx.x;
  double s = 0.0;
  for (int r = threadIdx.x; r < n_rows; r += blockDim.x) s += (double)in[(size_t)r * ncol + c];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[c] = (float)sh[0];
}

static unsigned blocks_for(long long n, int t) { return (unsigned)((n + t - 1) / t); }

int gate_fwd(int ns, int m1, int m2, int n_nodes, const float* in, float* out, cudaStream_t st) {
  if (n_nodes == 0) return B200_OK;
  const long long tot = (long long)n_nodes * (ns + 3 * m1 + 5 * m2);
  gate_fwd_kernel<<<blocks_for(tot, 256), 256, 0, st>>>(ns, m1, m2, n_nodes, in, out);
  return check_launch("gate_fwd");
}
int gate_bwd(int ns, int m1, int m2, int n_nodes, const float* in, const float* dout, float* din, cudaStream_t st) {
  if (n_nodes == 0) return B200_OK;
  const long long tot = (long long)n_nodes * (ns + m1 + m2 + 3 * m1 + 5 * m2);
  gate_bwd_kernel<<<blocks_for(tot, 256), 256, 0, st>>>(ns, m1, m2, n_nodes, in, dout, din);
  return check_launch("gate_bwd");
}
int gate_bwd2(int ns, int m1, int m2, int n_nodes, const float* in, const float* dout, const float* a, float* dx,
              float* dg, cudaStream_t st) {
  if (n_nodes == 0) return B200_OK;
  const long long tin = (long long)n_nodes * (ns + m1 + m2 + 3 * m1 + 5 * m2);
  const long long tout = (long long)n_nodes * (ns + 3 * m1 + 5 * m2);
  gate_bwd2_dx_kernel<<<blocks_for(tin, 256), 256, 0, st>>>(ns, m1, m2, n_nodes, in, dout, a, dx);
  gate_bwd2_dg_kernel<<<blocks_for(tout, 256), 256, 0, st>>>(ns, m1, m2, n_nodes, in, a, dg);
  return check_launch("gate_bwd2");
}
int embed_fwd(int n_nodes, int n_species, int dim, const int* types, const float* W, float* x, cudaStream_t st) {
  if (n_nodes == 0) return B200_OK;
  embed_fwd_kernel<<<blocks_for((long long)n_nodes * dim, 256), 256, 0, st>>>(n_nodes, n_species, dim, types, W, x);
  return check_launch("embed_fwd");
}
int rescale_fwd(int n_nodes, const int* types, const float* scaled, const float* scale, const float* shift,
                int species_wise, float* out, cudaStream_t st) {
  if (n_nodes == 0) return B200_OK;
  rescale_fwd_kernel<<<blocks_for(n_nodes, 256), 256, 0, st>>>(n_nodes, types, scaled, scale, shift, species_wise, out);
  return check_launch("rescale_fwd");
}
int colsum(int n_rows, int ncol, const float* in, float* out, cudaStream_t st) {
  B200_REQUIRE(ncol >= 1 && ncol <= 64, "colsum: ncol out of range");
  colsum_kernel<<<ncol, 1024, 0, st>>>(n_rows, ncol, in, out);
  return check_launch("colsum");
}

// ---- halo exchange: pack the rows a peer asked for / add the rows a peer returned ------------------------------------
// pair_e3gnn_parallel.cpp:803-866 (pack_forward_init / unpack_forward_init: index_select into the send buffer) and
// :867-933 (pack_reverse / unpack_reverse: index_add of the returned gradient rows into their owners).
// One thread per float4 (or float when the width is not a multiple of 4) of the packed buffer.
template <class V>
__global__ void halo_pack_kernel(long long total, int wv, const V* __restrict__ rows, const int* __restrict__ idx,
                                 V* __restrict__ out) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= total) return;
  const long long r = i / wv;
  const int c = (int)(i - r * wv);
  out[i] = rows[(size_t)__ldg(idx + r) * wv + c];
}
// the same owner row can be wanted by several peers: RED.ADD, like the reference's index_add_
__global__ void halo_unpack_add_kernel(long long total, int width, float* __restrict__ rows, const int* __restrict__ idx,
                                       const float* __restrict__ in) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= total) return;
  const long long r = i / width;
  const int c = (int)(i - r * width);
  atomicAdd(rows + (size_t)__ldg(idx + r) * width + c, in[i]);
}

int halo_pack(long long m, int width, const float* rows, const int* idx, float* out, cudaStream_t st) {
  if (m == 0 || width == 0) return B200_OK;
  const bool v4 = width % 4 == 0 && ((uintptr_t)rows % 16 == 0) && ((uintptr_t)out % 16 == 0);
  if (v4) {
    const long long tot = m * (width / 4);
    halo_pack_kernel<float4><<<blocks_for(tot, 256), 256, 0, st>>>(tot, width / 4, reinterpret_cast<const float4*>(rows), idx,
                                                                   reinterpret_cast<float4*>(out));
  } else {
    const long long tot = m * width;
    halo_pack_kernel<float><<<blocks_for(tot, 256), 256, 0, st>>>(tot, width, rows, idx, out);
  }
  return check_launch("halo_pack");
}
int halo_unpack_add(long long m, int width, float* rows, const int* idx, const float* in, cudaStream_t st) {
  if (m == 0 || width == 0) return B200_OK;
  const long long tot = m * width;
  halo_unpack_add_kernel<<<blocks_for(tot, 256), 256, 0, st>>>(tot, width, rows, idx, in);
  return check_launch("halo_unpack_add");
}

}  // namespace b200
