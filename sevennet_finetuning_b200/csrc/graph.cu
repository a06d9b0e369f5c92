// Fused radius-graph construction (cell list, float64, integer outputs) and CSR transpose.
//
// Reference: sevenn/train/dataload.py:31-68 - ase.neighborlist.primitive_neighbor_list('ijDS', pbc, cell,
// pos, cutoff, self_interaction=True) followed by removal of the zero-length self pair (:40-47); LAMMPS
// analogue pair_e3gnn.cpp:155-182 (Rij < cutoff^2 in double).  Contract: every ordered (i, j, S) with
// |pos[j] - pos[i] + S.cell| < cutoff, S integer (0 along non-periodic axes), except (i == j, S == 0).
//
// The accept test uses one fixed float64 expression with no FMA contraction, so the edge set is a pure
// function of the inputs and can be compared bit for bit:
//     D_k = (pos[j,k] - pos[i,k]) + ((S0*cell[0,k] + S1*cell[1,k]) + S2*cell[2,k])
//     d2  = (D0*D0 + D1*D1) + D2*D2 ;  edge <=> d2 < cutoff*cutoff
// Output rows (one per centre atom i) are sorted by (j, S0, S1, S2): the list is the canonical
// lexicographic order and is already the dst-sorted CSR the convolution kernels want.
#include <math.h>

#include "common.cuh"

namespace b200 {

struct GridHost {      // computed on the host from the 3x3 cell
  double inv[9];       // frac = pos @ inv
  double cell[9];      // true cell (zero rows are fine) used for the shift vectors
  double h[3];         // heights of the binning cell along each axis
  int pbc[3];
  double cutoff, cut2;
  int max_bins;
};
struct GridDev {       // written by the setup kernel
  double lo[3], ext[3];
  int nb[3], R[3];
  int nbins;
};

struct GraphWs {       // layout of the caller-provided workspace
  GridDev* grid;
  int* bin_of;         // [N]
  int* off;            // [N,3] floor of the fractional coordinates (periodic axes)
  int* bin_count;      // [max_bins+1]
  int* bin_start;      // [max_bins+1]
  int* bin_cursor;     // [max_bins]
  int* bin_atoms;      // [N]
  int* deg;            // [N]
  unsigned long long* tmp_key;  // [E] filled in phase 2 (allocated lazily from the tail)
};

static size_t align_up(size_t v) { return (v + 255) / 256 * 256; }
static int graph_max_bins(int n) { return 2 * n + 64; }

size_t graph_workspace_bytes(int n) {
  const size_t nb = (size_t)graph_max_bins(n);
  return align_up(sizeof(GridDev)) + align_up(sizeof(int) * (size_t)n) * 3 + align_up(sizeof(int) * 3 * (size_t)n) +
         align_up(sizeof(int) * (nb + 1)) * 3 + 1024;
}

static GraphWs carve(void* ws, int n) {
  GraphWs g;
  char* p = (char*)ws;
  const size_t nb = (size_t)graph_max_bins(n);
  g.grid = (GridDev*)p; p += align_up(sizeof(GridDev));
  g.bin_of = (int*)p; p += align_up(sizeof(int) * (size_t)n);
  g.bin_atoms = (int*)p; p += align_up(sizeof(int) * (size_t)n);
  g.deg = (int*)p; p += align_up(sizeof(int) * (size_t)n);
  g.off = (int*)p; p += align_up(sizeof(int) * 3 * (size_t)n);
  g.bin_count = (int*)p; p += align_up(sizeof(int) * (nb + 1));
  g.bin_start = (int*)p; p += align_up(sizeof(int) * (nb + 1));
  g.bin_cursor = (int*)p; p += align_up(sizeof(int) * (nb + 1));
  g.tmp_key = nullptr;
  return g;
}

__device__ __forceinline__ void frac_of(const GridHost& G, const double* __restrict__ pos, int i, double f[3]) {
  const double x = pos[3 * (size_t)i], y = pos[3 * (size_t)i + 1], z = pos[3 * (size_t)i + 2];
#pragma unroll
  for (int k = 0; k < 3; ++k) f[k] = x * G.inv[k] + y * G.inv[3 + k] + z * G.inv[6 + k];
}

// single CTA: min/max of the fractional coordinates, then the bin grid
__global__ void grid_setup_kernel(GridHost G, int n, const double* __restrict__ pos, GridDev* out) {
  __shared__ double smin[3][512], smax[3][512];
  double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    double f[3];
    frac_of(G, pos, i, f);
#pragma unroll
    for (int k = 0; k < 3; ++k) { mn[k] = fmin(mn[k], f[k]); mx[k] = fmax(mx[k], f[k]); }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) { smin[k][threadIdx.x] = mn[k]; smax[k][threadIdx.x] = mx[k]; }
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o)
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        smin[k][threadIdx.x] = fmin(smin[k][threadIdx.x], smin[k][threadIdx.x + o]);
        smax[k][threadIdx.x] = fmax(smax[k][threadIdx.x], smax[k][threadIdx.x + o]);
      }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    GridDev g;
    const double rc = G.cutoff * (1.0 + 1e-6) + 1e-9;
    double width[3];
    for (int k = 0; k < 3; ++k) {
      if (G.pbc[k]) { g.lo[k] = 0.0; g.ext[k] = 1.0; }
      else { g.lo[k] = smin[k][0]; g.ext[k] = fmax(smax[k][0] - smin[k][0], 1e-9) * (1.0 + 1e-9) + 1e-12; }
      width[k] = G.h[k] * g.ext[k];                     // extent along the axis normal
      long long nb = (long long)floor(width[k] / rc);
      g.nb[k] = (int)max(1LL, min(nb, 1LL << 20));
    }
    // cap the total number of bins (coarser bins stay correct)
    while ((long long)g.nb[0] * g.nb[1] * g.nb[2] > (long long)G.max_bins) {
      int a = 0;
      if (g.nb[1] > g.nb[a]) a = 1;
      if (g.nb[2] > g.nb[a]) a = 2;
      g.nb[a] = max(1, g.nb[a] / 2);
    }
    for (int k = 0; k < 3; ++k) {
      const double bw = width[k] / g.nb[k];
      int R = (int)ceil(rc / bw);
      if (!G.pbc[k]) R = min(R, g.nb[k] - 1);
      g.R[k] = max(R, G.pbc[k] ? 1 : 0);
    }
    g.nbins = g.nb[0] * g.nb[1] * g.nb[2];
    *out = g;
  }
}

__global__ void bin_atoms_kernel(GridHost G, int n, const double* __restrict__ pos, const GridDev* __restrict__ gd,
                                 int* __restrict__ bin_of, int* __restrict__ off, int* __restrict__ bin_count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const GridDev g = *gd;
  double f[3];
  frac_of(G, pos, i, f);
  int bc[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    double w;
    int o = 0;
    if (G.pbc[k]) {
      const double fl = floor(f[k]);
      o = (int)fl;
      w = f[k] - fl;
    } else {
      w = (f[k] - g.lo[k]) / g.ext[k];
    }
    off[3 * (size_t)i + k] = o;
    int b = (int)(w * g.nb[k]);
    bc[k] = min(max(b, 0), g.nb[k] - 1);
  }
  const int b = (bc[0] * g.nb[1] + bc[1]) * g.nb[2] + bc[2];
  bin_of[i] = b;
  atomicAdd(bin_count + b, 1);
}

// exclusive scan of in[0..n) into out[0..n], out[n] = total; single CTA, chunked
__global__ void exclusive_scan_kernel(const int* __restrict__ in, int* __restrict__ out, const int* n_dev, int n_host) {
  __shared__ int sh[1024];
  __shared__ int carry;
  const int n = n_dev ? *n_dev : n_host;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 1024 * 4) {
    int v[4], s = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int idx = base + threadIdx.x * 4 + q;
      v[q] = idx < n ? in[idx] : 0;
      s += v[q];
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
      int t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
      __syncthreads();
      sh[threadIdx.x] += t;
      __syncthreads();
    }
    int run = carry + sh[threadIdx.x] - s;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int idx = base + threadIdx.x * 4 + q;
      if (idx < n) out[idx] = run;
      run += v[q];
    }
    __syncthreads();
    if (threadIdx.x == 1023) carry += sh[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[n] = carry;
}

__global__ void fill_bins_kernel(int n, const int* __restrict__ bin_of, const int* __restrict__ bin_start,
                                 int* __restrict__ cursor, int* __restrict__ bin_atoms) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = bin_of[i];
  bin_atoms[bin_start[b] + atomicAdd(cursor + b, 1)] = i;
}

__device__ __forceinline__ unsigned long long pack_key(int j, int s0, int s1, int s2) {
  return ((unsigned long long)(unsigned)j << 32) | ((unsigned long long)(unsigned)(s0 + 512) << 20) |
         ((unsigned long long)(unsigned)(s1 + 512) << 10) | (unsigned long long)(unsigned)(s2 + 512);
}

// One warp per centre atom.  FILL = false: count accepted pairs; FILL = true: write the row (sorted).
template <bool FILL>
__global__ void __launch_bounds__(256)
neighbor_kernel(GridHost G, int n, const double* __restrict__ pos, const GridDev* __restrict__ gd,
                const int* __restrict__ bin_of, const int* __restrict__ off, const int* __restrict__ bin_start,
                const int* __restrict__ bin_atoms, int* __restrict__ deg, const int* __restrict__ rowptr,
                unsigned long long* __restrict__ tmp_key, int* __restrict__ src, int* __restrict__ dst,
                int* __restrict__ shift, float* __restrict__ edge_vec, int* __restrict__ err) {
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (i >= n) return;
  const GridDev g = *gd;
  const int b = bin_of[i];
  const int bc[3] = {b / (g.nb[1] * g.nb[2]), (b / g.nb[2]) % g.nb[1], b % g.nb[2]};
  const double pix = pos[3 * (size_t)i], piy = pos[3 * (size_t)i + 1], piz = pos[3 * (size_t)i + 2];
  const int oi[3] = {off[3 * (size_t)i], off[3 * (size_t)i + 1], off[3 * (size_t)i + 2]};
  int count = 0;
  const int base = FILL ? rowptr[i] : 0;
  for (int d0 = -g.R[0]; d0 <= g.R[0]; ++d0) {
    int c0 = bc[0] + d0, S0 = 0;
    if (G.pbc[0]) { S0 = (c0 >= 0) ? c0 / g.nb[0] : -((-c0 + g.nb[0] - 1) / g.nb[0]); c0 -= S0 * g.nb[0]; }
    else if (c0 < 0 || c0 >= g.nb[0]) continue;
    for (int d1 = -g.R[1]; d1 <= g.R[1]; ++d1) {
      int c1 = bc[1] + d1, S1 = 0;
      if (G.pbc[1]) { S1 = (c1 >= 0) ? c1 / g.nb[1] : -((-c1 + g.nb[1] - 1) / g.nb[1]); c1 -= S1 * g.nb[1]; }
      else if (c1 < 0 || c1 >= g.nb[1]) continue;
      for (int d2 = -g.R[2]; d2 <= g.R[2]; ++d2) {
        int c2 = bc[2] + d2, S2 = 0;
        if (G.pbc[2]) { S2 = (c2 >= 0) ? c2 / g.nb[2] : -((-c2 + g.nb[2] - 1) / g.nb[2]); c2 -= S2 * g.nb[2]; }
        else if (c2 < 0 || c2 >= g.nb[2]) continue;
        const int cb = (c0 * g.nb[1] + c1) * g.nb[2] + c2;
        const int q0 = bin_start[cb], q1 = bin_start[cb + 1];
        for (int q = q0 + lane; q < q1 + ((32 - (q1 - q0) % 32) % 32); q += 32) {
          bool keep = false;
          int j = 0, s0 = 0, s1 = 0, s2 = 0;
          double D0 = 0, D1 = 0, D2 = 0;
          if (q < q1) {
            j = bin_atoms[q];
            s0 = G.pbc[0] ? S0 - off[3 * (size_t)j] + oi[0] : 0;
            s1 = G.pbc[1] ? S1 - off[3 * (size_t)j + 1] + oi[1] : 0;
            s2 = G.pbc[2] ? S2 - off[3 * (size_t)j + 2] + oi[2] : 0;
            const double f0 = (double)s0, f1 = (double)s1, f2 = (double)s2;
            const double sh0 = __dadd_rn(__dadd_rn(__dmul_rn(f0, G.cell[0]), __dmul_rn(f1, G.cell[3])), __dmul_rn(f2, G.cell[6]));
            const double sh1 = __dadd_rn(__dadd_rn(__dmul_rn(f0, G.cell[1]), __dmul_rn(f1, G.cell[4])), __dmul_rn(f2, G.cell[7]));
            const double sh2 = __dadd_rn(__dadd_rn(__dmul_rn(f0, G.cell[2]), __dmul_rn(f1, G.cell[5])), __dmul_rn(f2, G.cell[8]));
            D0 = __dadd_rn(__dadd_rn(pos[3 * (size_t)j], -pix), sh0);
            D1 = __dadd_rn(__dadd_rn(pos[3 * (size_t)j + 1], -piy), sh1);
            D2 = __dadd_rn(__dadd_rn(pos[3 * (size_t)j + 2], -piz), sh2);
            const double dd = __dadd_rn(__dadd_rn(__dmul_rn(D0, D0), __dmul_rn(D1, D1)), __dmul_rn(D2, D2));
            keep = dd < G.cut2 && !(i == j && s0 == 0 && s1 == 0 && s2 == 0);
          }
          const unsigned m = __ballot_sync(0xffffffffu, keep);
          // image shifts beyond the 10-bit fields of the packed sort key: flagged in the counting pass, whose result the
          // host reads back anyway (graph_count fails with B200_ERR_INVALID)
          if (!FILL && keep && err && (abs(s0) > 511 || abs(s1) > 511 || abs(s2) > 511)) atomicExch(err, 1);
          if (FILL && keep) {
            const int p = base + count + __popc(m & ((1u << lane) - 1u));
            tmp_key[p] = pack_key(j, s0, s1, s2);
          }
          count += __popc(m);
        }
      }
    }
  }
  if (!FILL) {
    if (lane == 0) deg[i] = count;
    return;
  }
  __syncwarp();
  // rank sort of the row by key (keys are unique), then recompute the geometry in canonical form
  for (int a = lane; a < count; a += 32) {
    const unsigned long long ka = tmp_key[base + a];
    int rank = 0;
    for (int c = 0; c < count; ++c) rank += tmp_key[base + c] < ka;
    const int j = (int)(ka >> 32);
    const int s0 = (int)((ka >> 20) & 1023) - 512, s1 = (int)((ka >> 10) & 1023) - 512, s2 = (int)(ka & 1023) - 512;
    const double f0 = (double)s0, f1 = (double)s1, f2 = (double)s2;
    const double sh0 = __dadd_rn(__dadd_rn(__dmul_rn(f0, G.cell[0]), __dmul_rn(f1, G.cell[3])), __dmul_rn(f2, G.cell[6]));
    const double sh1 = __dadd_rn(__dadd_rn(__dmul_rn(f0, G.cell[1]), __dmul_rn(f1, G.cell[4])), __dmul_rn(f2, G.cell[7]));
    const double sh2 = __dadd_rn(__dadd_rn(__dmul_rn(f0, G.cell[2]), __dmul_rn(f1, G.cell[5])), __dmul_rn(f2, G.cell[8]));
    const size_t e = (size_t)base + rank;
    src[e] = j;
    dst[e] = i;
    shift[3 * e] = s0; shift[3 * e + 1] = s1; shift[3 * e + 2] = s2;
    edge_vec[3 * e] = (float)__dadd_rn(__dadd_rn(pos[3 * (size_t)j], -pix), sh0);
    edge_vec[3 * e + 1] = (float)__dadd_rn(__dadd_rn(pos[3 * (size_t)j + 1], -piy), sh1);
    edge_vec[3 * e + 2] = (float)__dadd_rn(__dadd_rn(pos[3 * (size_t)j + 2], -piz), sh2);
  }
}

static int make_grid_host(const double* cell, const int* pbc, double cutoff, int n, GridHost& G) {
  const bool any = pbc[0] || pbc[1] || pbc[2];
  double c[9];
  for (int k = 0; k < 9; ++k) c[k] = cell[k];
  double det = c[0] * (c[4] * c[8] - c[5] * c[7]) - c[1] * (c[3] * c[8] - c[5] * c[6]) + c[2] * (c[3] * c[7] - c[4] * c[6]);
  const bool use_cell = any && fabs(det) > 1e-12;
  if (any && !use_cell) {
    set_error("radius_graph: periodic axes need a non-singular cell");
    return B200_ERR_INVALID;
  }
  double e[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  if (use_cell) for (int k = 0; k < 9; ++k) e[k] = c[k];
  else det = 1.0;
  // inverse: inv = adj(e)/det, frac = pos @ inv
  G.inv[0] = (e[4] * e[8] - e[5] * e[7]) / det; G.inv[1] = (e[2] * e[7] - e[1] * e[8]) / det; G.inv[2] = (e[1] * e[5] - e[2] * e[4]) / det;
  G.inv[3] = (e[5] * e[6] - e[3] * e[8]) / det; G.inv[4] = (e[0] * e[8] - e[2] * e[6]) / det; G.inv[5] = (e[2] * e[3] - e[0] * e[5]) / det;
  G.inv[6] = (e[3] * e[7] - e[4] * e[6]) / det; G.inv[7] = (e[1] * e[6] - e[0] * e[7]) / det; G.inv[8] = (e[0] * e[4] - e[1] * e[3]) / det;
  for (int k = 0; k < 3; ++k) {
    const double* a = e + 3 * ((k + 1) % 3);
    const double* b = e + 3 * ((k + 2) % 3);
    const double cx = a[1] * b[2] - a[2] * b[1], cy = a[2] * b[0] - a[0] * b[2], cz = a[0] * b[1] - a[1] * b[0];
    G.h[k] = fabs(det) / sqrt(cx * cx + cy * cy + cz * cz);
    G.pbc[k] = use_cell ? (pbc[k] != 0) : 0;
  }
  for (int k = 0; k < 9; ++k) G.cell[k] = use_cell ? c[k] : 0.0;
  G.cutoff = cutoff;
  G.cut2 = cutoff * cutoff;
  G.max_bins = graph_max_bins(n);
  return B200_OK;
}

int graph_count(int n, const double* pos, const double* cell_host, const int* pbc_host, double cutoff, void* ws,
                int* rowptr, long long* n_edges_host, cudaStream_t st) {
  B200_REQUIRE(n >= 0 && cutoff > 0, "radius_graph: bad arguments");
  if (n == 0) {
    cudaMemsetAsync(rowptr, 0, sizeof(int), st);
    *n_edges_host = 0;
    return B200_OK;
  }
  GridHost G;
  if (int rc = make_grid_host(cell_host, pbc_host, cutoff, n, G)) return rc;
  GraphWs w = carve(ws, n);
  grid_setup_kernel<<<1, 512, 0, st>>>(G, n, pos, w.grid);
  cudaMemsetAsync(w.bin_count, 0, sizeof(int) * (size_t)(G.max_bins + 1), st);
  cudaMemsetAsync(w.bin_cursor, 0, sizeof(int) * (size_t)(G.max_bins + 1), st);
  const int nb = (n + 255) / 256;
  bin_atoms_kernel<<<nb, 256, 0, st>>>(G, n, pos, w.grid, w.bin_of, w.off, w.bin_count);
  exclusive_scan_kernel<<<1, 1024, 0, st>>>(w.bin_count, w.bin_start, &w.grid->nbins, 0);
  fill_bins_kernel<<<nb, 256, 0, st>>>(n, w.bin_of, w.bin_start, w.bin_cursor, w.bin_atoms);
  const unsigned nwb = (unsigned)(((long long)n * 32 + 255) / 256);
  int* err = w.bin_cursor;   // free again after fill_bins_kernel: the overflow flag of the packed sort key
  cudaMemsetAsync(err, 0, sizeof(int), st);
  neighbor_kernel<false><<<nwb, 256, 0, st>>>(G, n, pos, w.grid, w.bin_of, w.off, w.bin_start, w.bin_atoms, w.deg, nullptr,
                                              nullptr, nullptr, nullptr, nullptr, nullptr, err);
  exclusive_scan_kernel<<<1, 1024, 0, st>>>(w.deg, rowptr, nullptr, n);
  int total = 0, overflow = 0;
  cudaMemcpyAsync(&total, rowptr + n, sizeof(int), cudaMemcpyDeviceToHost, st);
  cudaMemcpyAsync(&overflow, err, sizeof(int), cudaMemcpyDeviceToHost, st);
  cudaStreamSynchronize(st);
  *n_edges_host = total;
  B200_REQUIRE(overflow == 0, "radius_graph: an edge needs a periodic image shift beyond +-511 cells (cell much thinner than the "
                              "cutoff, or atoms far outside the cell): not representable in the packed sort key");
  return check_launch("graph_count", 6);
}

int graph_fill(int n, const double* pos, const double* cell_host, const int* pbc_host, double cutoff, void* ws,
               const int* rowptr, int* src, int* dst, int* shift, float* edge_vec, unsigned long long* tmp_key,
               cudaStream_t st) {
  if (n == 0) return B200_OK;
  GridHost G;
  if (int rc = make_grid_host(cell_host, pbc_host, cutoff, n, G)) return rc;
  GraphWs w = carve(ws, n);
  const unsigned nwb = (unsigned)(((long long)n * 32 + 255) / 256);
  neighbor_kernel<true><<<nwb, 256, 0, st>>>(G, n, pos, w.grid, w.bin_of, w.off, w.bin_start, w.bin_atoms, nullptr, rowptr,
                                             tmp_key, src, dst, shift, edge_vec, nullptr);
  return check_launch("graph_fill");
}

// ---- CSR transpose: edge ids sorted by source, ties in increasing edge id ---------------------------
__global__ void histogram_kernel(long long E, const int* __restrict__ key, int* __restrict__ count) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e < E) atomicAdd(count + key[e], 1);
}
__global__ void scatter_by_key_kernel(long long E, const int* __restrict__ key, const int* __restrict__ start,
                                      int* __restrict__ cursor, int* __restrict__ out) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e < E) out[start[key[e]] + atomicAdd(cursor + key[e], 1)] = (int)e;
}
// warp per row: rank-sort the row's edge ids in place (via registers for rows <= 32*8, else insertion by lane 0)
__global__ void sort_rows_kernel(int n, const int* __restrict__ start, int* __restrict__ ids) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= n) return;
  const int b = start[r], cnt = start[r + 1] - b;
  if (cnt <= 256) {
    int v[8], rk[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int a = lane + 32 * q;
      v[q] = a < cnt ? ids[b + a] : 0x7fffffff;
      rk[q] = 0;
    }
    for (int c = 0; c < cnt; ++c) {
      const int x = ids[b + c];
#pragma unroll
      for (int q = 0; q < 8; ++q) rk[q] += x < v[q];
    }
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 8; ++q)
      if (lane + 32 * q < cnt) ids[b + rk[q]] = v[q];
  } else if (lane == 0) {
    for (int a = 1; a < cnt; ++a) {
      const int x = ids[b + a];
      int c = a - 1;
      while (c >= 0 && ids[b + c] > x) { ids[b + c + 1] = ids[b + c]; --c; }
      ids[b + c + 1] = x;
    }
  }
}

size_t transpose_workspace_bytes(int n, long long E) { return align_up(sizeof(int) * ((size_t)n + 1)) * 2 + 256; }

int csr_transpose(int n, long long E, const int* src, void* ws, int* rrowptr, int* eid, cudaStream_t st) {
  int* count = (int*)ws;
  int* cursor = (int*)((char*)ws + align_up(sizeof(int) * ((size_t)n + 1)));
  cudaMemsetAsync(count, 0, sizeof(int) * ((size_t)n + 1), st);
  cudaMemsetAsync(cursor, 0, sizeof(int) * ((size_t)n + 1), st);
  if (E > 0) histogram_kernel<<<(unsigned)((E + 255) / 256), 256, 0, st>>>(E, src, count);
  exclusive_scan_kernel<<<1, 1024, 0, st>>>(count, rrowptr, nullptr, n);
  if (E > 0) {
    scatter_by_key_kernel<<<(unsigned)((E + 255) / 256), 256, 0, st>>>(E, src, rrowptr, cursor, eid);
    sort_rows_kernel<<<(unsigned)(((long long)n * 32 + 255) / 256), 256, 0, st>>>(n, rrowptr, eid);
  }
  return check_launch("csr_transpose", E > 0 ? 4 : 1);
}

// ---- spatial decomposition: ownership / ghost classification of every atom of the cell (parallel.py) -------------------
// pair_e3gnn_parallel.cpp:277-314 takes LAMMPS' brick decomposition as given; a stand-alone caller has to make it.  One
// thread per atom: fractional coordinates, owner brick, distance to this rank's brick (periodic images included), and the
// cell-list bin of the atom.  key = class * n_bins + bin with class 0 = owned by `rank`, 1 + owner = ghost of `rank` owned
// by `owner`, n_ranks + 1 = not needed: ONE stable sort of the keys gives the rank's atoms in cell-list order followed by
// its ghosts grouped by owner, and a histogram of the classes gives every count.  Fully periodic cells only.
struct PartitionArgs {
  double inv[9];      // inverse cell (frac = pos @ inv)
  int grid[3];        // bricks per axis
  int mine[3];        // this rank's brick
  double margin[3];   // cutoff in units of the cell height per axis
  int nb[3];          // cell-list bins per axis
};
__global__ void partition_keys_kernel(long long n, const double* __restrict__ pos, PartitionArgs a, long long* __restrict__ key) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double p0 = pos[3 * i], p1 = pos[3 * i + 1], p2 = pos[3 * i + 2];
  int b[3], c[3];
  bool near = true, local = true;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    double f = p0 * a.inv[k] + p1 * a.inv[3 + k] + p2 * a.inv[6 + k];
    f -= floor(f);
    b[k] = min(max((int)floor(f * a.grid[k]), 0), a.grid[k] - 1);
    c[k] = min(max((int)(f * a.nb[k]), 0), a.nb[k] - 1);
    local = local && b[k] == a.mine[k];
    const double lo = (double)a.mine[k] / a.grid[k], hi = (double)(a.mine[k] + 1) / a.grid[k];
    double d = fmax(fmax(lo - f, f - hi), 0.0);
    d = fmin(d, fmax(fmax(lo - (f - 1.0), (f - 1.0) - hi), 0.0));
    d = fmin(d, fmax(fmax(lo - (f + 1.0), (f + 1.0) - hi), 0.0));
    near = near && d <= a.margin[k];
  }
  const int owner = (b[0] * a.grid[1] + b[1]) * a.grid[2] + b[2];
  const int world = a.grid[0] * a.grid[1] * a.grid[2];
  const long long cls = local ? 0 : (near ? 1 + owner : world + 1);
  const long long bins = (long long)a.nb[0] * a.nb[1] * a.nb[2];
  key[i] = cls * bins + ((long long)c[0] * a.nb[1] + c[1]) * a.nb[2] + c[2];
}

int partition_keys(long long n, const double* pos, const double* inv_cell_host, const int* grid, const int* mine,
                   const double* margin, const int* nbins, long long* key, cudaStream_t st) {
  if (n == 0) return B200_OK;
  PartitionArgs a;
  for (int k = 0; k < 9; ++k) a.inv[k] = inv_cell_host[k];
  for (int k = 0; k < 3; ++k) { a.grid[k] = grid[k]; a.mine[k] = mine[k]; a.margin[k] = margin[k]; a.nb[k] = nbins[k]; }
  partition_keys_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, pos, a, key);
  return check_launch("partition_keys");
}

// ---- allocation-free variant of csr_transpose -----------------------------------------------------------
// The two per-node counter arrays live in a file-scope device array instead of a caller workspace, so the call makes
// no cudaMalloc / cudaFree (CUDA-graph capture, serving loops, harnesses that forbid allocation).  Every GPU that
// loads the library holds the array (2 x 1 MiB + padding) until exit.  All calls on a device share it: calls that
// overlap on the GPU (two streams, two threads) race.  It keeps the previous call's contents, so both counter arrays
// are cleared (memset on the caller's stream, not an allocation) at the start of every call.
constexpr int kStaticTransposeMaxNodes = 262143;   // (n + 1) counters per array <= 2^18
constexpr size_t kStaticTransposeBytes = 2 * (((size_t)kStaticTransposeMaxNodes + 1) * sizeof(int)) + 256;
static __device__ __align__(256) unsigned char g_transpose_scratch[kStaticTransposeBytes];

// In host code `g_transpose_scratch` is only a stand-in for the symbol: the device address is looked up once per device
// (with that device current) and cached.
static int transpose_scratch_ptr(void** out) {
  static void* ptr[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    set_error("csr_transpose_static: cudaGetDevice failed");
    return B200_ERR_CUDA;
  }
  if (!ptr[dev]) {
    void* p = nullptr;
    const cudaError_t rc = cudaGetSymbolAddress(&p, g_transpose_scratch);
    if (rc != cudaSuccess || !p) {
      set_error("csr_transpose_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(rc));
      return B200_ERR_CUDA;
    }
    ptr[dev] = p;
  }
  *out = ptr[dev];
  return B200_OK;
}

}  // namespace b200

extern "C" int b200_csr_transpose_static_max_nodes(void) { return b200::kStaticTransposeMaxNodes; }

extern "C" int b200_csr_transpose_static(int n, int64_t E, const int32_t* src, int32_t* rrowptr, int32_t* eid,
                                         void* stream) {
  using namespace b200;
  B200_REQUIRE(n >= 0 && E >= 0 && rrowptr && (E == 0 || (src && eid)), "csr_transpose_static: null or negative argument");
  B200_REQUIRE(n <= kStaticTransposeMaxNodes,
               "csr_transpose_static: %d nodes exceed the static scratch limit of %d; use b200_csr_transpose with a "
               "caller workspace", n, kStaticTransposeMaxNodes);
  B200_REQUIRE(E < (1LL << 31), "csr_transpose_static: more than 2^31 edges");
  B200_REQUIRE(transpose_workspace_bytes(n, E) <= kStaticTransposeBytes, "csr_transpose_static: scratch too small");
  void* ws = nullptr;
  const int rc = transpose_scratch_ptr(&ws);
  if (rc != B200_OK) return rc;
  return csr_transpose(n, E, src, ws, rrowptr, eid, (cudaStream_t)stream);
}

extern "C" int b200_partition_keys(int64_t n_atoms, const double* pos, const double* inv_cell_host, const int* grid,
                                   const int* my_brick, const double* margin, const int* cell_bins, int64_t* key, void* stream) {
  B200_REQUIRE(n_atoms >= 0 && inv_cell_host && grid && my_brick && margin && cell_bins && (n_atoms == 0 || (pos && key)),
               "partition_keys: null argument");
  for (int k = 0; k < 3; ++k)
    B200_REQUIRE(grid[k] >= 1 && my_brick[k] >= 0 && my_brick[k] < grid[k] && cell_bins[k] >= 1, "partition_keys: bad brick grid");
  return b200::partition_keys(n_atoms, pos, inv_cell_host, grid, my_brick, margin, cell_bins, (long long*)key,
                              (cudaStream_t)stream);
}
