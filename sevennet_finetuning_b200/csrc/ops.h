// Internal C++ launchers shared by the C-ABI (capi.cu) and the whole-model engine (engine.cu).
#pragma once
#include "common.cuh"

namespace b200 {

void fill_conv_desc(int m0, int m1, int m2, int lout, ConvDesc& d);

int edge_embed_fwd(const EdgeParams& p, long long E, const float* vec, const float* coeffs, float* len, float* Y,
                   float* emb, cudaStream_t st);
int edge_embed_bwd(const EdgeParams& p, long long E, const float* vec, const float* coeffs, const float* dY,
                   const float* demb, float* dvec, cudaStream_t st);
int edge_embed_bwd2(const EdgeParams& p, long long E, const float* vec, const float* coeffs, const float* gY,
                    const float* gemb, const float* a, float* dvec2, float* dgY, float* dgemb, float* dk, cudaStream_t st);
int edge_embed_dk(const EdgeParams& p, long long E, const float* vec, const float* coeffs, const float* gemb, float* dk,
                  cudaStream_t st);
int edge_embed_bwd_full(const EdgeParams& p, long long E, const float* vec, const float* coeffs, const float* dY,
                        const float* demb, const float* dr, float* dvec, cudaStream_t st);
int edge_embed_bwd_r(const EdgeParams& p, long long E, const float* vec, const float* dY, const float* dr, float* dvec,
                     cudaStream_t st);
int force_virial(int n_atoms, const int* rowptr, const int* rrowptr, const int* eid_by_src, const float* vec,
                 const float* dvec, float* force, float* virial_atom, cudaStream_t st);

int radial_mlp_fwd(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                   float* w, cudaStream_t st);
int radial_mlp_bwd(const MlpDesc& d, long long E, const float* emb, const float* W0, const float* W1, const float* W2,
                   const float* dw, float* demb, int accumulate, cudaStream_t st);

int conv_fwd(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
             const float* w, float denom, float* out, cudaStream_t st);
// second-order pass of the training path: per-path scale of the weights (in) and of the weight gradient (out), NULL = 1;
// dw_mode 0: d weight overwritten, 1: added into, 2: not computed (dw may be nullptr)
int conv_fwd_scaled(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
                    const float* w, float denom, float* out, const float* path_scale, cudaStream_t st);
int conv_bwd_scaled(const ConvDesc& d, int n_dst, int n_src, int n_edges, const int* rowptr, const int* src, const float* x,
                    const float* Y, const float* w, const float* gout, float denom, float* dw, float* dY, int accumulate,
                    float* dx, const float* path_scale, int dw_mode, cudaStream_t st);
int conv_bwd_x(const ConvDesc& d, int n_src, const int* rrowptr, const int* eid, const int* dst, const float* gout,
               const float* Y, const float* w, float denom, float* dx, cudaStream_t st);
int conv_bwd_edge_x(const ConvDesc& d, int n_dst, int n_src, int n_edges, const int* rowptr, const int* src, const float* x,
                    const float* Y, const float* w, const float* gout, float denom, float* dw, float* dY, int accumulate,
                    float* dx, cudaStream_t st);
// per-edge weights interpolated from the tabulated radial function inside the kernels (no [E, wn] tensors)
int conv_fwd_tab(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
                 const float* len, const float* table, float inv_h, int n_grid, float denom, float* out, cudaStream_t st);
int conv_bwd_tab(const ConvDesc& d, int n_dst, int n_src, int n_edges, const int* rowptr, const int* src, const float* x,
                 const float* Y, const float* len, const float* table, float inv_h, int n_grid, const float* gout, float denom,
                 float* dY, int accumulate_dY, float* dr, int accumulate_dr, float* dx, cudaStream_t st);
int conv_bwd_edge(const ConvDesc& d, int n_dst, int n_edges, const int* rowptr, const int* src, const float* x,
                  const float* Y, const float* w, const float* gout, float denom, float* dw, float* dY, int accumulate,
                  cudaStream_t st);

// tabulated radial weights (radial_table.cu): the MD path's replacement of the per-edge MLP evaluation
struct RadialTableGrid {
  double h;   // grid spacing
  int n;      // number of intervals; nodes 0..n
};
RadialTableGrid radial_table_grid(const EdgeParams& p);
size_t radial_table_floats(const EdgeParams& p, const MlpDesc& d);
bool radial_table_supported(const MlpDesc& d);
int radial_table_build(const EdgeParams& p, const MlpDesc& d, const float* coeffs, const float* W0, const float* W1,
                       const float* W2, float* T, cudaStream_t st);
// packed table (12 bytes per weight) read by the channel-pair convolution kernels (conv_pair.cu)
constexpr float kTabCDScale = 4096.0f;
size_t radial_table_packed_bytes(const EdgeParams& p, const MlpDesc& d);
int radial_table_build_packed(const EdgeParams& p, const MlpDesc& d, const float* coeffs, const float* W0, const float* W1,
                              const float* W2, void* T, cudaStream_t st);
bool conv_pair_supported(const ConvDesc& d);
int conv_pair_interleave(const ConvDesc& d, long long n_rows, float* x, int inverse, cudaStream_t st);
int conv_pair_fwd(const ConvDesc& d, int n_dst, const int* rowptr, const int* src, const float* x, const float* Y,
                  const float* len, const void* packed_table, float inv_h, int n_grid, float denom, float* out, cudaStream_t st);
// d Y and d E / d r are ADDED into dY / dr (the caller zeroes them once per reverse pass); d x is overwritten
int conv_pair_bwd(const ConvDesc& d, int n_dst, int n_src, const int* rowptr, const int* src, const float* x, const float* Y,
                  const float* len, const void* packed_table, float inv_h, int n_grid, const float* gout, float denom,
                  float* dY, float* dr, float* dx, cudaStream_t st);
size_t radial_table_order_bytes(long long E);
int radial_table_order(const EdgeParams& p, long long E, const float* len, void* ws, int* order, cudaStream_t st);   // edge ids sorted by interval
// order: edge ids sorted by interval (nullptr = natural order)
int radial_table_fwd(const EdgeParams& p, const MlpDesc& d, long long E, const float* len, const int* order, const float* T,
                     float* w, cudaStream_t st);
int radial_table_bwd(const EdgeParams& p, const MlpDesc& d, long long E, const float* len, const int* order, const float* T,
                     const float* dw, float* dr, int accumulate, cudaStream_t st);


int irreps_linear(const LinDesc& d, int n_nodes, const float* x, const float* W, float* y, int transpose, int accumulate,
                  cudaStream_t st, float** pack_cache = nullptr);
// layout bit 0 / bit 1: the l-blocks of the input / output rows are component-major ([2l+1][mul] instead of [mul][2l+1]),
// the row layout of the channel-pair convolution kernels (conv_pair.cu)
constexpr int kLinXcm = 1, kLinYcm = 2;
bool irreps_linear_tc_supported(const LinDesc& d, int transpose, int layout = 0);
int irreps_linear_tc(const LinDesc& d, int n_nodes, const float* x, const float* W, float* y, int transpose, int accumulate,
                     cudaStream_t st, int layout = 0, float** pack_cache = nullptr);
int irreps_linear_wgrad(const LinDesc& d, int n_nodes, const float* x, const float* dy, float* dW, int accumulate,
                        cudaStream_t st);

// train-mode radial weights from per-interval cubic coefficients P[n_grid, 4, wn] (table_train.cu)
int table_interp(int m, int wn, long long E, float inv_h, int n_grid, const float* P, const float* r, const int* order,
                 const float* coef, float* out, cudaStream_t st);
int table_dot(int m, int wn, long long E, float inv_h, int n_grid, const float* P, const float* r, const int* order,
              const float* g, float* out, cudaStream_t st);
int table_scatter(int m, int wn, long long E, float inv_h, int n_grid, const float* g, const float* coef, const float* r,
                  const int* order, float* G, cudaStream_t st);

// table_scatter(m) and table_dot(md) of the same g in one pass over it
int table_scatter_dot(int m, int md, int wn, long long E, float inv_h, int n_grid, const float* P, const float* g,
                      const float* coef, const float* r, const int* order, float* G, float* dot, cudaStream_t st);
int table_hermite(int n_grid, int wn, double h, const double* Z, const double* dZ, float* P, cudaStream_t st);
int table_hermite_adj(int n_grid, int wn, double h, const float* G, double* gZ, double* gdZ, cudaStream_t st);

int gate_fwd(int ns, int m1, int m2, int n_nodes, const float* in, float* out, cudaStream_t st);
int gate_bwd(int ns, int m1, int m2, int n_nodes, const float* in, const float* dout, float* din, cudaStream_t st);
int gate_bwd2(int ns, int m1, int m2, int n_nodes, const float* in, const float* dout, const float* a, float* dx,
              float* dg, cudaStream_t st);
int embed_fwd(int n_nodes, int n_species, int dim, const int* types, const float* W, float* x, cudaStream_t st);
int rescale_fwd(int n_nodes, const int* types, const float* scaled, const float* scale, const float* shift,
                int species_wise, float* out, cudaStream_t st);
int colsum(int n_rows, int ncol, const float* in, float* out, cudaStream_t st);
int halo_pack(long long m, int width, const float* rows, const int* idx, float* out, cudaStream_t st);
int halo_unpack_add(long long m, int width, float* rows, const int* idx, const float* in, cudaStream_t st);

size_t graph_workspace_bytes(int n);
int graph_count(int n, const double* pos, const double* cell_host, const int* pbc_host, double cutoff, void* ws,
                int* rowptr, long long* n_edges_host, cudaStream_t st);
int graph_fill(int n, const double* pos, const double* cell_host, const int* pbc_host, double cutoff, void* ws,
               const int* rowptr, int* src, int* dst, int* shift, float* edge_vec, unsigned long long* tmp_key,
               cudaStream_t st);
size_t transpose_workspace_bytes(int n, long long E);
int csr_transpose(int n, long long E, const int* src, void* ws, int* rrowptr, int* eid, cudaStream_t st);

}  // namespace b200
