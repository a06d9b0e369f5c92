// extern "C" surface of libsevennet_b200.so (declared in include/sevennet_b200.h).
#include <stdarg.h>

#include <atomic>
#include <string.h>

#include "ops.h"

namespace b200 {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
static std::atomic<long long> g_launches{0};
int check_launch(const char* what, int n_launched) {
  g_launches += n_launched;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}
}  // namespace b200

using namespace b200;
#define ST(s) ((cudaStream_t)(s))

extern "C" {

const char* b200_last_error(void) { return g_err; }
int b200_version(void) { return 1; }
long long b200_launch_count(void) { return g_launches.load(); }

int b200_edge_embed_fwd(const EdgeParams* p, int64_t E, const float* vec, const float* coeffs, float* len, float* Y,
                        float* emb, void* stream) {
  B200_REQUIRE(p && (E == 0 || (vec && coeffs && Y && emb)), "edge_embed_fwd: null argument");
  return edge_embed_fwd(*p, E, vec, coeffs, len, Y, emb, ST(stream));
}
int b200_edge_embed_bwd(const EdgeParams* p, int64_t E, const float* vec, const float* coeffs, const float* dY,
                        const float* demb, float* dvec, void* stream) {
  B200_REQUIRE(p && (E == 0 || (vec && coeffs && dvec)), "edge_embed_bwd: null argument");
  return edge_embed_bwd(*p, E, vec, coeffs, dY, demb, dvec, ST(stream));
}
int b200_edge_embed_bwd2(const EdgeParams* p, int64_t E, const float* vec, const float* coeffs, const float* gY,
                         const float* gemb, const float* a, float* dvec2, float* dgY, float* dgemb, float* dk, void* stream) {
  B200_REQUIRE(p && (E == 0 || (vec && coeffs && a && dvec2)), "edge_embed_bwd2: null argument");
  return edge_embed_bwd2(*p, E, vec, coeffs, gY, gemb, a, dvec2, dgY, dgemb, dk, ST(stream));
}
int b200_edge_embed_dk(const EdgeParams* p, int64_t E, const float* vec, const float* coeffs, const float* gemb, float* dk,
                       void* stream) {
  B200_REQUIRE(p && dk && (E == 0 || (vec && coeffs && gemb)), "edge_embed_dk: null argument");
  return edge_embed_dk(*p, E, vec, coeffs, gemb, dk, ST(stream));
}
size_t b200_radial_table_floats(const EdgeParams* p, const MlpDesc* d) { return (p && d) ? radial_table_floats(*p, *d) : 0; }
int b200_radial_table_build(const EdgeParams* p, const MlpDesc* d, const float* coeffs, const float* W0, const float* W1,
                            const float* W2, float* table, void* stream) {
  B200_REQUIRE(p && d && coeffs && W0 && W1 && W2 && table, "radial_table_build: null argument");
  return radial_table_build(*p, *d, coeffs, W0, W1, W2, table, ST(stream));
}
int b200_radial_table_fwd(const EdgeParams* p, const MlpDesc* d, int64_t E, const float* len, const int32_t* order,
                          const float* table, float* w, void* stream) {
  B200_REQUIRE(p && d && (E == 0 || (len && table && w)), "radial_table_fwd: null argument");
  return radial_table_fwd(*p, *d, E, len, order, table, w, ST(stream));
}
int b200_radial_table_bwd(const EdgeParams* p, const MlpDesc* d, int64_t E, const float* len, const int32_t* order,
                          const float* table, const float* dw, float* dr, int accumulate, void* stream) {
  B200_REQUIRE(p && d && (E == 0 || (len && table && dw && dr)), "radial_table_bwd: null argument");
  return radial_table_bwd(*p, *d, E, len, order, table, dw, dr, accumulate, ST(stream));
}
int b200_edge_embed_bwd_r(const EdgeParams* p, int64_t E, const float* vec, const float* dY, const float* dr, float* dvec,
                          void* stream) {
  B200_REQUIRE(p && (E == 0 || (vec && dvec)), "edge_embed_bwd_r: null argument");
  return edge_embed_bwd_r(*p, E, vec, dY, dr, dvec, ST(stream));
}
int b200_radial_mlp_fwd(const MlpDesc* d, int64_t E, const float* emb, const float* W0, const float* W1, const float* W2,
                        float* w, void* stream) {
  B200_REQUIRE(d && W0 && W1 && W2 && (E == 0 || (emb && w)), "radial_mlp_fwd: null argument");
  return radial_mlp_fwd(*d, E, emb, W0, W1, W2, w, ST(stream));
}
int b200_radial_mlp_bwd(const MlpDesc* d, int64_t E, const float* emb, const float* W0, const float* W1, const float* W2,
                        const float* dw, float* demb, int accumulate, void* stream) {
  B200_REQUIRE(d && W0 && W1 && W2 && (E == 0 || (emb && dw && demb)), "radial_mlp_bwd: null argument");
  return radial_mlp_bwd(*d, E, emb, W0, W1, W2, dw, demb, accumulate, ST(stream));
}
int b200_conv_dims(int m0, int m1, int m2, int lout, int* wn, int* odim) {
  B200_REQUIRE(lout == 0 || lout == 2, "conv_dims: lout must be 0 or 2");
  ConvDesc d;
  fill_conv_desc(m0, m1, m2, lout, d);
  if (wn) *wn = d.wn;
  if (odim) *odim = d.odim;
  return B200_OK;
}
int b200_conv_fwd(int m0, int m1, int m2, int lout, int n_dst, const int32_t* rowptr, const int32_t* src, const float* x,
                  const float* Y, const float* w, float denom, float* out, void* stream) {
  B200_REQUIRE(lout == 0 || lout == 2, "conv_fwd: lout must be 0 or 2");
  B200_REQUIRE(rowptr && out, "conv_fwd: null argument");
  ConvDesc d;
  fill_conv_desc(m0, m1, m2, lout, d);
  return conv_fwd(d, n_dst, rowptr, src, x, Y, w, denom, out, ST(stream));
}
int b200_conv_bwd_edge(int m0, int m1, int m2, int lout, int n_dst, int64_t E, const int32_t* rowptr, const int32_t* src,
                       const float* x, const float* Y, const float* w, const float* gout, float denom, float* dw,
                       float* dY, int accumulate, void* stream) {
  B200_REQUIRE(lout == 0 || lout == 2, "conv_bwd_edge: lout must be 0 or 2");
  ConvDesc d;
  fill_conv_desc(m0, m1, m2, lout, d);
  return conv_bwd_edge(d, n_dst, (int)E, rowptr, src, x, Y, w, gout, denom, dw, dY, accumulate, ST(stream));
}
int b200_conv_bwd(int m0, int m1, int m2, int lout, int n_dst, int n_src, int64_t E, const int32_t* rowptr, const int32_t* src,
                  const float* x, const float* Y, const float* w, const float* gout, float denom, float* dw, float* dY,
                  int accumulate, float* dx, void* stream) {
  B200_REQUIRE(lout == 0 || lout == 2, "conv_bwd: lout must be 0 or 2");
  B200_REQUIRE(dx != nullptr, "conv_bwd: d_x is NULL (use b200_conv_bwd_edge)");
  ConvDesc d;
  fill_conv_desc(m0, m1, m2, lout, d);
  return conv_bwd_edge_x(d, n_dst, n_src, (int)E, rowptr, src, x, Y, w, gout, denom, dw, dY, accumulate, dx, ST(stream));
}
int b200_conv_fwd_scaled(int m0, int m1, int m2, int lout, int n_dst, const int32_t* rowptr, const int32_t* src, const float* x,
                         const float* Y, const float* w, float denom, float* out, const float* path_scale, void* stream) {
  B200_REQUIRE(lout == 0 || lout == 2, "conv_fwd_scaled: lout must be 0 or 2");
  B200_REQUIRE(rowptr && out, "conv_fwd_scaled: null argument");
  ConvDesc d;
  fill_conv_desc(m0, m1, m2, lout, d);
  return conv_fwd_scaled(d, n_dst, rowptr, src, x, Y, w, denom, out, path_scale, ST(stream));
}
int b200_conv_bwd_scaled(int m0, int m1, int m2, int lout, int n_dst, int n_src, int64_t E, const int32_t* rowptr,
                         const int32_t* src, const float* x, const float* Y, const float* w, const float* gout, float denom,
                         float* dw, float* dY, int accumulate, float* dx, const float* path_scale, int dw_mode, void* stream) {
  B200_REQUIRE(lout == 0 || lout == 2, "conv_bwd_scaled: lout must be 0 or 2");
  B200_REQUIRE(dY != nullptr, "conv_bwd_scaled: d_edge_attr is NULL");   // d_x NULL: not computed
  ConvDesc d;
  fill_conv_desc(m0, m1, m2, lout, d);
  return conv_bwd_scaled(d, n_dst, n_src, (int)E, rowptr, src, x, Y, w, gout, denom, dw, dY, accumulate, dx, path_scale, dw_mode,
                         ST(stream));
}
int b200_conv_bwd_x(int m0, int m1, int m2, int lout, int n_src, const int32_t* rrowptr, const int32_t* eid,
                    const int32_t* dst, const float* gout, const float* Y, const float* w, float denom, float* dx,
                    void* stream) {
  B200_REQUIRE(lout == 0 || lout == 2, "conv_bwd_x: lout must be 0 or 2");
  ConvDesc d;
  fill_conv_desc(m0, m1, m2, lout, d);
  return conv_bwd_x(d, n_src, rrowptr, eid, dst, gout, Y, w, denom, dx, ST(stream));
}
int b200_irreps_linear(const LinDesc* d, int n, const float* x, const float* W, float* y, int transpose, int accumulate,
                       void* stream) {
  B200_REQUIRE(d && W && (n == 0 || (x && y)), "irreps_linear: null argument");
  return irreps_linear(*d, n, x, W, y, transpose, accumulate, ST(stream));
}
int b200_irreps_linear_wgrad(const LinDesc* d, int n, const float* x, const float* dy, float* dW, int accumulate,
                             void* stream) {
  B200_REQUIRE(d && dW, "irreps_linear_wgrad: null argument");
  return irreps_linear_wgrad(*d, n, x, dy, dW, accumulate, ST(stream));
}
int b200_gate_fwd(int ns, int m1, int m2, int n, const float* in, float* out, void* stream) {
  return gate_fwd(ns, m1, m2, n, in, out, ST(stream));
}
int b200_gate_bwd(int ns, int m1, int m2, int n, const float* in, const float* dout, float* din, void* stream) {
  return gate_bwd(ns, m1, m2, n, in, dout, din, ST(stream));
}
int b200_table_interp(int m, int wn, int64_t E, float inv_h, int n_grid, const float* P, const float* r, const int32_t* order,
                      const float* coef, float* out, void* stream) {
  B200_REQUIRE(E == 0 || (P && r && out), "table_interp: null argument");
  return table_interp(m, wn, E, inv_h, n_grid, P, r, order, coef, out, ST(stream));
}
int b200_table_dot(int m, int wn, int64_t E, float inv_h, int n_grid, const float* P, const float* r, const int32_t* order,
                   const float* g, float* out, void* stream) {
  B200_REQUIRE(E == 0 || (P && r && g && out), "table_dot: null argument");
  return table_dot(m, wn, E, inv_h, n_grid, P, r, order, g, out, ST(stream));
}
int b200_table_scatter(int m, int wn, int64_t E, float inv_h, int n_grid, const float* g, const float* coef, const float* r,
                       const int32_t* order, float* G, void* stream) {
  B200_REQUIRE(G && (E == 0 || (g && r)), "table_scatter: null argument");
  return table_scatter(m, wn, E, inv_h, n_grid, g, coef, r, order, G, ST(stream));
}
int b200_table_scatter_dot(int m, int md, int wn, int64_t E, float inv_h, int n_grid, const float* P, const float* g,
                           const float* coef, const float* r, const int32_t* order, float* G, float* dot, void* stream) {
  B200_REQUIRE(G && (E == 0 || (P && g && r && dot)), "table_scatter_dot: null argument");
  return table_scatter_dot(m, md, wn, E, inv_h, n_grid, P, g, coef, r, order, G, dot, ST(stream));
}
int b200_table_hermite(int n_grid, int wn, double h, const double* Z, const double* dZ, float* P, void* stream) {
  B200_REQUIRE(Z && dZ && P, "table_hermite: null argument");
  return table_hermite(n_grid, wn, h, Z, dZ, P, ST(stream));
}
int b200_table_hermite_adj(int n_grid, int wn, double h, const float* G, double* gZ, double* gdZ, void* stream) {
  B200_REQUIRE(G && gZ && gdZ, "table_hermite_adj: null argument");
  return table_hermite_adj(n_grid, wn, h, G, gZ, gdZ, ST(stream));
}
int b200_gate_bwd2(int ns, int m1, int m2, int n, const float* in, const float* dout, const float* a, float* dx, float* dg,
                   void* stream) {
  B200_REQUIRE(n == 0 || (in && dout && a && dx && dg), "gate_bwd2: null argument");
  return gate_bwd2(ns, m1, m2, n, in, dout, a, dx, dg, ST(stream));
}
int b200_embed_fwd(int n, int nsp, int dim, const int32_t* types, const float* W, float* x, void* stream) {
  return embed_fwd(n, nsp, dim, types, W, x, ST(stream));
}
int b200_rescale_fwd(int n, const int32_t* types, const float* scaled, const float* scale, const float* shift,
                     int species_wise, float* out, void* stream) {
  return rescale_fwd(n, types, scaled, scale, shift, species_wise, out, ST(stream));
}
int b200_halo_pack(int64_t n_send, int width, const float* rows, const int32_t* idx, float* out, void* stream) {
  B200_REQUIRE(n_send >= 0 && width >= 0 && (n_send == 0 || (rows && idx && out)), "halo_pack: null argument");
  return halo_pack(n_send, width, rows, idx, out, ST(stream));
}
int b200_halo_unpack_add(int64_t n_recv, int width, float* rows, const int32_t* idx, const float* in, void* stream) {
  B200_REQUIRE(n_recv >= 0 && width >= 0 && (n_recv == 0 || (rows && idx && in)), "halo_unpack_add: null argument");
  return halo_unpack_add(n_recv, width, rows, idx, in, ST(stream));
}
int b200_colsum(int n_rows, int ncol, const float* in, float* out, void* stream) {
  return colsum(n_rows, ncol, in, out, ST(stream));
}
int b200_force_virial(int n, const int32_t* rowptr, const int32_t* rrowptr, const int32_t* eid, const float* vec,
                      const float* dvec, float* force, float* virial_atom, void* stream) {
  return force_virial(n, rowptr, rrowptr, eid, vec, dvec, force, virial_atom, ST(stream));
}
size_t b200_graph_workspace_bytes(int n) { return graph_workspace_bytes(n); }
int b200_graph_count(int n, const double* pos, const double* cell_host, const int* pbc_host, double cutoff, void* ws,
                     int32_t* rowptr, int64_t* n_edges_host, void* stream) {
  B200_REQUIRE(cell_host && pbc_host && ws && rowptr && n_edges_host, "graph_count: null argument");
  long long e = 0;
  int rc = graph_count(n, pos, cell_host, pbc_host, cutoff, ws, rowptr, &e, ST(stream));
  *n_edges_host = e;
  return rc;
}
int b200_graph_fill(int n, const double* pos, const double* cell_host, const int* pbc_host, double cutoff, void* ws,
                    const int32_t* rowptr, int32_t* src, int32_t* dst, int32_t* shift, float* edge_vec,
                    uint64_t* tmp_keys, void* stream) {
  B200_REQUIRE(cell_host && pbc_host && ws && rowptr, "graph_fill: null argument");
  return graph_fill(n, pos, cell_host, pbc_host, cutoff, ws, rowptr, src, dst, shift, edge_vec,
                    (unsigned long long*)tmp_keys, ST(stream));
}
size_t b200_transpose_workspace_bytes(int n, int64_t E) { return transpose_workspace_bytes(n, E); }
int b200_csr_transpose(int n, int64_t E, const int32_t* src, void* ws, int32_t* rrowptr, int32_t* eid, void* stream) {
  B200_REQUIRE(ws && rrowptr, "csr_transpose: null argument");
  return csr_transpose(n, E, src, ws, rrowptr, eid, ST(stream));
}

}  // extern "C"
