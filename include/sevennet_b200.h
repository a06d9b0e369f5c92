/* sevennet_b200.h - C ABI of the B200-native SevenNet message-passing hot path.
 *
 * Drop-in boundary: the reference has no FFI of its own - its seam is the Python
 * nn.Module contract of sevenn.nn (forward(data: Dict[str, Tensor]) -> Dict[str, Tensor]) and, on the
 * native side, the TorchScript call made by pair_e3gnn.cpp / pair_e3gnn_parallel.cpp.  Every entry
 * point below is what a binding for one of those call sites needs; each cites the reference
 * interface it replaces (paths relative to the reference root).
 *
 * Conventions: plain pointers and sizes only; all array pointers are DEVICE pointers unless the name
 * ends in _host; fp32 features in the e3nn irreps-blocked layout [mul, 2l+1] (m fastest); int32
 * indices; `stream` is a cudaStream_t passed as void*; no allocation happens inside (callers pass
 * workspaces); every function returns B200_OK (0) or an error code, message via b200_last_error().
 * There is no CPU fallback: the library needs an sm_100a device.
 */
#ifndef SEVENNET_B200_H
#define SEVENNET_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_ERR_INVALID 1 /* bad argument / unsupported shape */
#define B200_ERR_CUDA 2    /* CUDA runtime reported an error  */

#define B200_CUTOFF_POLY 0  /* PolynomialCutoff, sevenn/nn/edge_embedding.py:119-145 */
#define B200_CUTOFF_XPLOR 1 /* XPLORCutoff,      sevenn/nn/edge_embedding.py:148-173 */

const char* b200_last_error(void);
int b200_version(void);
/* number of CUDA kernels this library has launched in this process (bench.py's gpu_launches) */
long long b200_launch_count(void);

/* ---- edge embedding: sevenn/nn/edge_embedding.py:202-230 (EdgeEmbedding.forward) ------------------ */
typedef struct EdgeParams {
  int num_basis;      /* BesselBasis.num_basis                       (:80-116) */
  int cutoff_kind;    /* B200_CUTOFF_*                                        */
  int normalize_sph;  /* SphericalEncoding(normalize=...)            (:177-198) */
  float r_cut;        /* cutoff_length                                         */
  float r_on;         /* XPLOR cutoff_on                                       */
  float poly_p;       /* PolynomialCutoff p                                    */
} EdgeParams;

/* edge_vec[E,3] -> edge_length[E] (may be NULL), edge_attr Y[E,9], edge_embedding emb[E,num_basis] */
int b200_edge_embed_fwd(const EdgeParams* p, int64_t n_edges, const float* edge_vec, const float* bessel_coeffs,
                        float* edge_length, float* edge_attr, float* edge_embedding, void* stream);
/* d edge_vec[E,3] = J^T (d edge_attr[E,9] (NULL ok), d edge_embedding[E,nb] (NULL ok)) */
int b200_edge_embed_bwd(const EdgeParams* p, int64_t n_edges, const float* edge_vec, const float* bessel_coeffs,
                        const float* d_edge_attr, const float* d_edge_embedding, float* d_edge_vec, void* stream);
/* d bessel_coeffs[nb] = sum_e d_edge_embedding[e, n] * d emb_n / d k_n (trainable frequencies, edge_embedding.py:95-101) */
int b200_edge_embed_dk(const EdgeParams* p, int64_t n_edges, const float* edge_vec, const float* bessel_coeffs,
                       const float* d_edge_embedding, float* d_coeffs, void* stream);
/* second-order pass: the reverse of b200_edge_embed_bwd (force_output.py:165-170 create_graph=self.training).
 * a[E,3] = dL/d(d_edge_vec).  Outputs: dL/d edge_vec (Hessian-vector product) [E,3]; dL/d(d_edge_attr) [E,9] and
 * dL/d(d_edge_embedding) [E,nb] (NULL ok); dL/d bessel_coeffs [nb] (NULL ok).  d_edge_attr / d_edge_embedding are the
 * inputs the first reverse pass was called with (NULL ok = zero). */
int b200_edge_embed_bwd2(const EdgeParams* p, int64_t n_edges, const float* edge_vec, const float* bessel_coeffs,
                         const float* d_edge_attr, const float* d_edge_embedding, const float* a, float* d_edge_vec,
                         float* d_d_edge_attr, float* d_d_edge_embedding, float* d_coeffs, void* stream);
/* same with dE/d|edge_vec| given directly (d_r[E], from the table-reading convolution kernels) instead of d edge_embedding */
int b200_edge_embed_bwd_r(const EdgeParams* p, int64_t n_edges, const float* edge_vec, const float* d_edge_attr,
                          const float* d_r, float* d_edge_vec, void* stream);

/* ---- radial MLP: sevenn/nn/convolution.py:97-106 (weight_nn = FullyConnectedNet([nb,H1,H2,wn])) ---- */
typedef struct MlpDesc {
  int nb, h1, h2, wn;
} MlpDesc;
/* weights in the reference's parameter layout: layer{k}.weight is [in, out] row-major, un-normalised */
int b200_radial_mlp_fwd(const MlpDesc* d, int64_t n_edges, const float* emb, const float* W0, const float* W1,
                        const float* W2, float* weight_out, void* stream);
int b200_radial_mlp_bwd(const MlpDesc* d, int64_t n_edges, const float* emb, const float* W0, const float* W1,
                        const float* W2, const float* d_weight, float* d_emb, int accumulate, void* stream);

/* ---- tabulated radial weights (MD path; same reference lines as the radial MLP) ----------------------------------
 * w_k(r) = weight_nn(BesselBasis(r) * cutoff(r))_k depends on the edge length only.  For fixed parameters it is evaluated
 * once in float64 on a uniform grid (spacing ~ r_cut / 4000, the XPLOR switching radius on a node) together with its exact
 * derivative and stored per interval as the coefficients of the cubic Hermite interpolant, table[n_intervals][4][wn]
 * (the slope is then as accurate as the value: no fp32 differences of neighbouring nodes).  Every edge evaluates its cubic
 * (b200_radial_table_fwd) and the reverse pass contracts d w with the derivative of the same spline
 * (b200_radial_table_bwd: d_r[E] (+)= sum_k d_weight[e,k] * d w_k/d r).  b200_model_compute uses this unless
 * B200_RADIAL_TABLE=0; training passes use the explicit MLP above. */
size_t b200_radial_table_floats(const EdgeParams* p, const MlpDesc* d);
int b200_radial_table_build(const EdgeParams* p, const MlpDesc* d, const float* bessel_coeffs, const float* W0,
                            const float* W1, const float* W2, float* table, void* stream);
/* order: edge ids sorted by grid interval (NULL = natural order); sorted visiting keeps the table rows in L1 */
int b200_radial_table_fwd(const EdgeParams* p, const MlpDesc* d, int64_t n_edges, const float* edge_length,
                          const int32_t* order, const float* table, float* weight, void* stream);
int b200_radial_table_bwd(const EdgeParams* p, const MlpDesc* d, int64_t n_edges, const float* edge_length,
                          const int32_t* order, const float* table, const float* d_weight, float* d_r, int accumulate,
                          void* stream);

/* ---- convolution: sevenn/nn/convolution.py:105-121 (TensorProduct 'uvu' + message_gather + /denominator) ----
 * x irreps m0x0e+m1x1e+m2x2e, filter 0e+1e+2e, output keeps l3 <= lout (0 or 2); instruction and slot
 * order as built by IrrepsConvolution.__init__ (:72-95).  Edges are given as a CSR sorted by
 * edge_index[0] (the aggregation destination, :112-113): rowptr[n_dst+1], src[E] = edge_index[1]. */
int b200_conv_dims(int m0, int m1, int m2, int lout, int* weight_numel, int* out_dim);
int b200_conv_fwd(int m0, int m1, int m2, int lout, int n_dst, const int32_t* rowptr, const int32_t* src,
                  const float* x, const float* edge_attr, const float* weight, float denominator, float* out,
                  void* stream);
/* d weight[E,wn] and d edge_attr[E,9] (component 0 untouched) from d out[n_dst,out_dim] */
int b200_conv_bwd_edge(int m0, int m1, int m2, int lout, int n_dst, int64_t n_edges, const int32_t* rowptr,
                       const int32_t* src, const float* x, const float* edge_attr, const float* weight,
                       const float* d_out, float denominator, float* d_weight, float* d_edge_attr, int accumulate,
                       void* stream);
/* the same pass with d x[n_src,x_dim] added (zeroed inside, accumulated with RED.ADD per edge: the sums are
 * order-nondeterministic like the reference's index_add scatter); replaces b200_conv_bwd_edge + b200_conv_bwd_x */
int b200_conv_bwd(int m0, int m1, int m2, int lout, int n_dst, int n_src, int64_t n_edges, const int32_t* rowptr,
                  const int32_t* src, const float* x, const float* edge_attr, const float* weight, const float* d_out,
                  float denominator, float* d_weight, float* d_edge_attr, int accumulate, float* d_x, void* stream);
/* Variants for the second-order pass of force-matching training (force_output.py:165-170 create_graph=self.training, then
 * trainer.py:173-186 loss.backward()): path_scale[15] (HOST array, canonical path order 000 011 022 101 110 111 112 121 122 202
 * 211 212 220 221 222; NULL = all 1) multiplies the weights of a path on the way in and its weight gradient on the way out -
 * the variation of edge_attr acts on the l2 > 0 paths only (scale 0 on 000, 101, 202), which would otherwise be a masked copy
 * of weight[E,wn] and of its gradient.  dw_mode 0: d_weight overwritten, 1: added into, 2: not computed (d_weight may be NULL).
 * d_x NULL: not computed. */
int b200_conv_fwd_scaled(int m0, int m1, int m2, int lout, int n_dst, const int32_t* rowptr, const int32_t* src,
                         const float* x, const float* edge_attr, const float* weight, float denominator, float* out,
                         const float* path_scale, void* stream);
int b200_conv_bwd_scaled(int m0, int m1, int m2, int lout, int n_dst, int n_src, int64_t n_edges, const int32_t* rowptr,
                         const int32_t* src, const float* x, const float* edge_attr, const float* weight, const float* d_out,
                         float denominator, float* d_weight, float* d_edge_attr, int accumulate, float* d_x,
                         const float* path_scale, int dw_mode, void* stream);
/* d x[n_src,x_dim]; rrowptr[n_src+1] / eid_by_src[E]: CSR of the same edges sorted by source, dst[E] = edge_index[0] */
int b200_conv_bwd_x(int m0, int m1, int m2, int lout, int n_src, const int32_t* rrowptr, const int32_t* eid_by_src,
                    const int32_t* dst, const float* d_out, const float* edge_attr, const float* weight,
                    float denominator, float* d_x, void* stream);


/* ---- per-node irreps linear: sevenn/nn/linear.py:46-49 (IrrepsLinear -> e3nn o3.Linear), also
 *      sevenn/nn/self_connection.py:60-62 (SelfConnectionLinearIntro) ------------------------------- */
typedef struct LinBlock {
  int d;        /* 2l+1 */
  int mul_in, mul_out;
  int in_off, out_off; /* offsets of the block in the input / output rows */
  int w_off;           /* offset of W[mul_in, mul_out] in the flat weight vector */
  float scale;         /* 1/sqrt(fan_in) */
} LinBlock;
typedef struct LinDesc {
  int nblk;
  LinBlock blk[4];
  int in_dim, out_dim;
} LinDesc;
/* y[n, out block] (+)= scale * x[n, in block] . W ;  transpose=1: x-gradient (roles of in/out swapped, W^T) */
int b200_irreps_linear(const LinDesc* d, int n_nodes, const float* x, const float* weight, float* y, int transpose,
                       int accumulate, void* stream);
/* dW (+)= scale * x^T dy  (flat, reference layout) */
int b200_irreps_linear_wgrad(const LinDesc* d, int n_nodes, const float* x, const float* dy, float* d_weight,
                             int accumulate, void* stream);

/* ---- gate: sevenn/nn/equivariant_gate.py:59-61 (e3nn Gate, silu on scalars and gates) --------------- */
/* in [n, ns+ng+3*m1+5*m2] = [scalars | gates | 1e | 2e], ng = m1+m2; out [n, ns+3*m1+5*m2] */
int b200_gate_fwd(int ns, int m1, int m2, int n_nodes, const float* in, float* out, void* stream);
int b200_gate_bwd(int ns, int m1, int m2, int n_nodes, const float* in, const float* d_out, float* d_in, void* stream);
/* second-order pass (the reverse of b200_gate_bwd; sevenn/nn/force_output.py:165-170 create_graph=self.training makes
 * the fine-tuning loss differentiate the forces): a = dL/d(d_in) [n, in_dim] -> dL/d in (dx) and dL/d(d_out) (dg) */
int b200_gate_bwd2(int ns, int m1, int m2, int n_nodes, const float* in, const float* d_out, const float* a, float* dx,
                   float* dg, void* stream);

/* ---- train-mode radial weights from a table: sevenn/nn/convolution.py:97-106 (weight = weight_nn(edge_embedding)) under
 *      sevenn/train/trainer.py:173-186.  weight_nn is a function of the edge length alone; the caller evaluates it on the
 *      n_grid + 1 grid points r = i h (differentiably, in its own framework) and passes per-interval cubic coefficients
 *      P[n_grid, 4, wn] = (a, b, c, d): w(t) = a + b t + c t^2 + d t^3, t = r * inv_h - i (inv_h = 1 / h as the fp32 number the grid points were generated with: r_i = i / inv_h).  m = order of the derivative in r
 *      (0..3; 4 gives zeros).  `order`: edge ids sorted by interval (NULL = natural order; locality only).
 *      The three maps are closed under differentiation (csrc/table_train.cu). --------------------------------------- */
/* out[e, k] = coef[e] * d^m w_k / d r^m (r[e])      (coef NULL = 1) */
int b200_table_interp(int m, int wn, int64_t n_edges, float inv_h, int n_grid, const float* P, const float* r,
                      const int32_t* order, const float* coef, float* out, void* stream);
/* out[e] = sum_k g[e, k] * d^m w_k / d r^m (r[e]) */
int b200_table_dot(int m, int wn, int64_t n_edges, float inv_h, int n_grid, const float* P, const float* r, const int32_t* order,
                   const float* g, float* out, void* stream);
/* G[n_grid, 4, wn] = d (sum_e coef[e] sum_k g[e, k] d^m w_k / d r^m (r[e])) / d P     (overwritten; coef NULL = 1) */
int b200_table_scatter(int m, int wn, int64_t n_edges, float inv_h, int n_grid, const float* g, const float* coef, const float* r,
                       const int32_t* order, float* G, void* stream);

/* b200_table_scatter(m, ...) and b200_table_dot(md, ...) of the same g in ONE pass over it (the two results of a reverse pass
 * that read the [E, wn] gradient): G as above, dot[e] = sum_k g[e, k] d^md w_k / d r^md (r[e]) (coef does not enter dot) */
int b200_table_scatter_dot(int m, int md, int wn, int64_t n_edges, float inv_h, int n_grid, const float* P, const float* g,
                           const float* coef, const float* r, const int32_t* order, float* G, float* dot, void* stream);
/* Z, dZ [n_grid + 1, wn] float64: weight_nn and its derivative in r on the grid points r_i = i h  ->  P[n_grid, 4, wn]
 * (cubic Hermite coefficients per interval; the differences of neighbouring values are formed in float64) */
int b200_table_hermite(int n_grid, int wn, double h, const double* Z, const double* dZ, float* P, void* stream);
/* the adjoint: G = d L / d P  ->  d L / d Z, d L / d dZ (float64, [n_grid + 1, wn]) */
int b200_table_hermite_adj(int n_grid, int wn, double h, const float* G, double* gZ, double* gdZ, void* stream);

/* ---- node embedding: sevenn/nn/node_embedding.py:39-48 + nn/linear.py:37-44 (one-hot . W / sqrt(n_species)) */
int b200_embed_fwd(int n_nodes, int n_species, int dim, const int32_t* types, const float* W, float* x, void* stream);

/* ---- rescale + reduce: sevenn/nn/scale.py:67-73, nn/linear.py:76-90 --------------------------------- */
int b200_rescale_fwd(int n_nodes, const int32_t* types, const float* scaled, const float* scale, const float* shift,
                     int species_wise, float* atomic_energy, void* stream);
/* out[c] = sum_n in[n, c], c < ncol <= 8 (fp64 accumulation, deterministic) */
int b200_colsum(int n_rows, int ncol, const float* in, float* out, void* stream);
/* Brick decomposition of a fully periodic cell for the spatially decomposed call (pair_e3gnn_parallel.cpp:277-314 takes
 * LAMMPS' decomposition as given): key[i] = class * n_bins + cell_list_bin, class 0 = owned by this brick, 1 + owner = ghost
 * (within margin = cutoff / cell height of the brick, periodic images included) owned by brick `owner`, n_bricks + 1 = not
 * needed.  A stable sort of the keys orders the rank's atoms (cell-list order) and its ghosts (grouped by owner). */
int b200_partition_keys(int64_t n_atoms, const double* pos, const double* inv_cell_host, const int* grid, const int* my_brick,
                        const double* margin, const int* cell_bins, int64_t* key, void* stream);
/* Halo exchange of the spatially decomposed call: pack_forward_init / unpack_forward_init (index_select of the rows a
 * peer asked for, pair_e3gnn_parallel.cpp:803-866) and pack_reverse / unpack_reverse (index_add of the returned gradient
 * rows into their owners, :867-933).  rows: [n, width] fp32; idx: int32 row numbers; out / in: packed [m, width]. */
int b200_halo_pack(int64_t n_send, int width, const float* rows, const int32_t* idx, float* out, void* stream);
int b200_halo_unpack_add(int64_t n_recv, int width, float* rows, const int32_t* idx, const float* in, void* stream);

/* ---- forces / virial: sevenn/nn/force_output.py:158-215 (ForceStressOutputFromEdge) ----------------- */
int b200_force_virial(int n_atoms, const int32_t* rowptr, const int32_t* rrowptr, const int32_t* eid_by_src,
                      const float* edge_vec, const float* d_edge_vec, float* force, float* virial_atom, void* stream);

/* ---- radius graph: sevenn/train/dataload.py:31-68 (ase primitive_neighbor_list 'ijDS' + self-pair filter),
 *      LAMMPS analogue pair_e3gnn.cpp:155-182.  Two-phase: count (returns E, fills rowptr), then fill. ---- */
size_t b200_graph_workspace_bytes(int n_atoms);
int b200_graph_count(int n_atoms, const double* pos, const double* cell_host, const int* pbc_host, double cutoff,
                     void* workspace, int32_t* rowptr, int64_t* n_edges_host, void* stream);
int b200_graph_fill(int n_atoms, const double* pos, const double* cell_host, const int* pbc_host, double cutoff,
                    void* workspace, const int32_t* rowptr, int32_t* src, int32_t* dst, int32_t* shift,
                    float* edge_vec, uint64_t* tmp_keys /* [E] scratch */, void* stream);
/* CSR of the same edge list sorted by source (stable in edge id) */
size_t b200_transpose_workspace_bytes(int n_nodes, int64_t n_edges);
int b200_csr_transpose(int n_nodes, int64_t n_edges, const int32_t* src, void* workspace, int32_t* rrowptr,
                       int32_t* eid_by_src, void* stream);
/* Allocation-free variant of b200_csr_transpose.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, serving loops).  The scratch (two per-node
 * counter arrays) is a static device array inside the library, held by every GPU that loads it (about 2 MiB); the call
 * allocates nothing.  Limit: n_nodes <= b200_csr_transpose_static_max_nodes() (262,143), otherwise B200_ERR_INVALID.
 * All calls on a device share the scratch: calls that overlap on the GPU race; back-to-back calls on one stream are fine
 * (the counters are cleared at the start of every call).  Same outputs as b200_csr_transpose, bit for bit. */
int b200_csr_transpose_static_max_nodes(void);
int b200_csr_transpose_static(int n_nodes, int64_t n_edges, const int32_t* src, int32_t* rrowptr, int32_t* eid_by_src,
                              void* stream);

/* ---- whole-model E+F+S: the call PairE3GNN::compute makes (pair_e3gnn.cpp:205-266) and
 *      SevenNetCalculator.calculate (sevennet_calculator.py:119-157) ---------------------------------- */
typedef struct B200ModelConfig {
  int n_layers;
  int n_species;
  int lmax;            /* 2 */
  int mul[8][3];       /* node irreps entering layer t: mul[t][l]; mul[n_layers] = irreps after the last gate */
  int hidden[2];       /* radial MLP hidden widths */
  int readout_hidden;  /* reduce_input_to_hidden width */
  EdgeParams edge;
  int species_wise_rescale;
} B200ModelConfig;
typedef struct B200Model B200Model;
/* weights_host: the reference state dict flattened in named_parameters() order */
int b200_model_create(const B200ModelConfig* cfg, B200Model** out);
int b200_model_set_param(B200Model* m, const char* name, const float* data_host, int64_t numel);
void b200_model_destroy(B200Model* m);
size_t b200_model_workspace_bytes(const B200Model* m, int n_atoms, int64_t n_edges);
/* edge list must be sorted by dst (rowptr); outputs: energy[1], atomic_energy[n], force[n,3], virial[6]
 * (sum over edges of [rx fx, ry fy, rz fz, rx fy, ry fz, rz fx]; stress = -virial / volume) */
int b200_model_compute(B200Model* m, int n_atoms, int64_t n_edges, const int32_t* types, const int32_t* rowptr,
                       const int32_t* src, const int32_t* dst, const float* edge_vec, void* workspace,
                       size_t workspace_bytes, float* energy, float* atomic_energy, float* force, float* virial,
                       void* stream);

/* ---- spatially decomposed force call: PairE3GNNParallel::compute (pair_e3gnn_parallel.cpp:347-454) ----------
 * A rank owns n_local atoms and sees n_ghost ghosts (rows n_local .. n_local+n_ghost-1 of `types`, targets of
 * `src` only; every edge's centre `dst` is local, as pair_e3gnn_parallel.cpp:277-314 builds them).
 * rowptr has n_local+n_ghost+1 entries (ghost rows empty).  `halo` is called on the calling thread, between
 * kernels enqueued on `stream`, where the reference does its LAMMPS comm (:397-400 forward, :446-453 reverse):
 *   B200_HALO_FWD_X  rows[n_local+n_ghost, width]: fill the ghost rows from their owners   (layers >= 1)
 *   B200_HALO_BWD_X  rows[...]: add each ghost row into its owner's local row               (layers >= 1)
 *   B200_HALO_FORCE  rows = force[n_local+n_ghost, 3]: add ghost rows into the owners' rows (LAMMPS reverse_comm)
 * The hook must enqueue its work stream-ordered with `stream` and return 0.  force has n_local+n_ghost rows. */
#define B200_HALO_FWD_X 0
#define B200_HALO_BWD_X 1
#define B200_HALO_FORCE 2
typedef int (*b200_halo_fn)(void* user, int phase, int layer, float* rows, int width);
size_t b200_model_workspace_bytes_halo(const B200Model* m, int n_local, int n_ghost, int64_t n_edges);
int b200_model_compute_halo(B200Model* m, int n_local, int n_ghost, int64_t n_edges, const int32_t* types,
                            const int32_t* rowptr, const int32_t* src, const int32_t* dst, const float* edge_vec,
                            void* workspace, size_t workspace_bytes, float* energy, float* atomic_energy, float* force,
                            float* virial, b200_halo_fn halo, void* halo_user, void* stream);

/* optional per-category device timing of b200_model_compute (CUDA events on the compute stream).
 * categories: 0 edge embedding+CSR, 1 radial MLP fwd, 2 conv fwd, 3 node linears/gate/readout,
 * 4 radial MLP bwd, 5 conv bwd (weights, harmonics), 6 conv bwd (x), 7 force/virial, 8 halo exchange */
#define B200_TIMING_CATEGORIES 9
int b200_model_set_timing(B200Model* m, int enable);
int b200_model_get_timing(B200Model* m, float* ms_out, int n_out);

#ifdef __cplusplus
}
#endif
#endif
