// Whole-model energy + forces + virial: the call PairE3GNN::compute (pair_e3gnn.cpp:205-266) and
// SevenNetCalculator.calculate (sevennet_calculator.py:119-157) make into the model, with a hand-written
// reverse pass instead of autograd (ForceStressOutputFromEdge semantics, nn/force_output.py:158-215).
//
// Layer order follows sevenn/model_build.py:219-437 and nn/interaction_blocks.py:49-84:
//   edge_embedding -> onehot embedding -> for t: sc_intro | self_interaction_1 -> convolution ->
//   self_interaction_2 -> (+ self connection) -> gate;  readout 2 linears -> rescale -> sum.
#include <stdlib.h>

#include <map>
#include <string>
#include <vector>

#include "ops.h"

namespace b200 {

void fill_conv_desc(int m0, int m1, int m2, int lout, ConvDesc& d) {
  d.m0 = m0; d.m1 = m1; d.m2 = m2; d.lout = lout;
  d.xdim = m0 + 3 * m1 + 5 * m2;
  static const int L[kNumPaths][3] = {{0, 0, 0}, {0, 1, 1}, {0, 2, 2}, {1, 0, 1}, {1, 1, 0}, {1, 1, 1}, {1, 1, 2}, {1, 2, 1},
                                      {1, 2, 2}, {2, 0, 2}, {2, 1, 1}, {2, 1, 2}, {2, 2, 0}, {2, 2, 1}, {2, 2, 2}};
  const int mul[3] = {m0, m1, m2};
  int woff = 0;
  for (int p = 0; p < kNumPaths; ++p) {
    d.woff[p] = -1;
    d.ooff[p] = -1;
    if (mul[L[p][0]] > 0 && L[p][2] <= lout) {
      d.woff[p] = woff;
      woff += mul[L[p][0]];
    }
  }
  d.wn = woff;
  int ooff = 0;
  for (int l3 = 0; l3 <= lout; ++l3)
    for (int p = 0; p < kNumPaths; ++p)
      if (d.woff[p] >= 0 && L[p][2] == l3) {
        d.ooff[p] = ooff;
        ooff += mul[L[p][0]] * (2 * l3 + 1);
      }
  d.odim = ooff;
}

// irreps linear between two (m0,m1,m2)-shaped irreps; blocks exist where both sides have the l
static LinDesc make_linear(const int in[3], const int out[3]) {
  LinDesc d{};
  int io = 0, oo = 0, wo = 0;
  for (int l = 0; l < 3; ++l) {
    const int D = 2 * l + 1;
    if (in[l] > 0 && out[l] > 0) {
      LinBlock& b = d.blk[d.nblk++];
      b.d = D; b.mul_in = in[l]; b.mul_out = out[l]; b.in_off = io; b.out_off = oo; b.w_off = wo;
      b.scale = 1.0f / sqrtf((float)in[l]);
      wo += in[l] * out[l];
    }
    io += in[l] * D;
    oo += out[l] * D;
  }
  d.in_dim = io;
  d.out_dim = oo;
  return d;
}
static int lin_numel(const LinDesc& d) {
  int n = 0;
  for (int i = 0; i < d.nblk; ++i) n += d.blk[i].mul_in * d.blk[i].mul_out;
  return n;
}
static bool lin_covers_input(const LinDesc& d) {
  int covered = 0;
  for (int i = 0; i < d.nblk; ++i) covered += d.blk[i].mul_in * d.blk[i].d;
  return covered == d.in_dim;
}
static bool lin_covers_output(const LinDesc& d) {
  int covered = 0;
  for (int i = 0; i < d.nblk; ++i) covered += d.blk[i].mul_out * d.blk[i].d;
  return covered == d.out_dim;
}

struct Layer {
  int in[3], out[3];      // node irreps entering / leaving the layer
  int gin[3];             // gate input irreps: (ns+ng, m1, m2)
  int msg[3];             // convolution output multiplicities per l
  ConvDesc conv;
  MlpDesc mlp;
  LinDesc sc, si1, si2;
  int lout;
  bool table;   // per-edge weights interpolated from the tabulated radial function (radial_table.cu)
  bool tabconv; // ... inside the convolution kernels: neither the weights nor their gradient are materialised
  bool si1_cm;  // ... and self_interaction_1 produces / consumes that row layout directly
  bool pair;    // ... by the channel-pair kernels (conv_pair.cu): packed table, component-major source rows
};

}  // namespace b200

using namespace b200;

struct B200Model {
  B200ModelConfig cfg;
  std::vector<Layer> layers;
  LinDesc r1, r2;
  std::map<std::string, std::pair<float*, long long>> params;  // name -> (device ptr, numel)
  std::map<std::string, long long> expected;
  std::vector<float> denom;  // host copies of {t}_convolution.denominator
  // optional per-category device timing (CUDA events on the compute stream)
  bool timing = false;
  bool split_bwd_x = false;   // B200_SPLIT_BWD_X=1: d x[src] by the separate source-sorted kernel (deterministic sums)
  std::vector<float*> tables; // per layer: tabulated radial weights (device), rebuilt when a parameter changes
  bool tables_dirty = true;
  // hi / lo operand packs of the tensor-core linears, one per (weight, transposed?): built at the first call after a
  // parameter change, reused by every call after it
  std::map<std::pair<const float*, int>, float*> packs;
  std::vector<cudaEvent_t> ev;
  std::vector<int> ev_cat;
  size_t ev_used = 0;
};

enum { CAT_EDGE = 0, CAT_MLP_FWD, CAT_CONV_FWD, CAT_NODE, CAT_MLP_BWD, CAT_CONV_BWD_EDGE, CAT_CONV_BWD_X, CAT_FORCE, CAT_HALO, CAT_COUNT };

static void tick(B200Model* m, int cat, cudaStream_t st) {
  if (!m->timing) return;
  if (m->ev_used + 2 > m->ev.size()) {
    for (int k = 0; k < 64; ++k) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      m->ev.push_back(e);
      m->ev_cat.push_back(-1);
    }
  }
  m->ev_cat[m->ev_used] = cat;
  cudaEventRecord(m->ev[m->ev_used++], st);
}
// OP(cat, call): run one launcher, bracketed by events when timing is on
#define OP(cat, call)                 \
  do {                                \
    tick(m, cat, st);                 \
    int rc_ = (call);                 \
    tick(m, -1, st);                  \
    if (rc_ != B200_OK) return rc_;   \
  } while (0)

extern "C" int b200_model_set_timing(B200Model* m, int enable) {
  B200_REQUIRE(m, "model_set_timing: null model");
  m->timing = enable != 0;
  m->ev_used = 0;
  return B200_OK;
}
// accumulated device milliseconds per category since set_timing(1); synchronises the device
extern "C" int b200_model_get_timing(B200Model* m, float* ms_out, int n_out) {
  B200_REQUIRE(m && ms_out && n_out >= CAT_COUNT, "model_get_timing: need %d output slots", (int)CAT_COUNT);
  cudaDeviceSynchronize();
  for (int k = 0; k < n_out; ++k) ms_out[k] = 0.f;
  for (size_t k = 0; k + 1 < m->ev_used; k += 2) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, m->ev[k], m->ev[k + 1]);
    ms_out[m->ev_cat[k]] += ms;
  }
  m->ev_used = 0;
  return B200_OK;
}

// slot of the operand pack of (weight, transposed?) in the model's cache
static float** PACK(B200Model* m, const float* W, int transpose) { return &m->packs[{W, transpose}]; }

static float* P(B200Model* m, const std::string& name) {
  auto it = m->params.find(name);
  return it == m->params.end() ? nullptr : it->second.first;
}

extern "C" int b200_model_create(const B200ModelConfig* cfg, B200Model** out) {
  B200_REQUIRE(cfg && out, "model_create: null argument");
  B200_REQUIRE(cfg->n_layers >= 1 && cfg->n_layers <= 7 && cfg->lmax == 2, "model_create: 1..7 layers, lmax 2");
  B200Model* m = new B200Model();
  m->cfg = *cfg;
  const int nl = cfg->n_layers;
  m->denom.assign(nl, 1.0f);
  for (int t = 0; t < nl; ++t) {
    Layer L{};
    for (int l = 0; l < 3; ++l) { L.in[l] = cfg->mul[t][l]; L.out[l] = cfg->mul[t + 1][l]; }
    L.lout = (L.out[1] + L.out[2] > 0) ? 2 : 0;
    L.gin[0] = L.out[0] + L.out[1] + L.out[2]; L.gin[1] = L.out[1]; L.gin[2] = L.out[2];
    fill_conv_desc(L.in[0], L.in[1], L.in[2], L.lout, L.conv);
    static const int L3[kNumPaths] = {0, 1, 2, 1, 0, 1, 2, 1, 2, 2, 1, 2, 0, 1, 2};
    static const int L1[kNumPaths] = {0, 0, 0, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2};
    L.msg[0] = L.msg[1] = L.msg[2] = 0;
    for (int p = 0; p < kNumPaths; ++p)
      if (L.conv.woff[p] >= 0) L.msg[L3[p]] += L.in[L1[p]];
    L.mlp = MlpDesc{cfg->edge.num_basis, cfg->hidden[0], cfg->hidden[1], L.conv.wn};
    L.sc = make_linear(L.in, L.gin);
    L.si1 = make_linear(L.in, L.in);
    L.si2 = make_linear(L.msg, L.gin);
    L.table = radial_table_supported(L.mlp) && !(getenv("B200_RADIAL_TABLE") && atoi(getenv("B200_RADIAL_TABLE")) == 0);
    L.tabconv = L.table && !(getenv("B200_TABLE_IN_CONV") && atoi(getenv("B200_TABLE_IN_CONV")) == 0);
    L.pair = L.tabconv && conv_pair_supported(L.conv) && !(getenv("B200_CONV_PAIR") && atoi(getenv("B200_CONV_PAIR")) == 0);
    // self_interaction_1 writes (and its transposed product reads) the component-major rows of the pair kernels itself
    // (linear_tc.cu) instead of a transposition pass either way; B200_LINEAR_CM=0: the separate passes
    L.si1_cm = L.pair && irreps_linear_tc_supported(L.si1, 0, kLinYcm) && irreps_linear_tc_supported(L.si1, 1, kLinXcm) &&
               !(getenv("B200_LINEAR_CM") && atoi(getenv("B200_LINEAR_CM")) == 0);
    m->layers.push_back(L);
    m->split_bwd_x = getenv("B200_SPLIT_BWD_X") && atoi(getenv("B200_SPLIT_BWD_X")) != 0;
    const std::string s = std::to_string(t);
    m->expected[s + "_self_connection_intro.linear.weight"] = lin_numel(L.sc);
    m->expected[s + "_self_interaction_1.linear.weight"] = lin_numel(L.si1);
    m->expected[s + "_self_interaction_2.linear.weight"] = lin_numel(L.si2);
    m->expected[s + "_convolution.weight_nn.layer0.weight"] = (long long)L.mlp.nb * L.mlp.h1;
    m->expected[s + "_convolution.weight_nn.layer1.weight"] = (long long)L.mlp.h1 * L.mlp.h2;
    m->expected[s + "_convolution.weight_nn.layer2.weight"] = (long long)L.mlp.h2 * L.mlp.wn;
    m->expected[s + "_convolution.denominator"] = 1;
  }
  const int last[3] = {cfg->mul[nl][0], cfg->mul[nl][1], cfg->mul[nl][2]};
  const int hid[3] = {cfg->readout_hidden, 0, 0}, one[3] = {1, 0, 0};
  m->r1 = make_linear(last, hid);
  m->r2 = make_linear(hid, one);
  m->expected["onehot_to_feature_x.linear.weight"] = (long long)cfg->n_species * cfg->mul[0][0];
  m->expected["reduce_input_to_hidden.linear.weight"] = lin_numel(m->r1);
  m->expected["reduce_hidden_to_energy.linear.weight"] = lin_numel(m->r2);
  const int nsc = cfg->species_wise_rescale ? cfg->n_species : 1;
  m->expected["rescale_atomic_energy.shift"] = nsc;
  m->expected["rescale_atomic_energy.scale"] = nsc;
  m->expected["edge_embedding.basis_function.coeffs"] = cfg->edge.num_basis;
  *out = m;
  return B200_OK;
}

extern "C" int b200_model_set_param(B200Model* m, const char* name, const float* data_host, int64_t numel) {
  B200_REQUIRE(m && name && data_host, "model_set_param: null argument");
  auto it = m->expected.find(name);
  B200_REQUIRE(it != m->expected.end(), "model_set_param: unknown parameter '%s'", name);
  B200_REQUIRE(it->second == numel, "model_set_param: '%s' expects %lld values, got %lld", name, it->second,
               (long long)numel);
  auto& slot = m->params[name];
  if (!slot.first) {
    if (cudaMalloc(&slot.first, sizeof(float) * (size_t)numel) != cudaSuccess) {
      set_error("model_set_param: cudaMalloc failed");
      return B200_ERR_CUDA;
    }
    slot.second = numel;
  }
  if (cudaMemcpy(slot.first, data_host, sizeof(float) * (size_t)numel, cudaMemcpyHostToDevice) != cudaSuccess) {
    set_error("model_set_param: cudaMemcpy failed");
    return B200_ERR_CUDA;
  }
  m->tables_dirty = true;
  if (!m->packs.empty()) {
    cudaDeviceSynchronize();
    for (auto& kv : m->packs) cudaFree(kv.second);
    m->packs.clear();
  }
  const std::string nm(name);
  const std::string suffix = "_convolution.denominator";
  if (nm.size() > suffix.size() && nm.compare(nm.size() - suffix.size(), suffix.size(), suffix) == 0)
    m->denom[std::stoi(nm)] = data_host[0];
  return B200_OK;
}

extern "C" void b200_model_destroy(B200Model* m) {
  if (!m) return;
  for (auto& kv : m->params) cudaFree(kv.second.first);
  for (auto t : m->tables) cudaFree(t);
  for (auto& kv : m->packs) cudaFree(kv.second);
  for (auto e : m->ev) cudaEventDestroy(e);
  delete m;
}

namespace {
struct Plan {
  size_t off = 0;
  char* base = nullptr;
  template <class T> T* take(size_t n) {
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += (n * sizeof(T) + 255) / 256 * 256;
    return p;
  }
};
struct Buffers {
  float *Y, *emb, *len, *dY, *demb, *ddr, *dvec, *dw, *x[2], *conv_out, *g_x[2], *g_gin, *g_conv, *h, *s, *g_h, *g_s, *virial_atom;
  std::vector<float*> w, xs, gin;
  int *rrowptr, *eid, *rorder;
  void *tws, *tws2;
};
void plan_buffers(const B200Model* m, int NL, int NA, long long E, Plan& p, Buffers& b) {
  const int N = NL;
  int maxx = 0, maxg = 0, maxmsg = 0, maxwn = 0;
  for (const Layer& L : m->layers) {
    maxx = std::max(maxx, L.conv.xdim);
    maxg = std::max(maxg, L.sc.out_dim);
    maxmsg = std::max(maxmsg, L.conv.odim);
    if (!L.tabconv) maxwn = std::max(maxwn, L.conv.wn);   // these layers materialise [E, wn] weights and their gradient
  }
  maxx = std::max(maxx, m->r1.in_dim);
  const int nb = m->cfg.edge.num_basis;
  b.Y = p.take<float>((size_t)E * 9);
  b.emb = p.take<float>((size_t)E * nb);
  b.len = p.take<float>((size_t)E);
  b.dY = p.take<float>((size_t)E * 9);
  b.demb = p.take<float>((size_t)E * nb);
  b.ddr = p.take<float>((size_t)E);
  b.dvec = p.take<float>((size_t)E * 3);
  b.dw = p.take<float>((size_t)E * maxwn);
  for (int k = 0; k < 2; ++k) b.x[k] = p.take<float>((size_t)NA * maxx);
  for (int k = 0; k < 2; ++k) b.g_x[k] = p.take<float>((size_t)N * maxx);
  b.conv_out = p.take<float>((size_t)NA * maxmsg);
  b.g_conv = p.take<float>((size_t)N * maxmsg);
  b.g_gin = p.take<float>((size_t)N * maxg);
  b.h = p.take<float>((size_t)N * m->r1.out_dim);
  b.g_h = p.take<float>((size_t)N * m->r1.out_dim);
  b.s = p.take<float>((size_t)N);
  b.g_s = p.take<float>((size_t)N);
  b.virial_atom = p.take<float>((size_t)NA * 6);
  b.w.clear(); b.xs.clear(); b.gin.clear();
  for (const Layer& L : m->layers) {
    b.w.push_back(p.take<float>(L.tabconv ? 0 : (size_t)E * L.conv.wn));
    b.xs.push_back(p.take<float>((size_t)NA * L.conv.xdim));
    b.gin.push_back(p.take<float>((size_t)N * L.sc.out_dim));
  }
  b.rrowptr = p.take<int>((size_t)NA + 1);
  b.eid = p.take<int>((size_t)E);
  b.tws = p.take<char>(transpose_workspace_bytes(NA, E));
  // edges sorted by the interval of the tabulated radial functions (radial_table.cu)
  b.rorder = p.take<int>((size_t)E);
  b.tws2 = p.take<char>(radial_table_order_bytes(E));
}

__global__ void rescale_bwd_kernel(int n, const int* __restrict__ types, const float* __restrict__ scale, int species_wise,
                                   float* __restrict__ g) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) g[i] = scale[species_wise ? types[i] : 0];
}
}  // namespace

extern "C" size_t b200_model_workspace_bytes_halo(const B200Model* m, int n_local, int n_ghost, int64_t n_edges) {
  Plan p;
  Buffers b;
  plan_buffers(m, n_local, n_local + n_ghost, n_edges, p, b);
  return p.off + 256;
}
extern "C" size_t b200_model_workspace_bytes(const B200Model* m, int n_atoms, int64_t n_edges) {
  return b200_model_workspace_bytes_halo(m, n_atoms, 0, n_edges);
}

#define TRY(call)                 \
  do {                            \
    int rc_ = (call);             \
    if (rc_ != B200_OK) return rc_; \
  } while (0)

// halo hook: called on the compute stream's host thread at the exchange points (see header)
#define HALO(phase, layer, rows, width)                                                   \
  do {                                                                                    \
    if (halo) {                                                                           \
      tick(m, CAT_HALO, st);                                                              \
      int rc_ = halo(halo_user, phase, layer, rows, width);                               \
      tick(m, -1, st);                                                                    \
      if (rc_ != 0) {                                                                     \
        set_error("model_compute: halo callback failed (phase %d, layer %d)", phase, layer); \
        return B200_ERR_INVALID;                                                          \
      }                                                                                   \
    }                                                                                     \
  } while (0)

extern "C" int b200_model_compute_halo(B200Model* m, int N, int n_ghost, int64_t E, const int32_t* types,
                                       const int32_t* rowptr, const int32_t* src, const int32_t* dst,
                                       const float* edge_vec, void* workspace, size_t workspace_bytes, float* energy,
                                       float* atomic_energy, float* force, float* virial, b200_halo_fn halo,
                                       void* halo_user, void* stream) {
  B200_REQUIRE(m, "model_compute: null model");
  B200_REQUIRE(N >= 0 && n_ghost >= 0 && (n_ghost == 0 || halo), "model_compute: ghosts need a halo callback");
  for (auto& kv : m->expected)
    B200_REQUIRE(m->params.count(kv.first), "model_compute: parameter '%s' was never set", kv.first.c_str());
  B200_REQUIRE(E < (1LL << 31), "model_compute: more than 2^31 edges per device");
  cudaStream_t st = (cudaStream_t)stream;
  const int NA = N + n_ghost;   // rows that can be edge sources: local atoms then ghosts
  Plan p;
  p.base = (char*)(((uintptr_t)workspace + 255) / 256 * 256);
  Buffers b;
  plan_buffers(m, N, NA, E, p, b);
  B200_REQUIRE(p.off + 256 <= workspace_bytes, "model_compute: workspace too small (%zu needed)", p.off + 256);
  const B200ModelConfig& c = m->cfg;
  const float* coeffs = P(m, "edge_embedding.basis_function.coeffs");
  const int nl = c.n_layers;

  // tabulated radial functions: rebuilt (float64, a few hundred microseconds per layer) after a parameter change
  if (m->tables_dirty) {
    if (m->tables.empty()) m->tables.assign(nl, nullptr);
    for (int t = 0; t < nl; ++t) {
      const Layer& L = m->layers[t];
      if (!L.table) continue;
      const std::string s = std::to_string(t);
      const size_t tbytes = L.pair ? radial_table_packed_bytes(c.edge, L.mlp) : sizeof(float) * radial_table_floats(c.edge, L.mlp);
      if (!m->tables[t] && cudaMalloc(&m->tables[t], tbytes) != cudaSuccess) {
        set_error("model_compute: cudaMalloc of the radial table failed");
        return B200_ERR_CUDA;
      }
      if (L.pair)
        OP(CAT_MLP_FWD, radial_table_build_packed(c.edge, L.mlp, coeffs, P(m, s + "_convolution.weight_nn.layer0.weight"),
                                                  P(m, s + "_convolution.weight_nn.layer1.weight"),
                                                  P(m, s + "_convolution.weight_nn.layer2.weight"), m->tables[t], st));
      else
        OP(CAT_MLP_FWD, radial_table_build(c.edge, L.mlp, coeffs, P(m, s + "_convolution.weight_nn.layer0.weight"),
                                           P(m, s + "_convolution.weight_nn.layer1.weight"),
                                           P(m, s + "_convolution.weight_nn.layer2.weight"), m->tables[t], st));
    }
    m->tables_dirty = false;
  }

  // ---------------- forward ----------------
  OP(CAT_EDGE, edge_embed_fwd(c.edge, E, edge_vec, coeffs, b.len, b.Y, b.emb, st));
  OP(CAT_EDGE, csr_transpose(NA, E, src, b.tws, b.rrowptr, b.eid, st));
  bool any_table = false;
  for (const Layer& L : m->layers) any_table = any_table || (L.table && !L.tabconv);
  if (any_table) {   // counting sort of the edges by table interval, once per call for all layers and both passes
    OP(CAT_EDGE, radial_table_order(c.edge, E, b.len, b.tws2, b.rorder, st));
  }
  float* x = b.x[0];
  float* xn = b.x[1];
  OP(CAT_NODE, embed_fwd(NA, c.n_species, c.mul[0][0], types, P(m, "onehot_to_feature_x.linear.weight"), x, st));
  const std::vector<float>& denom = m->denom;
  // a node linear with the named weight; its tensor-core operand pack is kept in the model between calls
  auto lin = [&](const LinDesc& d, int n, const float* xin, const std::string& name, float* yout, int transpose, int accumulate) {
    const float* W = P(m, name);
    return irreps_linear(d, n, xin, W, yout, transpose, accumulate, st, PACK(m, W, transpose));
  };
  for (int t = 0; t < nl; ++t) {
    const Layer& L = m->layers[t];
    const std::string s = std::to_string(t);
    if (!lin_covers_output(L.sc)) cudaMemsetAsync(b.gin[t], 0, sizeof(float) * (size_t)N * L.sc.out_dim, st);
    OP(CAT_NODE, lin(L.sc, N, x, s + "_self_connection_intro.linear.weight", b.gin[t], 0, 0));
    // layer 0: ghost rows are recomputed from their types (model_build.py:134-172 "ghost_" modules);
    // later layers: the owners' rows are fetched through the halo hook
    // the channel-pair kernels read component-major rows; ghost rows arrive from their owners already transposed
    const int n_si1 = t == 0 ? NA : N;
    const bool cm = L.si1_cm && n_si1 >= 64 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
    if (cm) {
      OP(CAT_NODE, irreps_linear_tc(L.si1, n_si1, x, P(m, s + "_self_interaction_1.linear.weight"), b.xs[t], 0, 0, st, kLinYcm,
                                    PACK(m, P(m, s + "_self_interaction_1.linear.weight"), 0)));
    } else {
      OP(CAT_NODE, lin(L.si1, n_si1, x, s + "_self_interaction_1.linear.weight", b.xs[t], 0, 0));
      if (L.pair) OP(CAT_NODE, conv_pair_interleave(L.conv, n_si1, b.xs[t], 0, st));
    }
    if (t > 0) HALO(B200_HALO_FWD_X, t, b.xs[t], L.conv.xdim);
    const float* W0 = P(m, s + "_convolution.weight_nn.layer0.weight");
    const float* W1 = P(m, s + "_convolution.weight_nn.layer1.weight");
    const float* W2 = P(m, s + "_convolution.weight_nn.layer2.weight");
    {
      if (L.pair) {
        const RadialTableGrid tg = radial_table_grid(c.edge);
        OP(CAT_CONV_FWD, conv_pair_fwd(L.conv, N, rowptr, src, b.xs[t], b.Y, b.len, m->tables[t], (float)(1.0 / tg.h), tg.n, denom[t],
                                       b.conv_out, st));
      } else if (L.tabconv) {
        const RadialTableGrid tg = radial_table_grid(c.edge);
        OP(CAT_CONV_FWD, conv_fwd_tab(L.conv, N, rowptr, src, b.xs[t], b.Y, b.len, m->tables[t], (float)(1.0 / tg.h), tg.n, denom[t],
                                      b.conv_out, st));
      } else {
        if (L.table) OP(CAT_MLP_FWD, radial_table_fwd(c.edge, L.mlp, E, b.len, b.rorder, m->tables[t], b.w[t], st));
        else OP(CAT_MLP_FWD, radial_mlp_fwd(L.mlp, E, b.emb, W0, W1, W2, b.w[t], st));
        OP(CAT_CONV_FWD, conv_fwd(L.conv, N, rowptr, src, b.xs[t], b.Y, b.w[t], denom[t], b.conv_out, st));
      }
    }
    OP(CAT_NODE, lin(L.si2, N, b.conv_out, s + "_self_interaction_2.linear.weight", b.gin[t], 0, 1));
    OP(CAT_NODE, gate_fwd(L.out[0], L.out[1], L.out[2], N, b.gin[t], xn, st));
    std::swap(x, xn);
  }
  OP(CAT_NODE, lin(m->r1, N, x, "reduce_input_to_hidden.linear.weight", b.h, 0, 0));
  OP(CAT_NODE, lin(m->r2, N, b.h, "reduce_hidden_to_energy.linear.weight", b.s, 0, 0));
  OP(CAT_NODE, rescale_fwd(N, types, b.s, P(m, "rescale_atomic_energy.scale"), P(m, "rescale_atomic_energy.shift"),
                  c.species_wise_rescale, atomic_energy, st));
  OP(CAT_NODE, colsum(N, 1, atomic_energy, energy, st));

  // ---------------- reverse pass: dE/d edge_vec ----------------
  if (N > 0)
    rescale_bwd_kernel<<<(N + 255) / 256, 256, 0, st>>>(N, types, P(m, "rescale_atomic_energy.scale"),
                                                        c.species_wise_rescale, b.g_s);
  OP(CAT_NODE, lin(m->r2, N, b.g_s, "reduce_hidden_to_energy.linear.weight", b.g_h, 1, 0));
  float* g = b.g_x[0];
  float* gp = b.g_x[1];
  OP(CAT_NODE, lin(m->r1, N, b.g_h, "reduce_input_to_hidden.linear.weight", g, 1, 0));
  bool dy_written = false, ddr_written = false, demb_written = false;   // which per-edge gradient buffers hold a sum already
  {   // the channel-pair kernels ADD their d Y and d E / d r into the per-edge buffers
    bool any_pair = false;
    for (const Layer& L : m->layers) any_pair = any_pair || L.pair;
    if (any_pair && E > 0) {
      cudaMemsetAsync(b.dY, 0, sizeof(float) * 9 * (size_t)E, st);
      cudaMemsetAsync(b.ddr, 0, sizeof(float) * (size_t)E, st);
      dy_written = ddr_written = true;
    }
  }
  for (int t = nl - 1; t >= 0; --t) {
    const Layer& L = m->layers[t];
    const std::string s = std::to_string(t);
    OP(CAT_NODE, gate_bwd(L.out[0], L.out[1], L.out[2], N, b.gin[t], g, b.g_gin, st));
    OP(CAT_NODE, lin(L.si2, N, b.g_gin, s + "_self_interaction_2.linear.weight", b.g_conv, 1, 0));
    const float* W0 = P(m, s + "_convolution.weight_nn.layer0.weight");
    const float* W1 = P(m, s + "_convolution.weight_nn.layer1.weight");
    const float* W2 = P(m, s + "_convolution.weight_nn.layer2.weight");
    // features entering layer t > 0 depend on the positions through the earlier layers: d x is needed there
    float* dxs = t > 0 ? b.conv_out : nullptr;
    if (L.pair) {
      const RadialTableGrid tg = radial_table_grid(c.edge);
      OP(CAT_CONV_BWD_EDGE, conv_pair_bwd(L.conv, N, NA, rowptr, src, b.xs[t], b.Y, b.len, m->tables[t], (float)(1.0 / tg.h), tg.n,
                                          b.g_conv, denom[t], b.dY, b.ddr, dxs, st));
    } else if (L.tabconv) {
      // d Y, d E / d r and (t > 0) d x[src] in one pass; the weight gradient is contracted with d w / d r in registers
      const RadialTableGrid tg = radial_table_grid(c.edge);
      OP(CAT_CONV_BWD_EDGE, conv_bwd_tab(L.conv, N, NA, (int)E, rowptr, src, b.xs[t], b.Y, b.len, m->tables[t], (float)(1.0 / tg.h),
                                         tg.n, b.g_conv, denom[t], b.dY, dy_written ? 1 : 0, b.ddr, ddr_written ? 1 : 0,
                                         dxs, st));   // split_bwd_x does not apply: there is no [E, wn] weight tensor to re-read
      dy_written = ddr_written = true;
    } else {
      // d weight, d Y and (t > 0) d x[src] in one pass: the source rows are accumulated with RED.ADD
      OP(CAT_CONV_BWD_EDGE, conv_bwd_edge_x(L.conv, N, NA, (int)E, rowptr, src, b.xs[t], b.Y, b.w[t], b.g_conv, denom[t], b.dw,
                                            b.dY, dy_written ? 1 : 0, m->split_bwd_x ? nullptr : dxs, st));
      dy_written = true;
      if (L.table) {   // d E / d r directly: joins the per-edge radial gradient of the table-in-convolution layers (b.ddr)
        OP(CAT_MLP_BWD, radial_table_bwd(c.edge, L.mlp, E, b.len, b.rorder, m->tables[t], b.dw, b.ddr, ddr_written ? 1 : 0, st));
        ddr_written = true;
      } else {
        OP(CAT_MLP_BWD, radial_mlp_bwd(L.mlp, E, b.emb, W0, W1, W2, b.dw, b.demb, demb_written ? 1 : 0, st));
        demb_written = true;
      }
    }
    if (t > 0) {
      if (!lin_covers_input(L.sc)) cudaMemsetAsync(gp, 0, sizeof(float) * (size_t)N * L.sc.in_dim, st);
      OP(CAT_NODE, lin(L.sc, N, b.g_gin, s + "_self_connection_intro.linear.weight", gp, 1, 0));
      if (!L.tabconv && m->split_bwd_x)
        OP(CAT_CONV_BWD_X, conv_bwd_x(L.conv, NA, b.rrowptr, b.eid, dst, b.g_conv, b.Y, b.w[t], denom[t], b.conv_out, st));
      HALO(B200_HALO_BWD_X, t, b.conv_out, L.conv.xdim);   // ghost rows are added into their owners' rows
      if (L.si1_cm && N >= 64 && (reinterpret_cast<uintptr_t>(b.conv_out) & 15) == 0) {
        OP(CAT_NODE, irreps_linear_tc(L.si1, N, b.conv_out, P(m, s + "_self_interaction_1.linear.weight"), gp, 1, 1, st, kLinXcm,
                                    PACK(m, P(m, s + "_self_interaction_1.linear.weight"), 1)));
      } else {
        if (L.pair) OP(CAT_NODE, conv_pair_interleave(L.conv, N, b.conv_out, 1, st));   // back to the standard row layout
        OP(CAT_NODE, lin(L.si1, N, b.conv_out, s + "_self_interaction_1.linear.weight", gp, 1, 1));
      }
      std::swap(g, gp);
    }
  }
  OP(CAT_EDGE, edge_embed_bwd_full(c.edge, E, edge_vec, coeffs, b.dY, demb_written ? b.demb : nullptr,
                                   ddr_written ? b.ddr : nullptr, b.dvec, st));
  OP(CAT_FORCE, force_virial(NA, rowptr, b.rrowptr, b.eid, edge_vec, b.dvec, force, b.virial_atom, st));
  HALO(B200_HALO_FORCE, -1, force, 3);
  OP(CAT_NODE, colsum(NA, 6, b.virial_atom, virial, st));   // per-source rows (force_output.py:186-213): ghosts included
  return check_launch("model_compute", N > 0 ? 1 : 0);
}

extern "C" int b200_model_compute(B200Model* m, int N, int64_t E, const int32_t* types, const int32_t* rowptr,
                                  const int32_t* src, const int32_t* dst, const float* edge_vec, void* workspace,
                                  size_t workspace_bytes, float* energy, float* atomic_energy, float* force,
                                  float* virial, void* stream) {
  return b200_model_compute_halo(m, N, 0, E, types, rowptr, src, dst, edge_vec, workspace, workspace_bytes, energy,
                                 atomic_energy, force, virial, nullptr, nullptr, stream);
}
