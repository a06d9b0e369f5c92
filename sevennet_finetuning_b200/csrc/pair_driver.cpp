// Stand-alone native caller of the C-ABI: what a LAMMPS pair style would do per MD step, without LAMMPS, libtorch or
// Python on the path.  It mirrors PairE3GNN::compute (sevenn/pair_e3gnn/pair_e3gnn.cpp:72-275):
//   positions / cell / types  ->  neighbour list within the cutoff (pair_e3gnn.cpp:155-182; here the CUDA cell list)
//   -> model call (pair_e3gnn.cpp:205-215)  ->  eng_vdwl, f[i], virial[6] copied back (pair_e3gnn.cpp:231-256).
// LAMMPS itself is not available in this repository's environment, so this driver is the C++ side the parity tests run
// (tests/test_gpu_model.py::test_native_pair_driver).
//
//   pair_driver <structure.bin> <weights.bin> <out.bin> [repeat]
// structure.bin: int32 n_atoms, int32 pbc[3], float64 cutoff, float64 cell[9], float64 pos[n*3], int32 types[n]
// weights.bin:   int32 n_layers, n_species, mul[(n_layers+1)*3], hidden[2], readout_hidden, cutoff_kind, species_wise, num_basis;
//                float32 r_cut, r_on, poly_p; int32 n_params; then per parameter: int32 name_len, name, int64 numel, float32[numel]
// out.bin:       int64 n_edges, float64 ms_per_call, float32 energy, float32 atomic_energy[n], float32 force[n*3], float32 virial[6]
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/sevennet_b200.h"

#define CK(call)                                                                      \
  do {                                                                                \
    int rc_ = (call);                                                                 \
    if (rc_ != 0) {                                                                   \
      fprintf(stderr, "pair_driver: %s failed (%d): %s\n", #call, rc_, b200_last_error()); \
      return 2;                                                                       \
    }                                                                                 \
  } while (0)
#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      fprintf(stderr, "pair_driver: %s: %s\n", #call, cudaGetErrorString(e_));            \
      return 3;                                                                           \
    }                                                                                     \
  } while (0)

template <class T> static bool rd(FILE* f, T* p, size_t n = 1) { return fread(p, sizeof(T), n, f) == n; }

int main(int argc, char** argv) {
  if (argc < 4) {
    fprintf(stderr, "usage: pair_driver structure.bin weights.bin out.bin [repeat]\n");
    return 1;
  }
  const int repeat = argc > 4 ? atoi(argv[4]) : 1;
  // ---- structure (what LAMMPS hands a pair style: x, type, box)
  FILE* f = fopen(argv[1], "rb");
  if (!f) { perror(argv[1]); return 1; }
  int32_t n = 0, pbc[3];
  double cutoff, cell[9];
  if (!rd(f, &n) || !rd(f, pbc, 3) || !rd(f, &cutoff) || !rd(f, cell, 9)) { fprintf(stderr, "bad structure file\n"); return 1; }
  std::vector<double> pos((size_t)n * 3);
  std::vector<int32_t> types(n);
  if (!rd(f, pos.data(), pos.size()) || !rd(f, types.data(), types.size())) { fprintf(stderr, "bad structure file\n"); return 1; }
  fclose(f);
  // ---- model (what init_style() does once: metadata of the deployed model + its parameters)
  f = fopen(argv[2], "rb");
  if (!f) { perror(argv[2]); return 1; }
  B200ModelConfig cfg;
  memset(&cfg, 0, sizeof(cfg));
  int32_t kind = 0, nparams = 0;
  if (!rd(f, &cfg.n_layers) || !rd(f, &cfg.n_species)) { fprintf(stderr, "bad weights file\n"); return 1; }
  cfg.lmax = 2;
  for (int t = 0; t <= cfg.n_layers; ++t)
    if (!rd(f, cfg.mul[t], 3)) return 1;
  if (!rd(f, cfg.hidden, 2) || !rd(f, &cfg.readout_hidden) || !rd(f, &kind) || !rd(f, &cfg.species_wise_rescale) ||
      !rd(f, &cfg.edge.num_basis))
    return 1;
  if (!rd(f, &cfg.edge.r_cut) || !rd(f, &cfg.edge.r_on) || !rd(f, &cfg.edge.poly_p) || !rd(f, &nparams)) return 1;
  cfg.edge.cutoff_kind = kind;
  cfg.edge.normalize_sph = 1;
  B200Model* model = nullptr;
  CK(b200_model_create(&cfg, &model));
  for (int k = 0; k < nparams; ++k) {
    int32_t len = 0;
    int64_t numel = 0;
    if (!rd(f, &len)) return 1;
    std::string name(len, '\0');
    if (!rd(f, &name[0], len) || !rd(f, &numel)) return 1;
    std::vector<float> w(numel);
    if (!rd(f, w.data(), w.size())) return 1;
    CK(b200_model_set_param(model, name.c_str(), w.data(), numel));
  }
  fclose(f);

  cudaStream_t st;
  CU(cudaStreamCreate(&st));
  double* d_pos;
  int32_t *d_types, *d_rowptr;
  void* d_gws;
  CU(cudaMalloc(&d_pos, sizeof(double) * pos.size()));
  CU(cudaMalloc(&d_types, sizeof(int32_t) * n));
  CU(cudaMalloc(&d_rowptr, sizeof(int32_t) * (n + 1)));
  CU(cudaMalloc(&d_gws, b200_graph_workspace_bytes(n)));
  CU(cudaMemcpy(d_types, types.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice));
  float *d_e, *d_ae, *d_f, *d_vir;
  CU(cudaMalloc(&d_e, sizeof(float)));
  CU(cudaMalloc(&d_ae, sizeof(float) * n));
  CU(cudaMalloc(&d_f, sizeof(float) * n * 3));
  CU(cudaMalloc(&d_vir, sizeof(float) * 6));
  int32_t *d_src = nullptr, *d_dst = nullptr, *d_shift = nullptr;
  float* d_vec = nullptr;
  uint64_t* d_keys = nullptr;
  void* d_ws = nullptr;
  size_t cap_edges = 0, cap_ws = 0;
  int64_t E = 0;
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0));
  CU(cudaEventCreate(&e1));
  float ms = 0.f;
  // ---- compute(): every MD step
  for (int it = 0; it < repeat + 1; ++it) {   // first pass = warm-up / allocation
    if (it == 1) CU(cudaEventRecord(e0, st));
    CU(cudaMemcpyAsync(d_pos, pos.data(), sizeof(double) * pos.size(), cudaMemcpyHostToDevice, st));
    CK(b200_graph_count(n, d_pos, cell, pbc, cutoff, d_gws, d_rowptr, &E, st));
    if ((size_t)E > cap_edges) {   // grow-only device buffers, as a pair style keeps them between steps
      cudaFree(d_src); cudaFree(d_dst); cudaFree(d_shift); cudaFree(d_vec); cudaFree(d_keys);
      cap_edges = (size_t)E + (size_t)E / 8 + 16;
      CU(cudaMalloc(&d_src, sizeof(int32_t) * cap_edges));
      CU(cudaMalloc(&d_dst, sizeof(int32_t) * cap_edges));
      CU(cudaMalloc(&d_shift, sizeof(int32_t) * 3 * cap_edges));
      CU(cudaMalloc(&d_vec, sizeof(float) * 3 * cap_edges));
      CU(cudaMalloc(&d_keys, sizeof(uint64_t) * cap_edges));
    }
    CK(b200_graph_fill(n, d_pos, cell, pbc, cutoff, d_gws, d_rowptr, d_src, d_dst, d_shift, d_vec, d_keys, st));
    const size_t need = b200_model_workspace_bytes(model, n, E);
    if (need > cap_ws) {
      cudaFree(d_ws);
      cap_ws = need + need / 8;
      CU(cudaMalloc(&d_ws, cap_ws));
    }
    CK(b200_model_compute(model, n, E, d_types, d_rowptr, d_src, d_dst, d_vec, d_ws, cap_ws, d_e, d_ae, d_f, d_vir, st));
  }
  CU(cudaEventRecord(e1, st));
  CU(cudaStreamSynchronize(st));
  if (repeat > 0) CU(cudaEventElapsedTime(&ms, e0, e1));
  float energy, vir[6];
  std::vector<float> ae(n), force((size_t)n * 3);
  CU(cudaMemcpy(&energy, d_e, sizeof(float), cudaMemcpyDeviceToHost));
  CU(cudaMemcpy(ae.data(), d_ae, sizeof(float) * n, cudaMemcpyDeviceToHost));
  CU(cudaMemcpy(force.data(), d_f, sizeof(float) * n * 3, cudaMemcpyDeviceToHost));
  CU(cudaMemcpy(vir, d_vir, sizeof(float) * 6, cudaMemcpyDeviceToHost));
  f = fopen(argv[3], "wb");
  if (!f) { perror(argv[3]); return 1; }
  const double ms_per_call = repeat > 0 ? (double)ms / repeat : 0.0;
  fwrite(&E, sizeof(E), 1, f);
  fwrite(&ms_per_call, sizeof(double), 1, f);
  fwrite(&energy, sizeof(float), 1, f);
  fwrite(ae.data(), sizeof(float), ae.size(), f);
  fwrite(force.data(), sizeof(float), force.size(), f);
  fwrite(vir, sizeof(float), 6, f);
  fclose(f);
  printf("pair_driver: %d atoms, %lld edges, E = %.6f eV, %.3f ms per call\n", n, (long long)E, energy, ms_per_call);
  b200_model_destroy(model);
  return 0;
}
