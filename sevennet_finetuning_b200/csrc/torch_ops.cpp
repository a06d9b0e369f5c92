// TorchScript-visible operator over the C-ABI: `torch.ops.b200.sevennet_efs` is registered with the
// dispatcher from C++, so a scripted + saved module that calls it (sevennet_finetuning_b200/deploy.py, the counterpart of
// sevenn/scripts/deploy.py:15-51) loads and runs in any libtorch process that has dlopen'ed this library - which is all
// PairE3GNN::compute (sevenn/pair_e3gnn/pair_e3gnn.cpp:205-266) needs: it feeds the dict {x, pos, edge_index, pbc_shift,
// cell_lattice_vectors, cell_volume, num_atoms} and reads inferred_total_energy / inferred_force / inferred_stress /
// atomic_energy.  No Python, no autograd: forces and stress come from the engine's hand-written reverse pass.
//
// Built by sevennet_finetuning_b200/build.py into libb200_torch_ops.so (g++ against the torch headers, linked with
// libsevennet_b200.so).
#include <ATen/ATen.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/library.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/sevennet_b200.h"

namespace {

struct Entry {
  B200Model* model = nullptr;
  int64_t version = -1;
};
std::mutex g_mu;
std::map<std::pair<const void*, int>, Entry> g_models;   // (flat weight buffer, device) -> engine handle

void check(int rc, const char* what) {
  TORCH_CHECK(rc == B200_OK, "b200::sevennet_efs: ", what, " failed (status ", rc, "): ", b200_last_error());
}

// cfg_i = [n_layers, n_species, lmax, hidden0, hidden1, readout_hidden, species_wise_rescale, num_basis, cutoff_kind,
//          normalize_sph, mul[0][0..2], ..., mul[n_layers][0..2]];  cfg_f = [r_cut, r_on, poly_p]
B200Model* get_model(const at::Tensor& flat, at::IntArrayRef offsets, const std::vector<std::string>& names,
                     at::IntArrayRef cfg_i, at::ArrayRef<double> cfg_f) {
  const int dev = flat.get_device();
  std::lock_guard<std::mutex> lock(g_mu);
  Entry& e = g_models[{flat.data_ptr(), dev}];
  const int64_t version = (int64_t)flat._version();
  if (e.model && e.version == version) return e.model;
  TORCH_CHECK(cfg_i.size() >= 10 && cfg_f.size() == 3, "b200::sevennet_efs: malformed configuration");
  const int nl = (int)cfg_i[0];
  TORCH_CHECK(nl >= 1 && nl <= 7 && (int64_t)cfg_i.size() == 10 + 3 * (nl + 1), "b200::sevennet_efs: malformed configuration");
  TORCH_CHECK(offsets.size() == names.size() + 1, "b200::sevennet_efs: offsets / names mismatch");
  if (!e.model) {
    B200ModelConfig c{};
    c.n_layers = nl; c.n_species = (int)cfg_i[1]; c.lmax = (int)cfg_i[2];
    c.hidden[0] = (int)cfg_i[3]; c.hidden[1] = (int)cfg_i[4]; c.readout_hidden = (int)cfg_i[5];
    c.species_wise_rescale = (int)cfg_i[6];
    c.edge.num_basis = (int)cfg_i[7]; c.edge.cutoff_kind = (int)cfg_i[8]; c.edge.normalize_sph = (int)cfg_i[9];
    c.edge.r_cut = (float)cfg_f[0]; c.edge.r_on = (float)cfg_f[1]; c.edge.poly_p = (float)cfg_f[2];
    for (int t = 0; t <= nl; ++t)
      for (int l = 0; l < 3; ++l) c.mul[t][l] = (int)cfg_i[10 + 3 * t + l];
    check(b200_model_create(&c, &e.model), "model_create");
  }
  const at::Tensor host = flat.detach().to(at::kCPU, at::kFloat).contiguous();
  for (size_t k = 0; k < names.size(); ++k)
    check(b200_model_set_param(e.model, names[k].c_str(), host.data_ptr<float>() + offsets[k], offsets[k + 1] - offsets[k]),
          names[k].c_str());
  e.version = version;
  return e.model;
}

// -> [inferred_total_energy (scalar), atomic_energy [N,1], inferred_force [N,3], inferred_stress [6]]
std::vector<at::Tensor> sevennet_efs(const at::Tensor& x, const at::Tensor& pos, const at::Tensor& edge_index,
                                     const at::Tensor& pbc_shift, const at::Tensor& cell, const at::Tensor& cell_volume,
                                     const at::Tensor& flat, at::IntArrayRef offsets, std::vector<std::string> names,
                                     at::IntArrayRef cfg_i, at::ArrayRef<double> cfg_f) {
  TORCH_CHECK(pos.is_cuda() && flat.is_cuda(), "b200::sevennet_efs: CUDA tensors only (there is no CPU path)");
  TORCH_CHECK(edge_index.dim() == 2 && edge_index.size(0) == 2, "b200::sevennet_efs: edge_index must be [2, E]");
  const c10::cuda::CUDAGuard guard(pos.device());
  B200Model* m = get_model(flat, offsets, names, cfg_i, cfg_f);
  const auto dev = pos.device();
  const int64_t N = pos.size(0), E = edge_index.size(1);
  at::Tensor dst = edge_index.select(0, 0).to(dev), src = edge_index.select(0, 1).to(dev), shift = pbc_shift.to(dev);
  // the engine wants the edges sorted by their centre atom (LAMMPS lists and ase lists are; a shuffled list is re-sorted)
  if (E > 1 && dst.slice(0, 1).lt(dst.slice(0, 0, E - 1)).any().item<bool>()) {
    const at::Tensor perm = at::argsort(dst, /*stable=*/true, 0, false);
    dst = dst.index_select(0, perm); src = src.index_select(0, perm); shift = shift.index_select(0, perm);
  }
  const at::Tensor p32 = pos.detach().to(at::kFloat);
  // edge_embedding.py:62-75 at zero strain
  const at::Tensor vec = (p32.index_select(0, src) - p32.index_select(0, dst) + at::matmul(shift.to(at::kFloat), cell.to(dev, at::kFloat).view({3, 3}))).contiguous();
  at::Tensor rowptr = at::zeros({N + 1}, at::TensorOptions().dtype(at::kInt).device(dev));
  if (E > 0) rowptr.slice(0, 1).copy_(at::cumsum(at::bincount(dst, {}, N), 0));
  const at::Tensor dst32 = dst.to(at::kInt).contiguous(), src32 = src.to(at::kInt).contiguous(), types = x.to(dev, at::kInt).contiguous();
  const size_t ws_bytes = b200_model_workspace_bytes(m, (int)N, E);
  at::Tensor ws = at::empty({(int64_t)ws_bytes}, at::TensorOptions().dtype(at::kByte).device(dev));
  const auto f32 = at::TensorOptions().dtype(at::kFloat).device(dev);
  at::Tensor energy = at::empty({1}, f32), ae = at::empty({N, 1}, f32), force = at::empty({N, 3}, f32), virial = at::empty({6}, f32);
  check(b200_model_compute(m, (int)N, E, types.data_ptr<int>(), rowptr.data_ptr<int>(), src32.data_ptr<int>(),
                           dst32.data_ptr<int>(), vec.data_ptr<float>(), ws.data_ptr(), ws_bytes, energy.data_ptr<float>(),
                           ae.data_ptr<float>(), force.data_ptr<float>(), virial.data_ptr<float>(),
                           (void*)c10::cuda::getCurrentCUDAStream().stream()),
        "model_compute");
  // force_output.py:98-128: stress = -(dE/d strain) / volume, order [xx, yy, zz, xy, yz, zx], volume clamped at 1e-3
  const at::Tensor stress = -virial / at::clamp_min(cell_volume.to(dev, at::kFloat).reshape({}), 1e-3);
  return {energy.reshape({}), ae, force, stress};
}

}  // namespace

TORCH_LIBRARY(b200, m) {
  m.def("sevennet_efs(Tensor x, Tensor pos, Tensor edge_index, Tensor pbc_shift, Tensor cell, Tensor cell_volume, Tensor flat, "
        "int[] offsets, str[] names, int[] cfg_i, float[] cfg_f) -> Tensor[]");
}
TORCH_LIBRARY_IMPL(b200, CUDA, m) { m.impl("sevennet_efs", &sevennet_efs); }
// no autograd through the operator (the reverse pass is inside the engine): outputs never require grad, whatever the inputs do
TORCH_LIBRARY_IMPL(b200, Autograd, m) { m.impl("sevennet_efs", torch::CppFunction::makeFallthrough()); }
