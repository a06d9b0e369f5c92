"""TorchScript deployment of the B200 path: the counterpart of ``sevenn/scripts/deploy.py:15-51``.

``deploy(state_dict, config, fname)`` writes a scripted module whose forward takes the dict ``PairE3GNN::compute`` feeds
(``sevenn/pair_e3gnn/pair_e3gnn.cpp:205-212``: x, pos, edge_index, pbc_shift, cell_lattice_vectors, cell_volume, num_atoms) and
returns it with ``inferred_total_energy``, ``atomic_energy``, ``inferred_force``, ``inferred_stress`` (``:230-266``), plus the
reference's extra files (``chemical_symbols_to_index``, ``cutoff``, ``num_species``, ``model_type``, ``version``, ``dtype``,
``time``).  The whole model is ONE dispatcher operator, ``torch.ops.b200.sevennet_efs`` (``csrc/torch_ops.cpp``, registered from
C++), so the file loads in any libtorch process that has loaded ``libb200_torch_ops.so`` first - from C++:
``dlopen("libb200_torch_ops.so", RTLD_NOW | RTLD_GLOBAL); torch::jit::load(path)``; from Python:
``torch.ops.load_library(ops_library()); torch.jit.load(path)``.  No Python at run time, no autograd: forces and stress come from
the engine's hand-written reverse pass (``pos.requires_grad`` is ignored).
"""
import os
from datetime import datetime
from typing import Dict, List

import torch

from .model import SEVENNET0_CONFIG, edge_params_from_config
from . import ops as _ops

HERE = os.path.dirname(os.path.abspath(__file__))


def ops_library():
    """Path of the operator library (built by ``build.py``); loads it into this process."""
    path = os.path.join(HERE, 'libb200_torch_ops.so')
    if not os.path.exists(path):
        raise RuntimeError(f'{path} is missing: run `python -m sevennet_finetuning_b200.build`')
    torch.ops.load_library(path)
    return path


class DeployedSevenNet(torch.nn.Module):
    """Scriptable whole-model module over ``torch.ops.b200.sevennet_efs``.  Parameters live in one flat buffer (`flat`) with the
    reference's parameter names and offsets beside it; the operator builds its engine handle from them on first use."""

    def __init__(self, state_dict, config=None):
        super().__init__()
        cfg = dict(SEVENNET0_CONFIG)
        cfg.update(config or {})
        if cfg.get('self_connection_type', 'linear') != 'linear':
            raise ValueError("deploy: the engine implements the 'linear' self connection")
        names = [k for k in state_dict]
        flats = [torch.as_tensor(state_dict[k], dtype=torch.float32).reshape(-1) for k in names]
        off = [0]
        for f in flats:
            off.append(off[-1] + int(f.numel()))
        self.register_buffer('flat', torch.cat(flats))
        self.names: List[str] = names
        self.offsets: List[int] = off
        ep = edge_params_from_config(cfg)
        nl = int(cfg['num_convolution_layer'])
        muls: List[int] = []
        for ir in cfg['irreps_manual']:
            muls += [int(m) for m in _ops.muls_of(ir)]
        self.cfg_i: List[int] = [nl, int(cfg['num_species']), int(cfg['lmax']), int(cfg['weight_nn_hidden_neurons'][0]),
                                 int(cfg['weight_nn_hidden_neurons'][1]), int(_ops.muls_of(cfg['irreps_manual'][-1])[0] // 2),
                                 int(bool(cfg['use_species_wise_shift_scale'])), int(ep.num_basis), int(ep.cutoff_kind),
                                 int(ep.normalize_sph)] + muls
        self.cfg_f: List[float] = [float(ep.r_cut), float(ep.r_on), float(ep.poly_p)]

    def forward(self, data: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        out = torch.ops.b200.sevennet_efs(data['x'], data['pos'].detach(), data['edge_index'], data['pbc_shift'],
                                          data['cell_lattice_vectors'], data['cell_volume'], self.flat, self.offsets,
                                          self.names, self.cfg_i, self.cfg_f)
        data['inferred_total_energy'] = out[0]
        data['atomic_energy'] = out[1]
        data['inferred_force'] = out[2]
        data['inferred_stress'] = out[3]
        return data


def deploy(state_dict, config, fname, chemical_symbols=None):
    """Script, freeze and save (``scripts/deploy.py:30-51``).  ``chemical_symbols``: the model's species in index order (the
    reference derives it from ``config['_type_map']``); returns the file name."""
    ops_library()
    model = DeployedSevenNet(state_dict, config).eval()
    scripted = torch.jit.freeze(torch.jit.script(model))
    cfg = dict(SEVENNET0_CONFIG)
    cfg.update(config or {})
    if chemical_symbols is None:
        from .synthetic import SEVENNET0_SPECIES
        chemical_symbols = SEVENNET0_SPECIES
    md = {'chemical_symbols_to_index': ' '.join(chemical_symbols) + ' ', 'cutoff': str(cfg['cutoff']),
          'num_species': str(cfg['num_species']), 'model_type': 'E3_equivariant_model', 'version': 'b200-0.1',
          'dtype': 'single', 'time': datetime.now().strftime('%Y-%m-%d')}
    if not fname.endswith('.pt'):
        fname += '.pt'
    torch.jit.save(scripted, fname, _extra_files=md)
    return fname
