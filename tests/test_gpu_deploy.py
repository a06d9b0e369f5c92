"""TorchScript deployment (sevennet_finetuning_b200/deploy.py, csrc/torch_ops.cpp): the saved file runs in a process that has
only torch and the operator library - what a libtorch caller (PairE3GNN::compute) has - and reproduces the golden vectors."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_CHILD = r'''
import json, sys
import numpy as np
import torch
lib, pt, npz = sys.argv[1:4]
assert 'sevennet_finetuning_b200' not in sys.modules
torch.ops.load_library(lib)
extra = {'chemical_symbols_to_index': '', 'cutoff': '', 'num_species': ''}
m = torch.jit.load(pt, map_location='cuda:0', _extra_files=extra)
g = np.load(npz)
dev = torch.device('cuda:0')
res = {}
for shuffle in (False, True):
    ei, sh = g['edge_index'], g['shift']
    if shuffle:
        p = np.random.default_rng(0).permutation(ei.shape[1])
        ei, sh = ei[:, p], sh[p]
    # the dict of pair_e3gnn.cpp:205-212
    d = {'x': torch.tensor(g['types'], dtype=torch.long, device=dev),
         'pos': torch.tensor(g['pos'], dtype=torch.float32, device=dev).requires_grad_(True),
         'edge_index': torch.tensor(ei, dtype=torch.long, device=dev), 'pbc_shift': torch.tensor(sh, dtype=torch.float32, device=dev),
         'num_atoms': torch.tensor(len(g['pos'])), 'cell_lattice_vectors': torch.tensor(g['cell'], dtype=torch.float32, device=dev),
         'cell_volume': torch.tensor(abs(np.linalg.det(g['cell'])), dtype=torch.float32, device=dev)}
    o = m(d)
    res[str(shuffle)] = dict(e=float(o['inferred_total_energy'].cpu()), f=o['inferred_force'].cpu().numpy().tolist(),
                             s=o['inferred_stress'].cpu().numpy().tolist(), ae=o['atomic_energy'].cpu().numpy().reshape(-1).tolist())
res['extra'] = {k: (v.decode() if isinstance(v, bytes) else v) for k, v in extra.items()}
print('RESULT' + json.dumps(res))
'''


def test_deployed_file_runs_without_the_python_package(weights, tmp_path):
    from sevennet_finetuning_b200 import deploy as dp
    lib = dp.ops_library()
    pt = dp.deploy(weights, None, str(tmp_path / 'deployed_serial_b200'))
    assert pt.endswith('.pt') and os.path.getsize(pt) > 3_000_000
    for name in ('si216.npz', 'hfo2_96.npz'):
        npz = os.path.join(ROOT, 'tests', 'golden', name)
        r = subprocess.run([sys.executable, '-c', _CHILD, lib, pt, npz], capture_output=True, text=True, timeout=600,
                           cwd=str(tmp_path), env={k: v for k, v in os.environ.items() if k != 'PYTHONPATH'})
        assert r.returncode == 0, r.stderr[-3000:]
        res = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith('RESULT')][0][6:])
        g = np.load(npz)
        n = len(g['pos'])
        for key in ('False', 'True'):
            o = res[key]
            etol = max(1e-6, 3.0 * float(np.spacing(np.float32(abs(o["e"])))) / n)
            assert abs(o['e'] - float(g['ref_energy'])) / n <= etol and abs(o['e'] - float(g['o64_energy'])) / n <= etol
            # fp32 positions in the dict (as LAMMPS passes them): 4e-5 eV/A, see test_deployed_serial_form_matches_golden
            assert np.abs(np.array(o['f']) - g['ref_force']).max() <= 4e-5 and np.abs(np.array(o['f']) - g['o64_force']).max() <= 4e-5
            assert np.abs(np.array(o['s']) - g['o64_stress']).max() <= 1e-4 * np.abs(g['o64_stress']).max()
            assert np.abs(np.array(o['ae']) - g['o64_atomic_energy'].reshape(-1)).max() <= 2e-5
        assert res['extra']['cutoff'] == '5.0' and res['extra']['num_species'] == '89'
        assert res['extra']['chemical_symbols_to_index'].split()[:3] == ['Ac', 'Ag', 'Al']
