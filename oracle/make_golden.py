"""TEST INFRASTRUCTURE - generate golden fixtures from the reference's own artefacts.

Run in the build container (where ``/root/reference`` exists):

    python -m oracle.make_golden            # writes tests/golden/* and oracle/_ref/*
    python -m oracle.make_golden --dump     # also prints the frozen graph listing

What it executes is the reference itself: the scripted+frozen SevenNet-0 graphs produced
by ``sevenn/scripts/deploy.py`` and shipped under
``sevenn/pretrained_potentials/SevenNet_0__11July2024/`` (``import sevenn`` is impossible
here: e3nn / ase / torch_geometric are absent).  Outputs:

  tests/golden/sevennet0_weights.npz   state dict under the reference's parameter names
                                       (from example_inputs/fine_tuning/estimate_Fisher/opt_params_sevenn.pt)
  tests/golden/cg_tables.npz           constants c22..c29 (coupling tables) + misc constants
  tests/golden/si216.npz               C1 inputs + reference E / per-atom E / F / stress (+fp64 oracle)
  tests/golden/hfo2_96.npz             triclinic HfO2 (example_inputs/md_serial_example/res.dat) same fields
  tests/golden/si216_segments.npz      per-segment x rows from deployed_parallel_{0..4}.pt
  oracle/_ref/*.pt                     copies of the deployed graphs (git-ignored; travel to
                                       the GPU box) for the timed CPU baseline
"""
import os
import shutil
import sys

import numpy as np
import torch

REF = '/root/reference'
POT = f'{REF}/sevenn/pretrained_potentials/SevenNet_0__11July2024'
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(ROOT, 'tests', 'golden')


def load_weights():
    sd = torch.load(f'{REF}/example_inputs/fine_tuning/estimate_Fisher/opt_params_sevenn.pt',
                    map_location='cpu')
    return {k: v.detach().float().numpy() for k, v in sd.items()}


def read_lammps_data(path):
    lines = open(path).read().splitlines()
    n = int([l for l in lines if l.strip().endswith('atoms')][0].split()[0])
    def pick(tag):
        return [float(t) for t in [l for l in lines if l.strip().endswith(tag)][0].split()[:-len(tag.split())]]
    xlo, xhi = pick('xlo xhi'); ylo, yhi = pick('ylo yhi'); zlo, zhi = pick('zlo zhi')
    xy, xz, yz = pick('xy xz yz')
    cell = np.array([[xhi - xlo, 0, 0], [xy, yhi - ylo, 0], [xz, yz, zhi - zlo]])
    k = [i for i, l in enumerate(lines) if l.strip() == 'Atoms'][0]
    rows = [l.split() for l in lines[k + 1:] if len(l.split()) >= 5][:n]
    types = np.array([int(r[1]) for r in rows])
    pos = np.array([[float(t) for t in r[2:5]] for r in rows])
    return pos, cell, types


def run_serial(ts, pos, cell, types, cutoff=5.0):
    from oracle.graph import radius_graph
    ei, S, D = radius_graph(pos, cell, [True] * 3, cutoff)
    d = {'x': torch.tensor(types), 'pos': torch.tensor(pos, dtype=torch.float32).requires_grad_(True),
         'edge_index': torch.tensor(ei), 'pbc_shift': torch.tensor(S, dtype=torch.float32),
         'cell_lattice_vectors': torch.tensor(cell, dtype=torch.float32),
         'cell_volume': torch.tensor(abs(np.linalg.det(cell)), dtype=torch.float32),
         'num_atoms': torch.tensor(len(pos)), 'batch': torch.zeros(0, dtype=torch.long)}
    o = ts(d)
    return ei, S, D, {k: o[k].detach().numpy() for k in
                      ('inferred_total_energy', 'atomic_energy', 'inferred_force', 'inferred_stress')}


def run_oracle64(w, pos, cell, types, ei, D):
    from oracle.refmodel import build_model
    torch.set_default_dtype(torch.float64)
    m = build_model(dict(shift=w['rescale_atomic_energy.shift'].tolist(),
                         scale=w['rescale_atomic_energy.scale'].tolist(),
                         conv_denominator=float(w['0_convolution.denominator'][0]))).double()
    m.load_state_dict({k: torch.tensor(v) for k, v in w.items()})
    m.eval(); m.set_is_batch_data(False)
    d = {'x': torch.tensor(types), 'edge_index': torch.tensor(ei),
         'edge_vec': torch.tensor(D, dtype=torch.float64).requires_grad_(True),
         'num_atoms': torch.tensor(len(pos)),
         'cell_volume': torch.tensor(abs(np.linalg.det(cell)), dtype=torch.float64),
         'batch': torch.zeros(0, dtype=torch.long)}
    o = m(d)
    torch.set_default_dtype(torch.float32)
    return {k: o[k].detach().numpy() for k in
            ('inferred_total_energy', 'atomic_energy', 'inferred_force', 'inferred_stress')}


def save_case(name, ts, w, pos, cell, types):
    ei, S, D, o = run_serial(ts, pos, cell, types)
    o64 = run_oracle64(w, pos, cell, types, ei, D)
    np.savez_compressed(
        os.path.join(GOLD, name), pos=pos, cell=cell, types=types, n_edges=ei.shape[1],
        edge_index=ei.astype(np.int32), shift=S.astype(np.int8),
        ref_energy=o['inferred_total_energy'], ref_atomic_energy=o['atomic_energy'],
        ref_force=o['inferred_force'], ref_stress=o['inferred_stress'],
        o64_energy=o64['inferred_total_energy'], o64_atomic_energy=o64['atomic_energy'],
        o64_force=o64['inferred_force'], o64_stress=o64['inferred_stress'])
    print(name, 'N', len(pos), 'E', ei.shape[1], 'energy', float(o['inferred_total_energy']),
          '|ref32-o64| F', np.abs(o['inferred_force'] - o64['inferred_force']).max())
    return ei, S, D


def main():
    from sevennet_finetuning_b200.synthetic import SPECIES_INDEX, si_diamond
    os.makedirs(GOLD, exist_ok=True)
    os.makedirs(os.path.join(HERE, '_ref'), exist_ok=True)
    for f in ['serial_model/deployed_serial.pt'] + [f'parallel_model/deployed_parallel_{i}.pt' for i in range(5)]:
        shutil.copyfile(f'{POT}/{f}', os.path.join(HERE, '_ref', os.path.basename(f)))
    ts = torch.jit.load(f'{POT}/serial_model/deployed_serial.pt', map_location='cpu').eval()
    if '--dump' in sys.argv:
        print(ts.code)
    w = load_weights()
    np.savez_compressed(os.path.join(GOLD, 'sevennet0_weights.npz'), **w)
    c = ts.code_with_constants[1].const_mapping
    np.savez_compressed(os.path.join(GOLD, 'cg_tables.npz'),
                        **{f'{a}{b}{d}': c[k].numpy() for k, (a, b, d) in
                           {'c22': (1, 1, 1), 'c23': (1, 1, 2), 'c24': (1, 2, 1), 'c25': (1, 2, 2),
                            'c26': (2, 1, 1), 'c27': (2, 1, 2), 'c28': (2, 2, 1), 'c29': (2, 2, 2)}.items()},
                        silu_2mom=c['c5'].numpy(), bessel_coeffs=c['c0'].numpy(), denominator=c['c9'].numpy())
    # frozen-graph weights == state-dict weights (sanity)
    assert np.allclose(c['c4'].numpy() * np.sqrt(8.0), w['0_convolution.weight_nn.layer0.weight'], atol=1e-6)

    pos, cell, pbc, sym = si_diamond()
    types = np.array([SPECIES_INDEX[s] for s in sym])
    ei, S, D = save_case('si216.npz', ts, w, pos, cell, types)

    hpos, hcell, htypes = read_lammps_data(f'{REF}/example_inputs/md_serial_example/res.dat')
    htypes = np.array([SPECIES_INDEX['Hf'], SPECIES_INDEX['O']])[htypes - 1]
    save_case('hfo2_96.npz', ts, w, hpos, hcell, htypes)

    # per-segment intermediates from the parallel deployment (no ghosts), SURVEY appendix A
    segs = [torch.jit.load(f'{POT}/parallel_model/deployed_parallel_{i}.pt', map_location='cpu').eval()
            for i in range(5)]
    d = {'x': torch.tensor(types), 'x_ghost': torch.zeros(0, dtype=torch.long),
         'edge_index': torch.tensor(ei), 'edge_vec': torch.tensor(D, dtype=torch.float32).requires_grad_(True),
         'num_atoms': torch.tensor([len(pos)]), 'nlocal': torch.tensor([len(pos)])}
    rows = np.arange(0, len(pos), 9)  # 24 atoms
    xs = {}
    out = segs[0](d)
    xs['x_after_seg0'] = out['x'].detach().numpy()[rows]
    for i in range(1, 5):
        out['x_ghost'] = torch.zeros(0, out['x'].shape[1])
        out = segs[i](out)
        if i < 4:
            xs[f'x_after_seg{i}'] = out['x'].detach().numpy()[rows]
    xs['energy'] = out['inferred_total_energy'].detach().numpy()
    xs['atomic_energy'] = out['atomic_energy'].detach().numpy()
    np.savez_compressed(os.path.join(GOLD, 'si216_segments.npz'), rows=rows, **xs)
    print('segments energy', float(xs['energy']))


if __name__ == '__main__':
    main()
