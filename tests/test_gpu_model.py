"""Whole-model E/F/S parity of the C-ABI engine against the reference's frozen-graph outputs
(tests/golden/*.npz, produced by oracle/make_golden.py from deployed_serial.pt) and against
the float64 oracle.  Tolerances (north_star): |dE|/N <= 1e-6 eV/atom; forces: the reference's own
fp32 graph sits 1-2e-5 eV/A from the fp64 evaluation, so max|dF| <= 1e-5 eV/A is asserted against
the fp64 oracle and <= 4e-5 against the fp32 reference; stress rel <= 1e-4 of max|sigma|."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def model(weights):
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from sevennet_finetuning_b200.model import B200SevenNet
    return B200SevenNet(state_dict=weights)


@pytest.mark.parametrize('name', ['si216.npz', 'hfo2_96.npz'])
def test_efs_matches_reference_graph(model, gold, name):
    g = gold(name)
    dev = model.device
    out = model.calculate(torch.tensor(g['pos'], dtype=torch.float64, device=dev), g['cell'], [True] * 3,
                          torch.tensor(g['types'], dtype=torch.int32, device=dev))
    n = len(g['pos'])
    e = float(out['energy'].item())
    f = out['forces'].cpu().numpy()
    s = out['stress'].cpu().numpy()
    ae = out['atomic_energy'].cpu().numpy()
    print(name, 'E', e, 'ref', float(g['ref_energy']), 'o64', float(g['o64_energy']),
          'dF32', np.abs(f - g['ref_force']).max(), 'dF64', np.abs(f - g['o64_force']).max(),
          'dS', np.abs(s - g['o64_stress']).max(), 'Smax', np.abs(g['o64_stress']).max())
    assert abs(e - float(g['o64_energy'])) / n <= 1e-6
    assert abs(e - float(g['ref_energy'])) / n <= 1e-6
    assert np.abs(ae - g['o64_atomic_energy'].reshape(-1)).max() <= 1e-5
    assert np.abs(f - g['o64_force']).max() <= 1e-5
    assert np.abs(f - g['ref_force']).max() <= 4e-5
    assert np.abs(s - g['o64_stress']).max() <= 1e-4 * np.abs(g['o64_stress']).max()
    assert np.abs(s - g['ref_stress']).max() <= 2e-4 * np.abs(g['ref_stress']).max()


def test_efs_fused_engine_matches_reference_graph(weights, gold, monkeypatch):
    """same parity gates with the fused tcgen05 convolution kernels switched on (B200_FUSED=1, csrc/fused.cu)"""
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from sevennet_finetuning_b200.model import B200SevenNet
    monkeypatch.setenv('B200_FUSED', '1')
    m = B200SevenNet(state_dict=weights)
    for name in ('si216.npz', 'hfo2_96.npz'):
        g = gold(name)
        out = m.calculate(torch.tensor(g['pos'], dtype=torch.float64, device=m.device), g['cell'], [True] * 3,
                          torch.tensor(g['types'], dtype=torch.int32, device=m.device))
        n = len(g['pos'])
        e, f, s = float(out['energy'].item()), out['forces'].cpu().numpy(), out['stress'].cpu().numpy()
        assert abs(e - float(g['o64_energy'])) / n <= 1e-6
        assert np.abs(f - g['o64_force']).max() <= 1e-5
        assert np.abs(s - g['o64_stress']).max() <= 1e-4 * np.abs(g['o64_stress']).max()


def test_full_size_properties(model):
    """C3-like cell at reduced size (20k atoms): size-independent properties - net force is zero,
    energy is invariant under a rigid translation and a permutation of the atoms."""
    from sevennet_finetuning_b200.synthetic import SPECIES_INDEX, amorphous_cell
    pos, cell, pbc, sym = amorphous_cell(20000, seed=3)
    types = np.array([SPECIES_INDEX[s] for s in sym], np.int32)
    dev = model.device
    o1 = model.calculate(torch.tensor(pos, device=dev), cell, pbc, torch.tensor(types, device=dev))
    f = o1['forces'].double().cpu().numpy()
    e1 = float(o1['energy'].item())
    assert np.isfinite(f).all() and np.isfinite(e1)
    assert np.abs(f.sum(0)).max() < 2e-2 and np.abs(f).max() > 1e-2
    perm = np.random.default_rng(0).permutation(len(pos))
    o2 = model.calculate(torch.tensor(pos[perm] + np.array([0.37, -1.2, 2.5]), device=dev), cell, pbc,
                         torch.tensor(types[perm], device=dev))
    e2 = float(o2['energy'].item())
    assert abs(e1 - e2) / len(pos) < 1e-6
    assert np.abs(o2['forces'].double().cpu().numpy() - f[perm]).max() < 2e-4
    assert o1['graph']['src'].shape[0] == o2['graph']['src'].shape[0]


def test_cell_order_evaluation_matches_caller_order(model):
    """large systems are evaluated in cell-list atom order; per-atom results come back in the caller's order"""
    from sevennet_finetuning_b200.synthetic import SPECIES_INDEX, amorphous_cell
    pos, cell, pbc, sym = amorphous_cell(6000, seed=5)
    types = np.array([SPECIES_INDEX[s] for s in sym], np.int32)
    dev = model.device
    assert len(pos) >= model.SORT_MIN_ATOMS
    args = (torch.tensor(pos, device=dev), cell, pbc, torch.tensor(types, device=dev))
    a = model.calculate(*args, sort_atoms=True)
    b = model.calculate(*args, sort_atoms=False)
    assert 'perm' in a['graph'] and 'perm' not in b['graph']
    perm = a['graph']['perm']
    assert sorted(perm.tolist()) == list(range(len(pos)))
    # same edge set: map the internal numbering back
    ea = torch.stack([perm[a['graph']['dst'].long()], perm[a['graph']['src'].long()]]).cpu().numpy()
    eb = torch.stack([b['graph']['dst'].long(), b['graph']['src'].long()]).cpu().numpy()
    assert ea.shape == eb.shape
    ka = np.sort(ea[0].astype(np.int64) * len(pos) + ea[1])
    kb = np.sort(eb[0].astype(np.int64) * len(pos) + eb[1])
    assert (ka == kb).all()
    assert abs(float(a['energy']) - float(b['energy'])) / len(pos) <= 1e-6
    assert float((a['atomic_energy'] - b['atomic_energy']).abs().max()) <= 1e-5
    assert float((a['forces'] - b['forces']).abs().max()) <= 6e-5   # two fp32 summation orders, forces up to 10 eV/A
    assert float((a['stress'] - b['stress']).abs().max()) <= 1e-4 * float(b['stress'].abs().max())
