"""The oracle (oracle/) against the reference's own frozen-graph outputs (tests/golden/)."""
import numpy as np
import torch

from oracle import wigner
from oracle.graph import radius_graph, radius_graph_bruteforce
from oracle.refmodel import build_model


def _model(weights, dtype=torch.float32):
    m = build_model(dict(shift=weights['rescale_atomic_energy.shift'].tolist(),
                         scale=weights['rescale_atomic_energy.scale'].tolist(),
                         conv_denominator=float(weights['0_convolution.denominator'][0])))
    m.load_state_dict({k: torch.tensor(v) for k, v in weights.items()})
    m = m.to(dtype).eval()
    m.set_is_batch_data(False)
    return m


def test_w3j_matches_frozen_graph_tables(gold):
    g = gold('cg_tables.npz')
    for key in ['111', '112', '121', '122', '211', '212', '221', '222']:
        l1, l2, l3 = (int(c) for c in key)
        C = wigner.w3j(l1, l2, l3) * np.sqrt(2 * l3 + 1)
        assert np.abs(C - g[key]).max() < 1e-6, key


def test_w3j_trivial_paths():
    # (0,l,l), (l,0,l) are identity/sqrt(2l+1); (l,l,0) likewise: SURVEY appendix B
    for l in (0, 1, 2):
        d = 2 * l + 1
        assert np.allclose(wigner.w3j(0, l, l)[0] * np.sqrt(d), np.eye(d), atol=1e-9)
        assert np.allclose(wigner.w3j(l, 0, l)[:, 0] * np.sqrt(d), np.eye(d), atol=1e-9)
        assert np.allclose(wigner.w3j(l, l, 0)[:, :, 0] * np.sqrt(d), np.eye(d), atol=1e-9)


def test_radius_graph_matches_golden_edge_sets(gold):
    for name in ('si216.npz', 'hfo2_96.npz'):
        g = gold(name)
        ei, S, D = radius_graph(g['pos'], g['cell'], [True] * 3, 5.0)
        assert ei.shape[1] == int(g['n_edges'])
        assert (ei == g['edge_index']).all() and (S == g['shift']).all()


def test_radius_graph_celllist_equals_bruteforce():
    rng = np.random.default_rng(3)
    cell = np.array([[7.0, 0, 0], [1.5, 6.5, 0], [-0.8, 1.1, 8.0]])
    pos = rng.uniform(-3, 12, (40, 3))  # deliberately not wrapped
    for pbc in ([True] * 3, [True, False, True], [False] * 3):
        a = radius_graph(pos, cell, pbc, 4.0)
        b = radius_graph_bruteforce(pos, cell, pbc, 4.0)
        assert a[0].shape == b[0].shape and (a[0] == b[0]).all() and (a[1] == b[1]).all()
        assert np.abs(a[2] - b[2]).max() == 0.0


def test_radius_graph_small_cell_self_images():
    # one atom in a 3 A cubic cell, cutoff 5: periodic self images only
    ei, S, D = radius_graph(np.zeros((1, 3)), np.eye(3) * 3.0, [True] * 3, 5.0)
    assert ei.shape[1] == 18 and (ei == 0).all() and not (S == 0).all(1).any()
    ei, S, D = radius_graph(np.zeros((0, 3)), np.eye(3) * 3.0, [True] * 3, 5.0)
    assert ei.shape == (2, 0)


def _run(m, g, dtype):
    ei, S, D = radius_graph(g['pos'], g['cell'], [True] * 3, 5.0)
    d = {'x': torch.tensor(g['types']), 'edge_index': torch.tensor(ei),
         'edge_vec': torch.tensor(D, dtype=dtype).requires_grad_(True),
         'num_atoms': torch.tensor(len(g['pos'])),
         'cell_volume': torch.tensor(abs(np.linalg.det(g['cell'])), dtype=dtype),
         'batch': torch.zeros(0, dtype=torch.long)}
    return m(d)


def test_oracle_model_matches_frozen_graph(weights, gold):
    m = _model(weights)
    for name in ('si216.npz', 'hfo2_96.npz'):
        g = gold(name)
        o = _run(m, g, torch.float32)
        n = len(g['pos'])
        assert abs(o['inferred_total_energy'].item() - float(g['ref_energy'])) / n <= 1e-6
        assert np.abs(o['atomic_energy'].detach().numpy() - g['ref_atomic_energy']).max() <= 1e-5
        # the frozen fp32 graph itself sits 1-2e-5 eV/A from the fp64 evaluation
        assert np.abs(o['inferred_force'].detach().numpy() - g['ref_force']).max() <= 4e-5
        s = o['inferred_stress'].detach().numpy()
        assert np.abs(s - g['ref_stress']).max() <= 1e-4 * np.abs(g['ref_stress']).max()


def test_oracle_segment_intermediates(weights, gold):
    """Layer-boundary features vs deployed_parallel_{0..3}.pt (post-gate, post next si1)."""
    g, seg = gold('si216.npz'), gold('si216_segments.npz')
    m = _model(weights)
    ei, S, D = radius_graph(g['pos'], g['cell'], [True] * 3, 5.0)
    d = {'x': torch.tensor(g['types']), 'edge_index': torch.tensor(ei),
         'edge_vec': torch.tensor(D, dtype=torch.float32), 'num_atoms': torch.tensor(216),
         'cell_volume': torch.tensor(1.0), 'batch': torch.zeros(0, dtype=torch.long)}
    rows = seg['rows']
    with torch.no_grad():
        for name, mod in m.named_children():
            if name == 'force_output':
                break
            d = mod(d)
            for t in range(1, 5):
                if name == f'{t}_self_interaction_1':
                    ref = seg[f'x_after_seg{t - 1}']
                    got = d['x'].numpy()[rows]
                    assert np.abs(got - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max()), name


def test_nequip_self_connection_matches_frozen_graph(gold):
    """oracle SelfConnectionIntro (FullyConnectedTensorProduct with the one-hot species vector, self_connection.py:11-38)
    against the reference example model's own frozen-graph arithmetic and weights (oracle/make_golden_fctp.py)."""
    from oracle.refnn import SelfConnectionIntro
    g = gold('fctp_hfo.npz')
    m = SelfConnectionIntro('4x0e+4x1e', '2x0e', '12x0e+4x1e')
    with torch.no_grad():
        m.fc_tensor_product.weight.copy_(torch.tensor(np.concatenate([g['w0'].reshape(-1), g['w1'].reshape(-1)])))
    onehot = torch.nn.functional.one_hot(torch.tensor(g['types']), 2)
    out = m({'x': torch.tensor(g['x']), 'node_attr': onehot})['self_cont_tmp']
    assert out.shape == (len(g['x']), 24)
    assert np.abs(out.detach().numpy() - g['out']).max() <= 1e-6
