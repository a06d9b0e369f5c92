"""The sevenn.nn twins (sevennet_finetuning_b200/nn.py): same module names / state dict as the reference, dict-in
dict-out forward, energy/forces/stress parity with the golden vectors, and first-order autograd (parameter
gradients of the energy) against the float64 oracle model."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _data(g, dev, shuffle=False):
    from oracle.graph import radius_graph
    ei, S, D = radius_graph(g['pos'], g['cell'], [True] * 3, 5.0)
    if shuffle:   # the reference does not promise any edge order
        p = np.random.default_rng(0).permutation(ei.shape[1])
        ei, D = ei[:, p], D[p]
    return {'x': torch.tensor(g['types'], dtype=torch.long, device=dev), 'edge_index': torch.tensor(ei, device=dev),
            'edge_vec': torch.tensor(D, dtype=torch.float32, device=dev).requires_grad_(True),
            'num_atoms': torch.tensor(len(g['pos']), device=dev),
            'cell_volume': torch.tensor(abs(np.linalg.det(g['cell'])), dtype=torch.float32, device=dev),
            'batch': torch.zeros(len(g['pos']), dtype=torch.long, device=dev)}, ei, D


def _oracle(weights):
    from oracle.refmodel import build_model
    m = build_model(dict(shift=weights['rescale_atomic_energy.shift'].tolist(), scale=weights['rescale_atomic_energy.scale'].tolist(),
                         conv_denominator=float(weights['0_convolution.denominator'][0])))
    m.load_state_dict({k: torch.tensor(v) for k, v in weights.items()})
    return m.double()


@pytest.fixture(scope='module')
def twin(weights):
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from sevennet_finetuning_b200 import nn as b200nn
    m = b200nn.build_E3_equivariant_model(dict(shift=weights['rescale_atomic_energy.shift'].tolist(),
                                               scale=weights['rescale_atomic_energy.scale'].tolist(),
                                               conv_denominator=float(weights['0_convolution.denominator'][0])))
    missing = m.load_state_dict({k: torch.tensor(v) for k, v in weights.items()}, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return m.cuda()


@pytest.mark.parametrize('name,shuffle', [('si216.npz', False), ('hfo2_96.npz', True)])
def test_module_chain_efs(twin, gold, name, shuffle):
    g = gold(name)
    dev = torch.device('cuda:0')
    twin.eval()
    twin.set_is_batch_data(False)
    data, ei, D = _data(g, dev, shuffle)
    out = twin(data)
    n = len(g['pos'])
    e = float(out['inferred_total_energy'])
    f = out['inferred_force'].detach().cpu().numpy()
    s = out['inferred_stress'].detach().cpu().numpy()
    assert abs(e - float(g['o64_energy'])) / n <= 1e-6
    assert np.abs(out['atomic_energy'].detach().cpu().numpy().reshape(-1) - g['o64_atomic_energy'].reshape(-1)).max() <= 1e-5
    assert np.abs(f - g['o64_force']).max() <= 1e-5
    assert np.abs(s - g['o64_stress']).max() <= 1e-4 * np.abs(g['o64_stress']).max()
    assert 'self_cont_tmp' not in out and '_b200_graph' not in out
    # batched flavour of the same call
    twin.set_is_batch_data(True)
    data, _, _ = _data(g, dev, shuffle)
    data['cell_volume'] = data['cell_volume'].view(1)
    out = twin(data)
    assert out['inferred_total_energy'].shape == (1,) and out['inferred_stress'].shape == (1, 6)
    assert abs(float(out['inferred_total_energy'][0]) - e) / n <= 2e-6   # fp32 sum order (1 ulp of 1158 eV = 1.2e-4)


def test_parameter_gradients_of_energy(twin, gold, weights):
    """d E_total / d theta for every parameter, against autograd through the float64 oracle model"""
    g = gold('hfo2_96.npz')
    dev = torch.device('cuda:0')
    twin.train()
    twin.set_is_batch_data(False)
    for p in twin.parameters():
        p.requires_grad_(True)
        p.grad = None
    data, ei, D = _data(g, dev)
    chain = [m for k, m in twin._modules.items() if k != 'force_output']
    for m in chain:
        data = m(data)
    data['inferred_total_energy'].backward()
    torch.set_default_dtype(torch.float64)   # the oracle builds its tables and one-hot rows in the default dtype
    try:
        ref = _oracle(weights)
        ref.set_is_batch_data(False)
        for p in ref.parameters():
            p.requires_grad_(True)
        d = {'x': torch.tensor(g['types']), 'edge_index': torch.tensor(ei),
             'edge_vec': torch.tensor(D, dtype=torch.float64).requires_grad_(True), 'num_atoms': torch.tensor(len(g['pos'])),
             'cell_volume': torch.tensor(abs(np.linalg.det(g['cell']))), 'batch': torch.zeros(0, dtype=torch.long)}
        for k, m in ref._modules.items():
            if k != 'force_output':
                d = m(d)
        d['inferred_total_energy'].backward()
    finally:
        torch.set_default_dtype(torch.float32)
    refg = dict(ref.named_parameters())
    worst = 0.0
    for k, p in twin.named_parameters():
        assert p.grad is not None, k
        a, b = p.grad.double().cpu().reshape(-1), refg[k].grad.reshape(-1)
        rel = float((a - b).abs().max() / max(float(b.abs().max()), 1e-12))
        worst = max(worst, rel)
        assert rel < 5e-4, (k, rel)
    print('worst relative parameter-gradient error', worst)
    twin.eval()


def test_force_output_refuses_second_order(twin, gold):
    g = gold('si216.npz')
    twin.train()
    twin.set_is_batch_data(False)
    data, _, _ = _data(g, torch.device('cuda:0'))
    with pytest.raises(NotImplementedError):
        twin(data)
    twin.eval()
