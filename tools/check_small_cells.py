"""|dE|/N and max|dF| of the engine (tabulated radial weights vs tensor-core MLP) against the float64 oracle over random tiny
triclinic cells (the self-image edge case of tests/test_gpu_model.py), to tell rounding noise from a systematic offset."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np, torch
from sevennet_finetuning_b200.model import B200SevenNet
import test_gpu_model as tm

w = dict(np.load(os.path.join(ROOT, 'tests/golden/sevennet0_weights.npz')))
dev = torch.device('cuda:0')
models = {}
for mode in ('1', '0'):
    os.environ['B200_RADIAL_TABLE'] = mode
    models[mode] = B200SevenNet(state_dict=w, device=dev)
res = {'1': [], '0': []}
for seed in range(24):
    rng = np.random.default_rng(seed)
    cell = np.array([[3.3, 0.0, 0.0], [0.4, 3.1, 0.0], [-0.3, 0.5, 3.6]]) * rng.uniform(0.95, 1.3)
    pos = np.array([[0.1, 0.2, 0.0], [1.7, 1.5, 1.9]]) + rng.normal(scale=0.08, size=(2, 3))
    types = rng.integers(0, 89, 2)
    ei, e64, f64, ae64 = tm._oracle_efs(w, pos, cell, [True] * 3, types)
    for mode, m in models.items():
        out = m.calculate(torch.tensor(pos, device=dev), cell, [True] * 3, torch.tensor(types, dtype=torch.int32, device=dev))
        res[mode].append(((float(out['energy']) - e64) / 2, np.abs(out['forces'].cpu().numpy() - f64).max(), abs(e64)))
for mode in ('1', '0'):
    a = np.array(res[mode])
    print('table' if mode == '1' else 'mlp  ', 'dE/N: mean %.2e rms %.2e max %.2e | max dF %.2e | over 1e-6: %d of %d' % (
        a[:, 0].mean(), np.sqrt((a[:, 0] ** 2).mean()), np.abs(a[:, 0]).max(), a[:, 1].max(), (np.abs(a[:, 0]) > 1e-6).sum(), len(a)))
    print('   ', ' '.join('%+.1e' % v for v in a[:, 0]))
