"""sevennet_finetuning_b200/stats.py on CPU: the dataset statistics against literal restatements of the reference's
per-structure loops (sevenn/train/dataset.py:262-395; Ridge from scikit-learn, as the reference uses) and the error
recorder against per-value formulas (sevenn/error_recorder.py:66-232), incl. the world_size-2 reduction over gloo."""
import math
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sevennet_finetuning_b200 import finetune as ft
from sevennet_finetuning_b200 import stats

NSP = 7


def _graph(n, e, seed):
    r = np.random.default_rng(seed)
    x = r.integers(0, NSP - 2, n)             # the last two species never occur
    ei = r.integers(0, n - 1, (2, e))         # atom n-1 has no neighbours
    ei = ei[:, np.argsort(ei[0], kind='stable')]
    ref_e = np.array([-1.0, -3.5, 0.7, -8.0, 2.0, 0.0, 0.0])
    return {'x': torch.tensor(x), 'edge_index': torch.tensor(ei), 'edge_vec': torch.tensor(r.normal(size=(e, 3)), dtype=torch.float32),
            'cell_volume': 100.0 + seed, 'num_atoms': n, 'total_energy': float(ref_e[x].sum() + 0.01 * r.normal()),
            'force_of_atoms': torch.tensor(r.normal(size=(n, 3)) * (1 + x[:, None]), dtype=torch.float32),
            'stress': torch.tensor(r.normal(size=6), dtype=torch.float32)}


def test_dataset_statistics_match_reference_loops():
    from sklearn.linear_model import Ridge
    graphs = [_graph(5 + s % 4, 20 + s, s) for s in range(40)]
    st = stats.DatasetStatistics(graphs, NSP)
    # dataset.py:344-354
    n_neigh = []
    for g in graphs:
        n_neigh.extend(np.unique(g['edge_index'][0].numpy(), return_counts=True)[1])
    assert abs(st.get_avg_num_neigh() - np.average(n_neigh)) < 1e-12
    # dataset.py:311-322
    fl = []
    for g in graphs:
        fl.extend(g['force_of_atoms'].reshape(-1).tolist())
    assert abs(st.get_force_rms() - float(torch.sqrt(torch.mean(torch.pow(torch.Tensor(fl), 2))))) < 1e-6
    # dataset.py:324-342
    atomx = torch.concat([g['x'] for g in graphs])
    force = torch.concat([g['force_of_atoms'] for g in graphs])
    rms = torch.zeros((NSP, 3))
    rms.scatter_reduce_(0, atomx.repeat_interleave(3, 0).reshape(force.shape), force.square(), reduce='mean', include_self=False)
    assert torch.allclose(st.get_species_wise_force_rms(), torch.sqrt(rms.mean(dim=1)), atol=1e-6)
    # dataset.py:262-276
    want = float(torch.mean(torch.Tensor([g['total_energy'] / g['num_atoms'] for g in graphs])))
    assert abs(st.get_per_atom_energy_mean() - want) < 1e-6
    # dataset.py:278-309
    c = np.zeros((len(graphs), NSP))
    for i, g in enumerate(graphs):
        c[i] = np.bincount(g['x'].numpy(), minlength=NSP)
    y = np.array([g['total_energy'] for g in graphs], dtype=np.float32)
    zero = np.all(c == 0, axis=0)
    full = np.zeros(NSP)
    full[~zero] = Ridge(alpha=0.1, fit_intercept=False).fit(c[:, ~zero], y).coef_
    got = st.get_species_ref_energy_by_linear_comb()
    assert np.abs(got - full).max() < 1e-4 and np.all(got[-2:] == 0.0)
    assert np.abs(got[:5] - np.array([-1.0, -3.5, 0.7, -8.0, 2.0])).max() < 0.05
    # dataset.py:357-395
    v = torch.cat([g['force_of_atoms'].reshape(-1) for g in graphs])
    s = st.get_statistics('force_of_atoms')
    assert abs(s['mean'] - float(v.mean())) < 1e-6 and abs(s['std'] - float(v.std())) < 1e-5
    assert abs(s['median'] - float(v.median())) < 1e-6 and s['max'] == float(v.max()) and s['min'] == float(v.min())
    sh, sc = st.shift_scale('elemwise_reference_energies', 'elemwise_force_rms')
    assert len(sh) == NSP and len(sc) == NSP


def _fake_output(seed, n_struct=3):
    r = np.random.default_rng(seed)
    na = r.integers(3, 7, n_struct)
    n = int(na.sum())
    out = {'num_atoms': torch.tensor(na), 'total_energy': torch.tensor(r.normal(size=n_struct), dtype=torch.float32),
           'inferred_total_energy': torch.tensor(r.normal(size=n_struct), dtype=torch.float32),
           'force_of_atoms': torch.tensor(r.normal(size=(n, 3)), dtype=torch.float32),
           'inferred_force': torch.tensor(r.normal(size=(n, 3)), dtype=torch.float32),
           'stress': torch.tensor(r.normal(size=(n_struct, 6)), dtype=torch.float32),
           'inferred_stress': torch.tensor(r.normal(size=(n_struct, 6)), dtype=torch.float32)}
    out['stress'][0] = float('nan')          # an unlabelled structure (error_recorder.py:118-121)
    return out


def _naive(outs):
    """every metric of the default list over all batches, by hand"""
    e2, f2, s2, fm, fv, sc = [], [], [], [], [], []
    for o in outs:
        e2 += (((o['total_energy'] - o['inferred_total_energy']) / o['num_atoms']) ** 2).tolist()
        d = (o['force_of_atoms'] - o['inferred_force']).double()
        f2 += (d ** 2).sum(1).tolist()
        fm += d.abs().reshape(-1).tolist()
        fv += (d ** 2).sum(1).sqrt().tolist()
        ds = ((o['stress'] - o['inferred_stress']).double() * ft.TO_KB)
        ok = ~torch.isnan(ds).any(1)
        s2 += (ds[ok] ** 2).sum(1).tolist()
        sc += (ds[ok] ** 2).reshape(-1).tolist()
    return {'Energy_RMSE (eV/atom)': math.sqrt(np.mean(e2)), 'Force_RMSE (eV/A)': math.sqrt(np.mean(f2)),
            'Stress_RMSE (kbar)': math.sqrt(np.mean(s2)), 'Force_MAE (eV/A)': float(np.mean(fm)),
            'Force_VectorMAE (eV/A)': float(np.mean(fv)), 'Stress_ComponentRMSE (kbar)': math.sqrt(np.mean(sc))}


_METRICS = (('Energy', 'RMSE'), ('Force', 'RMSE'), ('Stress', 'RMSE'), ('Force', 'MAE'), ('Force', 'VectorMAE'),
            ('Stress', 'ComponentRMSE'))


def test_error_recorder_single_process():
    outs = [_fake_output(s) for s in range(4)]
    rec = stats.ErrorRecorder(_METRICS)
    for o in outs:
        rec.update(o)
    got, want = rec.epoch_forward(), _naive(outs)
    assert set(got) == set(want)
    for k in want:
        assert abs(got[k] - want[k]) <= 1e-6 * max(1.0, abs(want[k])), k
    assert all(math.isnan(v) for v in rec.epoch_forward().values())      # reset after an epoch


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    rec = stats.ErrorRecorder(_METRICS)
    for s in range(4):
        if s % world == rank:
            rec.update(_fake_output(s))
    q.put((rank, rec.epoch_forward()))
    dist.barrier()
    dist.destroy_process_group()


def test_error_recorder_reduces_over_ranks_gloo_world2():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = dict(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = _naive([_fake_output(s) for s in range(4)])
    for r in range(2):
        for k in want:
            assert abs(res[r][k] - want[k]) <= 1e-6 * max(1.0, abs(want[k])), (r, k)
