"""Host logic of the fine-tuning loop (sevennet_finetuning_b200/finetune.py) on CPU: loss definitions against the
reference formulas (sevenn/train/loss.py), EWC known answer, collate, DistributedSampler-style sharding, and the
rehearsal iteration under DDP with gloo, world_size 2 (the model there is a stand-in: the B200 modules have no CPU path)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sevennet_finetuning_b200 import finetune as ft


def _graph(n, e, seed):
    r = np.random.default_rng(seed)
    return {'x': torch.tensor(r.integers(0, 89, n)), 'edge_index': torch.tensor(r.integers(0, n, (2, e))),
            'edge_vec': torch.tensor(r.normal(size=(e, 3)), dtype=torch.float32), 'cell_volume': 100.0 + seed, 'num_atoms': n,
            'total_energy': -5.0 * n, 'force_of_atoms': torch.tensor(r.normal(size=(n, 3)), dtype=torch.float32),
            'stress': torch.tensor(r.normal(size=6), dtype=torch.float32)}


def test_collate_offsets_and_keys():
    gs = [_graph(5, 12, 0), _graph(7, 20, 1), _graph(3, 4, 2)]
    b = ft.collate(gs)
    assert b['x'].shape == (15,) and b['edge_index'].shape == (2, 36) and b['edge_vec'].shape == (36, 3)
    assert b['batch'].tolist() == [0] * 5 + [1] * 7 + [2] * 3
    assert b['num_atoms'].tolist() == [5, 7, 3] and b['stress'].shape == (3, 6) and b['total_energy'].shape == (3,)
    # edges of structure k point inside structure k's node range
    lo = torch.tensor([0] * 12 + [5] * 20 + [12] * 4)
    hi = torch.tensor([5] * 12 + [12] * 20 + [15] * 4)
    assert bool(((b['edge_index'] >= lo) & (b['edge_index'] < hi)).all())
    assert torch.equal(b['edge_index'][:, 12:32] - 5, gs[1]['edge_index'])
    assert ft.batch_bytes(b) == sum(v.numel() * v.element_size() for v in b.values() if torch.is_tensor(v))
    assert '_b200_dst_sorted' not in b      # random edge order: no promise made


def test_loss_definitions_match_reference_formulas():
    """PerAtomEnergyLoss / ForceLoss / StressLoss / EWCLoss (sevenn/train/loss.py:96-265) and the weights of
    get_loss_functions_from_config (:268-307): total = E/atom + w_f F + w_s S[kbar] + lambda/2 EWC."""
    torch.manual_seed(0)
    out = {'inferred_total_energy': torch.randn(3), 'total_energy': torch.randn(3), 'num_atoms': torch.tensor([5, 7, 3]),
           'inferred_force': torch.randn(15, 3), 'force_of_atoms': torch.randn(15, 3),
           'inferred_stress': torch.randn(3, 6) * 1e-3, 'stress': torch.randn(3, 6) * 1e-3}
    out['force_of_atoms'][4] = float('nan')     # unlabeled entries are dropped (loss.py:83-90)
    model = torch.nn.Linear(4, 2)
    fisher = {k: torch.rand_like(v) for k, v in model.named_parameters()}
    opt = {k: v.detach() + 0.1 for k, v in model.named_parameters()}
    lf = ft.get_loss_functions('huber', {'delta': 0.01}, 1.0, True, 0.01, fisher, opt, ewc_lambda=1e5, device='cpu')
    assert [type(l).__name__ for l, _ in lf] == ['PerAtomEnergyLoss', 'ForceLoss', 'StressLoss', 'EWCLoss']
    assert [w for _, w in lf] == [1.0, 1.0, 0.01, 5e4]
    hub = torch.nn.HuberLoss(delta=0.01)
    keep = ~torch.isnan(out['force_of_atoms'].reshape(-1))
    want = (hub(out['inferred_total_energy'] / out['num_atoms'], out['total_energy'] / out['num_atoms'])
            + hub(out['inferred_force'].reshape(-1)[keep], out['force_of_atoms'].reshape(-1)[keep])
            + 0.01 * hub(out['inferred_stress'].reshape(-1) * 1602.1766208, out['stress'].reshape(-1) * 1602.1766208)
            + 5e4 * sum((fisher[k] * 0.01).sum() for k in fisher))
    got = ft.total_loss(lf, out, model, 'cpu')
    assert got.shape == (1,) and abs(float(got) - float(want)) <= 1e-6 * abs(float(want))


def test_per_structure_data_weights():
    """load_dataset_with_weights (loss.py:23-93,172-175): mean(criterion * weight), the weight of a structure repeated over
    its components (energy 1, stress 6) or over its atoms' components (forces)."""
    graphs = [_graph(4 + s, 10, s) for s in range(3)]
    for s, g in enumerate(graphs):
        g['data_weight'] = {'per_atom_energy': 1.0 + s, 'force_of_atoms': 0.5 * (s + 1), 'stress': 2.0 - 0.5 * s}
    graphs[1]['stress'] = torch.full((6,), float('nan'))          # an unlabelled stress: dropped together with its weights
    b = ft.to_device(ft.collate(graphs), 'cpu')
    r = np.random.default_rng(1)
    n = int(b['num_atoms'].sum())
    b['inferred_total_energy'] = torch.tensor(r.normal(size=3), dtype=torch.float32)
    b['inferred_force'] = torch.tensor(r.normal(size=(n, 3)), dtype=torch.float32)
    b['inferred_stress'] = torch.tensor(r.normal(size=(3, 6)), dtype=torch.float32)
    lf = ft.get_loss_functions('huber', {'delta': 0.01}, 1.0, True, 0.01, use_weight=True)
    h = torch.nn.HuberLoss(reduction='none', delta=0.01)
    na = b['num_atoms'].float()
    we = torch.tensor([1.0, 2.0, 3.0])
    want_e = (h(b['inferred_total_energy'] / na, b['total_energy'] / na) * we).mean()
    wf = torch.tensor([0.5, 1.0, 1.5])[b['batch']][:, None].expand(n, 3)
    want_f = (h(b['inferred_force'], b['force_of_atoms']) * wf).mean()
    ws = torch.tensor([2.0, 1.5, 1.0])[:, None].expand(3, 6)
    ok = ~torch.isnan(b['stress'])
    want_s = (h(b['inferred_stress'] * ft.TO_KB, b['stress'] * ft.TO_KB)[ok] * ws[ok]).mean()
    got = [float(ld.get_loss(b)) for ld, _ in lf]
    assert np.allclose(got, [float(want_e), float(want_f), float(want_s)], rtol=1e-6)
    with pytest.raises(AssertionError):
        ft.PerAtomEnergyLoss(torch.nn.MSELoss(), use_weight=True)   # weighted losses need reduction='none' (loss.py:27-30)


def test_shard_indices_is_distributed_sampler_split():
    for n, w in [(100, 8), (37, 4), (5, 8)]:
        parts = [ft.shard_indices(n, r, w, seed=3, epoch=1) for r in range(w)]
        assert len({len(p) for p in parts}) == 1 and len(parts[0]) == -(-n // w)
        assert set(sum(parts, [])) == set(range(n))
        from torch.utils.data.distributed import DistributedSampler
        for r in range(w):
            s = DistributedSampler(list(range(n)), num_replicas=w, rank=r, shuffle=True, seed=3)
            s.set_epoch(1)
            assert list(s) == parts[r]


class _Toy(torch.nn.Module):
    """dict in / dict out stand-in with an energy that depends on edge_vec, so that forces need create_graph."""

    def __init__(self):
        super().__init__()
        self.w = torch.nn.Parameter(torch.tensor([0.3, -0.2, 0.5]))
        self.b = torch.nn.Parameter(torch.tensor([0.1]))

    def forward(self, data):
        v = data['edge_vec']
        e_edge = torch.tanh((v * self.w).sum(-1)) + self.b * (v ** 2).sum(-1)
        nb = data['num_atoms'].shape[0]
        eb = data['batch'][data['edge_index'][0]]
        data['inferred_total_energy'] = torch.zeros(nb).index_add(0, eb, e_edge)
        (fij,) = torch.autograd.grad(e_edge.sum(), v, create_graph=True)
        n = data['x'].shape[0]
        data['inferred_force'] = torch.zeros(n, 3).index_add(0, data['edge_index'][0], fij).index_add(0, data['edge_index'][1], -fij)
        vir = torch.cat([v * fij, (v[:, 0] * fij[:, 1])[:, None], (v[:, 1] * fij[:, 2])[:, None], (v[:, 2] * fij[:, 0])[:, None]], 1)
        data['inferred_stress'] = -torch.zeros(nb, 6).index_add(0, eb, vir) / data['cell_volume'][:, None]
        return data


def _losses(model):
    sd = {k: v.detach().clone() for k, v in model.named_parameters()}
    return ft.get_loss_functions('huber', {'delta': 0.01}, 1.0, True, 0.01, {k: torch.ones_like(v) * 1e-3 for k, v in sd.items()},
                                 sd, ewc_lambda=1e5, device='cpu')


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    torch.manual_seed(0)
    model = _Toy()
    tr = ft.RehearsalTrainer(model, _losses(model), 'sgd', {'lr': 0.05}, device='cpu', distributed=True, local_rank=rank)
    graphs = [_graph(6, 15, s) for s in range(8)]
    mem = [_graph(6, 15, 100 + s) for s in range(8)]
    mine = ft.shard_indices(8, rank, world, shuffle=False)
    b = ft.to_device(ft.collate([graphs[i] for i in mine]), 'cpu')
    m = ft.to_device(ft.collate([mem[i] for i in mine]), 'cpu')
    l, lm = tr.run_one_step(b, m)
    q.put((rank, [p.detach().tolist() for p in tr.model.module.parameters()], float(l), float(lm)))   # plain lists: no fd passing
    dist.barrier()
    dist.destroy_process_group()


def test_rehearsal_step_under_ddp_gloo_world2():
    """Two ranks, each with half of the batch: parameters stay identical across ranks and equal a single process that
    averages the two ranks' gradients by hand, for both optimizer steps of the iteration (trainer.py:185-210)."""
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted([q.get(timeout=120) for _ in ps], key=lambda t: t[0])
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    for a, b in zip(res[0][1], res[1][1]):
        assert a == b
    # single-process restatement: average of per-rank gradients, accumulated across the two backward calls
    torch.manual_seed(0)
    model = _Toy()
    lf = _losses(model)
    graphs = [_graph(6, 15, s) for s in range(8)]
    mem = [_graph(6, 15, 100 + s) for s in range(8)]
    opt = torch.optim.SGD(model.parameters(), lr=0.05)
    acc = [torch.zeros_like(p) for p in model.parameters()]
    for src in (graphs, mem):
        gs = []
        for r in range(2):
            d = ft.to_device(ft.collate([src[i] for i in ft.shard_indices(8, r, 2, shuffle=False)]), 'cpu')
            loss = ft.total_loss(lf, model(d), model, 'cpu')
            gs.append(torch.autograd.grad(loss, list(model.parameters())))
        for k, p in enumerate(model.parameters()):
            acc[k] = acc[k] + 0.5 * (gs[0][k] + gs[1][k])
            p.grad = acc[k].clone()
        opt.step()
    for a, b in zip(res[0][1], model.parameters()):
        assert torch.allclose(torch.tensor(a), b.detach(), rtol=1e-5, atol=1e-7), (a, b)
