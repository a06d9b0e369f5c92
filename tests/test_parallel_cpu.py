"""Host-side bookkeeping of the spatially decomposed path (parallel.py) on CPU: ownership, ghost shells,
forward fetch and reverse add - with the in-process transport (4 bricks) and with torch.distributed gloo
(world_size 2).  The checker for "every neighbour is present" is the float64 oracle graph."""
import os
import threading

import numpy as np
import torch

from sevennet_finetuning_b200.parallel import DistComm, HaloPlan, ThreadComm, brick_grid


def _cell(n=600, seed=5, tric=True):
    rng = np.random.default_rng(seed)
    cell = np.diag([21.0, 17.0, 13.0])
    if tric:
        cell = cell + np.array([[0, 0, 0], [2.5, 0, 0], [-1.5, 3.0, 0]])
    pos = rng.random((n, 3)) @ cell + rng.normal(0, 3.0, (n, 3))   # some atoms outside the cell: wrapping is exercised
    return pos, cell, [True, True, True]


def _check_rank(plan, pos, cell, pbc, results, rank):
    from oracle.graph import radius_graph
    ei, S, D = radius_graph(pos, cell, pbc, 5.0)
    loc = set(plan.local_ids.tolist())
    sub = set(plan.subset.tolist())
    assert len(sub) == plan.n_local + plan.n_ghost
    need = set(ei[1][np.isin(ei[0], list(loc))].tolist())
    assert need <= sub, 'a neighbour of a local atom is missing from the ghost shell'
    # forward fetch: rows carry a function of the global id
    n = plan.n_local + plan.n_ghost
    rows = torch.zeros(n, 3, dtype=torch.float64)
    gid = plan.local_ids.double()
    rows[:plan.n_local] = torch.stack([gid, gid * 2 + 1, torch.sin(gid)], 1)
    plan.fetch_ghost_rows(rows)
    gg = plan.ghost_ids.double()
    assert torch.equal(rows[plan.n_local:], torch.stack([gg, gg * 2 + 1, torch.sin(gg)], 1))
    # reverse add: every ghost row holds (1, gid); an owner ends with (#ranks that see it as a ghost)
    back = torch.zeros(n, 2, dtype=torch.float64)
    back[plan.n_local:, 0] = 1.0
    back[plan.n_local:, 1] = gg
    plan.return_ghost_rows(back)
    results[rank] = (plan.local_ids.clone(), back[:plan.n_local].clone(), plan.ghost_ids.clone())


def _verify_global(results, n_atoms):
    owners = torch.cat([r[0] for r in results])
    assert sorted(owners.tolist()) == list(range(n_atoms)), 'bricks do not partition the atoms'
    seen = torch.zeros(n_atoms)
    for r in results:
        seen.index_add_(0, r[2], torch.ones(len(r[2])))
    for ids, back, _ in results:
        assert torch.equal(back[:, 0], seen[ids].double())
        assert torch.equal(back[:, 1], (seen[ids] * ids).double())


def test_brick_grid():
    assert brick_grid(1) == (1, 1, 1) and brick_grid(2) == (2, 1, 1)
    assert brick_grid(4) == (2, 2, 1) and brick_grid(8) == (2, 2, 2)


def test_halo_plan_threads_four_bricks():
    pos, cell, pbc = _cell()
    world, grid = 4, (2, 2, 1)
    comms = ThreadComm.create(world)
    results, errs = [None] * world, []

    def run(r):
        try:
            plan = HaloPlan(torch.tensor(pos), cell, pbc, 5.0, grid, r, comms[r]).connect()
            _check_rank(plan, pos, cell, pbc, results, r)
        except Exception as ex:   # pragma: no cover
            errs.append(ex)
            comms[r].sh.barrier.abort()
    th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    _verify_global(results, len(pos))


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        pos, cell, pbc = _cell(tric=False)
        plan = HaloPlan(torch.tensor(pos), cell, pbc, 5.0, brick_grid(world), rank, DistComm()).connect()
        res = [None] * world
        _check_rank(plan, pos, cell, pbc, res, rank)
        q.put((rank, res[rank]))
    finally:
        dist.destroy_process_group()


def test_halo_plan_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    ps = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    got = dict(q.get(timeout=240) for _ in range(2))
    [p.join(timeout=60) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    _verify_global([got[0], got[1]], 600)
