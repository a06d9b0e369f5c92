#!/usr/bin/env python
"""bench.py - SevenNet-0 energy+forces+stress atom-steps/s on B200 (BASELINE.json metric).

Workload at N=1: config 3 of BASELINE.json - one MD force call (radius graph + E + F + virial) on a
100,000-atom synthetic amorphous cell (~5M edges), SevenNet-0 weights, fp32.  A "step" is one full
force call with the graph rebuilt.  N>1: config 4 - ONE cell of N x 131,072 atoms (1,048,576 at N=8) cut
into N bricks (2x1x1 / 2x2x1 / 2x2x2), one rank per brick, ghost features exchanged per interaction layer
over NCCL (sevennet_finetuning_b200/parallel.py); weak scaling (fixed atoms per GPU), value = atoms of the
whole cell x steps / max-over-ranks time.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--atoms A] [--impl reference] [--workload md|finetune]

`--workload finetune` measures the other half of the metric: config 2 (N = 1) / config 5 (N > 1, DDP over NCCL),
structures/s of the reference's rehearsal + EWC iteration (sevenn/train/trainer.py:173-210) over the B200 module twins.
The default MD line at N = 1 also carries that measurement under the key `finetune`.

One JSON line on stdout (rank 0).  `value` = device-resident inputs; `e2e` = host buffers in, host
results out, through the public API (B200SevenNet.calculate).  `--impl reference` times the
reference's own frozen TorchScript graph (oracle/_ref/deployed_serial.pt) on the host CPU cores on a
bounded sample of the same workload.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = 'SevenNet-0 E+F+S atom-steps/s'
UNIT = 'atom-steps/s'
CAT_NAMES = ['edge_embed+csr', 'radial_mlp_fwd', 'conv_fwd', 'node_linear_gate', 'radial_mlp_bwd', 'conv_bwd_edge',
             'conv_bwd_x', 'force_virial', 'halo_exchange']


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        return dict(hbm=float(p['hbm_gbs']), tf=float(p['bf16_tflops_sustained']), src='measured')
    except Exception:
        return dict(hbm=6650.0, tf=1400.0, src='fallback')


_JSON_OUT = None


class ClockSampler:
    """SM clock and throttle reasons during the timed region (B200_PROFILING.md recipe), read through NVML in-process (the
    counters nvidia-smi prints; spawning nvidia-smi from a process that maps >100 GB of device memory stalls the launching
    thread for tens of milliseconds per sample).  One sample per timed step, taken by the MAIN thread right after it has
    queued the step - the GPU is executing that step then.  A sampling thread beside the launching thread was the first
    version: its NVML queries collided with the kernel launches in the driver and cost single timed steps 10-300 ms
    (2 runs of 6, always the second step; none of 6 without the thread)."""

    def __init__(self, index):
        self.rows, self.index = [], index
        self.nv = None
        try:
            import pynvml as nv
            nv.nvmlInit()
            self.h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.mx = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
            self.get_reasons = getattr(nv, 'nvmlDeviceGetCurrentClocksEventReasons', None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
            self.get_reasons(self.h)
            self.bits = [(getattr(nv, 'nvmlClocksEventReasonHwSlowdown', 0x8), 0), (getattr(nv, 'nvmlClocksEventReasonHwThermalSlowdown', 0x40), 1),
                         (getattr(nv, 'nvmlClocksEventReasonSwThermalSlowdown', 0x20), 2), (getattr(nv, 'nvmlClocksEventReasonSwPowerCap', 0x4), 3)]
            self.nv = nv
        except Exception:
            self.nv = None

    def sample(self):
        if self.nv is not None:
            try:
                r = self.get_reasons(self.h)
                flags = ['Not Active'] * 4
                for bit, k in self.bits:
                    if r & bit:
                        flags[k] = 'Active'
                self.rows.append([str(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)), str(self.mx)] + flags)
                return
            except Exception:
                self.nv = None
        q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')
        try:
            out = subprocess.run(['nvidia-smi', f'--id={self.index}', f'--query-gpu={q}', '--format=csv,noheader,nounits'],
                                 capture_output=True, text=True, timeout=5).stdout.strip()
            if out:
                self.rows.append([t.strip() for t in out.split(',')])
        except Exception:
            pass

    def summary(self):
        if not self.rows:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['unavailable'])
        sm = sorted(float(r[0]) for r in self.rows)
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for k, n in enumerate(names) if any(r[2 + k].lower().startswith('active') for r in self.rows)]
        return dict(sm_mhz=sm[len(sm) // 2], sm_max_mhz=float(self.rows[0][1]), reasons=reasons, samples=len(sm))


def make_cell(n_atoms, seed=0):
    from sevennet_finetuning_b200.synthetic import SPECIES_INDEX, amorphous_cell
    pos, cell, pbc, sym = amorphous_cell(n_atoms, seed=seed)
    types = np.array([SPECIES_INDEX[s] for s in sym], np.int32)
    return pos, cell, pbc, types


def load_weights():
    return dict(np.load(os.path.join(ROOT, 'tests', 'golden', 'sevennet0_weights.npz')))


def _reference_callable(pos, cell, pbc, types, device='cpu'):
    """The reference's own implementation of the path: its frozen TorchScript graph (oracle/_ref/deployed_serial.pt,
    kind 'reference'), else the oracle restatement (kind 'port').  Returns (call, kind); call() -> (E, F, S, n_edges, ei)."""
    from oracle.graph import radius_graph
    path = os.path.join(ROOT, 'oracle', '_ref', 'deployed_serial.pt')
    if os.path.exists(path):
        ts = torch.jit.load(path, map_location=device).eval()

        def call():
            ei, S, D = radius_graph(pos, cell, pbc, 5.0)
            d = {'x': torch.tensor(types, dtype=torch.long, device=device),
                 'pos': torch.tensor(pos, dtype=torch.float32, device=device).requires_grad_(True),
                 'edge_index': torch.tensor(ei, device=device), 'pbc_shift': torch.tensor(S, dtype=torch.float32, device=device),
                 'cell_lattice_vectors': torch.tensor(cell, dtype=torch.float32, device=device),
                 'cell_volume': torch.tensor(abs(np.linalg.det(cell)), dtype=torch.float32, device=device),
                 'num_atoms': torch.tensor(len(pos)), 'batch': torch.zeros(0, dtype=torch.long)}
            o = ts(d)
            return (float(o['inferred_total_energy'].detach()), o['inferred_force'].detach().cpu().numpy(),
                    o['inferred_stress'].detach().cpu().numpy().reshape(-1), ei.shape[1], ei)
        return call, 'reference'
    # the restated oracle (port) when the frozen graph did not travel
    from oracle.refmodel import build_model
    w = load_weights()
    m = build_model(dict(shift=w['rescale_atomic_energy.shift'].tolist(), scale=w['rescale_atomic_energy.scale'].tolist(),
                         conv_denominator=float(w['0_convolution.denominator'][0])))
    m.load_state_dict({k: torch.tensor(v) for k, v in w.items()})
    m.eval()
    m.set_is_batch_data(False)

    def call():
        ei, S, D = radius_graph(pos, cell, pbc, 5.0)
        o = m({'x': torch.tensor(types, dtype=torch.long), 'edge_index': torch.tensor(ei),
               'edge_vec': torch.tensor(D, dtype=torch.float32).requires_grad_(True),
               'num_atoms': torch.tensor(len(pos)), 'cell_volume': torch.tensor(abs(np.linalg.det(cell)), dtype=torch.float32),
               'batch': torch.zeros(0, dtype=torch.long)})
        return (float(o['inferred_total_energy'].detach()), o['inferred_force'].detach().numpy(),
                o['inferred_stress'].detach().numpy().reshape(-1), ei.shape[1], ei)
    return call, 'port'


def reference_cpu_time(n_atoms, steps, warmup, seed=1, full_atoms=None):
    """The reference's own PyTorch CPU path on the host cores, on a bounded SAMPLE of the workload: an n_atoms cell of the
    same generator; the per-atom rate is what is reported (`extrapolated`: the full-size cell is never run on the CPU)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    pos, cell, pbc, types = make_cell(n_atoms, seed)
    call, kind = _reference_callable(pos, cell, pbc, types)
    for _ in range(warmup):
        call()
    t0 = time.perf_counter()
    for _ in range(steps):
        ne = call()[3]
    dt = (time.perf_counter() - t0) / steps
    return dict(value=n_atoms / dt, unit=UNIT, cores=cores, kind=kind, extrapolated=True, sample_atoms=n_atoms, sample_edges=int(ne),
                full_atoms=full_atoms,
                sample=f'{n_atoms}-atom amorphous cell of the same generator (NOT the full-size cell: per-atom rate, extrapolated), '
                       f'{ne} edges, {steps} timed calls ({dt:.2f} s/call), graph via the numpy restatement'), dt


def si216_cpu_time(steps=3, warmup=1):
    """BASELINE config 1 at its full size: Si-216 on the host cores through the same reference graph (not extrapolated)."""
    from sevennet_finetuning_b200.synthetic import SPECIES_INDEX, si_diamond
    pos, cell, pbc, sym = si_diamond()
    types = np.array([SPECIES_INDEX[x] for x in sym], np.int32)
    torch.set_num_threads(os.cpu_count() or 1)
    call, kind = _reference_callable(pos, cell, pbc, types)
    for _ in range(warmup):
        call()
    t0 = time.perf_counter()
    for _ in range(steps):
        e, _, _, ne, _ = call()
    dt = (time.perf_counter() - t0) / steps
    return dict(config='config 1: Si-216 diamond supercell, full size', value=len(pos) / dt, unit=UNIT, s_per_call=dt, n_edges=int(ne),
                energy=e, cores=os.cpu_count() or 1, kind=kind, extrapolated=False)


def parity_vs_reference(model, n_atoms, seed=1):
    """The benchmarked workload's generator at the size the reference arm evaluates: CUDA engine against the reference's frozen
    fp32 graph on the CPU (tolerances of tests/test_gpu_model.py: 1e-6 eV/atom; 4e-5 eV/A relative to max(1, |F|max) against
    the fp32 graph, which itself sits 1-2e-5 from float64; stress 2e-4 of max|sigma|)."""
    pos, cell, pbc, types = make_cell(n_atoms, seed)
    call, kind = _reference_callable(pos, cell, pbc, types)
    e, f, sg, ne, ei = call()
    dev = model.device
    out = model.calculate(torch.tensor(pos, dtype=torch.float64, device=dev), cell, pbc, torch.tensor(types, dtype=torch.int32, device=dev),
                          sort_atoms=False)
    same = bool(out['graph']['src'].shape[0] == ne and (out['graph']['src'].cpu().numpy() == ei[1]).all()
                and (out['graph']['dst'].cpu().numpy() == ei[0]).all())
    de = abs(float(out['energy'].item()) - e) / n_atoms
    df = float(np.abs(out['forces'].cpu().numpy() - f).max())
    fmax = float(np.abs(f).max())
    ds = float(np.abs(out['stress'].cpu().numpy().reshape(-1) - sg).max() / max(np.abs(sg).max(), 1e-30))
    ok = bool(same and de <= 1e-6 and df <= 4e-5 * max(1.0, fmax) and ds <= 2e-4)
    return dict(against=f'{kind}: frozen TorchScript graph, fp32, CPU', n_atoms=n_atoms, n_edges=int(ne), edge_set_equal=same,
                dE_per_atom=de, max_dF=df, max_F=fmax, rel_dS=ds, within_tolerance=ok)


def parity_parallel(model, par, dev, n_atoms=20000):
    """N > 1: the decomposed call over NCCL against the single-GPU engine on the same cell (every rank evaluates the whole
    cell alone and compares its own atoms); max over ranks."""
    import torch.distributed as dist
    pos, cell, pbc, types = make_cell(n_atoms, seed=2)
    pos_d = torch.tensor(pos, dtype=torch.float64, device=dev)
    types_d = torch.tensor(types, dtype=torch.int32, device=dev)
    ref = model.calculate(pos_d, cell, pbc, types_d)
    out = par.calculate(pos_d, cell, pbc, types_d)
    ids = out['local_ids'].long()
    df = (out['forces'] - ref['forces'][ids]).abs().max().double()
    de = (out['energy'].double() - ref['energy'].double()).abs().max() / n_atoms
    ds = ((out['stress'] - ref['stress']).abs().max() / ref['stress'].abs().max()).double()
    ne = torch.tensor([float(out['n_edges'])], device=dev, dtype=torch.float64)
    t = torch.stack([de.reshape(()), df.reshape(()), ds.reshape(())])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(ne)
    de, df, ds = (float(v) for v in t.cpu())
    return dict(against='single-GPU engine on the whole cell (same kernels, one rank)', transport='NCCL', n_atoms=n_atoms,
                edge_count_equal=bool(int(ne.item()) == int(ref['graph']['src'].shape[0])), dE_per_atom=de, max_dF=df, rel_dS=ds,
                within_tolerance=bool(de <= 1e-6 and df <= 6e-5 and ds <= 1e-4))


def reference_gpu_time(dev, n_atoms, steps=3, warmup=2):
    """the reference's frozen TorchScript graph with map_location='cuda' on this B200 ("the one to beat"), at the
    largest size of the same generator that fits (it materialises [E, 3136] messages and keeps them for autograd)."""
    from sevennet_finetuning_b200 import ops
    path = os.path.join(ROOT, 'oracle', '_ref', 'deployed_serial.pt')
    if not os.path.exists(path):
        return dict(unavailable='oracle/_ref/deployed_serial.pt did not travel')
    for fn in ('_jit_override_can_fuse_on_gpu', '_jit_set_texpr_fuser_enabled', '_jit_set_nvfuser_enabled'):
        try:   # as pair_e3gnn.cpp:324-329 does: no JIT fusers
            getattr(torch._C, fn)(False)
        except Exception:
            pass
    ts = torch.jit.load(path, map_location=dev).eval()
    while n_atoms >= 500:
        try:
            pos, cell, pbc, types = make_cell(n_atoms, seed=1)
            pos_d = torch.tensor(pos, dtype=torch.float64, device=dev)
            x = torch.tensor(types, dtype=torch.long, device=dev)
            cell_t = torch.tensor(cell, dtype=torch.float32, device=dev)
            vol = torch.tensor(abs(np.linalg.det(cell)), dtype=torch.float32, device=dev)

            def call():
                g = ops.radius_graph(pos_d, cell, pbc, 5.0)   # graph by the CUDA cell list (the reference would use ASE on the CPU)
                d = {'x': x, 'pos': pos_d.float().requires_grad_(True),
                     'edge_index': torch.stack([g['dst'].long(), g['src'].long()]), 'pbc_shift': g['shift'].float(),
                     'cell_lattice_vectors': cell_t, 'cell_volume': vol, 'num_atoms': torch.tensor(n_atoms),
                     'batch': torch.zeros(0, dtype=torch.long)}
                o = ts(d)
                return o['inferred_total_energy'], o['inferred_force'], int(g['src'].shape[0])
            for _ in range(warmup):
                call()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                e, f, ne = call()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            return dict(value=n_atoms / (ms * 1e-3), unit=UNIT, n_atoms=n_atoms, n_edges=ne, ms_per_call=ms, steps=steps,
                        kind='reference: frozen TorchScript graph (deployed_serial.pt), map_location=cuda, fp32, JIT fusers off',
                        peak_mem_gb=torch.cuda.max_memory_allocated(dev) / 1e9, energy=float(e.detach()))
        except torch.cuda.OutOfMemoryError:
            torch.cuda.empty_cache()
            n_atoms //= 2
        except Exception as ex:
            return dict(unavailable=repr(ex)[:300])
    return dict(unavailable='out of memory at every size tried')


FT_METRIC = 'SevenNet-0 fine-tune structures/s'
FT_UNIT = 'structures/s'
# byte model of one structure-step: the second-order pass runs, per interaction layer, 1 + 3
# forward-type and 1 + 3 backward-type tensor-product kernels (higher_order.ConvB.backward), each moving the fused-model
# bytes of SURVEY 8(d): 4 x (8,292 E + 43,136 N) + 4 x (16,092 E + 43,136 N)
FT_BYTES_E, FT_BYTES_N = 4 * 8292.0 + 4 * 16092.0, 8 * 43136.0


def finetune_cfg(world, batch):
    if world == 1:
        return {'workload': f'config 2: SevenNet-0 fine-tuning step (Huber E/F/S loss + rehearsal replay batch + EWC Fisher penalty, '
                            f'Adam), batches of {batch} synthetic ~100-atom bulk crystals + {batch} replay structures',
                'parallelism': 'single', 'l2': 'inputs larger than L2'}
    return {'workload': f'config 5: data-parallel SevenNet-0 fine-tuning with rehearsal + EWC, {batch} + {batch} synthetic '
                        f'~100-atom bulk crystals per GPU and step, gradient all-reduce over NCCL (DDP)',
            'parallelism': f'dp{world}', 'l2': 'inputs larger than L2'}


def finetune_reference_cpu(n_struct, steps, warmup):
    """The reference's fine-tuning iteration (sevenn/train/trainer.py:173-210) on the host cores through the oracle
    restatement of its modules (the reference package itself needs e3nn and cannot be imported; its frozen TorchScript
    graphs have no parameters to train) on a bounded sample: n_struct + n_struct structures per step."""
    from oracle.graph import radius_graph
    from oracle.refmodel import build_model
    from sevennet_finetuning_b200 import finetune as ft
    from sevennet_finetuning_b200.synthetic import SPECIES_INDEX, bulk_batch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    w = load_weights()
    m = build_model(dict(shift=w['rescale_atomic_energy.shift'].tolist(), scale=w['rescale_atomic_energy.scale'].tolist(),
                         conv_denominator=float(w['0_convolution.denominator'][0])))
    sd = {k: torch.tensor(v) for k, v in w.items()}
    m.load_state_dict(sd)
    m.train()
    m.set_is_batch_data(True)

    def graphs(seed):
        out = []
        for pos, cell, pbc, sym in bulk_batch(n_struct, seed=seed):
            ei, S, D = radius_graph(pos, cell, pbc, 5.0)
            n = len(pos)
            out.append({'x': torch.tensor([SPECIES_INDEX[t] for t in sym]), 'edge_index': torch.tensor(ei),
                        'edge_vec': torch.tensor(D, dtype=torch.float32), 'cell_volume': abs(float(np.linalg.det(cell))),
                        'num_atoms': n, 'total_energy': -5.0 * n, 'force_of_atoms': torch.zeros(n, 3), 'stress': torch.zeros(6)})
        return out
    tb, mb = ft.collate(graphs(11)), ft.collate(graphs(12))
    lf = ft.get_loss_functions('huber', {'delta': 0.01}, 1.0, True, 0.01, {k: torch.full_like(v, 1e-3) for k, v in sd.items()},
                               {k: v.clone() for k, v in sd.items()}, ewc_lambda=1e5, device='cpu')
    opt = torch.optim.Adam([p for p in m.parameters() if p.requires_grad], lr=1e-6)

    def step():
        opt.zero_grad()
        for k, b in enumerate((tb, mb)):
            d = {kk: (v.clone() if torch.is_tensor(v) else v) for kk, v in b.items()}
            d['edge_vec'].requires_grad_(True)
            loss = ft.total_loss(lf, m(d), m, 'cpu')
            loss.backward()
            opt.step()
        return float(loss.detach())
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    ne = int(tb['edge_index'].shape[1] + mb['edge_index'].shape[1])
    return dict(value=2 * n_struct / dt, unit=FT_UNIT, cores=cores, kind='port',
                sample=f'{n_struct} + {n_struct} structures per step of the same generator ({int(tb["x"].shape[0] + mb["x"].shape[0])} '
                       f'atoms, {ne} edges), {steps} timed steps ({dt:.2f} s/step), oracle restatement of the reference modules'), dt


def finetune_bench(a, rank, world, local, warmup, cpu_baseline=True, clocks=True):
    """Config 2 (N = 1) / config 5 (N > 1): structures/s of the rehearsal + EWC iteration over the B200 module twins."""
    import torch.distributed as dist
    from sevennet_finetuning_b200 import _lib, finetune as ft, nn as b200nn, ops
    from sevennet_finetuning_b200.model import B200SevenNet
    from sevennet_finetuning_b200.synthetic import bulk_batch
    dev = torch.device('cuda', local)
    L = _lib.lib()
    w = load_weights()
    sd = {k: torch.tensor(v) for k, v in w.items()}
    cfg = dict(shift=w['rescale_atomic_energy.shift'].tolist(), scale=w['rescale_atomic_energy.scale'].tolist(),
               conv_denominator=float(w['0_convolution.denominator'][0]))
    model = b200nn.build_E3_equivariant_model(cfg)
    model.load_state_dict(sd)
    engine = B200SevenNet(state_dict=w, device=dev)
    B = a.batch
    # two different (train, replay) batch pairs per rank, alternated, so no step re-reads the previous step's inputs
    pairs = []
    for k in range(2):
        tg = ft.synthetic_graphs(bulk_batch(B, seed=100 + 10 * rank + k), 5.0, dev, label_model=engine, noise=0.002, seed=k)
        mg = ft.synthetic_graphs(bulk_batch(B, seed=500 + 10 * rank + k), 5.0, dev, label_model=engine, noise=0.002, seed=k)
        pairs.append((ft.collate(tg, pin=True), ft.collate(mg, pin=True)))
    del engine
    torch.cuda.empty_cache()
    # EWC: Fisher / theta* of the shape the reference ships (fisher_sevenn.pt / opt_params_sevenn.pt); synthetic Fisher
    g = torch.Generator().manual_seed(7)
    fisher = {k: torch.rand(v.shape, generator=g) * 1e-3 for k, v in sd.items()}
    lf = ft.get_loss_functions('huber', {'delta': 0.01}, 1.0, True, 0.01, fisher, {k: v.clone() for k, v in sd.items()},
                               ewc_lambda=1e5, device=dev)
    tr = ft.RehearsalTrainer(model, lf, 'adam', {'lr': 1e-6}, device=dev, distributed=world > 1, local_rank=local)
    dev_pairs = [(ft.to_device(t, dev), ft.to_device(m, dev)) for t, m in pairs]
    it = {'k': 0}

    def step_resident():
        t, m = dev_pairs[it['k'] % 2]
        it['k'] += 1
        t['edge_vec'].grad = m['edge_vec'].grad = None
        return tr.run_one_step(dict(t), dict(m))   # the modules rewrite keys of the dict they are given (x -> features)

    def step_e2e():
        t, m = pairs[it['k'] % 2]
        it['k'] += 1
        l, lm = tr.run_one_step(ft.to_device(t, dev), ft.to_device(m, dev))
        return float(l), float(lm)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            out = fn()
            if cs:
                cs.sample()   # the step is queued, the GPU is executing it
        e1.record()
        barrier()
        wall = time.perf_counter() - t0
        t = torch.tensor([max(e0.elapsed_time(e1) * 1e-3, wall if fn is step_e2e else 0.0)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), out

    cs = ClockSampler(local) if clocks else None
    for _ in range(warmup):
        step_resident()
        if cs:
            cs.sample()
    launches0 = L.b200_launch_count()
    torch.cuda.synchronize()
    if cs:
        cs.rows = []
    t_res, _ = timed(step_resident, a.steps)
    launches = L.b200_launch_count() - launches0
    for _ in range(2):
        step_e2e()
    t_e2e, losses = timed(step_e2e, a.steps)
    # which kernel family dominates: per-op CUDA events over two extra (untimed) steps
    ops.enable_timing(True)
    step_resident()
    step_resident()
    per_op = {k: v[0] / 2 for k, v in ops.timing_summary().items()}
    ops.enable_timing(False)
    n_struct = 2 * B * world
    E = sum(int(t['edge_index'].shape[1] + m['edge_index'].shape[1]) for t, m in pairs) / 2
    N = sum(int(t['x'].shape[0] + m['x'].shape[0]) for t, m in pairs) / 2
    pk = peaks()
    step_ms = t_res / a.steps * 1e3
    byt = FT_BYTES_E * E + FT_BYTES_N * N
    dom = max(per_op, key=per_op.get)
    conv_ms = sum(v for k, v in per_op.items() if k.split()[0] in ('conv_fwd', 'conv_bwd')) or 1e-9
    roof = dict(kernel='conv_fwd+conv_bwd (uvu tensor product + scatter and their reverse / second-order passes)', bound='hbm',
                achieved=byt / (conv_ms * 1e-3) / 1e9, peak=pk['hbm'], unit='GB/s', frac=byt / (conv_ms * 1e-3) / 1e9 / pk['hbm'],
                traffic=None, peak_source=pk['src'], ms_per_step=conv_ms, slowest_op=dom)
    line = dict(metric=FT_METRIC, value=n_struct * a.steps / t_res, unit=FT_UNIT, n_gpus=world, steps=a.steps, warmup=warmup,
                ms_per_step=step_ms, higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f32', data='synthetic',
                config=finetune_cfg(world, B), n_atoms_per_step=N, n_edges_per_step=E,
                e2e=dict(value=n_struct * a.steps / t_e2e, unit=FT_UNIT,
                         h2d_bytes_per_step=int(sum(ft.batch_bytes(t) + ft.batch_bytes(m) for t, m in pairs) / 2),
                         d2h_bytes_per_step=8, ms_per_step=t_e2e / a.steps * 1e3),
                gpu_launches=int(launches), roofline=roof,
                step_roofline=dict(bound='hbm', algorithmic_bytes=byt, achieved=byt / (step_ms * 1e-3) / 1e9, peak=pk['hbm'],
                                   unit='GB/s', frac=byt / (step_ms * 1e-3) / 1e9 / pk['hbm']),
                op_ms_per_step=per_op, last_losses=list(losses), clocks=cs.summary() if clocks else None)
    if rank == 0 and cpu_baseline and world == 1:
        line['cpu_baseline'], _ = finetune_reference_cpu(2, 1, 1)
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--atoms', type=int, default=100_000)
    ap.add_argument('--atoms-per-gpu', type=int, default=131_072, help='N>1: the cell holds N x this many atoms')
    ap.add_argument('--impl', default='b200')
    ap.add_argument('--cpu-atoms', type=int, default=600)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--workload', default='md', choices=['md', 'finetune'],
                    help="md: configs 3/4 (atom-steps/s, the default line); finetune: configs 2/5 (structures/s)")
    ap.add_argument('--batch', type=int, default=32, help='finetune: structures per train batch (and per replay batch)')
    ap.add_argument('--no-finetune', action='store_true', help='md line without the appended fine-tune measurement')
    ap.add_argument('--strong-atoms', type=int, default=1_048_576, help='N>1: size of the one cell of the strong-scaling view')
    ap.add_argument('--no-alt', action='store_true', help='skip the second scaling view (scaling_base / strong_scaling)')
    ap.add_argument('--ref-gpu-atoms', type=int, default=8000, help='N=1: size tried first for the reference graph on the GPU')
    a = ap.parse_args()
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    warmup = max(a.warmup, 3) if a.impl != 'reference' else a.warmup
    n_total = a.atoms if world == 1 else a.atoms_per_gpu * world
    if world == 1:
        cfg = {'workload': f'config 3: SevenNet-0 single-structure MD force call, {a.atoms}-atom synthetic amorphous cell '
                           '(Si:O 1:2, ~50 edges/atom, 5 A cutoff), graph rebuilt every step',
               'parallelism': 'single', 'l2': 'inputs larger than L2'}
    else:
        from sevennet_finetuning_b200.parallel import brick_grid
        cfg = {'workload': f'config 4: SevenNet-0 spatially decomposed MD force call, one {n_total}-atom synthetic amorphous '
                           f'cell (Si:O 1:2, ~50 edges/atom, 5 A cutoff) in {"x".join(map(str, brick_grid(world)))} bricks, '
                           'ghost shell of one cutoff, per-layer halo exchange over NCCL, partition + graph rebuilt every step',
               'parallelism': f'spatial{world}', 'l2': 'inputs larger than L2'}

    if a.impl == 'reference':
        if rank != 0:
            return
        if a.workload == 'finetune':
            steps = max(1, min(a.steps, 2))
            cb, dt = finetune_reference_cpu(2, steps, min(a.warmup, 1))
            print(json.dumps(dict(metric=FT_METRIC, value=cb['value'], unit=FT_UNIT, n_gpus=a.gpus, steps=steps, warmup=min(a.warmup, 1),
                                  ms_per_step=dt * 1e3, higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f32',
                                  data='synthetic', config=finetune_cfg(world, a.batch), impl='reference', cpu_baseline=cb,
                                  e2e=dict(value=cb['value'], unit=FT_UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                                  gpu_launches=0)))
            return
        steps = max(1, min(a.steps, 3))
        cb, dt = reference_cpu_time(max(a.cpu_atoms, 3000), steps, min(a.warmup, 1), full_atoms=n_total)
        cfg = dict(cfg, reference_sample=cb['sample'])
        line = dict(metric=METRIC, value=cb['value'], unit=UNIT, n_gpus=a.gpus, steps=steps, warmup=min(a.warmup, 1),
                    ms_per_step=dt * 1e3, higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f32',
                    data='synthetic', config=cfg, impl='reference', cpu_baseline=cb,
                    e2e=dict(value=cb['value'], unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0,
                    extrapolated=True)
        try:
            line['si216'] = si216_cpu_time()
        except Exception as ex:
            line['si216'] = {'error': repr(ex)[:200]}
        print(json.dumps(line))
        return

    import torch.distributed as dist
    from sevennet_finetuning_b200 import _lib, ops
    from sevennet_finetuning_b200.model import B200SevenNet
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        if os.environ.get('NCCL_DEBUG', 'VERSION').upper() == 'VERSION':
            os.environ['NCCL_DEBUG'] = 'WARN'
        # NCCL writes its version banner straight to file descriptor 1: keep stdout to the one JSON line by pointing
        # fd 1 at stderr for the rest of the run and printing the line through a duplicate of the original stdout
        global _JSON_OUT
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), 'w')
        os.dup2(2, 1)
        dist.init_process_group('nccl', device_id=dev)
    if a.workload == 'finetune':
        line = finetune_bench(a, rank, world, local, warmup, cpu_baseline=not a.no_cpu_baseline)
        if rank == 0:
            print(json.dumps(line), file=_JSON_OUT or sys.stdout, flush=True)
        if world > 1:
            dist.destroy_process_group()
        return
    L = _lib.lib()
    model = B200SevenNet(state_dict=load_weights(), device=dev)
    pos, cell, pbc, types = make_cell(n_total, seed=0)
    par = None
    if world > 1:
        from sevennet_finetuning_b200.parallel import ParallelSevenNet
        par = ParallelSevenNet(model)
    pos_d = torch.tensor(pos, dtype=torch.float64, device=dev)
    types_d = torch.tensor(types, dtype=torch.int32, device=dev)
    pos_pin = torch.tensor(pos, dtype=torch.float64).pin_memory()
    types_pin = torch.tensor(types, dtype=torch.int32).pin_memory()

    calc = model.calculate if par is None else par.calculate

    def step_resident():
        return calc(pos_d, cell, pbc, types_d)

    def step_e2e():
        p = pos_pin.to(dev, non_blocking=True)
        t = types_pin.to(dev, non_blocking=True)
        o = calc(p, cell, pbc, t)
        return o['energy'].cpu(), o['forces'].cpu(), o['stress'].cpu()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        evs = []
        for _ in range(steps):
            out = fn()
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            evs.append(ev)
            cs.sample()   # the step is queued, the GPU is executing it
        e1.record()
        barrier()
        wall = time.perf_counter() - t0
        ms = e0.elapsed_time(e1)
        timed.per_step = [round((e0 if k == 0 else evs[k - 1]).elapsed_time(evs[k]), 3) for k in range(steps)]
        # host-synchronous pieces (edge count readback, D2H) sit between the events as well; use the larger
        t = torch.tensor([max(ms * 1e-3, wall if fn is step_e2e else ms * 1e-3)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), out

    # parity of the benchmarked workload before anything is timed
    try:
        parity = parity_vs_reference(model, a.cpu_atoms) if par is None else parity_parallel(model, par, dev)
    except Exception as ex:
        parity = {'error': repr(ex)[:300]}
    # the clock sampler and the engine's per-category events are switched on BEFORE the warm-up steps: the first NVML
    # queries and the first event creations otherwise land in the timed steps; what they record during warm-up is discarded
    cs = ClockSampler(local)
    L.b200_model_set_timing(model._h, 1)
    for _ in range(warmup):
        out = step_resident()
        cs.sample()
    if par is None:
        n_edges, n_own = int(out['graph']['src'].shape[0]), n_total
    else:
        cnt = torch.tensor([out['n_edges'], out['n_local'], out['n_ghost']], dtype=torch.int64, device=dev)
        per_rank = [torch.zeros_like(cnt) for _ in range(world)]
        dist.all_gather(per_rank, cnt)
        per_rank = torch.stack(per_rank).cpu().numpy()
        n_edges, n_own = int(per_rank[:, 0].sum()), int(out['n_local'])
        cfg['per_rank'] = {'edges': per_rank[:, 0].tolist(), 'local': per_rank[:, 1].tolist(), 'ghost': per_rank[:, 2].tolist()}
    launches0 = L.b200_launch_count()
    torch.cuda.synchronize()
    L.b200_model_set_timing(model._h, 1)   # resets the accumulated events of the warm-up
    cs.rows = []
    t_res, out = timed(step_resident, a.steps)
    per_step_res = list(timed.per_step)
    cat = (C.c_float * 9)()
    L.b200_model_get_timing(model._h, cat, 9)
    L.b200_model_set_timing(model._h, 0)
    launches = L.b200_launch_count() - launches0
    for _ in range(2):
        step_e2e()
    t_e2e, _ = timed(step_e2e, a.steps)
    clocks = cs.summary()
    N, E = n_total, n_edges
    value = N * a.steps / t_res
    e2e_v = N * a.steps / t_e2e
    pk = peaks()
    cat_ms = {n: float(cat[k]) / a.steps for k, n in enumerate(CAT_NAMES)}
    step_ms = t_res / a.steps * 1e3
    # dominant kernel category -> roofline (algorithmic bytes/flops per step: SURVEY 8d)
    # per-kernel figures are rank 0's (its own edges / atoms); the step figure is the whole job over N x peak
    Ek, Nk = (E, N) if par is None else (int(out['n_edges']), int(out['n_local']))
    # SURVEY 8(d) byte model per category and step (fused model: per-edge weights never counted); the edge-gradient
    # pass now also accumulates d x[src] (b200_conv_bwd), so it carries the whole backward figure
    split = os.environ.get('B200_SPLIT_BWD_X', '0') not in ('', '0')
    conv_bytes = {'conv_fwd': 8292.0 * Ek + 43136.0 * Nk,
                  'conv_bwd_edge': (16092.0 - (4 * 1920.0 if split else 0.0)) * Ek + 43136.0 * Nk,
                  'conv_bwd_x': 4 * 1920.0 * Ek}
    mlp_flops = {'radial_mlp_fwd': 2.0 * (29184 + 3 * 66048 + 18944) * Ek, 'radial_mlp_bwd': 2.0 * (29184 + 3 * 66048 + 18944) * Ek}
    # HBM bytes the radial kernels must move in the un-fused layout: the per-edge weights [E, wn] written (fwd) / their
    # gradient read (bwd) once, wn summed over the five layers = 3488
    mlp_bytes = {'radial_mlp_fwd': 4.0 * 3488 * Ek, 'radial_mlp_bwd': 4.0 * 3488 * Ek}
    dom = max((k for k in cat_ms if k != 'halo_exchange'), key=cat_ms.get)
    if dom in mlp_flops:
        ach = mlp_flops[dom] / (cat_ms[dom] * 1e-3) / 1e12
        roof = dict(kernel=dom, bound='tensor', achieved=ach, peak=pk['tf'], unit='TFLOP/s', frac=ach / pk['tf'], traffic=None,
                    note='algorithmic fp32 flops (3xTF32 executes 2-3x as many tf32 flops; tf32 peak is half the bf16 figure)',
                    hbm_side=dict(achieved=mlp_bytes[dom] / (cat_ms[dom] * 1e-3) / 1e9, peak=pk['hbm'], unit='GB/s',
                                  frac=mlp_bytes[dom] / (cat_ms[dom] * 1e-3) / 1e9 / pk['hbm']))
    else:
        byt = conv_bytes.get(dom, 24384.0 * Ek + 86272.0 * Nk)
        ach = byt / (cat_ms[dom] * 1e-3) / 1e9
        roof = dict(kernel=dom, bound='hbm', achieved=ach, peak=pk['hbm'], unit='GB/s', frac=ach / pk['hbm'], traffic=None)
    # measured DRAM traffic of the dominant category (sum over its launches of one step) from the committed ncu capture
    try:
        tr = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json')))
        if world == 1 and tr.get('n_edges') == E and dom in tr.get('dram_bytes_per_step', {}):
            roof['traffic'] = tr['dram_bytes_per_step'][dom]
            roof['traffic_source'] = tr.get('source')
    except Exception:
        pass
    roof['peak_source'] = pk['src']
    roof['ms_per_step'] = cat_ms[dom]
    step_bytes = 24384.0 * E + 86272.0 * N
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=a.steps, warmup=warmup, ms_per_step=step_ms,
                higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f32', data='synthetic', config=cfg,
                n_atoms=N, n_edges=E,
                e2e=dict(value=e2e_v, unit=UNIT, h2d_bytes_per_step=int(pos_pin.numel() * 8 + types_pin.numel() * 4),
                         d2h_bytes_per_step=int(4 + n_own * 12 + 24), ms_per_step=t_e2e / a.steps * 1e3),
                gpu_launches=int(launches), roofline=roof,
                step_roofline=dict(bound='hbm', algorithmic_bytes=step_bytes, achieved=step_bytes / (step_ms * 1e-3) / 1e9,
                                   peak=pk['hbm'] * world, unit='GB/s',
                                   frac=step_bytes / (step_ms * 1e-3) / 1e9 / (pk['hbm'] * world)),
                kernel_ms_per_step=cat_ms, ms_each_step=per_step_res, clocks=clocks, parity=parity)
    # second scaling views (3 steps each after 1 warm-up): N = 1 at the per-GPU size of the multi-GPU runs; N > 1 strong
    # scaling of ONE 1,048,576-atom cell next to the weak-scaling `value`
    try:
        n_alt = a.atoms_per_gpu if world == 1 else a.strong_atoms
        if n_alt and n_alt != n_total and not a.no_alt:
            model._ws = None
            out = None
            torch.cuda.empty_cache()
            pos2, cell2, pbc2, types2 = make_cell(n_alt, seed=0)
            pos2_d = torch.tensor(pos2, dtype=torch.float64, device=dev)
            types2_d = torch.tensor(types2, dtype=torch.int32, device=dev)
            fn2 = lambda: calc(pos2_d, cell2, pbc2, types2_d)   # noqa: E731
            fn2()
            t_alt, out = timed(fn2, 3)
            key = 'scaling_base' if world == 1 else 'strong_scaling'
            line[key] = dict(n_atoms=n_alt, value=n_alt * 3 / t_alt, unit=UNIT, ms_per_step=t_alt / 3 * 1e3, steps=3,
                             note=('N = 1 at the per-GPU cell size of the N > 1 runs' if world == 1 else
                                   f'strong scaling: one {n_alt}-atom cell on {world} GPUs'))
            pos2_d = types2_d = None
        elif world > 1 and n_alt == n_total:
            line['strong_scaling'] = dict(n_atoms=n_total, value=value, unit=UNIT, ms_per_step=step_ms, steps=a.steps,
                                          note='the weak-scaling run at this N is the strong-scaling cell')
    except Exception as ex:
        line['scaling_base' if world == 1 else 'strong_scaling'] = {'error': repr(ex)[:300]}
    if not a.no_finetune:
        # second half of BASELINE.json's metric ("...; fine-tune structures/s"): config 2 (N = 1) / config 5 (N > 1: data
        # parallel, DDP over NCCL, 32 + 32 structures per GPU) in the same run, after the MD workspace is released (full
        # line: python bench.py --workload finetune)
        model._ws = None
        out = pos_d = types_d = None
        torch.cuda.empty_cache()
        try:
            fa = argparse.Namespace(**vars(a))
            fa.steps = min(a.steps, 5)
            fl = finetune_bench(fa, rank, world, local, 3, cpu_baseline=not a.no_cpu_baseline, clocks=False)
            line['finetune'] = {k: fl[k] for k in ('metric', 'value', 'unit', 'ms_per_step', 'steps', 'config', 'e2e', 'gpu_launches',
                                                   'roofline', 'step_roofline', 'op_ms_per_step', 'cpu_baseline') if k in fl}
        except Exception as ex:   # the MD line stands on its own
            line['finetune'] = {'error': repr(ex)}
    if rank == 0:
        if not a.no_cpu_baseline and world == 1:
            line['cpu_baseline'], _ = reference_cpu_time(a.cpu_atoms, 2, 1, full_atoms=n_total)
            model._ws = None
            torch.cuda.empty_cache()
            line['reference_gpu'] = reference_gpu_time(dev, a.ref_gpu_atoms)
        print(json.dumps(line), file=_JSON_OUT or sys.stdout, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
