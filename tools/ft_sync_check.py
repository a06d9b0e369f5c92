"""Lists the host synchronisations of one fine-tuning iteration (torch.cuda.set_sync_debug_mode('warn')) and times the
iteration with the CPU clock next to the CUDA-event clock: a gap between the two is launch overhead the GPU waits for."""
import os, sys, time, warnings
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from sevennet_finetuning_b200 import finetune as ft, nn as b200nn
from sevennet_finetuning_b200.synthetic import bulk_batch

dev = torch.device('cuda:0')
w = bench.load_weights()
sd = {k: torch.tensor(v) for k, v in w.items()}
model = b200nn.build_E3_equivariant_model(dict(shift=w['rescale_atomic_energy.shift'].tolist(), scale=w['rescale_atomic_energy.scale'].tolist(),
                                               conv_denominator=float(w['0_convolution.denominator'][0])))
model.load_state_dict(sd)


def graphs(seed):
    gs = ft.synthetic_graphs(bulk_batch(32, seed=seed), 5.0, dev)
    for g in gs:
        n = g['num_atoms']
        g.update(total_energy=-5.0 * n, force_of_atoms=torch.zeros(n, 3), stress=torch.zeros(6))
    return gs


t, m = ft.collate(graphs(100), dev), ft.collate(graphs(500), dev)
lf = ft.get_loss_functions('huber', {'delta': 0.01}, 1.0, True, 0.01, {k: torch.full_like(v, 1e-3) for k, v in sd.items()},
                           {k: v.clone() for k, v in sd.items()}, ewc_lambda=1e5, device=dev)
tr = ft.RehearsalTrainer(model, lf, 'adam', {'lr': 1e-6}, device=dev)
for _ in range(3):
    tr.run_one_step(dict(t), dict(m))
torch.cuda.synchronize()
torch.cuda.set_sync_debug_mode('warn')
with warnings.catch_warnings(record=True) as rec:
    warnings.simplefilter('always')
    tr.run_one_step(dict(t), dict(m))
torch.cuda.set_sync_debug_mode('default')
torch.cuda.synchronize()
print('host synchronisations in one iteration:', len(rec))
seen = {}
for r in rec:
    key = f'{r.filename}:{r.lineno}'
    seen[key] = seen.get(key, 0) + 1
for k, v in sorted(seen.items(), key=lambda kv: -kv[1]):
    print(v, k)
for _ in range(3):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    tr.run_one_step(dict(t), dict(m))
    t1 = time.perf_counter()
    e1.record()
    torch.cuda.synchronize()
    print('cpu issue time %.1f ms, gpu time %.1f ms' % ((t1 - t0) * 1e3, e0.elapsed_time(e1)))
