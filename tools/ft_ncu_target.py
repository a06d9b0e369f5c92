"""Target for ncu captures of the fine-tuning iteration: builds config 2's batches (labels = zeros, no engine calls),
runs 3 warm-up iterations, then ONE iteration between cudaProfilerStart/Stop (run ncu with --profile-from-start off)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
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
torch.cuda.profiler.start()
tr.run_one_step(dict(t), dict(m))
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print('edges', int(t['edge_index'].shape[1] + m['edge_index'].shape[1]), 'atoms', int(t['x'].shape[0] + m['x'].shape[0]))
