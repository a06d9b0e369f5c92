"""NCCL check of the spatially decomposed force call: run under torchrun on N GPUs; every rank also evaluates the
whole cell alone and compares its own atoms.  `gpurun --gpus 2 -- python -m torch.distributed.run --nnodes=1
--nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/check_parallel.py [atoms]`"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sevennet_finetuning_b200.model import B200SevenNet  # noqa: E402
from sevennet_finetuning_b200.parallel import ParallelSevenNet  # noqa: E402
from sevennet_finetuning_b200.synthetic import SPECIES_INDEX, amorphous_cell  # noqa: E402

rank, local = int(os.environ['RANK']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
w = dict(np.load(os.path.join(root, 'tests', 'golden', 'sevennet0_weights.npz')))
pos, cell, pbc, sym = amorphous_cell(n, seed=2)
types = torch.tensor(np.array([SPECIES_INDEX[s] for s in sym], np.int32), device=dev)
pos_d = torch.tensor(pos, device=dev)
model = B200SevenNet(state_dict=w, device=dev)
ref = model.calculate(pos_d, cell, pbc, types)
out = ParallelSevenNet(model).calculate(pos_d, cell, pbc, types)
ids = out['local_ids']
df = float((out['forces'] - ref['forces'][ids]).abs().max())
de = abs(float(out['energy'].item()) - float(ref['energy'].item())) / n
ds = float((out['stress'] - ref['stress']).abs().max() / ref['stress'].abs().max())
print(f'rank {rank}: local {out["n_local"]} ghost {out["n_ghost"]} edges {out["n_edges"]} |dE|/N {de:.2e} max|dF| {df:.2e} '
      f'rel dS {ds:.2e}', flush=True)
assert de <= 1e-6 and df <= 4e-5 and ds <= 1e-4
dist.barrier()
dist.destroy_process_group()
