"""Stand-alone timing of the fused convolution ops on a realistic graph (not a benchmark; for ncu / tuning).
usage: python tools/prof_fused.py [atoms] [reps]"""
import sys

import numpy as np
import torch

sys.path.insert(0, '.')
from sevennet_finetuning_b200 import ops  # noqa: E402
from sevennet_finetuning_b200._lib import EdgeParams  # noqa: E402
from sevennet_finetuning_b200.synthetic import amorphous_cell  # noqa: E402

dev = torch.device('cuda:0')
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
pos, cell, pbc, _ = amorphous_cell(n, seed=0)
g = ops.radius_graph(torch.tensor(pos, device=dev), cell, pbc, 5.0)
E = g['src'].shape[0]
torch.manual_seed(0)
p = EdgeParams(8, 1, 1, 5.0, 4.5, 6.0)
coeffs = torch.tensor([0.5079, 1.1099, 1.7265, 2.3767, 3.0392, 3.7201, 4.4142, 5.1186], device=dev)
_, Y, emb = ops.edge_embed_fwd(p, g['edge_vec'], coeffs)


def timeit(fn):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


for muls, lout, name in [((128, 64, 32), 2, 'mid'), ((128, 0, 0), 2, 'layer0'), ((128, 64, 32), 0, 'last')]:
    wn, od = ops.conv_dims(*muls, lout)
    xdim = muls[0] + 3 * muls[1] + 5 * muls[2]
    Ws = [torch.randn(8, 64, device=dev), torch.randn(64, 64, device=dev), torch.randn(64, wn, device=dev)]
    x = torch.randn(n, xdim, device=dev)
    go = torch.randn(n, od, device=dev)
    rad = ops.fused_radial(p, g['edge_vec'], coeffs, Ws[0], Ws[1])
    t_r = timeit(lambda: ops.fused_radial(p, g['edge_vec'], coeffs, Ws[0], Ws[1]))
    t_f = timeit(lambda: ops.fused_conv_fwd(muls, lout, n, g['dst'], g['src'], x, Y, rad, Ws[2], 36.0))
    t_b = timeit(lambda: ops.fused_conv_bwd(muls, lout, g['dst'], g['src'], x, Y, rad, Ws[2], go, 36.0))
    print(f'{name:7s} N={n} E={E}: radial {t_r:.3f} ms  fwd {t_f:.3f} ms  bwd {t_b:.3f} ms   '
          f'(per 1M edges: {t_r / E * 1e6:.3f} / {t_f / E * 1e6:.3f} / {t_b / E * 1e6:.3f})', flush=True)
