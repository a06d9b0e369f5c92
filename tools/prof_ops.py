"""Stand-alone launcher of single ops at C3-like sizes, for ncu captures (not a benchmark)."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, '.')
from sevennet_finetuning_b200 import ops  # noqa: E402

dev = torch.device('cuda:0')
which = sys.argv[1] if len(sys.argv) > 1 else 'radial'
E = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
torch.manual_seed(0)
if which == 'radial':
    wn = 960
    emb = torch.randn(E, 8, device=dev) * 0.3
    Ws = [torch.randn(8, 64, device=dev), torch.randn(64, 64, device=dev), torch.randn(64, wn, device=dev)]
    g = torch.randn(E, wn, device=dev)
    for _ in range(3):
        w = ops.radial_mlp_fwd(emb, *Ws)
        de = ops.radial_mlp_bwd(emb, *Ws, g)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    w = ops.radial_mlp_fwd(emb, *Ws)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    de = ops.radial_mlp_bwd(emb, *Ws, g)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f'radial fwd {1e3 * (t1 - t0):.3f} ms  bwd {1e3 * (t2 - t1):.3f} ms  E={E}')

    import ctypes as C
    from sevennet_finetuning_b200._lib import lib
    buf = (C.c_longlong * 64)()
    w = ops.radial_mlp_fwd(emb, *Ws)
    torch.cuda.synchronize()
    lib().b200_debug_tc(buf, 64)
    t = list(buf)
    print('fwd CTA timeline (clk rel. start): hidden done', t[1] - t[0])
    for mt in range(9):
        b = 2 + 5 * mt
        print(f' it{mt}: top {t[b]-t[0]} full_wait {t[b+1]-t[b] if mt<8 else 0} mma_issue {t[b+2]-t[b+1] if mt<8 else 0} '
              f'done_wait {t[b+3]-t[b+2] if mt>0 else 0} epilogue {t[b+4]-t[b+3] if mt>0 else 0}')
