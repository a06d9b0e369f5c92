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
    buf = (C.c_longlong * 128)()
    w = ops.radial_mlp_fwd(emb, *Ws)
    torch.cuda.synchronize()
    lib().b200_debug_tc(buf, 128)
    t = list(buf)
    z = t[0]
    print('fwd worker: h1 done', t[1] - z, 'z2 done', t[2] - z, 'h2 done', t[3] - z, 'end', t[30] - z)
    for mt in range(8):
        b = 4 + 3 * mt
        print(f'  slab{mt}: top {t[b]-z} mma_wait {t[b+1]-t[b]} tmem_ld {t[b+2]-t[b+1]} stores {(t[b+3] if mt < 7 else t[30])-t[b+2]}')
    print('fwd issuer: start', t[40] - z, 'h1 ready', t[41] - z, 'z2 done', t[42] - z, 'h2 ready', t[43] - z)
    for mt in range(8):
        b = 44 + 4 * mt
        print(f'  slab{mt}: top {t[b]-z} full_wait {t[b+1]-t[b]} empty_wait {t[b+2]-t[b+1]} issue {t[b+3]-t[b+2]}')

    def cta_stats(tag):
        cb = (C.c_longlong * (3 * 8192))()
        lib().b200_debug_tc_cta(cb, 3 * 8192)
        a = np.array(list(cb), dtype=np.int64).reshape(8192, 3)
        n = min(8192, (E + 127) // 128)
        a = a[:n]
        dur = a[:, 1] - a[:, 0]
        t0 = a[:, 0].min()
        print(f'{tag} CTAs: n={n} kernel span {(a[:,1].max()-t0)/1e3:.1f} us; CTA duration us mean {dur.mean()/1e3:.2f} min {dur.min()/1e3:.2f} '
              f'p50 {np.median(dur)/1e3:.2f} p90 {np.percentile(dur,90)/1e3:.2f} max {dur.max()/1e3:.2f}')
        for lo in range(0, n, 1000):
            print(f'  blocks {lo}-{lo+999}: start {(a[lo:lo+1000,0].min()-t0)/1e3:.1f} us mean dur {dur[lo:lo+1000].mean()/1e3:.2f} us')
        sm = a[:, 2]
        per_sm = np.bincount(sm, minlength=148)
        print('  CTAs per SM: min', per_sm.min(), 'max', per_sm.max(), ' busy us per SM: min', min(dur[sm == k].sum() for k in range(148)) / 1e3,
              'max', max(dur[sm == k].sum() for k in range(148)) / 1e3)

    cta_stats('fwd')
    de = ops.radial_mlp_bwd(emb, *Ws, g)
    torch.cuda.synchronize()
    lib().b200_debug_tc(buf, 128)
    t = list(buf)
    z = t[0]
    print('bwd worker: h1 done', t[1] - z, 'z2 done', t[2] - z, 'loop start', t[3] - z, 'loop end', t[31] - z, 'mma tail', t[32] - t[31],
          'dz2 done', t[33] - z, 'z1 done', t[34] - z, 'dh1 wait', t[35] - t[34], 'end', t[30] - z)
    for i in range(8):
        b = 4 + 3 * i
        print(f'  chunk{i}: top {t[b]-z} gload {t[b+1]-t[b]} stage_wait {t[b+2]-t[b+1]} split+arrive {(t[b+3] if i < 7 else t[b+2])-t[b+2]}')
    print('bwd issuer: z2 done', t[40] - z, 'loop end', t[41] - z, 'tail', t[42] - t[41])
    for i in range(8):
        b = 44 + 4 * i
        print(f'  chunk{i}: top {t[b]-z} fullB_wait {t[b+1]-t[b]} a_wait {t[b+2]-t[b+1]} issue {t[b+3]-t[b+2]}')
    print('bwd chunks 8..15 (issuer waits for its own commit): issue+exec+commit clk', [t[88 + k] - t[72 + k] for k in range(8)])
    cta_stats('bwd')
