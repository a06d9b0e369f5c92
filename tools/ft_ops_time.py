"""Times the dense (l = 0) linear kernels at the shapes the radial MLP takes in a fine-tuning batch (rows = edges)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sevennet_finetuning_b200 import ops

dev = torch.device('cuda:0')
n = int(sys.argv[1]) if len(sys.argv) > 1 else 107607


def t(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for a, b in [(8, 64), (64, 64), (64, 384), (64, 960), (64, 224)]:
    desc, numel = ops.make_lin_desc(f'{a}x0e', f'{b}x0e')
    x = torch.randn(n, a, device=dev)
    g = torch.randn(n, b, device=dev)
    W = torch.randn(numel, device=dev)
    fl = 2.0 * n * a * b
    r = [t(lambda: ops.irreps_linear(desc, x, W)), t(lambda: ops.irreps_linear(desc, g, W, transpose=True)),
         t(lambda: ops.irreps_linear_wgrad(desc, x, g, numel))]
    print(f'{a:4d}->{b:4d} rows {n}: fwd {r[0]:.3f} ms ({fl / r[0] / 1e9:.1f} TF/s)  transpose {r[1]:.3f} ms ({fl / r[1] / 1e9:.1f})  '
          f'wgrad {r[2]:.3f} ms ({fl / r[2] / 1e9:.1f})  mm {t(lambda: x @ W.view(a, b)):.3f} ms')
