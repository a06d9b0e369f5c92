"""timing of the irreps linear kernels at SevenNet-0 shapes, 100k nodes (tuning aid)"""
import sys
import torch
sys.path.insert(0, '.')
from sevennet_finetuning_b200 import ops
dev = torch.device('cuda:0')
n = 100000
tot = {0: 0.0, 1: 0.0}
for irin, irout, name in [('224x0e+384x1e+352x2e', '224x0e+64x1e+32x2e', 'si2 mid'), ('128x0e+64x1e+32x2e', '224x0e+64x1e+32x2e', 'sc mid'),
                          ('128x0e+64x1e+32x2e', '128x0e+64x1e+32x2e', 'si1 mid'), ('128x0e', '64x0e', 'readout')]:
    desc, numel = ops.make_lin_desc(irin, irout)
    x = torch.randn(n, desc.in_dim, device=dev)
    g = torch.randn(n, desc.out_dim, device=dev)
    W = torch.randn(numel, device=dev)
    for tr, inp in ((False, x), (True, g)):
        for _ in range(2):
            ops.irreps_linear(desc, inp, W, transpose=tr)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            ops.irreps_linear(desc, inp, W, transpose=tr)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        tot[int(tr)] += ms
        print(f'{name:8s} transpose={int(tr)}  {ms:.3f} ms', flush=True)
print('sum fwd %.3f  transposed %.3f' % (tot[0], tot[1]))
