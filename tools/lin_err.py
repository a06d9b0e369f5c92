"""error statistics of the irreps linear kernels vs float64 (tuning aid)"""
import os, sys
import numpy as np, torch
sys.path.insert(0, '.')
from sevennet_finetuning_b200 import ops
dev = torch.device('cuda:0')
torch.manual_seed(0)
for irin, irout in [('224x0e+384x1e+352x2e', '224x0e+64x1e+32x2e'), ('128x0e+64x1e+32x2e', '224x0e+64x1e+32x2e'), ('128x0e', '64x0e')]:
    desc, numel = ops.make_lin_desc(irin, irout)
    n = 4096
    x = torch.randn(n, desc.in_dim, dtype=torch.float64).abs() + 0.5   # same-sign terms expose a rounding bias
    W = torch.randn(numel, dtype=torch.float64).abs() + 0.5
    y = ops.irreps_linear(desc, x.float().to(dev), W.float().to(dev)).double().cpu()
    # float64 reference with the fp32-rounded inputs
    xr, Wr = x.float().double(), W.float().double()
    ref = torch.zeros(n, desc.out_dim, dtype=torch.float64)
    for i in range(desc.nblk):
        b = desc.blk[i]
        xb = xr[:, b.in_off:b.in_off + b.mul_in * b.d].reshape(n, b.mul_in, b.d)
        Wb = Wr[b.w_off:b.w_off + b.mul_in * b.mul_out].reshape(b.mul_in, b.mul_out)
        ref[:, b.out_off:b.out_off + b.mul_out * b.d] = (torch.einsum('nkm,ko->nom', xb, Wb) * b.scale).reshape(n, -1)
    rel = ((y - ref) / ref)
    print(irin, '->', irout, 'mean rel err %.3e  max |rel| %.3e' % (rel.mean().item(), rel.abs().max().item()), flush=True)
