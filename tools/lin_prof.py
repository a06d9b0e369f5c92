import ctypes as C, sys, torch
sys.path.insert(0, '.')
from sevennet_finetuning_b200 import ops
from sevennet_finetuning_b200._lib import lib
dev = torch.device('cuda:0')
n = 100000
desc, numel = ops.make_lin_desc('224x0e+384x1e+352x2e', '224x0e+64x1e+32x2e')
x = torch.randn(n, desc.in_dim, device=dev); W = torch.randn(numel, device=dev)
for _ in range(3): ops.irreps_linear(desc, x, W)
torch.cuda.synchronize()
out = (C.c_ulonglong * 32)()
L = lib()
print('rc', L.b200_debug_lin_prof(out))
it = max(1, out[6])
print('issuer', {nm: round(out[k] / it) for k, nm in enumerate(['wait mma_done(c-2)', 'B request + wait a_full', 'wait full_b', 'MMA issue + commit'])}, 'iters', out[6], 'sum', round(sum(out[k] for k in range(4)) / it))
it = max(1, out[14])
print('worker', {nm: round(out[8 + k] / it) for k, nm in enumerate(['wait mma_done', 'drain', 'stage+read+split+STS', 'fence+arrive', 'cp.async wait', 'worker barrier'])}, 'iters', out[14], 'sum', round(sum(out[8 + k] for k in range(6)) / it))
