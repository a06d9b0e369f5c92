"""Accuracy of the tabulated radial weights vs the tensor-core MLP kernel, both against the float64 oracle, with the real
SevenNet-0 weights (tests/golden/sevennet0_weights.npz)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import refnn
from sevennet_finetuning_b200 import ops
from sevennet_finetuning_b200._lib import B200_CUTOFF_XPLOR, EdgeParams

dev = torch.device('cuda:0')
w = dict(np.load('tests/golden/sevennet0_weights.npz'))
rng = np.random.default_rng(0)
r = rng.uniform(0.7, 4.9999, 20000).astype(np.float32)
p = EdgeParams(8, B200_CUTOFF_XPLOR, 1, 5.0, 4.5, 6.0)
co = torch.tensor(w['edge_embedding.basis_function.coeffs'])
for t in range(5):
    Ws = [torch.tensor(w[f'{t}_convolution.weight_nn.layer{k}.weight']) for k in range(3)]
    r64 = torch.tensor(r.astype(np.float64), requires_grad=True)
    k64 = co.double()
    emb = (2.0 / 5.0) * torch.sin(k64 * r64[:, None]) / r64[:, None]
    rc2, ron2, r2 = 25.0, 20.25, r64 * r64
    f = torch.where(r64 < 4.5, torch.ones_like(r64), (rc2 - r2) ** 2 * (rc2 + 2 * r2 - 3 * ron2) / (rc2 - ron2) ** 3)
    x = emb * f[:, None]
    c = 1.6791767923989418
    for k, W in enumerate(Ws):
        x = x @ (W.double() / np.sqrt(W.shape[0]))
        if k < 2:
            x = c * x * torch.sigmoid(x)
    w64 = x.detach()
    g = torch.tensor(rng.normal(size=w64.shape))
    (dr64,) = torch.autograd.grad((x * g).sum(), r64)
    Wd = [W.to(dev).contiguous() for W in Ws]
    ln = torch.tensor(r, device=dev)
    T = ops.radial_table_build(p, co.to(dev), *Wd)
    wt = ops.radial_table_fwd(p, Wd[2].shape[1], 8, 64, 64, ln, T).double().cpu()
    vec = torch.zeros(len(r), 3, device=dev); vec[:, 0] = ln
    _, _, emb32 = ops.edge_embed_fwd(p, vec, co.to(dev))
    wm = ops.radial_mlp_fwd(emb32, *Wd).double().cpu()
    drt = ops.radial_table_bwd(p, Wd[2].shape[1], 8, 64, 64, ln, T, g.float().to(dev)).double().cpu()
    demb = ops.radial_mlp_bwd(emb32, *Wd, g.float().to(dev).contiguous())
    dvec = ops.edge_embed_bwd(p, vec, co.to(dev), None, demb)
    drm = dvec[:, 0].double().cpu()
    sc = float(w64.abs().max())
    print(f'layer {t}: |w|max {sc:.3f}  table max err {float((wt - w64).abs().max()):.2e} rms {float((wt - w64).pow(2).mean().sqrt()):.2e} | '
          f'mlp kernel max err {float((wm - w64).abs().max()):.2e} rms {float((wm - w64).pow(2).mean().sqrt()):.2e} || d/dr: |max| {float(dr64.abs().max()):.1f} '
          f'table {float((drt - dr64).abs().max()):.2e}  mlp {float((drm - dr64).abs().max()):.2e}')
