"""Train-mode radial table (csrc/table_train.cu) and the *_scaled convolution passes against float64 restatements written
with plain tensor operations; the closed autograd family of higher_order.py against autograd of that restatement.

Reference semantics: sevenn/nn/convolution.py:97-106 (weight = weight_nn(edge_embedding)) under the create_graph force pass of
sevenn/nn/force_output.py:165-170 and loss.backward() of sevenn/train/trainer.py:173-186.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def dev():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    return torch.device('cuda:0')


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-30))


INV_H = 800.0


def _basis(m, t):
    z, o = torch.zeros_like(t), torch.ones_like(t)
    q = {0: [o, t, t * t, t ** 3], 1: [z, o * INV_H, 2 * t * INV_H, 3 * t * t * INV_H],
         2: [z, z, 2 * o * INV_H ** 2, 6 * t * INV_H ** 2], 3: [z, z, z, 6 * o * INV_H ** 3], 4: [z, z, z, z]}[m]
    return torch.stack(q, 1)   # [E, 4]


def _ref_interp(m, P, r):
    """float64: d^m/dr^m of the piecewise cubic, located like the kernel (fp32 product r * inv_h)."""
    n = P.shape[0]
    u = (r.float() * INV_H)
    i = u.to(torch.int64).clamp(0, n - 1)
    t = (r.float().double() * INV_H - i.double())
    return torch.einsum('ej,ejk->ek', _basis(m, t), P.double()[i]), i, t


def _case(dev, E=3000, n=4001, wn=96, seed=0):
    g = torch.Generator(device='cpu').manual_seed(seed)
    P = torch.randn(n, 4, wn, generator=g).to(dev)
    r = (torch.rand(E, generator=g) * 4.9 + 0.05).to(dev)
    order = torch.argsort((r * INV_H).to(torch.int32)).to(torch.int32)
    return P, r, order


@pytest.mark.parametrize('m', [0, 1, 2, 3, 4])
def test_table_interp_dot_scatter(dev, m):
    from sevennet_finetuning_b200 import ops
    P, r, order = _case(dev)
    E, wn = r.shape[0], P.shape[2]
    gen = torch.Generator(device='cpu').manual_seed(1)
    g = torch.randn(E, wn, generator=gen).to(dev)
    coef = torch.randn(E, generator=gen).to(dev)
    ref, i, t = _ref_interp(m, P, r)
    for od in (None, order):
        assert _rel(ops.table_interp(m, P, INV_H, r, od), ref) < 2e-6
        assert _rel(ops.table_interp(m, P, INV_H, r, od, coef), ref * coef.double().unsqueeze(1)) < 2e-6
        assert _rel(ops.table_dot(m, P, INV_H, r, g, od), (ref * g.double()).sum(1)) < 2e-5
        Gref = torch.zeros(P.shape, dtype=torch.float64, device=dev)
        Gref.index_add_(0, i, torch.einsum('ej,ek->ejk', _basis(m, t) * coef.double().unsqueeze(1), g.double()))
        G = ops.table_scatter(m, P.shape[0], INV_H, g, r, od, coef)
        assert _rel(G, Gref) < 2e-5
        md = min(m + 1, 4)
        G2, d2 = ops.table_scatter_dot(m, md, P, INV_H, g, r, od, coef)
        assert _rel(G2, Gref) < 2e-5
        assert _rel(d2, (_ref_interp(md, P, r)[0] * g.double()).sum(1)) < 2e-5


def test_table_edge_cases(dev):
    from sevennet_finetuning_b200 import ops
    P, r, _ = _case(dev, E=5, n=16, wn=8)
    empty = torch.empty(0, device=dev)
    assert ops.table_interp(0, P, INV_H, empty).shape == (0, 8)
    assert float(ops.table_scatter(0, 16, INV_H, torch.empty(0, 8, device=dev), empty).abs().max()) == 0.0
    # a length past the last grid point is evaluated on the last interval (t > 1), one at zero on the first
    rr = torch.tensor([0.0, 16.5 / INV_H], device=dev)
    out = ops.table_interp(0, P, INV_H, rr)
    t = 16.5 - 15.0
    ref1 = (P[15].double() * torch.tensor([1.0, t, t * t, t ** 3], dtype=torch.float64, device=dev).unsqueeze(1)).sum(0)
    assert _rel(out[0], P[0, 0]) < 1e-6 and _rel(out[1], ref1) < 5e-6
    with pytest.raises(RuntimeError):
        ops.table_interp(5, P, INV_H, rr)


def test_hermite_and_adjoint(dev):
    from sevennet_finetuning_b200 import ops
    gen = torch.Generator(device='cpu').manual_seed(2)
    n, wn, h = 57, 40, 1.25e-3
    Z = torch.randn(n + 1, wn, dtype=torch.float64, generator=gen).to(dev)
    dZ = torch.randn(n + 1, wn, dtype=torch.float64, generator=gen).to(dev)
    P = ops.table_hermite(Z, dZ, h)
    v0, v1, s0, s1 = Z[:-1], Z[1:], h * dZ[:-1], h * dZ[1:]
    ref = torch.stack([v0, s0, 3 * (v1 - v0) - 2 * s0 - s1, s0 + s1 - 2 * (v1 - v0)], 1)
    assert _rel(P, ref) < 1e-7
    # the interpolant reproduces the nodal values and slopes
    assert _rel(P[:, 0] + P[:, 1] + P[:, 2] + P[:, 3], v1) < 1e-6
    assert _rel(P[:, 1] + 2 * P[:, 2] + 3 * P[:, 3], s1) < 2e-6 * max(1.0, float(ref.abs().max() / s1.abs().max()))
    G = torch.randn(n, 4, wn, generator=gen).to(dev)
    gZ, gdZ = ops.table_hermite_adj(G, h)
    lhs = float((ref * G.double()).sum())
    rhs = float((Z * gZ).sum() + (dZ * gdZ).sum())
    assert abs(lhs - rhs) <= 1e-9 * max(abs(lhs), 1.0)


def test_table_family_second_order_matches_autograd(dev):
    """loss = f(w(P, r)) + |d f / d r|^2-like term: gradients with respect to P and r through TabInterp / TabDot / TabScatter,
    against autograd of the float64 restatement (gather + polynomial)."""
    from sevennet_finetuning_b200 import higher_order as ho
    P, r, order = _case(dev, E=700, n=4001, wn=32, seed=3)
    gen = torch.Generator(device='cpu').manual_seed(4)
    A = torch.randn(r.shape[0], 32, generator=gen).to(dev)

    def loss_of(w_fn, P_, r_):
        w = w_fn(P_, r_)
        e = (torch.tanh(w) * A.to(w.dtype)).sum()
        (dr,) = torch.autograd.grad(e, r_, create_graph=True)
        return e + (dr * dr).sum() * 1e-6, dr

    Pk, rk = P.clone().requires_grad_(True), r.clone().requires_grad_(True)
    lk, drk = loss_of(lambda P_, r_: ho.TabInterp.apply(P_, r_, None, 0, (INV_H, order)), Pk, rk)
    gPk, grk = torch.autograd.grad(lk, [Pk, rk])

    Pr = P.double().requires_grad_(True)
    rr = r.double().requires_grad_(True)
    n = P.shape[0]
    i = (r.float() * INV_H).to(torch.int64).clamp(0, n - 1)

    def ref_w(P_, r_):
        t = r_ * INV_H - i.double()
        return torch.einsum('ej,ejk->ek', torch.stack([torch.ones_like(t), t, t * t, t ** 3], 1), P_[i])

    lr, drr = loss_of(ref_w, Pr, rr)
    gPr, grr = torch.autograd.grad(lr, [Pr, rr])
    assert abs(float(lk) - float(lr)) <= 2e-5 * abs(float(lr))
    assert _rel(drk, drr) < 2e-5
    assert _rel(gPk, gPr) < 1e-4
    assert _rel(grk, grr) < 1e-4


@pytest.mark.parametrize('muls,lout', [((32, 16, 8), 2), ((32, 16, 8), 0), ((16, 0, 0), 2)])
def test_conv_scaled_passes(dev, muls, lout):
    """conv_fwd_scaled / conv_bwd_scaled == the plain passes on explicitly masked weights (w * m, then d w * m), with the
    weight gradient overwritten, added into, or left out."""
    from sevennet_finetuning_b200 import ops
    from sevennet_finetuning_b200 import higher_order as ho
    rng = np.random.default_rng(7)
    n, deg = 40, 9
    dst = np.repeat(np.arange(n), deg)
    src = rng.integers(0, n, n * deg)
    rowptr = torch.tensor(np.arange(n + 1) * deg, dtype=torch.int32, device=dev)
    src_t = torch.tensor(src, dtype=torch.int32, device=dev)
    wn, od = ops.conv_dims(*muls, lout)
    xdim = muls[0] + 3 * muls[1] + 5 * muls[2]
    E = n * deg
    x = torch.tensor(rng.normal(size=(n, xdim)), dtype=torch.float32, device=dev)
    Y = torch.tensor(rng.normal(size=(E, 9)), dtype=torch.float32, device=dev)
    Y[:, 0] = 1.0
    w = torch.tensor(rng.normal(size=(E, wn)), dtype=torch.float32, device=dev)
    g = torch.tensor(rng.normal(size=(n, od)), dtype=torch.float32, device=dev)
    scale = ho._y_scale()
    m = ho._y_mask((muls, lout), w)
    ref_out = ops.conv_fwd(muls, lout, rowptr, src_t, x, Y, w * m, 3.0)
    assert _rel(ops.conv_fwd_scaled(muls, lout, rowptr, src_t, x, Y, w, 3.0, scale), ref_out) < 1e-6
    rdw, rdY, rdx = ops.conv_bwd(muls, lout, n, rowptr, src_t, x, Y, w * m, g, 3.0)
    dw, dY, dx = ops.conv_bwd_scaled(muls, lout, n, rowptr, src_t, x, Y, w, g, 3.0, path_scale=scale)
    assert _rel(dw, rdw * m) < 1e-6 and _rel(dY, rdY) < 1e-5 and _rel(dx, rdx) < 1e-5
    base = torch.tensor(rng.normal(size=(E, wn)), dtype=torch.float32, device=dev)
    acc = base.clone()
    dw2, dY2, dx2 = ops.conv_bwd_scaled(muls, lout, n, rowptr, src_t, x, Y, w, g, 3.0, path_scale=scale, dw=acc, dw_mode=1,
                                        dY=dY.clone(), want_dx=False)
    assert dw2.data_ptr() == acc.data_ptr() and dx2 is None
    assert _rel(acc, base + rdw * m) < 1e-6 and _rel(dY2, 2 * rdY) < 1e-5
    dw3, dY3, dx3 = ops.conv_bwd_scaled(muls, lout, n, rowptr, src_t, x, Y, w, g, 3.0, dw_mode=2)
    rdw0, rdY0, rdx0 = ops.conv_bwd(muls, lout, n, rowptr, src_t, x, Y, w, g, 3.0)
    assert dw3 is None and _rel(dY3, rdY0) < 1e-5 and _rel(dx3, rdx0) < 1e-5
