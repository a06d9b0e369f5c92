"""GPU parity of every C-ABI op against the CPU oracle (oracle/), float64 on the oracle side.

Tolerances are relative to the magnitude of the reference tensor: fp32 kernels vs an fp64
restatement, rel <= 2e-5 unless stated.  Integer/index work (graph, CSR) must be bit-exact.
"""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).abs().max() / max(1.0, float(b.abs().max())))


@pytest.fixture(scope='module')
def dev():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    return torch.device('cuda:0')


def _graph_cases():
    from sevennet_finetuning_b200.synthetic import si_diamond
    rng = np.random.default_rng(5)
    cases = []
    p, c, b, _ = si_diamond()
    cases.append(('si216', p, c, [True] * 3, 5.0))
    tri = np.array([[7.0, 0, 0], [1.5, 6.5, 0], [-0.8, 1.1, 8.0]])
    cases.append(('triclinic_unwrapped', rng.uniform(-3, 12, (60, 3)), tri, [True] * 3, 4.0))
    cases.append(('mixed_pbc', rng.uniform(-3, 12, (60, 3)), tri, [True, False, True], 4.0))
    cases.append(('non_periodic', rng.uniform(0, 9, (50, 3)), np.zeros((3, 3)), [False] * 3, 3.0))
    cases.append(('one_atom_small_cell', np.zeros((1, 3)), np.eye(3) * 3.0, [True] * 3, 5.0))
    cases.append(('two_atoms_tiny_cell', np.array([[0.1, 0.2, 0.3], [1.0, 1.1, 0.9]]), np.eye(3) * 2.2, [True] * 3, 5.0))
    big = rng.uniform(0, 30, (3000, 3))
    cases.append(('random_3000', big, np.eye(3) * 30.0, [True] * 3, 5.0))
    return cases


def test_radius_graph_bit_exact(dev, gold):
    from oracle.graph import radius_graph as ref_graph
    from sevennet_finetuning_b200 import ops
    for name, pos, cell, pbc, rc in _graph_cases():
        ei, S, D = ref_graph(pos, cell, pbc, rc)
        g = ops.radius_graph(torch.tensor(pos, dtype=torch.float64, device=dev), cell, pbc, rc)
        E = ei.shape[1]
        assert g['src'].shape[0] == E, (name, g['src'].shape[0], E)
        assert (g['dst'].cpu().numpy() == ei[0]).all(), name
        assert (g['src'].cpu().numpy() == ei[1]).all(), name
        assert (g['shift'].cpu().numpy() == S).all(), name
        assert (g['edge_vec'].cpu().numpy() == D.astype(np.float32)).all(), name
        rp = g['rowptr'].cpu().numpy()
        assert rp[0] == 0 and rp[-1] == E and (np.diff(rp) == np.bincount(ei[0], minlength=len(pos))).all(), name
    # golden edge sets produced from the reference run
    for fn in ('si216.npz', 'hfo2_96.npz'):
        gd = gold(fn)
        g = ops.radius_graph(torch.tensor(gd['pos'], dtype=torch.float64, device=dev), gd['cell'], [True] * 3, 5.0)
        assert (g['dst'].cpu().numpy() == gd['edge_index'][0]).all() and (g['src'].cpu().numpy() == gd['edge_index'][1]).all()
        assert (g['shift'].cpu().numpy() == gd['shift']).all()


def test_radius_graph_empty(dev):
    from sevennet_finetuning_b200 import ops
    g = ops.radius_graph(torch.zeros(0, 3, dtype=torch.float64, device=dev), np.eye(3) * 3, [True] * 3, 5.0)
    assert g['src'].shape[0] == 0 and g['rowptr'].cpu().tolist() == [0]


def test_csr_transpose(dev):
    from sevennet_finetuning_b200 import ops
    rng = np.random.default_rng(0)
    n, E = 500, 20000
    src = rng.integers(0, n, E).astype(np.int32)
    src[:300] = 7  # one long row (> 256 entries)
    rr, eid = ops.csr_transpose(n, torch.tensor(src, device=dev))
    order = np.argsort(src, kind='stable')
    assert (eid.cpu().numpy() == order).all()
    assert (rr.cpu().numpy() == np.concatenate([[0], np.cumsum(np.bincount(src, minlength=n))])).all()


def _rand_edges(rng, n_nodes, n_edges):
    dst = np.sort(rng.integers(0, n_nodes, n_edges))
    src = rng.integers(0, n_nodes, n_edges)
    rowptr = np.concatenate([[0], np.cumsum(np.bincount(dst, minlength=n_nodes))])
    vec = rng.normal(size=(n_edges, 3)) * 1.5
    vec *= (rng.uniform(0.8, 4.95, (n_edges, 1)) / np.linalg.norm(vec, axis=1, keepdims=True))
    return dst, src, rowptr, vec


@pytest.mark.parametrize('kind', ['xplor', 'poly'])
def test_edge_embedding_fwd_bwd(dev, kind):
    from oracle import refnn
    from sevennet_finetuning_b200 import ops
    from sevennet_finetuning_b200._lib import EdgeParams
    rng = np.random.default_rng(1)
    _, _, _, vec = _rand_edges(rng, 10, 4000)
    coeffs = np.array([0.5079, 1.1099, 1.7265, 2.3767, 3.0392, 3.7201, 4.4142, 5.1186])
    env = refnn.XPLORCutoff(5.0, 4.5) if kind == 'xplor' else refnn.PolynomialCutoff(5.0, 6)
    mod = refnn.EdgeEmbedding(refnn.BesselBasis(5.0, 8), env, refnn.SphericalEncoding(2, 1, True)).double()
    mod.basis_function.coeffs.data = torch.tensor(coeffs)
    v64 = torch.tensor(vec, dtype=torch.float64, requires_grad=True)
    d = mod({'edge_vec': v64})
    gY = torch.tensor(rng.normal(size=(len(vec), 9)))
    gE = torch.tensor(rng.normal(size=(len(vec), 8)))
    (d['edge_attr'] * gY).sum().backward(retain_graph=True)
    gv_Y = v64.grad.clone(); v64.grad = None
    (d['edge_embedding'] * gE).sum().backward()
    gv_E = v64.grad.clone()
    p = EdgeParams(8, 1 if kind == 'xplor' else 0, 1, 5.0, 4.5, 6.0)
    vec32 = torch.tensor(vec, dtype=torch.float32, device=dev)
    c32 = torch.tensor(coeffs, dtype=torch.float32, device=dev)
    ln, Y, emb = ops.edge_embed_fwd(p, vec32, c32)
    assert _rel(ln, d['edge_length']) < 1e-6
    assert _rel(Y, d['edge_attr']) < 2e-6
    assert _rel(emb, d['edge_embedding']) < 5e-6
    gY32, gE32 = gY.float().to(dev).contiguous(), gE.float().to(dev).contiguous()
    assert _rel(ops.edge_embed_bwd(p, vec32, c32, gY32, None), gv_Y) < 1e-5
    assert _rel(ops.edge_embed_bwd(p, vec32, c32, None, gE32), gv_E) < 1e-5
    assert _rel(ops.edge_embed_bwd(p, vec32, c32, gY32, gE32), gv_Y + gv_E) < 1e-5


@pytest.mark.parametrize('path', ['tcgen05', 'ffma'])
@pytest.mark.parametrize('wn,E', [(960, 1000), (384, 130), (224, 64), (100, 77), (960, 128), (960, 129)])
def test_radial_mlp_fwd_bwd(dev, wn, E, path, monkeypatch):
    """both implementations of the radial MLP: tcgen05 3xTF32 (default) and the FFMA kernels"""
    from oracle import refnn
    from sevennet_finetuning_b200 import ops
    monkeypatch.setenv('B200_RADIAL_FFMA', '1' if path == 'ffma' else '0')
    torch.manual_seed(0)
    net = refnn.FullyConnectedNet([8, 64, 64, wn], 'silu').double()
    emb = (torch.randn(E, 8, dtype=torch.float64) * 0.3).requires_grad_(True)
    w = net(emb)
    g = torch.randn(E, wn, dtype=torch.float64)
    (w * g).sum().backward()
    Ws = [getattr(net, f'layer{k}').weight.detach().float().to(dev).contiguous() for k in range(3)]
    e32 = emb.detach().float().to(dev)
    w32 = ops.radial_mlp_fwd(e32, *Ws)
    assert _rel(w32, w) < 2e-5
    de = ops.radial_mlp_bwd(e32, *Ws, g.float().to(dev).contiguous())
    assert _rel(de, emb.grad) < 2e-5
    de2 = ops.radial_mlp_bwd(e32, *Ws, g.float().to(dev).contiguous(), demb=de.clone())
    assert _rel(de2, 2 * emb.grad) < 2e-5


CONV_CASES = [((128, 0, 0), 2, 'layer0'), ((128, 64, 32), 2, 'mid'), ((128, 64, 32), 0, 'last'),
              ((32, 16, 8), 2, 'small'), ((48, 40, 0), 2, 'no_l2'), ((20, 0, 12), 2, 'no_l1')]


@pytest.mark.parametrize('muls,lout,name', CONV_CASES)
def test_convolution_fwd_bwd(dev, muls, lout, name):
    from oracle import refnn
    from sevennet_finetuning_b200 import ops
    rng = np.random.default_rng(2)
    n, E = 37, 1500
    dst, src, rowptr, vec = _rand_edges(rng, n, E)
    irx = '+'.join(f'{m}x{l}e' for l, m in enumerate(muls) if m)
    tp = refnn.TensorProductUVU(irx, '1x0e+1x1e+1x2e', '+'.join(f'1x{l}e' for l in range(lout + 1))).double()
    wn, od = ops.conv_dims(*muls, lout)
    assert wn == tp.weight_numel and od == refnn.irreps_dim(tp.irreps_out)
    sph = refnn.SphericalEncoding(2, 1, True)
    Y = sph(torch.tensor(vec))
    x = torch.randn(n, refnn.irreps_dim(tp.irreps_in1), dtype=torch.float64, requires_grad=True)
    w = torch.randn(E, wn, dtype=torch.float64, requires_grad=True)
    Yr = Y.clone().requires_grad_(True)
    denom = 7.3
    msg = tp(x[torch.tensor(src)], Yr, w)
    out = refnn.message_gather(x, torch.tensor(dst), msg) / denom
    g = torch.randn_like(out)
    (out * g).sum().backward()
    d = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a)).to(dt).to(dev).contiguous()
    rp, s32, d32 = d(rowptr, torch.int32), d(src, torch.int32), d(dst, torch.int32)
    x32, Y32, w32, g32 = d(x.detach()), d(Y), d(w.detach()), d(g)
    o = ops.conv_fwd(muls, lout, rp, s32, x32, Y32, w32, denom)
    assert _rel(o, out) < 2e-5, name
    dw, dY = ops.conv_bwd_edge(muls, lout, rp, s32, x32, Y32, w32, g32, denom)
    assert _rel(dw, w.grad) < 2e-5, name
    assert _rel(dY[:, 1:], Yr.grad[:, 1:]) < 2e-5, name
    dw2, dY2 = ops.conv_bwd_edge(muls, lout, rp, s32, x32, Y32, w32, g32, denom, dY=dY.clone())
    assert _rel(dY2[:, 1:], 2 * Yr.grad[:, 1:]) < 2e-5, name
    rr, eid = ops.csr_transpose(n, s32)
    dx = ops.conv_bwd_x(muls, lout, n, rr, eid, d32, g32, Y32, w32, denom)
    assert _rel(dx, x.grad) < 2e-5, name
    # one pass for all three gradients (d x accumulated with RED.ADD)
    dw3, dY3, dx3 = ops.conv_bwd(muls, lout, n, rp, s32, x32, Y32, w32, g32, denom)
    assert _rel(dw3, w.grad) < 2e-5 and _rel(dY3[:, 1:], Yr.grad[:, 1:]) < 2e-5 and _rel(dx3, x.grad) < 2e-5, name


def test_convolution_empty_rows_and_long_rows(dev):
    """ragged input: atoms without neighbours, and a row longer than one staging chunk"""
    from oracle import refnn
    from sevennet_finetuning_b200 import ops
    rng = np.random.default_rng(3)
    n = 6
    dst = np.array([1] * 70 + [4] * 3)
    src = rng.integers(0, n, len(dst))
    rowptr = np.concatenate([[0], np.cumsum(np.bincount(dst, minlength=n))])
    vec = rng.normal(size=(len(dst), 3)) * 2
    muls, lout = (128, 64, 32), 2
    tp = refnn.TensorProductUVU('128x0e+64x1e+32x2e', '1x0e+1x1e+1x2e', '1x0e+1x1e+1x2e').double()
    Y = refnn.SphericalEncoding(2, 1, True)(torch.tensor(vec))
    x = torch.randn(n, 480, dtype=torch.float64)
    w = torch.randn(len(dst), 960, dtype=torch.float64)
    out = refnn.message_gather(x, torch.tensor(dst), tp(x[torch.tensor(src)], Y, w))
    d = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a)).to(dt).to(dev).contiguous()
    o = ops.conv_fwd(muls, lout, d(rowptr, torch.int32), d(src, torch.int32), d(x), d(Y), d(w), 1.0)
    assert _rel(o, out) < 2e-5
    assert float(o[0].abs().max()) == 0.0 and float(o[5].abs().max()) == 0.0


LIN_CASES = [('128x0e+64x1e+32x2e', '224x0e+64x1e+32x2e'), ('224x0e+384x1e+352x2e', '224x0e+64x1e+32x2e'),
             ('128x0e', '224x0e+64x1e+32x2e'), ('128x0e+64x1e+32x2e', '128x0e'), ('128x0e', '64x0e'), ('64x0e', '1x0e'),
             ('20x0e+7x1e+5x2e', '33x0e+70x1e+3x2e')]


@pytest.mark.parametrize('irin,irout', LIN_CASES)
def test_irreps_linear(dev, irin, irout):
    from oracle import refnn
    from sevennet_finetuning_b200 import ops
    torch.manual_seed(1)
    n = 131
    lin = refnn.O3Linear(irin, irout).double()
    x = torch.randn(n, refnn.irreps_dim(lin.irreps_in), dtype=torch.float64, requires_grad=True)
    y = lin(x)
    g = torch.randn_like(y)
    (y * g).sum().backward()
    desc, numel = ops.make_lin_desc(irin, irout)
    assert numel == lin.weight.numel()
    W = lin.weight.detach().float().to(dev)
    x32, g32 = x.detach().float().to(dev), g.float().to(dev)
    assert _rel(ops.irreps_linear(desc, x32, W), y) < 2e-5
    assert _rel(ops.irreps_linear(desc, g32, W, transpose=True), x.grad) < 2e-5
    assert _rel(ops.irreps_linear_wgrad(desc, x32, g32, numel), lin.weight.grad) < 2e-5
    y2 = ops.irreps_linear(desc, x32, W, out=ops.irreps_linear(desc, x32, W))
    assert _rel(y2, 2 * y) < 2e-5


@pytest.mark.parametrize('irin,irout,n', [('64x0e', '960x0e', 30011), ('8x0e', '64x0e', 5000),
                                          ('20x0e+7x1e+5x2e', '33x0e+70x1e+3x2e', 4099), ('128x0e+64x1e+32x2e', '224x0e+64x1e+32x2e', 2048)])
def test_irreps_linear_wgrad_many_rows(dev, irin, irout, n):
    """split-K weight gradient (rows = edges of a training batch for the radial MLP layers) against float64"""
    from oracle import refnn
    from sevennet_finetuning_b200 import ops
    torch.manual_seed(2)
    lin = refnn.O3Linear(irin, irout).double()
    x = torch.randn(n, refnn.irreps_dim(lin.irreps_in), dtype=torch.float64)
    y = lin(x)
    g = torch.randn_like(y)
    (y * g).sum().backward()
    desc, numel = ops.make_lin_desc(irin, irout)
    x32, g32 = x.float().to(dev), g.float().to(dev)
    dW = ops.irreps_linear_wgrad(desc, x32, g32, numel)
    assert _rel(dW, lin.weight.grad) < 2e-5
    assert torch.equal(dW, ops.irreps_linear_wgrad(desc, x32, g32, numel))   # fixed reduction order


@pytest.mark.parametrize('irx', ['128x0e+64x1e+32x2e', '128x0e', '16x0e+3x1e', '16x0e+5x2e'])
def test_gate(dev, irx):
    from oracle import refnn
    from sevennet_finetuning_b200 import ops
    torch.manual_seed(2)
    gate = refnn.EquivariantGate(irx)
    n = 77
    din = refnn.irreps_dim(gate.irreps_in)
    x = torch.randn(n, din, dtype=torch.float64, requires_grad=True)
    y = gate({'x': x})['x']
    g = torch.randn_like(y)
    (y * g).sum().backward()
    m = ops.muls_of(irx)
    x32 = x.detach().float().to(dev)
    assert _rel(ops.gate_fwd(m[0], m[1], m[2], x32), y) < 1e-5
    assert _rel(ops.gate_bwd(m[0], m[1], m[2], x32, g.float().to(dev)), x.grad) < 1e-5


FUSED_CASES = [((128, 64, 32), 2, 'mid', 1500), ((128, 0, 0), 2, 'layer0', 700), ((128, 64, 32), 0, 'last', 1500),
               ((64, 32, 32), 2, 'half', 333), ((128, 64, 32), 2, 'one_tile', 64), ((128, 64, 32), 2, 'tail', 129)]


@pytest.mark.parametrize('cutoff,wn', [('xplor', 960), ('xplor', 224), ('poly', 384)])
def test_radial_table_matches_float64_mlp(dev, cutoff, wn):
    """Tabulated radial weights (b200_radial_table_*): w(r) and sum_k dw_k dw_k/dr against the float64 oracle chain
    BesselBasis * cutoff -> FullyConnectedNet (edge_embedding.py:114-116,138-173; convolution.py:97-106)."""
    from oracle import refnn
    from sevennet_finetuning_b200 import ops
    from sevennet_finetuning_b200._lib import B200_CUTOFF_POLY, B200_CUTOFF_XPLOR, EdgeParams
    rng = np.random.default_rng(11)
    E = 4001
    r = np.concatenate([rng.uniform(0.6, 4.9999, E - 4), [4.5, 4.5 - 1e-6, 4.5 + 1e-6, 4.99999]])
    r = r.astype(np.float32).astype(np.float64)   # the kernel sees fp32 lengths
    torch.set_default_dtype(torch.float64)
    try:
        torch.manual_seed(3)
        env = refnn.XPLORCutoff(5.0, 4.5) if cutoff == 'xplor' else refnn.PolynomialCutoff(5.0, 6)
        basis = refnn.BesselBasis(5.0, 8).double()
        net = refnn.FullyConnectedNet([8, 64, 64, wn], 'silu').double()
        r64 = torch.tensor(r, requires_grad=True)
        w64 = net(basis(r64) * env(r64).unsqueeze(-1))
        g64 = torch.tensor(rng.normal(size=(E, wn)))
        (dr64,) = torch.autograd.grad((w64 * g64).sum(), r64)
    finally:
        torch.set_default_dtype(torch.float32)
    p = EdgeParams(8, B200_CUTOFF_XPLOR if cutoff == 'xplor' else B200_CUTOFF_POLY, 1, 5.0, 4.5, 6.0)
    Ws = [getattr(net, f'layer{k}').weight.detach().float().to(dev).contiguous() for k in range(3)]
    T = ops.radial_table_build(p, basis.coeffs.detach().float().to(dev), *Ws)
    ln = torch.tensor(r, dtype=torch.float32, device=dev)
    order = torch.argsort((ln * 800).int(), stable=True).to(torch.int32)
    w = ops.radial_table_fwd(p, wn, 8, 64, 64, ln, T)
    assert torch.equal(w, ops.radial_table_fwd(p, wn, 8, 64, 64, ln, T, order=order))   # visiting order is not arithmetic
    # fp32 rounding of r itself moves w by |dw/dr| * ulp(r): compare at that scale
    assert _rel(w, w64.detach()) < 5e-6
    dr = ops.radial_table_bwd(p, wn, 8, 64, 64, ln, T, g64.float().to(dev), order=order)
    assert _rel(dr, dr64) < 1e-5
    dr2 = ops.radial_table_bwd(p, wn, 8, 64, 64, ln, T, g64.float().to(dev), dr=dr.clone())
    assert _rel(dr2, 2 * dr64) < 1e-5
