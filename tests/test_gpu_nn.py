"""The sevenn.nn twins (sevennet_finetuning_b200/nn.py): same module names / state dict as the reference, dict-in
dict-out forward, energy/forces/stress parity with the golden vectors, and first-order autograd (parameter
gradients of the energy) against the float64 oracle model."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _data(g, dev, shuffle=False):
    from oracle.graph import radius_graph
    ei, S, D = radius_graph(g['pos'], g['cell'], [True] * 3, 5.0)
    if shuffle:   # the reference does not promise any edge order
        p = np.random.default_rng(0).permutation(ei.shape[1])
        ei, D = ei[:, p], D[p]
    return {'x': torch.tensor(g['types'], dtype=torch.long, device=dev), 'edge_index': torch.tensor(ei, device=dev),
            'edge_vec': torch.tensor(D, dtype=torch.float32, device=dev).requires_grad_(True),
            'num_atoms': torch.tensor(len(g['pos']), device=dev),
            'cell_volume': torch.tensor(abs(np.linalg.det(g['cell'])), dtype=torch.float32, device=dev),
            'batch': torch.zeros(len(g['pos']), dtype=torch.long, device=dev)}, ei, D


def _oracle(weights):
    from oracle.refmodel import build_model
    m = build_model(dict(shift=weights['rescale_atomic_energy.shift'].tolist(), scale=weights['rescale_atomic_energy.scale'].tolist(),
                         conv_denominator=float(weights['0_convolution.denominator'][0])))
    m.load_state_dict({k: torch.tensor(v) for k, v in weights.items()})
    return m.double()


@pytest.fixture(scope='module')
def twin(weights):
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from sevennet_finetuning_b200 import nn as b200nn
    m = b200nn.build_E3_equivariant_model(dict(shift=weights['rescale_atomic_energy.shift'].tolist(),
                                               scale=weights['rescale_atomic_energy.scale'].tolist(),
                                               conv_denominator=float(weights['0_convolution.denominator'][0])))
    missing = m.load_state_dict({k: torch.tensor(v) for k, v in weights.items()}, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return m.cuda()


@pytest.mark.parametrize('name,shuffle', [('si216.npz', False), ('hfo2_96.npz', True)])
def test_module_chain_efs(twin, gold, name, shuffle):
    g = gold(name)
    dev = torch.device('cuda:0')
    twin.eval()
    twin.set_is_batch_data(False)
    data, ei, D = _data(g, dev, shuffle)
    out = twin(data)
    n = len(g['pos'])
    e = float(out['inferred_total_energy'])
    f = out['inferred_force'].detach().cpu().numpy()
    s = out['inferred_stress'].detach().cpu().numpy()
    assert abs(e - float(g['o64_energy'])) / n <= 1e-6
    assert np.abs(out['atomic_energy'].detach().cpu().numpy().reshape(-1) - g['o64_atomic_energy'].reshape(-1)).max() <= 1e-5
    assert np.abs(f - g['o64_force']).max() <= 1e-5
    assert np.abs(s - g['o64_stress']).max() <= 1e-4 * np.abs(g['o64_stress']).max()
    assert 'self_cont_tmp' not in out and '_b200_graph' not in out
    # batched flavour of the same call
    twin.set_is_batch_data(True)
    data, _, _ = _data(g, dev, shuffle)
    data['cell_volume'] = data['cell_volume'].view(1)
    out = twin(data)
    assert out['inferred_total_energy'].shape == (1,) and out['inferred_stress'].shape == (1, 6)
    assert abs(float(out['inferred_total_energy'][0]) - e) / n <= 2e-6   # fp32 sum order (1 ulp of 1158 eV = 1.2e-4)


def test_parameter_gradients_of_energy(twin, gold, weights):
    """d E_total / d theta for every parameter, against autograd through the float64 oracle model"""
    g = gold('hfo2_96.npz')
    dev = torch.device('cuda:0')
    twin.train()
    twin.set_is_batch_data(False)
    for p in twin.parameters():
        p.requires_grad_(True)
        p.grad = None
    data, ei, D = _data(g, dev)
    chain = [m for k, m in twin._modules.items() if k != 'force_output']
    for m in chain:
        data = m(data)
    data['inferred_total_energy'].backward()
    torch.set_default_dtype(torch.float64)   # the oracle builds its tables and one-hot rows in the default dtype
    try:
        ref = _oracle(weights)
        ref.set_is_batch_data(False)
        for p in ref.parameters():
            p.requires_grad_(True)
        d = {'x': torch.tensor(g['types']), 'edge_index': torch.tensor(ei),
             'edge_vec': torch.tensor(D, dtype=torch.float64).requires_grad_(True), 'num_atoms': torch.tensor(len(g['pos'])),
             'cell_volume': torch.tensor(abs(np.linalg.det(g['cell']))), 'batch': torch.zeros(0, dtype=torch.long)}
        for k, m in ref._modules.items():
            if k != 'force_output':
                d = m(d)
        d['inferred_total_energy'].backward()
    finally:
        torch.set_default_dtype(torch.float32)
    refg = dict(ref.named_parameters())
    worst = 0.0
    for k, p in twin.named_parameters():
        assert p.grad is not None, k
        a, b = p.grad.double().cpu().reshape(-1), refg[k].grad.reshape(-1)
        rel = float((a - b).abs().max() / max(float(b.abs().max()), 1e-12))
        worst = max(worst, rel)
        assert rel < 5e-4, (k, rel)
    print('worst relative parameter-gradient error', worst)
    twin.eval()


def _finetune_loss(out, lab, n):
    """Huber(delta=0.01) on E/atom, forces, stress [kbar] with the shipped fine-tuning weights
    (example_inputs/fine_tuning/FT_w_reEWC/input_full.yaml:41-56; sevenn/train/loss.py:96-214,268-292)."""
    hub = torch.nn.HuberLoss(delta=0.01)
    le = hub(out['inferred_total_energy'].reshape(-1) / n, lab['e'].reshape(-1) / n)
    lf = hub(out['inferred_force'].reshape(-1), lab['f'].reshape(-1))
    ls = hub(out['inferred_stress'].reshape(-1) * 1602.1766208, lab['s'].reshape(-1) * 1602.1766208)
    return le + 1.0 * lf + 0.01 * ls


@pytest.mark.parametrize('train_table', [True, False])
def test_parameter_gradients_of_force_stress_loss(twin, gold, weights, train_table, monkeypatch):
    """The fine-tuning gradient: d L(E, F, S) / d theta with F = -dE/dr taken with create_graph=True
    (sevenn/nn/force_output.py:165-170, sevenn/train/trainer.py:173-186) - a reverse pass through the reverse pass of
    every kernel - against double-precision autograd through the oracle model.  train_table: weight_nn read from the per-step
    table over the edge length (csrc/table_train.cu, the default) or evaluated once per edge (B200_TRAIN_TABLE=0)."""
    import sevennet_finetuning_b200.nn as b200nn
    monkeypatch.setattr(b200nn, '_TRAIN_TABLE', train_table)
    g = gold('hfo2_96.npz')
    dev = torch.device('cuda:0')
    n = len(g['pos'])
    rng = np.random.default_rng(5)
    lab64 = {'e': torch.tensor(float(g['o64_energy']) + 0.3, dtype=torch.float64), 'f': torch.tensor(g['o64_force'] + rng.normal(0, 0.05, (n, 3))),
             's': torch.tensor(g['o64_stress'].reshape(-1) + rng.normal(0, 1e-3, 6))}
    twin.train()
    twin.set_is_batch_data(False)
    for p in twin.parameters():   # every parameter, including shift / scale / conv denominators
        p.requires_grad_(True)
        p.grad = None
    data, ei, D = _data(g, dev)
    out = twin(data)
    assert out['inferred_force'].requires_grad          # the forces stay in the graph in train() mode
    lab32 = {k: v.to(dev, torch.float32) for k, v in lab64.items()}
    loss = _finetune_loss(out, lab32, n)
    loss.backward()
    torch.set_default_dtype(torch.float64)
    try:
        ref = _oracle(weights)
        ref.train()
        ref.set_is_batch_data(False)
        for p in ref.parameters():
            p.requires_grad_(True)
        d = {'x': torch.tensor(g['types']), 'edge_index': torch.tensor(ei),
             'edge_vec': torch.tensor(D, dtype=torch.float64).requires_grad_(True), 'num_atoms': torch.tensor(n),
             'cell_volume': torch.tensor(abs(np.linalg.det(g['cell']))), 'batch': torch.zeros(0, dtype=torch.long)}
        o = ref(d)
        lref = _finetune_loss(o, lab64, n)
        lref.backward()
    finally:
        torch.set_default_dtype(torch.float32)
    assert abs(float(loss) - float(lref)) <= 1e-5 * abs(float(lref))
    refg = dict(ref.named_parameters())
    worst = 0.0
    for k, p in twin.named_parameters():
        if not p.requires_grad:
            continue
        assert p.grad is not None, k
        a, b = p.grad.double().cpu().reshape(-1), refg[k].grad.reshape(-1)
        rel = float((a - b).abs().max() / max(float(b.abs().max()), 1e-12))
        worst = max(worst, rel)
        assert rel < 1e-3, (k, rel)
    print('worst relative fine-tuning-gradient error', worst)
    twin.eval()


@pytest.mark.parametrize('muls,lout', [((8, 0, 0), 2), ((8, 4, 4), 2), ((8, 4, 4), 0)])
def test_convolution_family_second_order(muls, lout):
    """ConvF / ConvB (higher_order.py): gradient of a functional of the first-order gradients, against the oracle's
    tensor product + scatter in float64."""
    from oracle.refnn import TensorProductUVU, message_gather
    from sevennet_finetuning_b200 import higher_order as ho, ops
    dev = torch.device('cuda:0')
    rng = np.random.default_rng(3)
    n, E = 13, 90
    dst = np.sort(rng.integers(0, n, E))
    src = rng.integers(0, n, E)
    rowptr = np.concatenate([[0], np.cumsum(np.bincount(dst, minlength=n))])
    xdim = muls[0] + 3 * muls[1] + 5 * muls[2]
    wn, od = ops.conv_dims(*muls, lout)
    ir_x = '+'.join(f'{m}x{l}e' for l, m in enumerate(muls) if m)
    ir_o = '+'.join(f'1x{l}e' for l in range(lout + 1))
    x64 = torch.tensor(rng.normal(size=(n, xdim)), requires_grad=True)
    Yv64 = torch.tensor(rng.normal(size=(E, 8)), requires_grad=True)
    w64 = torch.tensor(rng.normal(size=(E, wn)), requires_grad=True)
    g64 = torch.tensor(rng.normal(size=(n, od)), requires_grad=True)
    cs = [torch.tensor(rng.normal(size=s)) for s in [(n, xdim), (E, 8), (E, wn)]]
    denom = 3.5

    def functional(conv, x, Yv, w, g, cs):
        Y = torch.cat([torch.ones_like(Yv[:, :1]), Yv], 1)
        out = conv(x, Y, w)
        dx, dYv, dw = torch.autograd.grad((out * g).sum(), [x, Yv, w], create_graph=True)
        S = (dx * cs[0]).sum() + (dYv * cs[1]).sum() + (dw * cs[2]).sum()
        return out, torch.autograd.grad(S, [x, Yv, w, g], allow_unused=True)

    torch.set_default_dtype(torch.float64)
    try:
        tp = TensorProductUVU(ir_x, '1x0e+1x1e+1x2e', ir_o).double()
        sd, dd = torch.tensor(src), torch.tensor(dst)
        o_ref, g_ref = functional(lambda x, Y, w: message_gather(x, dd, tp(x[sd], Y, w)) / denom, x64, Yv64, w64, g64, cs)
    finally:
        torch.set_default_dtype(torch.float32)
    graph = dict(rowptr_dst=torch.tensor(rowptr, dtype=torch.int32, device=dev), src=torch.tensor(src, dtype=torch.int32, device=dev))
    f32 = [t.detach().to(dev, torch.float32).requires_grad_(True) for t in (x64, Yv64, w64, g64)]
    o, gg = functional(lambda x, Y, w: ho.ConvF.apply(x, Y, w, graph, muls, lout, denom), *f32,
                       [c.to(dev, torch.float32) for c in cs])
    assert (o.double().cpu() - o_ref).abs().max() <= 2e-5 * o_ref.abs().max()
    for a, b, name in zip(gg, g_ref, 'x Y w g'.split()):
        if b is None or float(b.abs().max()) == 0.0:
            assert a is None or float(a.abs().max()) < 1e-6, name
            continue
        assert (a.double().cpu() - b).abs().max() <= 5e-5 * b.abs().max(), name


@pytest.mark.parametrize('cutoff', ['xplor', 'poly'])
def test_edge_embedding_second_order(cutoff):
    """EdgeF / EdgeB (b200_edge_embed_bwd2: dual-number evaluation of the embedding's own formulas): gradient of a
    functional of d/d edge_vec, with respect to edge_vec, the Bessel coefficients and the incoming gradients, against
    float64 autograd through the oracle's EdgeEmbedding (sevenn/nn/edge_embedding.py:114-116,138-145,163-173,190-198)."""
    from oracle import refnn
    from sevennet_finetuning_b200 import higher_order as ho
    from sevennet_finetuning_b200._lib import B200_CUTOFF_POLY, B200_CUTOFF_XPLOR, EdgeParams
    dev = torch.device('cuda:0')
    rng = np.random.default_rng(8)
    E = 257
    d = rng.normal(size=(E, 3))
    d = d / np.linalg.norm(d, axis=1, keepdims=True) * rng.uniform(0.8, 4.95, (E, 1))   # both branches of the envelope
    torch.set_default_dtype(torch.float64)
    try:
        env = refnn.XPLORCutoff(5.0, 4.5) if cutoff == 'xplor' else refnn.PolynomialCutoff(5.0, 6)
        ref = refnn.EdgeEmbedding(refnn.BesselBasis(5.0, 8), env, refnn.SphericalEncoding(2, 1, normalize=True)).double()
        v64 = torch.tensor(d, requires_grad=True)
        gY64 = torch.tensor(rng.normal(size=(E, 9)), requires_grad=True)
        ge64 = torch.tensor(rng.normal(size=(E, 8)), requires_grad=True)
        c64 = torch.tensor(rng.normal(size=(E, 3)))
        o = ref({'edge_vec': v64})
        k64 = ref.basis_function.coeffs
        (dv,) = torch.autograd.grad((o['edge_attr'] * gY64).sum() + (o['edge_embedding'] * ge64).sum(), [v64], create_graph=True)
        S = (dv * c64).sum() + (dv ** 2).sum()
        want = torch.autograd.grad(S, [v64, k64, gY64, ge64])
        dv_ref = dv.detach()
    finally:
        torch.set_default_dtype(torch.float32)
    p = EdgeParams(8, B200_CUTOFF_XPLOR if cutoff == 'xplor' else B200_CUTOFF_POLY, 1, 5.0, 4.5, 6.0)
    v = v64.detach().to(dev, torch.float32).requires_grad_(True)
    k = k64.detach().to(dev, torch.float32).requires_grad_(True)
    gY = gY64.detach().to(dev, torch.float32).requires_grad_(True)
    ge = ge64.detach().to(dev, torch.float32).requires_grad_(True)
    ln, Y, emb = ho.EdgeF.apply(v, k, p)
    (dv32,) = torch.autograd.grad((Y * gY).sum() + (emb * ge).sum(), [v], create_graph=True)
    S32 = (dv32 * c64.to(dev, torch.float32)).sum() + (dv32 ** 2).sum()
    got = torch.autograd.grad(S32, [v, k, gY, ge])
    assert (dv32.detach().double().cpu() - dv_ref).abs().max() <= 2e-5 * dv_ref.abs().max()
    for a, b, name in zip(got, want, ['edge_vec', 'coeffs', 'gY', 'gemb']):
        b = b.clone()
        if name == 'gY':
            b[:, 0] = 0     # Y[:, 0] is the constant 1: its incoming gradient is never read
        assert (a.double().cpu() - b).abs().max() <= 1e-4 * b.abs().max(), name


def test_linear_and_gate_families_second_order():
    """LinF / LinT / LinWG and GateF / GateB against float64 tensor-op restatements (e3nn Linear: x W / sqrt(fan_in) per
    l block; Gate: c*silu on scalars and gates, gates multiply the l > 0 channels)."""
    from sevennet_finetuning_b200 import higher_order as ho, ops
    dev = torch.device('cuda:0')
    rng = np.random.default_rng(4)
    n = 37
    mi, mo = (24, 16, 8), (24, 8, 8)
    desc, numel = ops.make_lin_desc('24x0e+16x1e+8x2e', '24x0e+8x1e+8x2e')
    din = sum(m * (2 * l + 1) for l, m in enumerate(mi))
    x64 = torch.tensor(rng.normal(size=(n, din)), requires_grad=True)
    W64 = torch.tensor(rng.normal(size=(numel,)), requires_grad=True)
    C = 1.6791767923989418

    def lin64(x, W):
        outs, io, wo = [], 0, 0
        for l in range(3):
            D = 2 * l + 1
            xb = x[:, io:io + mi[l] * D].reshape(-1, mi[l], D)
            Wb = W[wo:wo + mi[l] * mo[l]].reshape(mi[l], mo[l])
            outs.append(torch.einsum('nud,uv->nvd', xb, Wb).reshape(x.shape[0], -1) / mi[l] ** 0.5)
            io += mi[l] * D
            wo += mi[l] * mo[l]
        return torch.cat(outs, 1)

    def gate64(y):
        ns, m1, m2 = mo[0] - mo[1] - mo[2], mo[1], mo[2]
        s, t, v1, v2 = y[:, :ns], y[:, ns:ns + m1 + m2], y[:, mo[0]:mo[0] + 3 * m1], y[:, mo[0] + 3 * m1:]
        act = lambda z: C * z * torch.sigmoid(z)
        gt = act(t)
        return torch.cat([act(s), (v1.reshape(-1, m1, 3) * gt[:, :m1, None]).reshape(y.shape[0], -1),
                          (v2.reshape(-1, m2, 5) * gt[:, m1:, None]).reshape(y.shape[0], -1)], 1)

    ns, m1, m2 = mo[0] - mo[1] - mo[2], mo[1], mo[2]
    dout = ns + 3 * m1 + 5 * m2
    c1 = torch.tensor(rng.normal(size=(n, dout)))
    c2 = torch.tensor(rng.normal(size=(n, din)))

    def functional(f, x, W, c1, c2):
        out = f(x, W)
        (dx,) = torch.autograd.grad((out * c1).sum() + (out ** 2).sum(), [x], create_graph=True)
        S = (dx * c2).sum() + (dx ** 2).sum()
        return out, torch.autograd.grad(S, [x, W])

    o_ref, g_ref = functional(lambda x, W: gate64(lin64(x, W)), x64, W64, c1, c2)
    x32 = x64.detach().to(dev, torch.float32).requires_grad_(True)
    W32 = W64.detach().to(dev, torch.float32).requires_grad_(True)
    o, gg = functional(lambda x, W: ho.GateF.apply(ho.LinF.apply(x, W, desc, numel), ns, m1, m2), x32, W32,
                       c1.to(dev, torch.float32), c2.to(dev, torch.float32))
    assert (o.double().cpu() - o_ref).abs().max() <= 1e-5 * o_ref.abs().max()
    for a, b, name in zip(gg, g_ref, 'x W'.split()):
        assert (a.double().cpu() - b).abs().max() <= 1e-4 * b.abs().max(), name


def test_rehearsal_ewc_step(weights):
    """One iteration of the rehearsal loop (sevenn/train/trainer.py:173-210) on a batch of small synthetic crystals:
    batched forward, E/F/S + EWC loss, two optimizer steps; the loss of the same batch goes down."""
    from sevennet_finetuning_b200 import finetune as ft, nn as b200nn
    from sevennet_finetuning_b200.model import B200SevenNet
    from sevennet_finetuning_b200.synthetic import bulk_batch
    dev = torch.device('cuda:0')
    cfg = dict(shift=weights['rescale_atomic_energy.shift'].tolist(), scale=weights['rescale_atomic_energy.scale'].tolist(),
               conv_denominator=float(weights['0_convolution.denominator'][0]))
    model = b200nn.build_E3_equivariant_model(cfg)
    sd = {k: torch.tensor(v) for k, v in weights.items()}
    model.load_state_dict(sd)
    engine = B200SevenNet(state_dict=weights, device=dev)
    graphs = ft.synthetic_graphs(bulk_batch(4, seed=1, target_atoms=32), 5.0, dev, label_model=engine, noise=0.002, seed=1)
    mem = ft.synthetic_graphs(bulk_batch(4, seed=2, target_atoms=32), 5.0, dev, label_model=engine, noise=0.002, seed=2)
    fisher = {k: torch.ones_like(v) * 1e-3 for k, v in sd.items()}
    lf = ft.get_loss_functions('huber', {'delta': 0.01}, 1.0, True, 0.01, fisher, {k: v.clone() for k, v in sd.items()},
                               ewc_lambda=1e5, device=dev)
    tr = ft.RehearsalTrainer(model, lf, 'adam', {'lr': 1e-4}, device=dev)
    b, bm = ft.collate(graphs, dev), ft.collate(mem, dev)
    out = tr.model(ft.to_device(ft.collate(graphs), dev))
    assert out['inferred_total_energy'].shape == (4,) and out['inferred_stress'].shape == (4, 6)
    # batched train-mode forward agrees with the engine labels up to the label noise
    assert (out['inferred_total_energy'].detach().cpu() - b['total_energy'].cpu()).abs().max() < 0.5
    # directional check of the batched gradient (E/F/S terms; the EWC term is a plain quadratic, checked below):
    # L(theta - eps g) - L(theta) = -eps |g|^2 + O(eps^2), with eps sized for a 1 % drop
    lf_data = lf[:3]
    tr.model.train()
    tr.optimizer.zero_grad()
    l0 = ft.total_loss(lf_data, tr.model(b), tr.model, dev)
    l0.backward()
    params = [p for p in tr.model.parameters() if p.grad is not None]
    g2 = sum(float((p.grad ** 2).sum()) for p in params)
    eps = 0.01 * float(l0) / g2
    with torch.no_grad():
        for p in params:
            p -= eps * p.grad
    l1 = ft.total_loss(lf_data, tr.model(ft.collate(graphs, dev)), tr.model, dev)
    ratio = (float(l0) - float(l1)) / (eps * g2)
    assert 0.7 < ratio < 1.1, (float(l0), float(l1), ratio)
    # EWC known answer: theta moved by -eps g from theta*, Fisher = 1e-3 everywhere
    ewc = float(lf[3][0].get_loss(None, tr.model))
    assert abs(ewc - 1e-3 * eps * eps * g2) <= 2e-2 * ewc   # fp32 rounding of the tiny update theta - eps g
    # the full rehearsal iteration: train batch step + replay batch step (gradients accumulate, trainer.py:185-210)
    l2, m2 = tr.run_one_step(ft.collate(graphs, dev), ft.collate(mem, dev))
    assert torch.isfinite(l2).all() and torch.isfinite(m2).all()
    fish, opt, cnt = tr.compute_fisher_matrix([ft.collate(graphs[:1], dev), ft.collate(graphs[1:2], dev)])
    assert cnt == 2 and set(fish) == set(dict(tr.model.named_parameters()))
    assert all(torch.isfinite(v).all() and float(v.min()) >= 0 for v in fish.values())


def test_nequip_self_connection_twin(gold):
    """SelfConnectionIntro twin ('nequip' self connection): values against the frozen-graph golden
    vectors (multiplicities below the tensor-core kernel's minimum: FFMA kernel) and, at SevenNet-0 widths with all three
    l-blocks, values, first and second derivatives against the float64 oracle."""
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from oracle.refnn import SelfConnectionIntro as Ref
    from sevennet_finetuning_b200 import nn as b200nn
    dev = torch.device('cuda:0')
    g = gold('fctp_hfo.npz')
    m = b200nn.SelfConnectionIntro('4x0e+4x1e', '2x0e', '12x0e+4x1e').to(dev).eval()
    with torch.no_grad():
        m.fc_tensor_product.weight.copy_(torch.tensor(np.concatenate([g['w0'].reshape(-1), g['w1'].reshape(-1)])))
    t = torch.tensor(g['types'], device=dev)
    out = m({'x': torch.tensor(g['x'], device=dev), 'atom_type': t})['self_cont_tmp']
    assert np.abs(out.detach().cpu().numpy() - g['out']).max() <= 2e-6

    ir_x, ir_o, nsp, n = '128x0e+64x1e+32x2e', '224x0e+64x1e+32x2e', 5, 700
    gen = torch.Generator().manual_seed(3)
    x = torch.randn(n, 480, generator=gen, dtype=torch.float64)
    types = torch.randint(0, nsp, (n,), generator=gen)
    types[types == 3] = 1                                   # one species absent from the batch
    ref = Ref(ir_x, f'{nsp}x0e', ir_o).double()
    tw = b200nn.SelfConnectionIntro(ir_x, f'{nsp}x0e', ir_o).to(dev)
    tw.load_state_dict({k: v.float() for k, v in ref.state_dict().items()}, strict=True)
    probe = torch.randn(n, 576, generator=gen, dtype=torch.float64)
    v = torch.randn(n, 480, generator=gen, dtype=torch.float64)

    def run(mod, xx, tt, pr, vv, dt):
        mod.train()
        xx = xx.to(dt).requires_grad_(True)
        d = {'x': xx, 'atom_type': tt, 'node_attr': torch.nn.functional.one_hot(tt, nsp).to(dt)}
        y = mod(d)['self_cont_tmp']
        gx, = torch.autograd.grad((y * pr.to(dt)).sum() + (y * y).sum() * 0.5, xx, create_graph=True)
        W = mod.fc_tensor_product.weight
        gW, = torch.autograd.grad((gx * vv.to(dt)).sum(), W)          # second order: d/dW of a contraction of dL/dx
        return y.detach(), gx.detach(), gW

    y0, gx0, gW0 = run(ref, x, types, probe, v, torch.float64)
    y1, gx1, gW1 = run(tw, x.to(dev), types.to(dev), probe.to(dev), v.to(dev), torch.float32)
    for a, b_, name in ((y0, y1, 'y'), (gx0, gx1, 'dx'), (gW0, gW1, 'dW')):
        err = (a - b_.double().cpu()).abs().max() / a.abs().max()
        assert err <= 2e-5, (name, float(err))
    assert float(gW1.reshape(-1)[tw._sel[3].to(dev)].abs().max()) == 0.0   # the absent species' slice gets no gradient


def test_nequip_self_connection_model(gold, weights):
    """A whole 'nequip' self-connection model (self_connection_type='nequip', model_build.py:45-56) on HfO2-96 through the
    module chain: energy, forces, stress against the float64 oracle model with the same (seeded) parameters."""
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from oracle.refmodel import build_model
    from sevennet_finetuning_b200 import nn as b200nn
    cfg = dict(self_connection_type='nequip', num_species=4, num_convolution_layer=3,
               irreps_manual=['64x0e', '64x0e+32x1e+32x2e', '64x0e+32x1e+32x2e', '64x0e'],
               shift=[-1.0, -2.0, 0.5, 0.0], scale=[1.0, 1.5, 0.7, 1.0], conv_denominator=30.0)
    torch.manual_seed(11)
    ref = build_model(cfg).double()
    tw = b200nn.build_E3_equivariant_model(cfg)
    res = tw.load_state_dict({k: v.float() for k, v in ref.state_dict().items()}, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    assert any(k.endswith('self_connection_intro.fc_tensor_product.weight') for k in tw.state_dict())
    tw = tw.cuda().eval()
    tw.set_is_batch_data(False)
    ref.eval()
    ref.set_is_batch_data(False)
    g = gold('hfo2_96.npz')
    dev = torch.device('cuda:0')
    data, ei, D = _data(g, dev)
    types = torch.tensor(g['types']) % 2 + torch.arange(len(g['types'])) % 2 * 2    # four species
    data['x'] = types.to(dev)
    out = tw(data)
    d64 = {'x': types.clone(), 'edge_index': torch.tensor(ei), 'edge_vec': torch.tensor(D, dtype=torch.float32).double().requires_grad_(True),
           'num_atoms': torch.tensor(len(types)), 'cell_volume': torch.tensor(abs(np.linalg.det(g['cell'])), dtype=torch.float64),
           'batch': torch.zeros(len(types), dtype=torch.long)}
    torch.set_default_dtype(torch.float64)   # the oracle builds its tables and one-hot rows in the default dtype
    try:
        o64 = ref(d64)
    finally:
        torch.set_default_dtype(torch.float32)
    n = len(types)
    assert abs(float(out['inferred_total_energy']) - float(o64['inferred_total_energy'])) / n <= 2e-6
    fe = (out['inferred_force'].detach().cpu().double() - o64['inferred_force'].detach()).abs().max()
    assert fe <= 2e-5 * max(1.0, float(o64['inferred_force'].abs().max())), float(fe)
    se = (out['inferred_stress'].detach().cpu().double() - o64['inferred_stress'].detach()).abs().max()
    assert se <= 1e-4 * float(o64['inferred_stress'].abs().max()), float(se)


def test_substitute_into_reference_sequential(gold, weights):
    """nn.substitute: the reference-side container (oracle restatement of sevenn/nn/sequential.py with replace_module) keeps
    its identity, every hot-path module is swapped key by key, parameters and their requires_grad flags are carried over,
    and E/F/S match the golden vectors."""
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from sevennet_finetuning_b200 import nn as b200nn
    cfg = dict(shift=weights['rescale_atomic_energy.shift'].tolist(), scale=weights['rescale_atomic_energy.scale'].tolist(),
               conv_denominator=float(weights['0_convolution.denominator'][0]))
    ref = _oracle(weights).float().cuda()
    ref._modules['1_convolution'].denominator.requires_grad_(True)     # as train_denominator would leave it
    keys = list(ref._modules)
    out_model = b200nn.substitute(ref, cfg)
    assert out_model is ref and list(ref._modules) == keys
    for key, mod in ref._modules.items():
        assert type(mod).__module__.endswith('sevennet_finetuning_b200.nn'), key
    sd = ref.state_dict()
    assert set(sd) == set(weights)
    for k, v in weights.items():
        assert np.array_equal(sd[k].cpu().numpy().reshape(-1), np.asarray(v, np.float32).reshape(-1)), k
    grad = {k: p.requires_grad for k, p in ref.named_parameters()}
    assert grad['1_convolution.denominator'] and not grad['0_convolution.denominator']
    ref.eval()
    ref.set_is_batch_data(False)
    for name in ('si216.npz', 'hfo2_96.npz'):
        g = gold(name)
        data, _, _ = _data(g, torch.device('cuda:0'))
        out = ref(data)
        n = len(g['pos'])
        assert abs(float(out['inferred_total_energy']) - float(g['o64_energy'])) / n <= 1e-6
        assert np.abs(out['inferred_force'].detach().cpu().numpy() - g['o64_force']).max() <= 1e-5
        assert np.abs(out['inferred_force'].detach().cpu().numpy() - g['ref_force']).max() <= 4e-5
        s = out['inferred_stress'].detach().cpu().numpy()
        assert np.abs(s - g['o64_stress']).max() <= 1e-4 * np.abs(g['o64_stress']).max()


@pytest.mark.parametrize('name', ['si216.npz', 'hfo2_96.npz'])
def test_deployed_serial_form_matches_golden(twin, gold, name):
    """The deployed serial form of the model, assembled exactly as sevenn/scripts/deploy.py:22-23 does: EdgePreprocess
    prepended, force_output replaced by ForceStressOutput, inputs = the deployed dict (pos, pbc_shift, cell_lattice_vectors,
    cell_volume) - the same fp32 arithmetic from the same inputs as the reference's frozen graph that produced ref_*."""
    import copy
    from sevennet_finetuning_b200 import nn as b200nn
    g = gold(name)
    dev = torch.device('cuda:0')
    m = copy.deepcopy(twin)
    m.prepand_module('edge_preprocess', b200nn.EdgePreprocess(True))
    m.replace_module('force_output', b200nn.ForceStressOutput())
    assert list(m._modules)[0] == 'edge_preprocess' and type(m._modules['force_output']).__name__ == 'ForceStressOutput'
    m = m.to(dev).eval()
    n = len(g['pos'])

    def inputs(batched, shuffle):
        ei, sh = g['edge_index'], g['shift']
        if shuffle:
            p = np.random.default_rng(1).permutation(ei.shape[1])
            ei, sh = ei[:, p], sh[p]
        vol = torch.tensor(abs(np.linalg.det(g['cell'])), dtype=torch.float32, device=dev)
        return {'x': torch.tensor(g['types'], dtype=torch.long, device=dev),
                'pos': torch.tensor(g['pos'], dtype=torch.float32, device=dev).requires_grad_(True),
                'edge_index': torch.tensor(ei, dtype=torch.long, device=dev),
                'pbc_shift': torch.tensor(sh, dtype=torch.float32, device=dev),
                'cell_lattice_vectors': torch.tensor(g['cell'], dtype=torch.float32, device=dev).view((1, 3, 3) if batched else (3, 3)),
                'cell_volume': vol.view(1) if batched else vol,
                'num_atoms': torch.tensor([n], device=dev) if batched else torch.tensor(n, device=dev),
                'batch': torch.zeros(n if batched else 0, dtype=torch.long, device=dev)}

    for batched, shuffle in ((False, False), (False, True), (True, False)):
        m.set_is_batch_data(batched)
        out = m(inputs(batched, shuffle))
        e = float(out['inferred_total_energy'].reshape(-1)[0])
        f = out['inferred_force'].detach().cpu().numpy()
        s = out['inferred_stress'].detach().cpu().numpy().reshape(-1)
        # 1e-6 eV/atom, but never tighter than three fp32 ulps of the total energy (96 atoms at -972 eV: one ulp = 6.4e-7
        # eV/atom; Si-216 at -1158.7 eV: 1e-6 eV/atom is 1.8 ulp, and the shuffled / batched variants reach the total through a
        # different summation order - 1 run in 6 landed 2.16 ulp from the float64 value)
        etol = max(1e-6, 3.0 * float(np.spacing(np.float32(abs(e)))) / n)
        assert abs(e - float(g['ref_energy'])) / n <= etol and abs(e - float(g['o64_energy'])) / n <= etol
        # fp32 positions: an edge vector carries ~2e-6 A of rounding before anything is computed (the reference's frozen graph,
        # fed the same dict, sits 1-2e-5 eV/A from float64 for the same reason), hence 4e-5 against both
        assert np.abs(f - g['ref_force']).max() <= 4e-5 and np.abs(f - g['o64_force']).max() <= 4e-5
        assert np.abs(s - g['ref_stress']).max() <= 2e-4 * np.abs(g['ref_stress']).max()
        assert np.abs(s - g['o64_stress']).max() <= 1e-4 * np.abs(g['o64_stress']).max()
    # ForceOutput alone (no stress): same forces
    m.replace_module('force_output', b200nn.ForceOutput())
    m.set_is_batch_data(False)
    out = m(inputs(False, False))
    assert np.abs(out['inferred_force'].detach().cpu().numpy() - g['o64_force']).max() <= 4e-5 and 'inferred_stress' not in out


def test_batched_rehearsal_gradient_matches_float64_oracle(weights):
    """Config 2's iteration on a BATCH (4 structures of ~32 atoms): every parameter gradient of the trainer's own loss
    (RehearsalTrainer._fwd_loss: Huber E/F/S + EWC) against float64 double-backward through the oracle model in batch mode, and
    the gradient the second optimizer.step of the iteration sees - the reference does not zero the gradients between the train
    batch and the replay batch (sevenn/train/trainer.py:185-210), so it is grad_train(theta0) + grad_replay(theta1)."""
    from sevennet_finetuning_b200 import finetune as ft, nn as b200nn
    from sevennet_finetuning_b200.model import B200SevenNet
    from sevennet_finetuning_b200.synthetic import bulk_batch
    dev = torch.device('cuda:0')
    cfg = dict(shift=weights['rescale_atomic_energy.shift'].tolist(), scale=weights['rescale_atomic_energy.scale'].tolist(),
               conv_denominator=float(weights['0_convolution.denominator'][0]))
    sd = {k: torch.tensor(v) for k, v in weights.items()}
    engine = B200SevenNet(state_dict=weights, device=dev)
    graphs = ft.synthetic_graphs(bulk_batch(4, seed=11, target_atoms=32), 5.0, dev, label_model=engine, noise=0.02, seed=11)
    mem = ft.synthetic_graphs(bulk_batch(4, seed=12, target_atoms=32), 5.0, dev, label_model=engine, noise=0.02, seed=12)

    def run(model, device, dtype):
        """the iteration of RehearsalTrainer.run_one_step with the gradients recorded before each optimizer.step"""
        fisher = {k: torch.full_like(v, 1e-3).to(dtype) for k, v in sd.items()}
        star = {k: (v + 0.01 * torch.sin(torch.arange(v.numel(), dtype=torch.float64).reshape(v.shape) + 1.0)).to(dtype) for k, v in sd.items()}
        lf = ft.get_loss_functions('huber', {'delta': 0.01}, 1.0, True, 0.01, fisher, star, ewc_lambda=1e3, device=device)
        for p in model.parameters():
            p.requires_grad_(True)
        tr = ft.RehearsalTrainer(model, lf, 'sgd', {'lr': 1e-6}, device=device)
        seen = []
        step = tr.optimizer.step
        tr.optimizer.step = lambda *a, **k: (seen.append({n: p.grad.detach().double().cpu().clone() for n, p in tr.model.named_parameters()}), step(*a, **k))[1]

        def batch_of(gs):
            b = ft.collate(gs, device)
            if dtype == torch.float64:
                b = {k: (v.double() if torch.is_tensor(v) and v.is_floating_point() else v) for k, v in b.items()}
                b['edge_vec'] = b['edge_vec'].detach().requires_grad_(True)
            return b
        l, lm = tr.run_one_step(batch_of(graphs), batch_of(mem))
        return seen, float(l), float(lm)

    twin = b200nn.build_E3_equivariant_model(cfg)
    twin.load_state_dict(sd)
    got, l32, lm32 = run(twin, dev, torch.float32)
    torch.set_default_dtype(torch.float64)
    try:
        ref = _oracle(weights)
        want, l64, lm64 = run(ref, torch.device('cpu'), torch.float64)
    finally:
        torch.set_default_dtype(torch.float32)
    assert len(got) == 2 and len(want) == 2
    assert abs(l32 - l64) <= 1e-4 * abs(l64) and abs(lm32 - lm64) <= 1e-4 * abs(lm64)
    worst = [0.0, 0.0]
    for s in range(2):
        for k in want[s]:
            a, b = got[s][k].reshape(-1), want[s][k].reshape(-1)
            rel = float((a - b).abs().max() / max(float(b.abs().max()), 1e-12))
            worst[s] = max(worst[s], rel)
            assert rel < 1e-3, (s, k, rel)
    # the second step really saw the sum (not the replay gradient alone)
    k = '1_convolution.weight_nn.layer2.weight'
    assert float((got[1][k] - got[0][k]).abs().max()) > 0 and float((got[1][k]).abs().max()) > 0
    print('worst relative error: train-batch gradient', worst[0], 'accumulated gradient at the second step', worst[1])
