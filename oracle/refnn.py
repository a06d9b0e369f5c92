"""TEST INFRASTRUCTURE - dependency-free PyTorch restatement of the reference hot-path modules.

Not part of the product (see ``oracle/__init__.py``).  Each class restates the reference
module of the same name; e3nn internals are restated from the literal arithmetic of the
reference's frozen graph ``deployed_serial.pt`` (``m.code``; "dump" line numbers below refer
to that listing, reproduced by ``oracle/make_golden.py --dump``).

Feature layout everywhere: irreps-blocked ``[mul, 2l+1]`` (m fastest) per block, blocks
concatenated in irreps order - exactly the e3nn layout.

Pinned against the frozen graph to fp32 round-off by ``tests/test_oracle_golden.py``.
"""
import math
from collections import OrderedDict
from typing import Dict, List, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import wigner

SILU_2MOM = 1.6791767923989418  # e3nn normalize2mom(silu) constant c5 of the frozen graph

Irreps = List[Tuple[int, int, int]]  # (mul, l, parity)


def parse_irreps(s) -> Irreps:
    if not isinstance(s, str):
        return [tuple(t) for t in s]
    out = []
    for tok in s.replace(' ', '').split('+'):
        if not tok:
            continue
        mul, ir = tok.split('x') if 'x' in tok else ('1', tok)
        out.append((int(mul), int(ir[:-1]), 1 if ir[-1] == 'e' else -1))
    return out


def irreps_dim(ir: Irreps) -> int:
    return sum(m * (2 * l + 1) for m, l, _ in ir)


def irreps_str(ir: Irreps) -> str:
    return '+'.join(f'{m}x{l}{"e" if p == 1 else "o"}' for m, l, p in ir)


def simplify(ir: Irreps) -> Irreps:
    out = []
    for m, l, p in ir:
        if out and out[-1][1:] == (l, p):
            out[-1] = (out[-1][0] + m, l, p)
        elif m > 0:
            out.append((m, l, p))
    return out


def sh_irreps(lmax: int, parity: int) -> Irreps:
    return [(1, l, parity ** l) for l in range(lmax + 1)]


class O3Linear(nn.Module):
    """e3nn ``o3.Linear`` without biases (reference use: ``nn/linear.py:36``,
    ``nn/self_connection.py:57``).  Dump lines 88-118, 219-246: per output block
    ``x_l[N,mul_in,2l+1] . W[mul_in,mul_out] / sqrt(sum of mul_in feeding that output)``;
    flat weight = concat of W.flatten() over instructions (i_in outer, i_out inner)."""

    def __init__(self, irreps_in, irreps_out):
        super().__init__()
        self.irreps_in = parse_irreps(irreps_in)
        self.irreps_out = parse_irreps(irreps_out)
        self.instr = [(i, j) for i, (_, l1, p1) in enumerate(self.irreps_in)
                      for j, (_, l2, p2) in enumerate(self.irreps_out) if (l1, p1) == (l2, p2)]
        fan = {}
        for i, j in self.instr:
            fan[j] = fan.get(j, 0) + self.irreps_in[i][0]
        self.path_weight = [1.0 / math.sqrt(fan[j]) for _, j in self.instr]
        n = sum(self.irreps_in[i][0] * self.irreps_out[j][0] for i, j in self.instr)
        self.weight = nn.Parameter(torch.randn(n))

    def forward(self, x):
        N = x.shape[0]
        offs_in = np.cumsum([0] + [m * (2 * l + 1) for m, l, _ in self.irreps_in])
        outs = [None] * len(self.irreps_out)
        w_off = 0
        for (i, j), pw in zip(self.instr, self.path_weight):
            mi, l, _ = self.irreps_in[i]
            mo = self.irreps_out[j][0]
            W = self.weight[w_off:w_off + mi * mo].view(mi, mo)
            w_off += mi * mo
            xi = x[:, offs_in[i]:offs_in[i + 1]].reshape(N, mi, 2 * l + 1)
            y = torch.einsum('nim,io->nom', xi, W) * pw
            outs[j] = y if outs[j] is None else outs[j] + y
        res = []
        for j, (mo, l, _) in enumerate(self.irreps_out):
            if outs[j] is None:
                res.append(x.new_zeros(N, mo * (2 * l + 1)))
            else:
                res.append(outs[j].reshape(N, mo * (2 * l + 1)))
        return torch.cat(res, -1)


class IrrepsLinear(nn.Module):
    """``sevenn/nn/linear.py:14-49``."""

    def __init__(self, irreps_in, irreps_out, data_key_in, data_key_out=None, is_embed=False,
                 **kw):
        super().__init__()
        self.key_input = data_key_in
        self.key_output = data_key_in if data_key_out is None else data_key_out
        self.linear = O3Linear(irreps_in, irreps_out)
        self._last_dim = irreps_dim(self.linear.irreps_in)
        if is_embed:  # nn/linear.py:37-44
            self.linear.weight = nn.Parameter(self.linear.weight / self.linear.path_weight[0])

    def forward(self, data):
        x = data[self.key_input].reshape(-1, self._last_dim)
        data[self.key_output] = self.linear(x)
        return data


class BesselBasis(nn.Module):
    """``sevenn/nn/edge_embedding.py:80-116``."""

    def __init__(self, cutoff_length, bessel_basis_num=8, trainable_coeff=True):
        super().__init__()
        self.num_basis = bessel_basis_num
        self.prefactor = 2.0 / cutoff_length
        c = torch.FloatTensor([n * math.pi / cutoff_length for n in range(1, bessel_basis_num + 1)])
        if trainable_coeff:
            self.coeffs = nn.Parameter(c)
        else:
            self.register_buffer('coeffs', c)

    def forward(self, r):
        ur = r.unsqueeze(-1)
        return self.prefactor * torch.sin(self.coeffs * ur) / ur


class PolynomialCutoff(nn.Module):
    """``sevenn/nn/edge_embedding.py:119-145``."""

    def __init__(self, cutoff_length, poly_cut_p_value=6):
        super().__init__()
        p = poly_cut_p_value
        self.cutoff_length, self.p = cutoff_length, p
        self.coeff_p0 = (p + 1.0) * (p + 2.0) / 2.0
        self.coeff_p1 = p * (p + 2.0)
        self.coeff_p2 = p * (p + 1.0) / 2.0

    def forward(self, r):
        r = r / self.cutoff_length
        return (1 - self.coeff_p0 * torch.pow(r, self.p) + self.coeff_p1 * torch.pow(r, self.p + 1.0)
                - self.coeff_p2 * torch.pow(r, self.p + 2.0))


class XPLORCutoff(nn.Module):
    """``sevenn/nn/edge_embedding.py:148-173``."""

    def __init__(self, cutoff_length, cutoff_on):
        super().__init__()
        self.r_on, self.r_cut = cutoff_on, cutoff_length

    def forward(self, r):
        r_sq = r * r
        r_on_sq, r_cut_sq = self.r_on * self.r_on, self.r_cut * self.r_cut
        return torch.where(
            r < self.r_on, 1.0,
            (r_cut_sq - r_sq) ** 2 * (r_cut_sq + 2 * r_sq - 3 * r_on_sq) / (r_cut_sq - r_on_sq) ** 3)


class SphericalEncoding(nn.Module):
    """``sevenn/nn/edge_embedding.py:177-198`` -> e3nn ``SphericalHarmonics(normalize,
    'component')``; dump lines 41-66."""

    def __init__(self, lmax, parity=-1, normalize=True):
        super().__init__()
        assert lmax <= 2
        self.lmax, self.normalize = lmax, normalize
        self.irreps_out = sh_irreps(lmax, parity)

    def forward(self, rvec):
        if self.normalize:
            rvec = rvec / torch.clamp_min(torch.norm(rvec, 2, -1, True), 1e-12)
        x, y, z = rvec[..., 0], rvec[..., 1], rvec[..., 2]
        out = [torch.ones_like(x)]
        if self.lmax >= 1:
            s3 = math.sqrt(3.0)
            out += [s3 * x, s3 * y, s3 * z]
        if self.lmax >= 2:
            s3, s5 = math.sqrt(3.0), math.sqrt(5.0)
            x2z2 = x * x + z * z
            out += [s5 * (s3 * x * z), s5 * (s3 * x * y), s5 * (y * y - 0.5 * x2z2),
                    s5 * (s3 * y * z), s5 * (0.5 * s3 * (z * z - x * x))]
        return torch.stack(out, -1)


class EdgeEmbedding(nn.Module):
    """``sevenn/nn/edge_embedding.py:202-230``."""

    def __init__(self, basis_module, cutoff_module, spherical_module):
        super().__init__()
        self.basis_function = basis_module
        self.cutoff_function = cutoff_module
        self.spherical = spherical_module

    def forward(self, data):
        rvec = data['edge_vec']
        r = torch.linalg.norm(rvec, dim=-1)
        data['edge_length'] = r
        data['edge_embedding'] = self.basis_function(r) * self.cutoff_function(r).unsqueeze(-1)
        data['edge_attr'] = self.spherical(rvec)
        return data


class EdgePreprocess(nn.Module):
    """``sevenn/nn/edge_embedding.py:13-77`` (strain trick for stress)."""

    def __init__(self, is_stress=True):
        super().__init__()
        self.is_stress = is_stress
        self._is_batch_data = True

    def forward(self, data):
        cell = data['cell_lattice_vectors'].view(-1, 3, 3) if self._is_batch_data \
            else data['cell_lattice_vectors'].view(3, 3)
        cell_shift, pos, batch = data['pbc_shift'], data['pos'], data['batch']
        if self.is_stress:
            if self._is_batch_data:
                nb = int(batch.max().item()) + 1
                strain = torch.zeros((nb, 3, 3), dtype=pos.dtype, device=pos.device, requires_grad=True)
                data['_strain'] = strain
                sym = 0.5 * (strain + strain.transpose(-1, -2))
                pos = pos + torch.bmm(pos.unsqueeze(-2), sym[batch]).squeeze(-2)
                cell = cell + torch.bmm(cell, sym)
            else:
                strain = torch.zeros((3, 3), dtype=pos.dtype, device=pos.device, requires_grad=True)
                data['_strain'] = strain
                sym = 0.5 * (strain + strain.transpose(-1, -2))
                pos = pos + torch.mm(pos, sym)
                cell = cell + torch.mm(cell, sym)
        i0, i1 = data['edge_index'][0], data['edge_index'][1]
        edge_vec = pos[i1] - pos[i0]
        if self._is_batch_data:
            edge_vec = edge_vec + torch.einsum('ni,nij->nj', cell_shift, cell[batch[i0]])
        else:
            edge_vec = edge_vec + torch.einsum('ni,ij->nj', cell_shift, cell)
        data['edge_vec'] = edge_vec
        data['edge_length'] = torch.linalg.norm(edge_vec, dim=-1)
        return data


def _act_2mom(name: str) -> float:
    if name == 'silu':
        return SILU_2MOM
    # deterministic Gauss-Hermite second moment (e3nn samples randomly; not pinned)
    f = {'tanh': np.tanh, 'abs': np.abs, 'relu': lambda t: np.maximum(t, 0),
         'sigmoid': lambda t: 1 / (1 + np.exp(-t))}[name]
    t, w = np.polynomial.hermite_e.hermegauss(200)
    return float(1.0 / math.sqrt((w * f(t) ** 2).sum() / w.sum()))


_ACT = {'silu': torch.nn.functional.silu, 'tanh': torch.tanh, 'abs': torch.abs,
        'relu': torch.relu, 'sigmoid': torch.sigmoid}


class _FCLayer(nn.Module):
    def __init__(self, h_in, h_out):
        super().__init__()
        self.weight = nn.Parameter(torch.randn(h_in, h_out))


class FullyConnectedNet(nn.Module):
    """e3nn ``nn.FullyConnectedNet`` (``sevenn/nn/convolution.py:97-100``); dump lines 119-125:
    hidden ``h = c * act(x W / sqrt(fan_in))``, last layer ``x W / sqrt(fan_in)``, no bias."""

    def __init__(self, hs: List[int], act='silu'):
        super().__init__()
        self.hs = list(hs)
        self.act_name = act
        self.act_cst = _act_2mom(act)
        for k, (a, b) in enumerate(zip(hs[:-1], hs[1:])):
            setattr(self, f'layer{k}', _FCLayer(a, b))

    def forward(self, x):
        n = len(self.hs) - 1
        for k in range(n):
            W = getattr(self, f'layer{k}').weight
            x = x @ (W / math.sqrt(W.shape[0]))
            if k < n - 1:
                x = _ACT[self.act_name](x) * self.act_cst
        return x


def tp_instructions(irreps_x: Irreps, irreps_filter: Irreps, irreps_out: Irreps):
    """Instruction list of ``IrrepsConvolution.__init__`` (``sevenn/nn/convolution.py:72-86``).
    Returns (instructions [(i_x, i_f, k_sorted)], irreps_mid sorted) where irreps_mid keeps
    one entry per instruction (e3nn ``Irreps.sort`` is stable, ordering by (l, parity))."""
    targets = {(l, p) for _, l, p in irreps_out}
    mid, ins = [], []
    for i, (mul, l1, p1) in enumerate(irreps_x):
        for j, (_, l2, p2) in enumerate(irreps_filter):
            for l3 in range(abs(l1 - l2), l1 + l2 + 1):
                if (l3, p1 * p2) in targets:
                    ins.append((i, j, len(mid)))
                    mid.append((mul, l3, p1 * p2))
    # e3nn Irrep ordering: by l, then 'e'/'o' such that (l, -p*(-1)^l) ascending
    order = sorted(range(len(mid)), key=lambda k: (mid[k][1], -mid[k][2] * (-1) ** mid[k][1]))
    inv = {k: pos for pos, k in enumerate(order)}
    return [(i, j, inv[k]) for i, j, k in ins], [mid[k] for k in order]


class TensorProductUVU(nn.Module):
    """e3nn ``o3.TensorProduct(..., 'uvu', shared_weights=False)`` with component
    normalisation and 'element' path normalisation (``sevenn/nn/convolution.py:88-95``).
    Dump lines 126-203, 446-575: ``out[e,u,:] = w[e,u] sum C[mx,msh,:] x[e,u,mx] Y[e,msh]``,
    ``C = w3j * sqrt(2 l_out + 1)`` (filter multiplicity is 1, each instruction owns its slot)."""

    def __init__(self, irreps_x, irreps_filter, irreps_out):
        super().__init__()
        self.irreps_in1 = parse_irreps(irreps_x)
        self.irreps_in2 = parse_irreps(irreps_filter)
        self.instr, self.irreps_out = tp_instructions(self.irreps_in1, self.irreps_in2,
                                                      parse_irreps(irreps_out))
        self.weight_numel = sum(self.irreps_in1[i][0] for i, _, _ in self.instr)
        for n, (i, j, k) in enumerate(self.instr):
            l1, l2, l3 = self.irreps_in1[i][1], self.irreps_in2[j][1], self.irreps_out[k][1]
            C = torch.tensor(wigner.w3j(l1, l2, l3) * math.sqrt(2 * l3 + 1), dtype=torch.float32)
            self.register_buffer(f'cg{n}', C, persistent=False)

    def forward(self, x, y, w):
        E = x.shape[0]
        o1 = np.cumsum([0] + [m * (2 * l + 1) for m, l, _ in self.irreps_in1])
        o2 = np.cumsum([0] + [m * (2 * l + 1) for m, l, _ in self.irreps_in2])
        outs = [None] * len(self.irreps_out)
        woff = 0
        for n, (i, j, k) in enumerate(self.instr):
            mul, l1, _ = self.irreps_in1[i]
            l2 = self.irreps_in2[j][1]
            xi = x[:, o1[i]:o1[i + 1]].reshape(E, mul, 2 * l1 + 1)
            yj = y[:, o2[j]:o2[j + 1]]
            C = getattr(self, f'cg{n}').to(x.dtype)
            wk = w[:, woff:woff + mul]
            woff += mul
            outs[k] = (torch.einsum('eua,eb,abc->euc', xi, yj, C) * wk.unsqueeze(-1)).reshape(E, -1)
        return torch.cat(outs, -1)


def message_gather(node_features, edge_dst, message):
    """``sevenn/nn/convolution.py:19-32``."""
    out = torch.zeros((len(node_features),) + tuple(message.shape[1:]), dtype=message.dtype,
                      device=message.device)
    return out.index_add_(0, edge_dst, message)


class IrrepsConvolution(nn.Module):
    """``sevenn/nn/convolution.py:36-123``."""

    def __init__(self, irreps_x, irreps_filter, irreps_out, weight_layer_input_to_hidden,
                 weight_layer_act='silu', denominator=1.0, train_denominator=False,
                 is_parallel=False, **kw):
        super().__init__()
        self.denominator = nn.Parameter(torch.FloatTensor([denominator]),
                                        requires_grad=train_denominator)
        self.is_parallel = is_parallel
        self.convolution = TensorProductUVU(irreps_x, irreps_filter, irreps_out)
        self.weight_nn = FullyConnectedNet(
            list(weight_layer_input_to_hidden) + [self.convolution.weight_numel], weight_layer_act)
        self._comm_size = irreps_dim(self.convolution.irreps_in1)

    def forward(self, data):
        weight = self.weight_nn(data['edge_embedding'])
        x = data['x']
        if self.is_parallel:
            x = torch.cat([x, data['x_ghost']])
        edge_src, edge_dst = data['edge_index'][1], data['edge_index'][0]
        message = self.convolution(x[edge_src], data['edge_attr'], weight)
        x = message_gather(x, edge_dst, message).div(self.denominator)
        if self.is_parallel:
            x = torch.tensor_split(x, data['nlocal'])[0]
        data['x'] = x
        return data


class SelfConnectionLinearIntro(nn.Module):
    """``sevenn/nn/self_connection.py:42-62``."""

    def __init__(self, irreps_x, irreps_out, **kw):
        super().__init__()
        self.linear = O3Linear(irreps_x, irreps_out)

    def forward(self, data):
        data['self_cont_tmp'] = self.linear(data['x'])
        return data


class SelfConnectionIntro(nn.Module):
    """``sevenn/nn/self_connection.py:11-38`` ('nequip' self connection): e3nn
    ``FullyConnectedTensorProduct(irreps_x, irreps_operand, irreps_out)`` of the node features with
    the one-hot species vector (``irreps_operand = '{n_species}x0e'``, ``interaction_blocks.py:43-52``).
    e3nn is not vendored; the arithmetic follows the frozen graph of the reference's own example
    model (``example_inputs/md_serial_example/deployed_serial.pt``, forward() lines 331-356):
    per l-block ``out[n,w,m] = sum_{u,v} x[n,u,m] onehot[n,v] W[u,v,w] / sqrt(mul_in n_species)``
    (the Wigner 3j of (l,0,l) and the component normalisation cancel), flat weight = concat of
    W.flatten() over the blocks in input order.  Pinned by tests/golden/fctp_hfo.npz."""

    def __init__(self, irreps_x, irreps_operand, irreps_out, data_key_x='x',
                 data_key_operand='node_attr', **kw):
        super().__init__()
        self.irreps_x, self.irreps_out = parse_irreps(irreps_x), parse_irreps(irreps_out)
        self.nsp = parse_irreps(irreps_operand)[0][0]
        self.key_x, self.key_operand = data_key_x, data_key_operand
        self.instr = [(i, j) for i, (_, l1, p1) in enumerate(self.irreps_x)
                      for j, (_, l2, p2) in enumerate(self.irreps_out) if (l1, p1) == (l2, p2)]
        fan = {}
        for i, j in self.instr:
            fan[j] = fan.get(j, 0) + self.irreps_x[i][0] * self.nsp
        self.path_weight = [1.0 / math.sqrt(fan[j]) for _, j in self.instr]
        n = sum(self.irreps_x[i][0] * self.nsp * self.irreps_out[j][0] for i, j in self.instr)
        holder = nn.Module()
        holder.weight = nn.Parameter(torch.randn(n))
        self.fc_tensor_product = holder

    def forward(self, data):
        x, onehot = data[self.key_x], data[self.key_operand].to(data[self.key_x].dtype)
        N = x.shape[0]
        offs_in = np.cumsum([0] + [m * (2 * l + 1) for m, l, _ in self.irreps_x])
        outs = [None] * len(self.irreps_out)
        w_off = 0
        for (i, j), pw in zip(self.instr, self.path_weight):
            mi, l, _ = self.irreps_x[i]
            mo = self.irreps_out[j][0]
            W = self.fc_tensor_product.weight[w_off:w_off + mi * self.nsp * mo].view(mi, self.nsp, mo)
            w_off += mi * self.nsp * mo
            xi = x[:, offs_in[i]:offs_in[i + 1]].reshape(N, mi, 2 * l + 1)
            y = torch.einsum('num,nv,uvw->nwm', xi, onehot, W) * pw
            outs[j] = y if outs[j] is None else outs[j] + y
        res = [x.new_zeros(N, mo * (2 * l + 1)) if outs[j] is None else outs[j].reshape(N, mo * (2 * l + 1))
               for j, (mo, l, _) in enumerate(self.irreps_out)]
        data['self_cont_tmp'] = torch.cat(res, -1)
        return data


class SelfConnectionOutro(nn.Module):
    """``sevenn/nn/self_connection.py:93-109``."""

    def forward(self, data):
        data['x'] = data['x'] + data['self_cont_tmp']
        del data['self_cont_tmp']
        return data


class EquivariantGate(nn.Module):
    """``sevenn/nn/equivariant_gate.py:13-61`` -> e3nn ``Gate``; dump lines 249-340: input
    ``[scalars | gates | gated]``, output ``[c act(scalars) | (c act(gates))[u] * gated[u,:]]``."""

    def __init__(self, irreps_x, act_scalar='silu', act_gate='silu'):
        super().__init__()
        irreps_x = parse_irreps(irreps_x)
        self.scalars = [t for t in irreps_x if t[1] == 0]
        self.gated = [t for t in irreps_x if t[1] > 0]
        self.gates = [(m, 0, 1) for m, _, _ in self.gated]
        self.irreps_in = simplify(self.scalars + self.gates + self.gated)
        self.irreps_out = simplify(self.scalars + self.gated)
        self.act_scalar, self.act_gate = act_scalar, act_gate
        self.cs, self.cg = _act_2mom(act_scalar), _act_2mom(act_gate)

    def get_gate_irreps_in(self):
        return self.irreps_in

    def forward(self, data):
        x = data['x']
        ns = sum(m for m, _, _ in self.scalars)
        ng = sum(m for m, _, _ in self.gates)
        s = _ACT[self.act_scalar](x[:, :ns]) * self.cs
        outs = [s]
        if ng:
            g = _ACT[self.act_gate](x[:, ns:ns + ng]) * self.cg
            off, goff = ns + ng, 0
            for m, l, _ in self.gated:
                d = 2 * l + 1
                blk = x[:, off:off + m * d].reshape(-1, m, d)
                outs.append((blk * g[:, goff:goff + m].unsqueeze(-1)).reshape(-1, m * d))
                off += m * d
                goff += m
        data['x'] = torch.cat(outs, -1)
        return data
