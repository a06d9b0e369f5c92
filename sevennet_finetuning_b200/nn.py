"""Drop-in twins of the reference's ``sevenn.nn`` modules on the B200 kernels.

Same class names, constructor arguments, parameter names/shapes and the same
``forward(data: Dict[str, Tensor]) -> Dict[str, Tensor]`` contract as the reference
(``sevenn/nn/*.py``, keys from ``sevenn/_keys.py:26-81``), so ``AtomGraphSequential.replace_module``
(``sevenn/nn/sequential.py:61-62``) can swap them in and ``load_state_dict`` of a reference checkpoint works
unchanged.  Every forward and every first-order backward is a kernel of libsevennet_b200.so reached through
``torch.autograd.Function``; per-edge intermediates (weights, messages) are recomputed in backward instead of
saved.  There is no CPU path: a CPU tensor raises.

Scope of autograd.  In ``eval()`` mode: first order, per-edge intermediates recomputed (the MD / calculator flavour).
In ``train()`` mode the modules switch to the Functions of ``higher_order.py``, which stay differentiable through their
own backward pass: ``ForceStressOutputFromEdge`` then calls ``autograd.grad(..., create_graph=True)`` exactly like the
reference (``sevenn/nn/force_output.py:165-170``) and a loss on forces / stress can be back-propagated to the parameters
(``sevenn/train/trainer.py:173-186``).  Every forward, reverse and second-order evaluation on that path is a kernel of
libsevennet_b200.so (the edge embedding's second derivatives are dual-number evaluations of its own formulas,
``b200_edge_embed_bwd2``); tensor ops remain only for the loss, the rescale/reduce scalars and the force scatter, as in
the reference.
"""
import math
import os
from collections import OrderedDict

import torch
import torch.nn as nn

from . import higher_order as ho
from . import ops
from ._lib import B200_CUTOFF_POLY, B200_CUTOFF_XPLOR, EdgeParams

C_SILU = 1.6791767923989418   # e3nn normalize2mom(silu)


# ---- graph cache: the kernels want edges sorted by destination (edge_index[0], convolution.py:112-113) ----------------
def _graph_of(data):
    ei = data['edge_index']
    g = data.get('_b200_graph')
    if g is not None and g['key'] == (ei.data_ptr(), ei.shape[1]):
        return g
    dst64, src64 = ei[0], ei[1]
    n = data['x'].shape[0] if data['x'].dim() > 0 else int(data['num_atoms'])
    if 'x_ghost' in data and data['x_ghost'].dim() > 0:
        n_src = n + data['x_ghost'].shape[0]
    else:
        n_src = n
    E = ei.shape[1]
    perm = None
    # a loader that built the edges itself can vouch for their order (finetune.collate) and spare the device->host
    # round trip of the check, which would otherwise stall the launch queue once per forward
    if E > 1 and data.get('_b200_dst_sorted') is not True and bool((dst64[1:] < dst64[:-1]).any()):
        perm = torch.argsort(dst64, stable=True)
        dst64, src64 = dst64[perm], src64[perm]
    dst, src = dst64.to(torch.int32).contiguous(), src64.to(torch.int32).contiguous()
    # CSR offsets of the sorted destinations (bincount would read its output size back from the device)
    rowptr = torch.searchsorted(dst64, torch.arange(n_src + 1, device=ei.device, dtype=dst64.dtype)).to(torch.int32)
    rrowptr, eid = ops.csr_transpose(n_src, src)
    g = dict(key=(ei.data_ptr(), ei.shape[1]), perm=perm, dst=dst, src=src, rowptr=rowptr, rrowptr=rrowptr, eid=eid,
             n_dst=n, n_src=n_src, rowptr_dst=rowptr[:n + 1].contiguous())
    data['_b200_graph'] = g
    return g


# ---- autograd functions ---------------------------------------------------------------------------------------------
class _EdgeEmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vec, coeffs, params):
        v = ops._f32(vec)
        ln, Y, emb = ops.edge_embed_fwd(params, v, ops._f32(coeffs))
        ctx.save_for_backward(v, coeffs.detach())
        ctx.params = params
        return ln, Y, emb

    @staticmethod
    def backward(ctx, g_len, gY, gemb):
        v, coeffs = ctx.saved_tensors
        p = ctx.params
        gY = None if gY is None else ops._f32(gY)
        gemb = None if gemb is None else ops._f32(gemb)
        dvec = None
        if ctx.needs_input_grad[0]:
            dvec = ops.edge_embed_bwd(p, v, ops._f32(coeffs), gY, gemb)
            if g_len is not None:
                dvec = dvec + g_len.unsqueeze(-1) * v / torch.linalg.norm(v, dim=-1, keepdim=True)
        dco = None
        if ctx.needs_input_grad[1] and gemb is not None:
            # d emb[e, n] / d k_n = (2 / r_c) cos(k_n r) f_cut(r)   (edge_embedding.py:114-116)
            dco = ops.edge_embed_dk(p, v, ops._f32(coeffs), gemb)
        return dvec, dco, None


def _num_structures(data):
    """Number of structures of a batch: the reference reads ``int(batch.max()) + 1`` (linear.py:79, force_output.py:204),
    a device->host synchronisation; the per-structure ``num_atoms`` vector of a collated batch carries the same number."""
    na = data.get('num_atoms')
    if torch.is_tensor(na) and na.dim() == 1 and na.numel() >= 1:
        return int(na.numel())
    return int(data['batch'].max()) + 1


# train() mode: weight_nn through a table over the edge length rebuilt from the current parameters every forward
# (csrc/table_train.cu) instead of one network evaluation per edge.  B200_TRAIN_TABLE=0 restores the per-edge network.
_TRAIN_TABLE = os.environ.get('B200_TRAIN_TABLE', '1') != '0'


def _second_order(module):
    """train() mode with grad enabled: the forces must stay differentiable (force_output.py:168 create_graph=self.training)."""
    return module.training and torch.is_grad_enabled()


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, W, desc, numel):
        x32, W32 = ops._f32(x), ops._f32(W)
        ctx.save_for_backward(x32, W32)
        ctx.desc, ctx.numel = desc, numel
        return ops.irreps_linear(desc, x32, W32)

    @staticmethod
    def backward(ctx, gy):
        x, W = ctx.saved_tensors
        gy = ops._f32(gy)
        dx = ops.irreps_linear(ctx.desc, gy, W, transpose=True) if ctx.needs_input_grad[0] else None
        dW = ops.irreps_linear_wgrad(ctx.desc, x, gy, ctx.numel) if ctx.needs_input_grad[1] else None
        return dx, dW, None, None


class _EmbedFn(torch.autograd.Function):
    """one_hot(type) . W / sqrt(n_species) as a row gather (node_embedding.py:39-48 + linear.py:37-44)."""

    @staticmethod
    def forward(ctx, types, W, n_species, dim):
        t32 = ops._i32(types)
        ctx.save_for_backward(t32)
        ctx.n_species, ctx.dim = n_species, dim
        return ops.embed_fwd(t32, ops._f32(W).view(n_species, dim), n_species)

    @staticmethod
    def backward(ctx, g):
        (t,) = ctx.saved_tensors
        dW = torch.zeros(ctx.n_species, ctx.dim, dtype=torch.float32, device=g.device)
        dW.index_add_(0, t.long(), ops._f32(g) / math.sqrt(ctx.n_species))
        return None, dW.reshape(-1), None, None


class _GateFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, ns, m1, m2):
        x32 = ops._f32(x)
        ctx.save_for_backward(x32)
        ctx.dims = (ns, m1, m2)
        return ops.gate_fwd(ns, m1, m2, x32)

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        return ops.gate_bwd(*ctx.dims, x, ops._f32(g)), None, None, None


class _ConvFn(torch.autograd.Function):
    """weight_nn + uvu tensor product + scatter / denominator (convolution.py:105-121); edges sorted by dst."""

    @staticmethod
    def forward(ctx, x, Y, emb, W0, W1, W2, denom, graph, muls, lout, denom_host):
        x32, Y32, e32 = ops._f32(x), ops._f32(Y), ops._f32(emb)
        Ws = [ops._f32(W) for W in (W0, W1, W2)]
        w = ops.radial_mlp_fwd(e32, *Ws)
        d = denom_host
        out = ops.conv_fwd(muls, lout, graph['rowptr'][:graph['n_dst'] + 1].contiguous(), graph['src'], x32, Y32, w, d)
        ctx.save_for_backward(x32, Y32, e32, *Ws, out)
        ctx.graph, ctx.muls, ctx.lout, ctx.denom = graph, muls, lout, d
        return out

    @staticmethod
    def backward(ctx, gout):
        x, Y, emb, W0, W1, W2, out = ctx.saved_tensors
        g, muls, lout, d = ctx.graph, ctx.muls, ctx.lout, ctx.denom
        gout = ops._f32(gout)
        rowptr = g['rowptr'][:g['n_dst'] + 1].contiguous()
        w = ops.radial_mlp_fwd(emb, W0, W1, W2)                       # recomputed, never saved
        if ctx.needs_input_grad[0]:
            dw, dY, dx = ops.conv_bwd(muls, lout, x.shape[0], rowptr, g['src'], x, Y, w, gout, d)
        else:
            (dw, dY), dx = ops.conv_bwd_edge(muls, lout, rowptr, g['src'], x, Y, w, gout, d), None
        dY[:, 0] = 0.0
        demb = ops.radial_mlp_bwd(emb, W0, W1, W2, dw) if ctx.needs_input_grad[2] else None
        dW0 = dW1 = dW2 = None
        if any(ctx.needs_input_grad[3:6]):
            # parameter gradients of the radial MLP over the edge axis with the library's own kernels: a dense layer is the l = 0
            # block of an irreps linear (b200_irreps_linear / its transpose / b200_irreps_linear_wgrad), c * silu a gate of scalars
            nb, h1n, h2n, wn = W0.shape[0], W0.shape[1], W1.shape[1], W2.shape[1]
            (d0, n0), (d1, n1), (d2, n2) = (ops.make_lin_desc(f'{a}x0e', f'{b}x0e') for a, b in ((nb, h1n), (h1n, h2n), (h2n, wn)))
            z1 = ops.irreps_linear(d0, emb, W0.reshape(-1))
            a1 = ops.gate_fwd(h1n, 0, 0, z1)
            z2 = ops.irreps_linear(d1, a1, W1.reshape(-1))
            a2 = ops.gate_fwd(h2n, 0, 0, z2)
            dW2 = ops.irreps_linear_wgrad(d2, a2, dw, n2).view_as(W2)
            dz2 = ops.gate_bwd(h2n, 0, 0, z2, ops.irreps_linear(d2, dw, W2.reshape(-1), transpose=True))
            dW1 = ops.irreps_linear_wgrad(d1, a1, dz2, n1).view_as(W1)
            dz1 = ops.gate_bwd(h1n, 0, 0, z1, ops.irreps_linear(d1, dz2, W1.reshape(-1), transpose=True))
            dW0 = ops.irreps_linear_wgrad(d0, emb, dz1, n0).view_as(W0)
        dden = -(out * gout).sum().reshape(1) / d if ctx.needs_input_grad[6] else None
        return dx, dY, demb, dW0, dW1, dW2, dden, None, None, None, None


# ---- modules ----------------------------------------------------------------------------------------------------------
class BesselBasis(nn.Module):
    """sevenn/nn/edge_embedding.py:80-116 (parameters only; evaluated inside the edge kernel)."""

    def __init__(self, cutoff_length, bessel_basis_num=8, trainable_coeff=True):
        super().__init__()
        self.num_basis, self.cutoff_length = bessel_basis_num, cutoff_length
        c = torch.tensor([n * math.pi / cutoff_length for n in range(1, bessel_basis_num + 1)], dtype=torch.float32)
        if trainable_coeff:
            self.coeffs = nn.Parameter(c)
        else:
            self.register_buffer('coeffs', c)


class PolynomialCutoff(nn.Module):
    """sevenn/nn/edge_embedding.py:119-145."""

    def __init__(self, cutoff_length, poly_cut_p_value=6):
        super().__init__()
        self.cutoff_length, self.p = cutoff_length, poly_cut_p_value


class XPLORCutoff(nn.Module):
    """sevenn/nn/edge_embedding.py:148-173."""

    def __init__(self, cutoff_length, cutoff_on):
        super().__init__()
        self.cutoff_length, self.cutoff_on = cutoff_length, cutoff_on


class SphericalEncoding(nn.Module):
    """sevenn/nn/edge_embedding.py:177-198 (lmax 2, component normalisation)."""

    def __init__(self, lmax, parity=-1, normalization='component', normalize=True):
        super().__init__()
        if lmax != 2 or normalization != 'component':
            raise ValueError('B200 path: lmax = 2 with component normalisation only')
        self.lmax, self.normalize = lmax, normalize


class EdgeEmbedding(nn.Module):
    """sevenn/nn/edge_embedding.py:202-230."""

    def __init__(self, basis_module, cutoff_module, spherical_module):
        super().__init__()
        self.basis_function, self.cutoff_function, self.spherical = basis_module, cutoff_module, spherical_module

    def _params(self):
        c, b = self.cutoff_function, self.basis_function
        if isinstance(c, XPLORCutoff):
            return EdgeParams(b.num_basis, B200_CUTOFF_XPLOR, int(self.spherical.normalize), float(c.cutoff_length),
                              float(c.cutoff_on), 6.0)
        return EdgeParams(b.num_basis, B200_CUTOFF_POLY, int(self.spherical.normalize), float(c.cutoff_length), 0.0, float(c.p))

    def radial_grid(self, device):
        """Grid of the train-mode radial table (csrc/table_train.cu): (inv_h, n_grid, B, dB) with B[i, n] = bessel_n(r_i) *
        cutoff(r_i) and dB = dB / dr on the n_grid + 1 points r_i = i / inv_h, float64, differentiable with respect to the
        Bessel frequencies.  edge_embedding.py:80-116 (2/r_c sin(k_n r)/r), :119-145 (polynomial), :148-173 (XPLOR).  The
        spacing is about r_c / 4000, adjusted so that the XPLOR switching radius is a grid point (the functions are smooth
        inside every interval); inv_h is an fp32 number: the kernels locate an edge with t = r * inv_h - i."""
        c, b = self.cutoff_function, self.basis_function
        rc = float(c.cutoff_length)
        h = rc / 4000.0
        r_on = float(c.cutoff_on) if isinstance(c, XPLORCutoff) else 0.0
        if 0.0 < r_on < rc and math.floor(r_on / h + 0.5) >= 1:
            h = r_on / math.floor(r_on / h + 0.5)
        inv_h = float(torch.tensor(1.0 / h, dtype=torch.float32))
        n_grid = int(math.ceil(rc * inv_h)) + 1
        # (first point 1e-6 instead of 0: sin(k r) / r; no edge is that short)
        r = (torch.arange(n_grid + 1, dtype=torch.float64, device=device) / inv_h).clamp_(min=1e-6)
        k = b.coeffs.to(torch.float64)
        ur, kr = r.unsqueeze(1), r.unsqueeze(1) * k.unsqueeze(0)
        bes = (2.0 / rc) * torch.sin(kr) / ur
        dbes = (2.0 / rc) * (k.unsqueeze(0) * torch.cos(kr) / ur - torch.sin(kr) / (ur * ur))
        if isinstance(c, XPLORCutoff):
            u, v, den = rc * rc - r * r, rc * rc + 2.0 * r * r - 3.0 * r_on * r_on, (rc * rc - r_on * r_on) ** 3
            inner = r < r_on
            env = torch.where(inner, torch.ones_like(r), u * u * v / den)
            denv = torch.where(inner, torch.zeros_like(r), (-4.0 * r * u * v + 4.0 * r * u * u) / den)
        else:
            pw, x = float(c.p), r / rc
            c0, c1, c2 = (pw + 1.0) * (pw + 2.0) / 2.0, pw * (pw + 2.0), pw * (pw + 1.0) / 2.0
            env = 1.0 - c0 * x ** pw + c1 * x ** (pw + 1.0) - c2 * x ** (pw + 2.0)
            denv = (-c0 * pw * x ** (pw - 1.0) + c1 * (pw + 1.0) * x ** pw - c2 * (pw + 2.0) * x ** (pw + 1.0)) / rc
        return inv_h, n_grid, bes * env.unsqueeze(1), dbes * env.unsqueeze(1) + bes * denv.unsqueeze(1)

    def forward(self, data):
        if _second_order(self):
            vec = data['edge_vec']
            if not vec.is_cuda:
                raise RuntimeError('sevennet_b200: CPU tensor passed to a CUDA-only module')
            if vec.dtype != torch.float32:
                vec = vec.to(torch.float32)
            ln, Y, emb = ho.EdgeF.apply(vec, self.basis_function.coeffs, self._params())
            if _TRAIN_TABLE:
                # the convolutions read weight_nn from a table over the edge length (IrrepsConvolution._table_weights)
                inv_h, n_grid, B, dB = self.radial_grid(vec.device)
                data['_b200_radial_table'] = {'inv_h': inv_h, 'n_grid': n_grid, 'B': B, 'dB': dB,
                                              'r': torch.linalg.vector_norm(vec, dim=1)}
        else:
            ln, Y, emb = _EdgeEmbedFn.apply(data['edge_vec'], self.basis_function.coeffs, self._params())
        data['edge_length'], data['edge_attr'], data['edge_embedding'] = ln, Y, emb
        return data


class OnehotEmbedding(nn.Module):
    """sevenn/nn/node_embedding.py:15-48: x (species index) -> one-hot x, node_attr, atom_type."""

    def __init__(self, num_classes, data_key_x='x', data_key_save='atom_type', data_key_additional='node_attr'):
        super().__init__()
        self.num_classes = num_classes
        self.key_x, self.key_save, self.key_additional_output = data_key_x, data_key_save, data_key_additional

    def forward(self, data):
        idx = data[self.key_x]
        if not idx.is_cuda:
            raise RuntimeError('sevennet_b200: CPU tensor passed to a CUDA-only module')
        data[self.key_save] = idx
        onehot = torch.nn.functional.one_hot(idx.long(), self.num_classes).to(torch.float32)
        data[self.key_x] = onehot
        if self.key_additional_output is not None:
            data[self.key_additional_output] = onehot
        return data


class IrrepsLinear(nn.Module):
    """sevenn/nn/linear.py:14-49; parameter ``linear.weight`` is the flat e3nn weight vector."""

    class _Holder(nn.Module):
        def __init__(self, numel):
            super().__init__()
            self.weight = nn.Parameter(torch.randn(numel))

    def __init__(self, irreps_in, irreps_out, data_key_in, data_key_out=None, is_embed=False, **e3nn_linear_params):
        super().__init__()
        if e3nn_linear_params.get('biases', False):
            raise ValueError('B200 path: IrrepsLinear with biases=True (use_bias_in_linear) is not supported')
        self.key_input = data_key_in
        self.key_output = data_key_in if data_key_out is None else data_key_out
        self.irreps_in, self.irreps_out, self.is_embed = irreps_in, irreps_out, is_embed
        self.desc, self.numel = ops.make_lin_desc(irreps_in, irreps_out)
        self.linear = IrrepsLinear._Holder(self.numel)
        if is_embed:   # the reference pre-divides the initial weight by the path weight (linear.py:37-44)
            with torch.no_grad():
                self.linear.weight.div_(1.0 / math.sqrt(ops.muls_of(irreps_in)[0]))

    def forward(self, data):
        x = data[self.key_input]
        if self.is_embed and 'atom_type' in data:
            # one-hot input: a row gather of W / sqrt(n_species) instead of a dense product with 89x wasted work
            nsp, dim = ops.muls_of(self.irreps_in)[0], ops.muls_of(self.irreps_out)[0]
            fn = ho.EmbedF if _second_order(self) else _EmbedFn
            data[self.key_output] = fn.apply(data['atom_type'], self.linear.weight, nsp, dim)
        else:
            fn = ho.LinF if _second_order(self) else _LinearFn
            data[self.key_output] = fn.apply(x, self.linear.weight, self.desc, self.numel)
        return data


class SelfConnectionLinearIntro(nn.Module):
    """sevenn/nn/self_connection.py:42-62."""

    def __init__(self, irreps_x, irreps_out, data_key_x='x', **kwargs):
        super().__init__()
        self.irreps_x, self.irreps_out, self.key_x = irreps_x, irreps_out, data_key_x
        self.desc, self.numel = ops.make_lin_desc(irreps_x, irreps_out)
        self.linear = IrrepsLinear._Holder(self.numel)

    def forward(self, data):
        fn = ho.LinF if _second_order(self) else _LinearFn
        data['self_cont_tmp'] = fn.apply(data[self.key_x], self.linear.weight, self.desc, self.numel)
        return data


def _species_groups(data):
    """[(species, node indices)] of the batch and the permutation that undoes their concatenation; one host sync per
    forward (cached in the dict, all layers share it)."""
    g = data.get('_b200_species_groups')
    if g is None:
        t = data['atom_type']
        order = torch.argsort(t, stable=True)
        counts = torch.bincount(t).tolist()
        groups, o = [], 0
        for sp, c in enumerate(counts):
            if c > 0:
                groups.append((sp, order[o:o + c]))
                o += c
        inv = torch.empty_like(order)
        inv[order] = torch.arange(order.numel(), device=order.device)
        g = data['_b200_species_groups'] = (groups, inv)
    return g


class SelfConnectionIntro(nn.Module):
    """sevenn/nn/self_connection.py:11-38 ('nequip' self connection): FullyConnectedTensorProduct of the node features
    with the one-hot species vector, parameter ``fc_tensor_product.weight`` = concat over the l-blocks of
    W[mul_in, n_species, mul_out].  With a one-hot operand this is an irreps linear whose weight matrix is chosen by the
    node's species, W[:, type_n, :] / sqrt(mul_in n_species): the nodes are grouped by species and each group runs the
    irreps-linear kernels (and their closed reverse / second-order family) with its slice of the weight."""

    def __init__(self, irreps_x, irreps_operand, irreps_out, data_key_x='x', data_key_operand='node_attr', **kwargs):
        super().__init__()
        self.irreps_x, self.irreps_out, self.key_x, self.key_operand = irreps_x, irreps_out, data_key_x, data_key_operand
        self.nsp = ops.muls_of(irreps_operand)[0]
        self.desc, self.numel = ops.make_lin_desc(irreps_x, irreps_out)
        self.fc_tensor_product = IrrepsLinear._Holder(self.numel * self.nsp)
        # position of W_b[u, s, w] in the flat FCTP weight for every position (b, u, w) of the linear's flat weight
        sel = torch.empty(self.nsp, self.numel, dtype=torch.long)
        for k in range(self.desc.nblk):
            b = self.desc.blk[k]
            u = torch.arange(b.mul_in).view(-1, 1)
            w = torch.arange(b.mul_out).view(1, -1)
            for sp in range(self.nsp):
                sel[sp, b.w_off:b.w_off + b.mul_in * b.mul_out] = (self.nsp * b.w_off + (u * self.nsp + sp) * b.mul_out + w).reshape(-1)
        self.register_buffer('_sel', sel, persistent=False)

    def forward(self, data):
        x = data[self.key_x]
        groups, inv = _species_groups(data)
        fn = ho.LinF if _second_order(self) else _LinearFn
        W = self.fc_tensor_product.weight * (1.0 / math.sqrt(self.nsp))
        ys = [fn.apply(x.index_select(0, idx), W.index_select(0, self._sel[sp]), self.desc, self.numel) for sp, idx in groups]
        data['self_cont_tmp'] = torch.cat(ys, 0).index_select(0, inv)
        return data


class SelfConnectionOutro(nn.Module):
    """sevenn/nn/self_connection.py:93-109."""

    def __init__(self, data_key_x='x'):
        super().__init__()
        self.key_x = data_key_x

    def forward(self, data):
        data[self.key_x] = data[self.key_x] + data['self_cont_tmp']
        del data['self_cont_tmp']
        return data


class _RadialNet(nn.Module):
    """parameter holder with e3nn FullyConnectedNet's names: layer{k}.weight [in, out]."""

    class _Layer(nn.Module):
        def __init__(self, a, b):
            super().__init__()
            self.weight = nn.Parameter(torch.randn(a, b))

    def __init__(self, hs):
        super().__init__()
        for k, (a, b) in enumerate(zip(hs[:-1], hs[1:])):
            setattr(self, f'layer{k}', _RadialNet._Layer(a, b))


class IrrepsConvolution(nn.Module):
    """sevenn/nn/convolution.py:36-123 (silu radial activation, 3-layer weight_nn, filter 0e+1e+2e)."""

    def __init__(self, irreps_x, irreps_filter, irreps_out, weight_layer_input_to_hidden, weight_layer_act='silu',
                 denominator=1.0, weight_shift=0.0, weight_scale=1.0, train_denominator=False, data_key_x='x',
                 data_key_filter='edge_attr', data_key_weight_input='edge_embedding', data_key_edge_idx='edge_index',
                 is_parallel=False):
        super().__init__()
        if weight_shift != 0.0 or weight_scale != 1.0:
            raise ValueError('B200 path: weight_shift / weight_scale other than 0 / 1 are not supported')
        if len(weight_layer_input_to_hidden) != 3:
            raise ValueError('B200 path: weight_nn with two hidden layers')
        self.muls = ops.muls_of(irreps_x)
        louts = [l for _, l in ops.parse_irreps(irreps_out)]
        self.lout = max(louts)
        self.weight_numel, self.out_dim = ops.conv_dims(*self.muls, self.lout)
        self.denominator = nn.Parameter(torch.tensor([float(denominator)]), requires_grad=train_denominator)
        self.weight_nn = _RadialNet(list(weight_layer_input_to_hidden) + [self.weight_numel])
        self.key_x, self.key_filter, self.key_weight_input, self.key_edge_idx = (
            data_key_x, data_key_filter, data_key_weight_input, data_key_edge_idx)
        self.is_parallel = is_parallel
        self._comm_size = self.muls[0] + 3 * self.muls[1] + 5 * self.muls[2]   # convolution.py:102
        self._mlp_descs = None

    def _denominator_host(self):
        # host copy of the denominator, refreshed only when the parameter changed (optimizer step, load_state_dict): reading
        # it every forward is a device->host synchronisation per layer
        ver = (self.denominator._version, self.denominator.data_ptr())
        if getattr(self, '_den_cache', None) is None or self._den_cache[0] != ver:
            self._den_cache = (ver, float(self.denominator.detach()))
        return self._den_cache[1]

    def _table_weights(self, tab, g):
        """weight_nn (convolution.py:97-106; e3nn FullyConnectedNet: W / sqrt(fan_in), normalize2mom silu) as a function of
        the edge length: value and r-derivative of the network on the grid points (forward mode, float64, plain tensor
        operations: differentiable to any order with respect to the weights and the Bessel frequencies), turned into
        per-interval cubic Hermite coefficients, then read per edge by the kernels of csrc/table_train.cu."""
        nn_ = self.weight_nn
        c = 1.6791767923989418   # normalize2mom(silu)
        z, dz = tab['B'], tab['dB']
        for W, last in ((nn_.layer0.weight, False), (nn_.layer1.weight, False), (nn_.layer2.weight, True)):
            W64 = W.to(torch.float64) / math.sqrt(W.shape[0])
            z, dz = z @ W64, dz @ W64
            if not last:
                sg = torch.sigmoid(z)
                z, dz = c * z * sg, c * sg * (1.0 + z * (1.0 - sg)) * dz
        P = ho.HermF.apply(z, dz, 1.0 / tab['inv_h'])
        key = None if g['perm'] is None else g['perm'].data_ptr()
        if tab.get('key', 0) != key:   # edge order of this graph: the same for every layer
            r = tab['r'] if g['perm'] is None else tab['r'].index_select(0, g['perm'])
            bins = (r.detach() * tab['inv_h']).to(torch.int32).clamp_(0, tab['n_grid'] - 1)
            tab['key'], tab['r_g'], tab['order'] = key, r, torch.argsort(bins).to(torch.int32)
        return ho.TabInterp.apply(P, tab['r_g'], None, 0, (tab['inv_h'], tab['order']))

    def forward(self, data):
        g = _graph_of(data)
        x = data[self.key_x]
        if self.is_parallel:
            x = torch.cat([x, data['x_ghost']])
        Y, emb = data[self.key_filter], data[self.key_weight_input]
        if g['perm'] is not None:
            Y, emb = Y.index_select(0, g['perm']), emb.index_select(0, g['perm'])
        nn_ = self.weight_nn
        if _second_order(self):
            tab = data.get('_b200_radial_table')
            if tab is not None:
                w = self._table_weights(tab, g)
            else:
                hs = [nn_.layer0.weight.shape[0], nn_.layer0.weight.shape[1], nn_.layer1.weight.shape[1], self.weight_numel]
                if self._mlp_descs is None:
                    self._mlp_descs = [ops.make_lin_desc(f'{a}x0e', f'{b}x0e') for a, b in zip(hs[:-1], hs[1:])]
                w = ho.radial_mlp(emb, nn_.layer0.weight, nn_.layer1.weight, nn_.layer2.weight, self._mlp_descs)
            if self.denominator.requires_grad:   # train_denominator: keep 1/denominator in the graph
                out = ho.ConvF.apply(x, Y, w, g, self.muls, self.lout, 1.0) / self.denominator
            else:
                out = ho.ConvF.apply(x, Y, w, g, self.muls, self.lout, self._denominator_host())
            data[self.key_x] = out
            return data
        out = _ConvFn.apply(x, Y, emb, nn_.layer0.weight, nn_.layer1.weight, nn_.layer2.weight, self.denominator, g,
                            self.muls, self.lout, self._denominator_host())
        data[self.key_x] = out   # rows of the local atoms only (convolution.py:119-121)
        return data


class EquivariantGate(nn.Module):
    """sevenn/nn/equivariant_gate.py:13-61 (silu on scalars and gates, even parity)."""

    def __init__(self, irreps_x, act_scalar_dict=None, act_gate_dict=None, data_key_x='x'):
        super().__init__()
        self.key_x = data_key_x
        self.muls = ops.muls_of(irreps_x)
        self.irreps_out = irreps_x

    def get_gate_irreps_in(self):
        m0, m1, m2 = self.muls
        parts = [f'{m0 + m1 + m2}x0e'] + ([f'{m1}x1e'] if m1 else []) + ([f'{m2}x2e'] if m2 else [])
        return '+'.join(parts)

    def forward(self, data):
        fn = ho.GateF if _second_order(self) else _GateFn
        data[self.key_x] = fn.apply(data[self.key_x], *self.muls)
        return data


class Rescale(nn.Module):
    """sevenn/nn/scale.py:12-38."""

    def __init__(self, shift, scale, data_key_in='scaled_atomic_energy', data_key_out='atomic_energy',
                 train_shift_scale=False, **kwargs):
        super().__init__()
        self.shift = nn.Parameter(torch.tensor([float(shift)]), requires_grad=train_shift_scale)
        self.scale = nn.Parameter(torch.tensor([float(scale)]), requires_grad=train_shift_scale)
        self.key_input, self.key_output = data_key_in, data_key_out

    def forward(self, data):
        data[self.key_output] = data[self.key_input] * self.scale + self.shift
        return data


class SpeciesWiseRescale(nn.Module):
    """sevenn/nn/scale.py:42-73."""

    def __init__(self, shift, scale, data_key_in='scaled_atomic_energy', data_key_out='atomic_energy',
                 data_key_indices='atom_type', train_shift_scale=False):
        super().__init__()
        self.shift = nn.Parameter(torch.tensor(shift, dtype=torch.float32), requires_grad=train_shift_scale)
        self.scale = nn.Parameter(torch.tensor(scale, dtype=torch.float32), requires_grad=train_shift_scale)
        self.key_input, self.key_output, self.key_indices = data_key_in, data_key_out, data_key_indices

    def forward(self, data):
        idx = data[self.key_indices].long()
        data[self.key_output] = data[self.key_input] * self.scale[idx].view(-1, 1) + self.shift[idx].view(-1, 1)
        return data


class AtomReduce(nn.Module):
    """sevenn/nn/linear.py:53-90."""

    def __init__(self, data_key_in, data_key_out, reduce='sum', constant=1.0):
        super().__init__()
        self.key_input, self.key_output, self.constant = data_key_in, data_key_out, constant
        self._is_batch_data = True

    def forward(self, data):
        src = data[self.key_input].squeeze(1)
        if self._is_batch_data:
            size = _num_structures(data)
            out = torch.zeros(size, dtype=src.dtype, device=src.device)
            out.scatter_reduce_(0, data['batch'], src, reduce='sum')
            data[self.key_output] = out * self.constant
        else:
            data[self.key_output] = torch.sum(src) * self.constant
        return data


class _EdgeVecFn(torch.autograd.Function):
    """edge_vec = pos[idx1] - pos[idx0] + shift @ cell at zero strain (sevenn/nn/edge_embedding.py:62-75) with the reverse pass
    on the CUDA force / virial kernel: d E / d pos is minus the force sum of b200_force_virial and d E / d strain the
    symmetrised sum of r (x) dE/dr (what autograd makes of the strain trick, :33-60).  First order (eval) only."""

    @staticmethod
    def forward(ctx, pos, strain, cell, shift, batch, g, batched):
        ei_dst, ei_src = g['dst'].long(), g['src'].long()          # destination-sorted edge list of the graph cache
        sh = shift if g['perm'] is None else shift.index_select(0, g['perm'])
        if batched:
            vec = pos[ei_src] - pos[ei_dst] + torch.einsum('ni,nij->nj', sh, cell[batch[ei_dst]])
        else:
            vec = pos[ei_src] - pos[ei_dst] + sh @ cell
        vec = vec.float().contiguous()
        ctx.g, ctx.batched, ctx.nb = g, batched, (strain.shape[0] if batched else 0)
        ctx.save_for_backward(vec, batch)
        if g['perm'] is not None:                                   # hand the vectors back in the caller's edge order
            out = torch.empty_like(vec)
            out[g['perm']] = vec
            return out
        return vec

    @staticmethod
    def backward(ctx, gvec):
        vec, batch = ctx.saved_tensors
        g = ctx.g
        gv = ops._f32(gvec)
        if g['perm'] is not None:
            gv = gv.index_select(0, g['perm']).contiguous()
        force, vir = ops.force_virial(g['rowptr'], g['rrowptr'], g['eid'], vec, gv)    # F = -dE/dpos; vir per source atom
        if ctx.batched:
            v6 = torch.zeros(ctx.nb, 6, dtype=vir.dtype, device=vir.device).index_add_(0, batch, vir[:batch.shape[0]])
        else:
            v6 = ops.colsum(vir).view(1, 6)
        # [xx, yy, zz, xy, yz, zx] -> symmetric 3x3 (sum_e r (x) g is symmetric for a rotation-invariant energy)
        idx = torch.tensor([0, 3, 5, 3, 1, 4, 5, 4, 2], device=v6.device)
        dstrain = v6[:, idx].view(-1, 3, 3)
        if not ctx.batched:
            dstrain = dstrain[0]
        return -force[:batch.shape[0]] if ctx.batched else -force, dstrain, None, None, None, None, None


class EdgePreprocess(nn.Module):
    """sevenn/nn/edge_embedding.py:13-77: edge vectors from positions, image shifts and the cell, with a zero strain tensor
    as a differentiable leaf (`_strain`) so that stress = -(dE/d strain)/V.  Prepended by scripts/deploy.py:22."""

    def __init__(self, is_stress=True):
        super().__init__()
        self.is_stress = is_stress
        self._is_batch_data = True

    def forward(self, data):
        pos, shift, batch = data['pos'], data['pbc_shift'], data['batch']
        cell = data['cell_lattice_vectors'].view(-1, 3, 3) if self._is_batch_data else data['cell_lattice_vectors'].view(3, 3)
        if self.training or not pos.is_cuda:
            # second-order (training) flavour: plain differentiable tensor algebra, exactly the reference's lines
            if self.is_stress:
                if self._is_batch_data:
                    strain = torch.zeros((_num_structures(data), 3, 3), dtype=pos.dtype, device=pos.device, requires_grad=True)
                    sym = 0.5 * (strain + strain.transpose(-1, -2))
                    pos = pos + torch.bmm(pos.unsqueeze(-2), sym[batch]).squeeze(-2)
                    cell = cell + torch.bmm(cell, sym)
                else:
                    strain = torch.zeros((3, 3), dtype=pos.dtype, device=pos.device, requires_grad=True)
                    sym = 0.5 * (strain + strain.transpose(-1, -2))
                    pos = pos + torch.mm(pos, sym)
                    cell = cell + torch.mm(cell, sym)
                data['_strain'] = strain
            i0, i1 = data['edge_index'][0], data['edge_index'][1]
            vec = pos[i1] - pos[i0]
            vec = vec + (torch.einsum('ni,nij->nj', shift, cell[batch[i0]]) if self._is_batch_data
                         else torch.einsum('ni,ij->nj', shift, cell))
        else:
            nb = _num_structures(data) if self._is_batch_data else 0
            strain = torch.zeros((nb, 3, 3) if self._is_batch_data else (3, 3), dtype=pos.dtype, device=pos.device,
                                 requires_grad=self.is_stress)
            if self.is_stress:
                data['_strain'] = strain
            vec = _EdgeVecFn.apply(pos, strain, cell, shift, batch, _graph_of(data), self._is_batch_data)
        data['edge_vec'] = vec
        data['edge_length'] = torch.linalg.norm(vec, dim=-1)
        return data


class ForceOutput(nn.Module):
    """sevenn/nn/force_output.py:12-48: F = -dE/dpos."""

    def __init__(self, data_key_pos='pos', data_key_energy='inferred_total_energy', data_key_force='inferred_force'):
        super().__init__()
        self.key_pos, self.key_energy, self.key_force = data_key_pos, data_key_energy, data_key_force

    def forward(self, data):
        (grad,) = torch.autograd.grad([data[self.key_energy].sum()], [data[self.key_pos]], create_graph=self.training,
                                      allow_unused=True)
        if grad is not None:
            data[self.key_force] = torch.neg(grad)
        return data


class ForceStressOutput(nn.Module):
    """sevenn/nn/force_output.py:52-130: F = -dE/dpos, stress = -(dE/d strain)/V in the order [xx, yy, zz, xy, yz, zx], volume
    clamped at 1e-3 (:98-102).  Swapped in by scripts/deploy.py:23; needs EdgePreprocess in front (`_strain`)."""

    def __init__(self, data_key_pos='pos', data_key_energy='inferred_total_energy', data_key_force='inferred_force',
                 data_key_stress='inferred_stress', data_key_cell_volume='cell_volume'):
        super().__init__()
        self.key_pos, self.key_energy, self.key_force = data_key_pos, data_key_energy, data_key_force
        self.key_stress, self.key_cell_volume = data_key_stress, data_key_cell_volume
        self._is_batch_data = True

    def forward(self, data):
        grad = torch.autograd.grad([data[self.key_energy].sum()], [data[self.key_pos], data['_strain']],
                                   create_graph=self.training, allow_unused=True)
        if grad[0] is not None:
            data[self.key_force] = torch.neg(grad[0])
        volume = data[self.key_cell_volume]
        vlim = 1e-3
        if self._is_batch_data:
            volume = torch.clamp(volume, min=vlim)
        elif float(volume) < vlim:
            volume = torch.tensor(vlim, device=volume.device)
        sgrad = grad[1]
        if sgrad is not None:
            if self._is_batch_data:
                s = torch.neg(sgrad) / volume.view(-1, 1, 1)
                data[self.key_stress] = torch.stack((s[:, 0, 0], s[:, 1, 1], s[:, 2, 2], s[:, 0, 1], s[:, 1, 2], s[:, 0, 2]), dim=1)
            else:
                s = torch.neg(sgrad) / volume
                data[self.key_stress] = torch.stack((s[0, 0], s[1, 1], s[2, 2], s[0, 1], s[1, 2], s[0, 2]))
        return data


class ForceStressOutputFromEdge(nn.Module):
    """sevenn/nn/force_output.py:133-215: forces and stress from dE/d edge_vec."""

    def __init__(self, data_key_edge='edge_vec', data_key_edge_idx='edge_index', data_key_energy='inferred_total_energy',
                 data_key_force='inferred_force', data_key_stress='inferred_stress', data_key_cell_volume='cell_volume'):
        super().__init__()
        self.key_edge, self.key_edge_idx, self.key_energy = data_key_edge, data_key_edge_idx, data_key_energy
        self.key_force, self.key_stress, self.key_cell_volume = data_key_force, data_key_stress, data_key_cell_volume
        self._is_batch_data = True

    def forward(self, data):
        energy = data[self.key_energy]
        rij = data[self.key_edge]
        if _second_order(self):
            with ho.force_pass():
                (fij,) = torch.autograd.grad([energy.sum()], [rij], create_graph=True, allow_unused=True)
        else:
            (fij,) = torch.autograd.grad([energy.sum()], [rij], create_graph=False, allow_unused=True)
        g = _graph_of(data)
        n = g['n_src']
        if fij is None:
            fij = torch.zeros_like(rij)
        if not _second_order(self) and rij.is_cuda:
            # first-order flavour (calculator, MD): the CUDA force / virial kernel on the destination-sorted edge list
            r_s, f_s = ops._f32(rij.detach()), ops._f32(fij)
            if g['perm'] is not None:
                r_s, f_s = r_s.index_select(0, g['perm']).contiguous(), f_s.index_select(0, g['perm']).contiguous()
            force, vir = ops.force_virial(g['rowptr'], g['rrowptr'], g['eid'], r_s, f_s)
            data[self.key_force] = force
            if self._is_batch_data:
                per = torch.zeros(_num_structures(data), 6, dtype=vir.dtype, device=vir.device).index_add_(0, data['batch'], vir[:data['batch'].shape[0]])
                data[self.key_stress] = -per / data[self.key_cell_volume].view(-1, 1)
            else:
                data[self.key_stress] = -ops.colsum(vir) / data[self.key_cell_volume]
            return data
        ei = data[self.key_edge_idx]
        force = torch.zeros(n, 3, dtype=fij.dtype, device=fij.device)
        force.index_add_(0, ei[0], fij)
        force.index_add_(0, ei[1], -fij)
        data[self.key_force] = force
        diag = rij * fij
        s12 = rij[:, 0] * fij[:, 1]
        s23 = rij[:, 1] * fij[:, 2]
        s31 = rij[:, 2] * fij[:, 0]
        vir = torch.cat([diag, s12.unsqueeze(-1), s23.unsqueeze(-1), s31.unsqueeze(-1)], dim=-1)
        if self._is_batch_data:
            batch = data['batch']
            nb = _num_structures(data)
            per = torch.zeros(nb, 6, dtype=vir.dtype, device=vir.device)
            per.index_add_(0, batch[ei[1]], vir)
            data[self.key_stress] = -per / data[self.key_cell_volume].view(-1, 1)
        else:
            data[self.key_stress] = -vir.sum(0) / data[self.key_cell_volume]
        return data


class AtomGraphSequential(nn.Sequential):
    """sevenn/nn/sequential.py:15-90."""

    def __init__(self, modules, cutoff=0.0, type_map=None):
        if not isinstance(modules, OrderedDict):
            modules = OrderedDict(modules)
        super().__init__(modules)
        self.cutoff, self.type_map = cutoff, type_map
        self._is_batch_data = True

    def set_is_batch_data(self, flag):
        for m in self:
            if hasattr(m, '_is_batch_data'):
                m._is_batch_data = flag
        self._is_batch_data = flag

    def prepand_module(self, key, module):
        # (torch >= 2.x keeps _modules in a plain dict: rebuild it with the new key first instead of move_to_end)
        rest = [(k, v) for k, v in self._modules.items() if k != key]
        self._modules.clear()
        self._modules[key] = module
        for k, v in rest:
            self._modules[k] = v

    def replace_module(self, key, module):
        self._modules.update({key: module})

    def delete_module_by_key(self, key):
        if key in self._modules:
            del self._modules[key]

    def forward(self, data):
        for m in self:
            data = m(data)
        data.pop('_b200_graph', None)
        return data


def build_E3_equivariant_model(config=None):
    """The module chain of ``sevenn/model_build.py:186-445`` ('nequip' interaction, 'linear' or 'nequip' self connection,
    ``irreps_manual``) with the reference's module names, built from the B200 twins."""
    from .model import SEVENNET0_CONFIG
    cfg = dict(SEVENNET0_CONFIG)
    cfg.update(config or {})
    if cfg.get('use_bias_in_linear', False):
        raise ValueError('B200 path: use_bias_in_linear is not supported')
    cutoff, nsp = cfg['cutoff'], cfg['num_species']
    man = cfg['irreps_manual']
    nconv = cfg['num_convolution_layer']
    cf = cfg['cutoff_function']
    env = (XPLORCutoff(cutoff, cf['cutoff_on']) if cf['name'].upper() == 'XPLOR'
           else PolynomialCutoff(cutoff, cf.get('poly_cut_p_value', 6)))
    L = OrderedDict()
    L['edge_embedding'] = EdgeEmbedding(BesselBasis(cutoff, cfg['bessel_basis_num']), env,
                                        SphericalEncoding(cfg['lmax'], 1, normalize=cfg.get('normalize_sph', True)))
    L['onehot_idx_to_onehot'] = OnehotEmbedding(nsp)
    irreps_x = man[0]
    L['onehot_to_feature_x'] = IrrepsLinear(f'{nsp}x0e', irreps_x, 'x', is_embed=True)
    den = cfg.get('conv_denominator', 1.0)
    den = den if isinstance(den, (list, tuple)) else [den] * nconv
    for t in range(nconv):
        irreps_out = man[t + 1]
        lmax_node = 0 if t == nconv - 1 else cfg['lmax']
        tp_out = '+'.join(f'1x{l}e' for l in range(lmax_node + 1))
        gate = EquivariantGate(irreps_out)
        gin = gate.get_gate_irreps_in()
        if cfg.get('self_connection_type', 'linear') == 'nequip':   # model_build.py:45-56
            L[f'{t}_self_connection_intro'] = SelfConnectionIntro(irreps_x, f'{nsp}x0e', gin)
        else:
            L[f'{t}_self_connection_intro'] = SelfConnectionLinearIntro(irreps_x, gin)
        L[f'{t}_self_interaction_1'] = IrrepsLinear(irreps_x, irreps_x, 'x')
        conv = IrrepsConvolution(irreps_x, '1x0e+1x1e+1x2e', tp_out,
                                 [cfg['bessel_basis_num']] + list(cfg['weight_nn_hidden_neurons']), 'silu', den[t],
                                 train_denominator=bool(cfg.get('train_denominator', False)))
        L[f'{t}_convolution'] = conv
        m = ops.muls_of(irreps_x)
        msg = {0: [0, 0, 0]}[0]
        for (l1, l2, l3) in [(a, b, c) for a in range(3) for b in range(3) for c in range(abs(a - b), min(a + b, lmax_node) + 1)]:
            if m[l1] > 0:
                msg[l3] += m[l1]
        irreps_msg = '+'.join(f'{msg[l]}x{l}e' for l in range(3) if msg[l] > 0)
        L[f'{t}_self_interaction_2'] = IrrepsLinear(irreps_msg, gin, 'x')
        L[f'{t}_self_connection_outro'] = SelfConnectionOutro()
        L[f'{t}_equivariant_gate'] = gate
        irreps_x = irreps_out
    mid = ops.muls_of(man[-1])[0]
    L['reduce_input_to_hidden'] = IrrepsLinear(irreps_x, f'{mid // 2}x0e', 'x')
    L['reduce_hidden_to_energy'] = IrrepsLinear(f'{mid // 2}x0e', '1x0e', 'x', 'scaled_atomic_energy')
    shift, scale = cfg.get('shift', 0.0), cfg.get('scale', 1.0)
    train_ss = bool(cfg.get('train_shift_scale', False))   # model_build.py:400-437
    if cfg['use_species_wise_shift_scale']:
        shift = shift if isinstance(shift, (list, tuple)) else [float(shift)] * nsp
        scale = scale if isinstance(scale, (list, tuple)) else [float(scale)] * nsp
        L['rescale_atomic_energy'] = SpeciesWiseRescale(list(shift), list(scale), train_shift_scale=train_ss)
    else:
        L['rescale_atomic_energy'] = Rescale(shift, scale, train_shift_scale=train_ss)
    L['reduce_total_enegy'] = AtomReduce('atomic_energy', 'inferred_total_energy')
    L['force_output'] = ForceStressOutputFromEdge()
    return AtomGraphSequential(L, cutoff)


_TWINS = ('EdgePreprocess', 'ForceOutput', 'ForceStressOutput', 'EdgeEmbedding', 'OnehotEmbedding', 'IrrepsLinear', 'SelfConnectionLinearIntro', 'SelfConnectionIntro', 'SelfConnectionOutro',
          'IrrepsConvolution', 'EquivariantGate', 'Rescale', 'SpeciesWiseRescale', 'AtomReduce', 'ForceStressOutputFromEdge')


def substitute(model, config=None):
    """Replace the hot-path modules of a reference ``AtomGraphSequential`` by their B200 twins, key by key
    (``replace_module``, ``sevenn/nn/sequential.py:61-62``), and copy the parameters (names and shapes are identical)."""
    twin = build_E3_equivariant_model(config)
    ref_sd = model.state_dict()
    twin_sd = twin.state_dict()
    replaced = [key for key in twin._modules if key in model._modules and type(model._modules[key]).__name__ in _TWINS]
    # every parameter of a replaced module must have a twin parameter (e.g. a `linear.bias` of use_bias_in_linear has none)
    orphan = [k for k in ref_sd if k.split('.')[0] in replaced and k not in twin_sd]
    if orphan:
        raise ValueError(f'B200 path: reference parameters without a twin (biases are not supported): {orphan}')
    twin.load_state_dict({k: v for k, v in ref_sd.items() if k in twin_sd}, strict=False)
    # trainability follows the module being replaced (train_denominator / train_shift_scale of the reference config)
    ref_grad = {k: p.requires_grad for k, p in model.named_parameters()}
    for k, p in twin.named_parameters():
        if k in ref_grad:
            p.requires_grad_(ref_grad[k])
    dev = next(model.parameters()).device
    for key in replaced:
        model.replace_module(key, twin._modules[key].to(dev))
    return model
