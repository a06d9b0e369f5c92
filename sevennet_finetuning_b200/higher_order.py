"""Autograd functions that stay differentiable through their own backward pass.

Force-matching fine-tuning differentiates the forces with respect to the parameters: the reference calls
``torch.autograd.grad(energy, edge_vec, create_graph=self.training)`` (``sevenn/nn/force_output.py:165-170``) and then
``total_loss.backward()`` on a loss of ``inferred_force`` / ``inferred_stress`` (``sevenn/train/trainer.py:173-186``), i.e.
a reverse pass through the reverse pass.  The kernels of libsevennet_b200.so are (multi)linear maps, so their
derivatives of every order are the same kernels with the arguments permuted:

* ``irreps_linear`` is bilinear in (x, W): {forward, transpose, weight-gradient} is closed under differentiation;
* the uvu tensor product + scatter is trilinear in (x[src], Y, w): {``conv_fwd``, ``conv_bwd``} is closed;
* the gate / c*silu activation is elementwise: ``gate_fwd`` -> ``gate_bwd`` -> ``gate_bwd2`` (second derivative of silu).

Each Function below saves its *inputs* (with their autograd history) and builds its backward from other Functions of
the same family, so ``create_graph=True`` works to any order the family supports.  No CPU path: CPU tensors raise in
``ops``.
"""
import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from . import ops


class force_pass:
    """Context of the first reverse pass (d E / d edge_vec, force_output.py:165-170): autograd asks every Function for
    all of its input gradients, but only the edge-vector branch is consumed there, so the weight-gradient kernels are
    skipped while it is active (they run in the loss's own reverse pass)."""
    active = False

    def __enter__(self):
        force_pass.active = True

    def __exit__(self, *a):
        force_pass.active = False


def _c(t):
    return t if (t.dtype == torch.float32 and t.is_contiguous()) else t.to(torch.float32).contiguous()


# ---- edge embedding: spherical harmonics + Bessel x cutoff (sevenn/nn/edge_embedding.py:220-230) -------------------------
class EdgeF(Function):
    """(edge_vec, bessel coeffs) -> (edge_length, edge_attr Y[E,9], edge_embedding[E,nb])."""

    @staticmethod
    def forward(ctx, vec, coeffs, params):
        ctx.save_for_backward(vec, coeffs)
        ctx.params = params
        ln, Y, emb = ops.edge_embed_fwd(params, _c(vec.detach()), _c(coeffs.detach()))
        ctx.mark_non_differentiable(ln)     # edge_length is informational: no module of the path consumes it
        return ln, Y, emb

    @staticmethod
    def backward(ctx, g_len, gY, gemb):
        vec, coeffs = ctx.saved_tensors
        dvec, dk = EdgeB.apply(vec, coeffs, gY, gemb, ctx.params)
        return (dvec if ctx.needs_input_grad[0] else None), (dk if ctx.needs_input_grad[1] else None), None


class EdgeB(Function):
    """First reverse pass of EdgeF: (d edge_vec [E,3], d coeffs [nb]) from (d Y, d emb)."""

    @staticmethod
    def forward(ctx, vec, coeffs, gY, gemb, params):
        ctx.save_for_backward(vec, coeffs, gY, gemb)
        ctx.params = params
        ctx.set_materialize_grads(False)   # an output nobody differentiates arrives as None, not as zeros to be tested on the host
        v, k = _c(vec.detach()), _c(coeffs.detach())
        gY_ = None if gY is None else _c(gY.detach())
        ge_ = None if gemb is None else _c(gemb.detach())
        dvec = ops.edge_embed_bwd(params, v, k, gY_, ge_)
        if ge_ is None or force_pass.active:   # the force pass consumes d edge_vec only
            dk = torch.zeros_like(k)
        else:
            dk = ops.edge_embed_dk(params, v, k, ge_)
        return dvec, dk

    @staticmethod
    @once_differentiable
    def backward(ctx, a_vec, a_k):
        vec, coeffs, gY, gemb = ctx.saved_tensors
        if a_k is not None and bool((a_k != 0).any()):
            raise NotImplementedError('sevennet_b200: second derivative through the Bessel-coefficient gradient')
        if a_vec is None:
            return None, None, None, None, None
        dvec2, dgY, dgemb, dk = ops.edge_embed_bwd2(ctx.params, _c(vec), _c(coeffs), None if gY is None else _c(gY),
                                                    None if gemb is None else _c(gemb), _c(a_vec),
                                                    want_dk=ctx.needs_input_grad[1] and gemb is not None)
        return (dvec2 if ctx.needs_input_grad[0] else None), dk, (dgY if gY is not None else None), \
               (dgemb if gemb is not None else None), None


# ---- per-l block-diagonal linear (sevenn/nn/linear.py:46-49; e3nn o3.Linear / FullyConnectedNet layer) ------------------
class LinF(Function):
    """y = L(x, W): rows x [n, in_dim], flat weight W."""

    @staticmethod
    def forward(ctx, x, W, desc, numel):
        ctx.save_for_backward(x, W)
        ctx.desc, ctx.numel = desc, numel
        return ops.irreps_linear(desc, _c(x.detach()), _c(W.detach()))

    @staticmethod
    def backward(ctx, g):
        x, W = ctx.saved_tensors
        dx = LinT.apply(g, W, ctx.desc, ctx.numel) if ctx.needs_input_grad[0] else None
        dW = (LinWG.apply(x, g, ctx.desc, ctx.numel).view_as(W)
              if ctx.needs_input_grad[1] and not force_pass.active else None)
        return dx, dW, None, None


class LinT(Function):
    """x-gradient of LinF: L^T(g, W) [n, in_dim]."""

    @staticmethod
    def forward(ctx, g, W, desc, numel):
        ctx.save_for_backward(g, W)
        ctx.desc, ctx.numel = desc, numel
        return ops.irreps_linear(desc, _c(g.detach()), _c(W.detach()), transpose=True)

    @staticmethod
    def backward(ctx, a):
        g, W = ctx.saved_tensors
        dg = LinF.apply(a, W, ctx.desc, ctx.numel) if ctx.needs_input_grad[0] else None
        dW = LinWG.apply(a, g, ctx.desc, ctx.numel).view_as(W) if ctx.needs_input_grad[1] else None
        return dg, dW, None, None


class LinWG(Function):
    """W-gradient of LinF: flat [numel] = scale * x^T g per block."""

    @staticmethod
    def forward(ctx, x, g, desc, numel):
        ctx.save_for_backward(x, g)
        ctx.desc, ctx.numel = desc, numel
        return ops.irreps_linear_wgrad(desc, _c(x.detach()), _c(g.detach()), numel)

    @staticmethod
    def backward(ctx, A):
        x, g = ctx.saved_tensors
        A = A.reshape(-1)
        dx = LinT.apply(g, A, ctx.desc, ctx.numel) if ctx.needs_input_grad[0] else None
        dg = LinF.apply(x, A, ctx.desc, ctx.numel) if ctx.needs_input_grad[1] else None
        return dx, dg, None, None


# ---- gate / activation (sevenn/nn/equivariant_gate.py:59-61; e3nn normalize2mom(silu)) ---------------------------------
class GateF(Function):
    @staticmethod
    def forward(ctx, x, ns, m1, m2):
        ctx.save_for_backward(x)
        ctx.dims = (ns, m1, m2)
        return ops.gate_fwd(ns, m1, m2, _c(x.detach()))

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        return GateB.apply(x, g, *ctx.dims), None, None, None


class GateB(Function):
    @staticmethod
    def forward(ctx, x, g, ns, m1, m2):
        ctx.save_for_backward(x, g)
        ctx.dims = (ns, m1, m2)
        return ops.gate_bwd(ns, m1, m2, _c(x.detach()), _c(g.detach()))

    @staticmethod
    @once_differentiable
    def backward(ctx, a):
        x, g = ctx.saved_tensors
        dx, dg = ops.gate_bwd2(*ctx.dims, _c(x), _c(g), _c(a))
        return dx, dg, None, None, None


# ---- uvu tensor product + scatter-sum / denominator (sevenn/nn/convolution.py:19-32,115-118) ---------------------------
class ConvF(Function):
    """out[n_dst, odim] = (1/denom) sum_{e in row i} TP_uvu(x[src_e], Y_e, w_e); edges sorted by destination."""

    @staticmethod
    def forward(ctx, x, Y, w, graph, muls, lout, denom):
        ctx.save_for_backward(x, Y, w)
        ctx.args = (graph, muls, lout, denom)
        rowptr = graph['rowptr_dst']
        return ops.conv_fwd(muls, lout, rowptr, graph['src'], _c(x.detach()), _c(Y.detach()), _c(w.detach()), denom)

    @staticmethod
    def backward(ctx, g):
        x, Y, w = ctx.saved_tensors
        dx, dY, dw = ConvB.apply(x, Y, w, g, *ctx.args)
        n = ctx.needs_input_grad
        return (dx if n[0] else None), (dY if n[1] else None), (dw if n[2] else None), None, None, None, None


class ConvB(Function):
    """(d x, d Y, d w) of ConvF for the output gradient g, in one kernel pass (b200_conv_bwd)."""

    @staticmethod
    def forward(ctx, x, Y, w, g, graph, muls, lout, denom):
        ctx.save_for_backward(x, Y, w, g)
        ctx.args = (graph, muls, lout, denom)
        dw, dY, dx = ops.conv_bwd(muls, lout, x.shape[0], graph['rowptr_dst'], graph['src'], _c(x.detach()),
                                  _c(Y.detach()), _c(w.detach()), _c(g.detach()), denom)
        return dx, dY, dw

    @staticmethod
    def backward(ctx, ax, aY, aw):
        x, Y, w, g = ctx.saved_tensors
        args = ctx.args
        need = ctx.needs_input_grad
        dx = dY = dw = dg = None

        def add(a, b):
            return b if a is None else a + b

        if not torch.is_grad_enabled():
            # the loss's own reverse pass (no third order asked for): the same three terms straight on the kernels, with the
            # filter mask as a per-path scale inside them, ONE weight-gradient array the calls add into, and no outputs
            # nobody reads - every [E, wn] tensor saved here is a pass over ~1 GB
            graph, muls, lout, denom = args
            rowptr, src = graph['rowptr_dst'], graph['src']
            xc, Yc, wc, gc = _c(x), _c(Y), _c(w), _c(g)
            mask = _y_scale((muls, lout))
            want_dw, first = need[2], True
            if ax is not None:
                axc = _c(ax)
                if need[3]:
                    dg = add(dg, ops.conv_fwd(muls, lout, rowptr, src, axc, Yc, wc, denom))
                if need[1] or need[2]:
                    dw, dY, _ = ops.conv_bwd_scaled(muls, lout, x.shape[0], rowptr, src, axc, Yc, wc, gc, denom,
                                                    dw_mode=0 if want_dw else 2, want_dx=False)
                    first = not want_dw
            if aY is not None:
                aYc = _c(aY)
                if need[3]:
                    dg = add(dg, ops.conv_fwd_scaled(muls, lout, rowptr, src, xc, aYc, wc, denom, mask))
                if need[0] or need[2]:
                    mode = 2 if not want_dw else (0 if first else 1)
                    tw, _, dx = ops.conv_bwd_scaled(muls, lout, x.shape[0], rowptr, src, xc, aYc, wc, gc, denom, path_scale=mask,
                                                    dw=dw, dw_mode=mode, want_dx=need[0])
                    if want_dw:
                        dw, first = tw, False
            if aw is not None:
                awc = _c(aw)
                if need[3]:
                    dg = add(dg, ops.conv_fwd(muls, lout, rowptr, src, xc, Yc, awc, denom))
                if need[0] or need[1]:
                    _, dY, tx = ops.conv_bwd_scaled(muls, lout, x.shape[0], rowptr, src, xc, Yc, awc, gc, denom, dw_mode=2,
                                                    dY=dY, want_dx=need[0])
                    dx = add(dx, tx)
            return (dx if need[0] else None), (dY if need[1] else None), (dw if need[2] else None), \
                   (dg if need[3] else None), None, None, None, None

        # S = <g, T(ax, Y, w) + T(x, aY, w) + T(x, Y, aw)> (T trilinear); every term is one of the two kernels
        if ax is not None:
            if need[3]:
                dg = add(dg, ConvF.apply(ax, Y, w, *args))
            if need[1] or need[2]:
                _, tY, tw = ConvB.apply(ax, Y, w, g, *args)
                dY, dw = add(dY, tY), add(dw, tw)
        if aY is not None:
            # the kernels take Y[:, 0] == 1 as given (the l = 0 harmonic is a constant, its paths read no Y), i.e. they
            # evaluate T0(x, w) + T1(x, Y[1:], w).  The variation aY acts on T1 only: mask the weight columns of the
            # l2 = 0 paths so that T0 drops out, and their weight-gradient columns likewise.
            m = _y_mask(muls_lout=(args[1], args[2]), like=w)
            wm = w * m
            if need[3]:
                dg = add(dg, ConvF.apply(x, aY, wm, *args))
            if need[0] or need[2]:
                tx, _, tw = ConvB.apply(x, aY, wm, g, *args)
                dx, dw = add(dx, tx), add(dw, tw * m)
        if aw is not None:
            if need[3]:
                dg = add(dg, ConvF.apply(x, Y, aw, *args))
            if need[0] or need[1]:
                tx, tY, _ = ConvB.apply(x, Y, aw, g, *args)
                dx, dY = add(dx, tx), add(dY, tY)
        return (dx if need[0] else None), (dY if need[1] else None), (dw if need[2] else None), \
               (dg if need[3] else None), None, None, None, None


_MASKS = {}
def _y_scale(muls_lout=None):
    """the mask of _y_mask as 15 per-path factors (canonical path order of csrc/common.cuh) for the *_scaled kernels."""
    from .cg import ALL_PATHS
    return [0.0 if l2 == 0 else 1.0 for (_, l2, _) in ALL_PATHS]


def _y_mask(muls_lout, like):
    """[wn] vector: 0 on the weight columns of the paths whose filter is the constant l2 = 0 harmonic, else 1."""
    from .cg import tp_paths
    key = (muls_lout, like.device)
    if key not in _MASKS:
        paths, wn, _ = tp_paths(*muls_lout[0], muls_lout[1])
        m = torch.ones(wn, dtype=torch.float32, device=like.device)
        for p in paths:
            if p['l2'] == 0:
                m[p['woff']:p['woff'] + p['mul']] = 0.0
        _MASKS[key] = m
    return _MASKS[key]


# ---- species embedding: one_hot(type) . W / sqrt(n_species) (node_embedding.py:39-48, linear.py:37-44) -----------------
class EmbedF(Function):
    @staticmethod
    def forward(ctx, types, W, n_species, dim):
        t32 = ops._i32(types)
        ctx.save_for_backward(t32)
        ctx.n_species, ctx.dim = n_species, dim
        return ops.embed_fwd(t32, _c(W.detach()).view(n_species, dim), n_species)

    @staticmethod
    def backward(ctx, g):
        (t,) = ctx.saved_tensors
        # linear in g; written with differentiable tensor ops so that a loss on forces can pass through it
        dW = torch.zeros(ctx.n_species, ctx.dim, dtype=g.dtype, device=g.device).index_add(
            0, t.long(), g * (1.0 / ctx.n_species ** 0.5))
        return None, dW.reshape(-1), None, None


# ---- train-mode radial weights from a table (csrc/table_train.cu; sevenn/nn/convolution.py:97-106) ------------------------
# P[n_grid, 4, wn]: per-interval cubic coefficients of weight_nn as a function of the edge length, built by the caller with
# differentiable tensor operations on the grid points.  tab = (inv_h, order): 1 / grid step (fp32-exact), edge ids sorted by interval (int32).
class HermF(Function):
    """(Z, dZ) float64 [n_grid + 1, wn] -> P float32 [n_grid, 4, wn] (linear; csrc/table_train.cu hermite_kernel)."""

    @staticmethod
    def forward(ctx, Z, dZ, h):
        ctx.h = h
        return ops.table_hermite(Z.detach().contiguous(), dZ.detach().contiguous(), h)

    @staticmethod
    def backward(ctx, G):
        gZ, gdZ = HermT.apply(G, ctx.h)
        return gZ, gdZ, None


class HermT(Function):
    """The adjoint of HermF."""

    @staticmethod
    def forward(ctx, G, h):
        ctx.h = h
        return ops.table_hermite_adj(_c(G.detach()), h)

    @staticmethod
    def backward(ctx, aZ, adZ):
        return HermF.apply(aZ, adZ, ctx.h), None


class TabInterp(Function):
    """[E, wn] = coef[e] * d^m w / d r^m (r[e])   (coef None = 1)."""

    @staticmethod
    def forward(ctx, P, r, coef, m, tab):
        ctx.save_for_backward(P, r, coef)
        ctx.m, ctx.tab = m, tab
        return ops.table_interp(m, _c(P.detach()), tab[0], _c(r.detach()), tab[1], None if coef is None else _c(coef.detach()))

    @staticmethod
    def backward(ctx, g):
        P, r, coef = ctx.saved_tensors
        dP = dr = dc = None
        n = ctx.needs_input_grad
        if not torch.is_grad_enabled() and n[0] and n[1] and not force_pass.active and not (coef is not None and n[2]):
            # the loss's own reverse pass: both results that read g in one pass over it
            dP, dr = ops.table_scatter_dot(ctx.m, ctx.m + 1, _c(P), ctx.tab[0], _c(g), _c(r), ctx.tab[1],
                                           None if coef is None else _c(coef))
            return dP, (dr if coef is None else dr * coef), None, None, None
        if ctx.needs_input_grad[0] and not force_pass.active:
            dP = TabScatter.apply(g, coef, r, ctx.m, P.shape[0], ctx.tab)
        if ctx.needs_input_grad[1]:
            dr = TabDot.apply(P, r, g, ctx.m + 1, ctx.tab)
            if coef is not None:
                dr = dr * coef
        if coef is not None and ctx.needs_input_grad[2]:
            dc = TabDot.apply(P, r, g, ctx.m, ctx.tab)
        return dP, dr, dc, None, None


class TabDot(Function):
    """[E] = sum_k g[e, k] d^m w_k / d r^m (r[e])."""

    @staticmethod
    def forward(ctx, P, r, g, m, tab):
        ctx.save_for_backward(P, r, g)
        ctx.m, ctx.tab = m, tab
        return ops.table_dot(m, _c(P.detach()), tab[0], _c(r.detach()), _c(g.detach()), tab[1])

    @staticmethod
    def backward(ctx, a):
        P, r, g = ctx.saved_tensors
        n = ctx.needs_input_grad
        if not torch.is_grad_enabled() and n[0] and n[1] and not force_pass.active:
            dP, dr = ops.table_scatter_dot(ctx.m, ctx.m + 1, _c(P), ctx.tab[0], _c(g), _c(r), ctx.tab[1], _c(a))
            dg = ops.table_interp(ctx.m, _c(P), ctx.tab[0], _c(r), ctx.tab[1], _c(a)) if n[2] else None
            return dP, a * dr, dg, None, None
        dP = TabScatter.apply(g, a, r, ctx.m, P.shape[0], ctx.tab) if ctx.needs_input_grad[0] and not force_pass.active else None
        dr = a * TabDot.apply(P, r, g, ctx.m + 1, ctx.tab) if ctx.needs_input_grad[1] else None
        dg = TabInterp.apply(P, r, a, ctx.m, ctx.tab) if ctx.needs_input_grad[2] else None
        return dP, dr, dg, None, None


class TabScatter(Function):
    """[n_grid, 4, wn] = d (sum_e coef[e] sum_k g[e, k] d^m w_k / d r^m (r[e])) / d P  (coef None = 1)."""

    @staticmethod
    def forward(ctx, g, coef, r, m, n_grid, tab):
        ctx.save_for_backward(g, coef, r)
        ctx.m, ctx.tab = m, tab
        return ops.table_scatter(m, n_grid, tab[0], _c(g.detach()), _c(r.detach()), tab[1],
                                 None if coef is None else _c(coef.detach()))

    @staticmethod
    def backward(ctx, A):
        g, coef, r = ctx.saved_tensors
        dg = dc = dr = None
        if ctx.needs_input_grad[0]:
            dg = TabInterp.apply(A, r, coef, ctx.m, ctx.tab)
        if coef is not None and ctx.needs_input_grad[1]:
            dc = TabDot.apply(A, r, g, ctx.m, ctx.tab)
        if ctx.needs_input_grad[2]:
            dr = TabDot.apply(A, r, g, ctx.m + 1, ctx.tab)
            if coef is not None:
                dr = dr * coef
        return dg, dc, dr, None, None, None


def radial_mlp(emb, W0, W1, W2, descs):
    """weight_nn (e3nn FullyConnectedNet [nb, h1, h2, wn], c*silu; sevenn/nn/convolution.py:97-106) as a chain of the
    closed families above: a dense layer is the l = 0 block of an irreps linear, the activation a gate with scalars only."""
    (d0, n0), (d1, n1), (d2, n2) = descs
    h = GateF.apply(LinF.apply(emb, W0.reshape(-1), d0, n0), W0.shape[1], 0, 0)
    h = GateF.apply(LinF.apply(h, W1.reshape(-1), d1, n1), W1.shape[1], 0, 0)
    return LinF.apply(h, W2.reshape(-1), d2, n2)
