"""Tensor-level wrappers over the C-ABI: torch only allocates device memory and supplies the
stream; every computation is a kernel of libsevennet_b200.so.  CPU tensors are an error."""
import ctypes as C
import math

import numpy as np
import torch

from . import _lib
from ._lib import EdgeParams, LinBlock, LinDesc, MlpDesc, check, lib


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# ---- optional per-op CUDA-event timing (bench.py: which kernel family dominates a fine-tuning step) ----------------------
_TIMING = None
_TIMING_SHAPES = False


def enable_timing(flag=True):
    global _TIMING
    _TIMING = {} if flag else None


def timing_summary():
    """{op name: (total ms, calls)} since enable_timing(); synchronises."""
    torch.cuda.synchronize()
    return {k: (sum(a.elapsed_time(b) for a, b in v), len(v)) for k, v in (_TIMING or {}).items()}


def _timed(fn):
    import functools

    @functools.wraps(fn)
    def wrap(*a, **k):
        if _TIMING is None:
            return fn(*a, **k)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn(*a, **k)
        e1.record()
        key = fn.__name__
        if _TIMING_SHAPES:
            key += ' ' + ' '.join('x'.join(map(str, t.shape)) for t in a if torch.is_tensor(t) and t.dim() == 2)[:60]
        _TIMING.setdefault(key, []).append((e0, e1))
        return out
    return wrap


def _p(t):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError('sevennet_b200: CPU tensor passed to a CUDA-only op (there is no CPU fallback)')
    if t.device.index != torch.cuda.current_device():
        # kernels launch on the current device's current stream (_stream): a tensor of another GPU would be an invalid pointer
        raise RuntimeError(f'sevennet_b200: tensor on {t.device} but the current CUDA device is {torch.cuda.current_device()}; '
                           'wrap the call in torch.cuda.device(tensor.device)')
    if not t.is_contiguous():
        raise RuntimeError('sevennet_b200: tensor must be contiguous')
    return C.c_void_p(t.data_ptr())


def _f32(t):
    return t.detach().to(torch.float32).contiguous()


def _i32(t):
    return t.detach().to(torch.int32).contiguous()


# ---- irreps helpers (only even-parity l <= 2 blocks, SevenNet's is_parity=False family) ----------
def parse_irreps(s):
    """'128x0e+64x1e+32x2e' (or a list of (mul, l, p)) -> [(mul, l)], entries kept in order."""
    if not isinstance(s, str):
        return [(int(t[0]), int(t[1])) for t in s]
    out = []
    for tok in str(s).replace(' ', '').split('+'):
        if tok:
            mul, ir = tok.split('x') if 'x' in tok else ('1', tok)
            out.append((int(mul), int(ir[:-1])))
    return out


def irreps_dim(ir):
    return sum(m * (2 * l + 1) for m, l in parse_irreps(ir))


def muls_of(ir):
    """Simplified multiplicities (m0, m1, m2); entries must be sorted by l."""
    m = [0, 0, 0]
    last = -1
    for mul, l in parse_irreps(ir):
        if l > 2 or l < last:
            raise ValueError(f'unsupported irreps for the B200 path: {ir}')
        last = l
        m[l] += mul
    return tuple(m)


def make_lin_desc(irreps_in, irreps_out):
    """e3nn o3.Linear instruction table for irreps whose l-blocks are contiguous (reference use)."""
    mi, mo = muls_of(irreps_in), muls_of(irreps_out)
    d = LinDesc()
    io = oo = wo = 0
    n = 0
    for l in range(3):
        D = 2 * l + 1
        if mi[l] > 0 and mo[l] > 0:
            b = d.blk[n]
            b.d, b.mul_in, b.mul_out, b.in_off, b.out_off, b.w_off = D, mi[l], mo[l], io, oo, wo
            b.scale = 1.0 / math.sqrt(mi[l])
            wo += mi[l] * mo[l]
            n += 1
        io += mi[l] * D
        oo += mo[l] * D
    d.nblk, d.in_dim, d.out_dim = n, io, oo
    return d, wo


def conv_dims(m0, m1, m2, lout):
    wn, od = C.c_int(), C.c_int()
    check(lib().b200_conv_dims(m0, m1, m2, lout, C.byref(wn), C.byref(od)), 'conv_dims')
    return wn.value, od.value


# ---- graph ------------------------------------------------------------------------------------------
def radius_graph(pos, cell, pbc, cutoff):
    """pos: float64 CUDA tensor [N,3]; cell: 3x3 array-like (host); pbc: 3 bools.
    Returns dict(rowptr[N+1], src[E], dst[E], shift[E,3] int32, edge_vec[E,3] f32), rows sorted."""
    L = lib()
    if pos.dtype != torch.float64:
        raise RuntimeError('radius_graph: positions must be float64 (reference: ase neighbour list in double)')
    pos = pos.contiguous()
    n = pos.shape[0]
    dev = pos.device
    cell_h = (C.c_double * 9)(*np.asarray(cell, np.float64).reshape(9))
    pbc_h = (C.c_int * 3)(*[int(bool(b)) for b in np.asarray(pbc).reshape(3)])
    ws = torch.empty(L.b200_graph_workspace_bytes(n), dtype=torch.uint8, device=dev)
    rowptr = torch.empty(n + 1, dtype=torch.int32, device=dev)
    ne = C.c_int64()
    check(L.b200_graph_count(n, _p(pos), cell_h, pbc_h, float(cutoff), _p(ws), _p(rowptr), C.byref(ne), _stream()),
          'graph_count')
    E = ne.value
    src = torch.empty(E, dtype=torch.int32, device=dev)
    dst = torch.empty(E, dtype=torch.int32, device=dev)
    shift = torch.empty(E, 3, dtype=torch.int32, device=dev)
    vec = torch.empty(E, 3, dtype=torch.float32, device=dev)
    tmp = torch.empty(E, dtype=torch.int64, device=dev)
    check(L.b200_graph_fill(n, _p(pos), cell_h, pbc_h, float(cutoff), _p(ws), _p(rowptr), _p(src), _p(dst), _p(shift),
                            _p(vec), _p(tmp), _stream()), 'graph_fill')
    return dict(rowptr=rowptr, src=src, dst=dst, shift=shift, edge_vec=vec)


def csr_transpose(n_nodes, src):
    L = lib()
    E = src.shape[0]
    dev = src.device
    ws = torch.empty(L.b200_transpose_workspace_bytes(n_nodes, E), dtype=torch.uint8, device=dev)
    rrowptr = torch.empty(n_nodes + 1, dtype=torch.int32, device=dev)
    eid = torch.empty(E, dtype=torch.int32, device=dev)
    check(L.b200_csr_transpose(n_nodes, E, _p(src), _p(ws), _p(rrowptr), _p(eid), _stream()), 'csr_transpose')
    return rrowptr, eid


# ---- edge / radial ------------------------------------------------------------------------------------
@_timed
def edge_embed_fwd(params, vec, coeffs):
    E = vec.shape[0]
    ln = torch.empty(E, dtype=torch.float32, device=vec.device)
    Y = torch.empty(E, 9, dtype=torch.float32, device=vec.device)
    emb = torch.empty(E, params.num_basis, dtype=torch.float32, device=vec.device)
    check(lib().b200_edge_embed_fwd(C.byref(params), E, _p(vec), _p(coeffs), _p(ln), _p(Y), _p(emb), _stream()),
          'edge_embed_fwd')
    return ln, Y, emb


@_timed
def edge_embed_bwd(params, vec, coeffs, dY, demb):
    E = vec.shape[0]
    dvec = torch.empty(E, 3, dtype=torch.float32, device=vec.device)
    check(lib().b200_edge_embed_bwd(C.byref(params), E, _p(vec), _p(coeffs), _p(dY), _p(demb), _p(dvec), _stream()),
          'edge_embed_bwd')
    return dvec


@_timed
def radial_mlp_fwd(emb, W0, W1, W2):
    E = emb.shape[0]
    d = MlpDesc(W0.shape[0], W0.shape[1], W1.shape[1], W2.shape[1])
    w = torch.empty(E, d.wn, dtype=torch.float32, device=emb.device)
    check(lib().b200_radial_mlp_fwd(C.byref(d), E, _p(emb), _p(W0), _p(W1), _p(W2), _p(w), _stream()), 'radial_mlp_fwd')
    return w


@_timed
def radial_mlp_bwd(emb, W0, W1, W2, dw, demb=None):
    E = emb.shape[0]
    d = MlpDesc(W0.shape[0], W0.shape[1], W1.shape[1], W2.shape[1])
    acc = demb is not None
    if demb is None:
        demb = torch.empty(E, d.nb, dtype=torch.float32, device=emb.device)
    check(lib().b200_radial_mlp_bwd(C.byref(d), E, _p(emb), _p(W0), _p(W1), _p(W2), _p(dw), _p(demb), int(acc),
                                    _stream()), 'radial_mlp_bwd')
    return demb


# ---- convolution ----------------------------------------------------------------------------------------
@_timed
def conv_fwd(muls, lout, rowptr, src, x, Y, w, denom):
    n = rowptr.shape[0] - 1
    _, od = conv_dims(*muls, lout)
    out = torch.empty(n, od, dtype=torch.float32, device=x.device)
    check(lib().b200_conv_fwd(*muls, lout, n, _p(rowptr), _p(src), _p(x), _p(Y), _p(w), float(denom), _p(out),
                              _stream()), 'conv_fwd')
    return out


@_timed
def conv_bwd_edge(muls, lout, rowptr, src, x, Y, w, gout, denom, dY=None):
    n = rowptr.shape[0] - 1
    E = src.shape[0]
    dw = torch.empty_like(w)
    acc = dY is not None
    if dY is None:
        dY = torch.zeros(E, 9, dtype=torch.float32, device=x.device)
    check(lib().b200_conv_bwd_edge(*muls, lout, n, E, _p(rowptr), _p(src), _p(x), _p(Y), _p(w), _p(gout), float(denom),
                                   _p(dw), _p(dY), int(acc), _stream()), 'conv_bwd_edge')
    return dw, dY


@_timed
def conv_bwd(muls, lout, n_src, rowptr, src, x, Y, w, gout, denom, dY=None):
    """d weight, d edge_attr and d x[n_src] in one pass (b200_conv_bwd)."""
    n = rowptr.shape[0] - 1
    E = src.shape[0]
    dw = torch.empty_like(w)
    acc = dY is not None
    if dY is None:
        dY = torch.zeros(E, 9, dtype=torch.float32, device=x.device)
    xdim = muls[0] + 3 * muls[1] + 5 * muls[2]
    dx = torch.empty(n_src, xdim, dtype=torch.float32, device=x.device)
    check(lib().b200_conv_bwd(*muls, lout, n, n_src, E, _p(rowptr), _p(src), _p(x), _p(Y), _p(w), _p(gout), float(denom),
                              _p(dw), _p(dY), int(acc), _p(dx), _stream()), 'conv_bwd')
    return dw, dY, dx


def _path_scale(path_scale):
    return None if path_scale is None else (C.c_float * 15)(*[float(v) for v in path_scale])


@_timed
def conv_fwd_scaled(muls, lout, rowptr, src, x, Y, w, denom, path_scale):
    """conv_fwd with the weights of path p multiplied by path_scale[p] (15 host floats, canonical path order)."""
    n = rowptr.shape[0] - 1
    _, od = conv_dims(*muls, lout)
    out = torch.empty(n, od, dtype=torch.float32, device=x.device)
    check(lib().b200_conv_fwd_scaled(*muls, lout, n, _p(rowptr), _p(src), _p(x), _p(Y), _p(w), float(denom), _p(out),
                                     _path_scale(path_scale), _stream()), 'conv_fwd_scaled')
    return out


@_timed
def conv_bwd_scaled(muls, lout, n_src, rowptr, src, x, Y, w, gout, denom, path_scale=None, dw=None, dw_mode=0, dY=None,
                    want_dx=True):
    """conv_bwd for the second-order pass: per-path scale of the weights and of their gradient; dw_mode 0 = new d weight,
    1 = added into `dw`, 2 = no weight gradient; dY given = added into; want_dx False = no d x.  -> (dw, dY, dx)."""
    n = rowptr.shape[0] - 1
    E = src.shape[0]
    if dw_mode == 0:
        dw = torch.empty_like(w)
    elif dw_mode == 2:
        dw = None
    acc = dY is not None
    if dY is None:
        dY = torch.zeros(E, 9, dtype=torch.float32, device=x.device)
    xdim = muls[0] + 3 * muls[1] + 5 * muls[2]
    dx = torch.empty(n_src, xdim, dtype=torch.float32, device=x.device) if want_dx else None
    check(lib().b200_conv_bwd_scaled(*muls, lout, n, n_src, E, _p(rowptr), _p(src), _p(x), _p(Y), _p(w), _p(gout), float(denom),
                                     _p(dw), _p(dY), int(acc), _p(dx), _path_scale(path_scale), int(dw_mode), _stream()),
          'conv_bwd_scaled')
    return dw, dY, dx


@_timed
def conv_bwd_x(muls, lout, n_src, rrowptr, eid, dst, gout, Y, w, denom):
    xdim = muls[0] + 3 * muls[1] + 5 * muls[2]
    dx = torch.empty(n_src, xdim, dtype=torch.float32, device=gout.device)
    check(lib().b200_conv_bwd_x(*muls, lout, n_src, _p(rrowptr), _p(eid), _p(dst), _p(gout), _p(Y), _p(w), float(denom),
                                _p(dx), _stream()), 'conv_bwd_x')
    return dx


# ---- node side --------------------------------------------------------------------------------------------
@_timed
def irreps_linear(desc, x, W, transpose=False, out=None):
    n = x.shape[0]
    width = desc.in_dim if transpose else desc.out_dim
    covered = sum((desc.blk[i].mul_in if transpose else desc.blk[i].mul_out) * desc.blk[i].d for i in range(desc.nblk))
    acc = out is not None
    if out is None:
        out = (torch.empty if covered == width else torch.zeros)(n, width, dtype=torch.float32, device=x.device)
    check(lib().b200_irreps_linear(C.byref(desc), n, _p(x), _p(W), _p(out), int(transpose), int(acc), _stream()),
          'irreps_linear')
    return out


@_timed
def irreps_linear_wgrad(desc, x, dy, numel):
    dW = torch.zeros(numel, dtype=torch.float32, device=x.device)
    check(lib().b200_irreps_linear_wgrad(C.byref(desc), x.shape[0], _p(x), _p(dy), _p(dW), 0, _stream()),
          'irreps_linear_wgrad')
    return dW


@_timed
def gate_fwd(ns, m1, m2, x):
    out = torch.empty(x.shape[0], ns + 3 * m1 + 5 * m2, dtype=torch.float32, device=x.device)
    check(lib().b200_gate_fwd(ns, m1, m2, x.shape[0], _p(x), _p(out), _stream()), 'gate_fwd')
    return out


@_timed
def gate_bwd(ns, m1, m2, x, gout):
    gin = torch.empty_like(x)
    check(lib().b200_gate_bwd(ns, m1, m2, x.shape[0], _p(x), _p(gout), _p(gin), _stream()), 'gate_bwd')
    return gin


def embed_fwd(types, W, n_species):
    n, dim = types.shape[0], W.shape[1]
    x = torch.empty(n, dim, dtype=torch.float32, device=W.device)
    check(lib().b200_embed_fwd(n, n_species, dim, _p(types), _p(W), _p(x), _stream()), 'embed_fwd')
    return x


def force_virial(rowptr, rrowptr, eid, vec, dvec):
    n = rowptr.shape[0] - 1
    force = torch.empty(n, 3, dtype=torch.float32, device=vec.device)
    vir = torch.empty(n, 6, dtype=torch.float32, device=vec.device)
    check(lib().b200_force_virial(n, _p(rowptr), _p(rrowptr), _p(eid), _p(vec), _p(dvec), _p(force), _p(vir), _stream()),
          'force_virial')
    return force, vir


def partition_keys(pos, inv_cell, grid, mine, margin, nbins):
    """Sort keys of the brick decomposition (see include/sevennet_b200.h); pos float64 [N,3] on the device."""
    n = pos.shape[0]
    key = torch.empty(n, dtype=torch.int64, device=pos.device)
    inv = (C.c_double * 9)(*[float(v) for v in np.asarray(inv_cell, np.float64).reshape(-1)])
    g3, m3 = (C.c_int * 3)(*[int(v) for v in grid]), (C.c_int * 3)(*[int(v) for v in mine])
    mg, nb = (C.c_double * 3)(*[float(v) for v in margin]), (C.c_int * 3)(*[int(v) for v in nbins])
    check(lib().b200_partition_keys(n, _p(pos), inv, g3, m3, mg, nb, _p(key), _stream()), 'partition_keys')
    return key


def halo_pack(rows, idx, out):
    """out[k] = rows[idx[k]] (rows the peers asked for); rows [n, width] fp32, idx int32, out [m, width]."""
    check(lib().b200_halo_pack(idx.shape[0], rows.shape[1], _p(rows), _p(idx), _p(out), _stream()), 'halo_pack')
    return out


def halo_unpack_add(rows, idx, packed):
    """rows[idx[k]] += packed[k] (gradient / force rows returned by the peers)."""
    check(lib().b200_halo_unpack_add(idx.shape[0], rows.shape[1], _p(rows), _p(idx), _p(packed), _stream()), 'halo_unpack_add')
    return rows


def colsum(x):
    out = torch.empty(x.shape[1], dtype=torch.float32, device=x.device)
    check(lib().b200_colsum(x.shape[0], x.shape[1], _p(x), _p(out), _stream()), 'colsum')
    return out


def edge_embed_bwd_r(params, vec, dY, dr):
    E = vec.shape[0]
    dvec = torch.empty(E, 3, dtype=torch.float32, device=vec.device)
    check(lib().b200_edge_embed_bwd_r(C.byref(params), E, _p(vec), _p(dY), _p(dr), _p(dvec), _stream()), 'edge_embed_bwd_r')
    return dvec


@_timed
def gate_bwd2(ns, m1, m2, x, gout, a):
    """Second-order pass of the gate: a = dL/d(gate_bwd output) -> (dL/dx, dL/d gout)."""
    dx = torch.empty_like(x)
    dg = torch.empty_like(gout)
    check(lib().b200_gate_bwd2(ns, m1, m2, x.shape[0], _p(x), _p(gout), _p(a), _p(dx), _p(dg), _stream()), 'gate_bwd2')
    return dx, dg


@_timed
def table_scatter_dot(m, md, P, inv_h, g, r, order=None, coef=None):
    """(table_scatter(m, g, coef), table_dot(md, P, g)) in one pass over g."""
    n_grid, _, wn = P.shape
    G = torch.empty(n_grid, 4, wn, dtype=torch.float32, device=r.device)
    dot = torch.empty(r.shape[0], dtype=torch.float32, device=r.device)
    check(lib().b200_table_scatter_dot(m, md, wn, r.shape[0], inv_h, n_grid, _p(P), _p(g), _p(coef), _p(r), _p(order), _p(G),
                                       _p(dot), _stream()), 'table_scatter_dot')
    return G, dot


@_timed
def table_hermite(Z, dZ, h):
    """Nodal values / r-derivatives [n_grid + 1, wn] (float64) -> cubic coefficients P[n_grid, 4, wn] (float32)."""
    n1, wn = Z.shape
    P = torch.empty(n1 - 1, 4, wn, dtype=torch.float32, device=Z.device)
    check(lib().b200_table_hermite(n1 - 1, wn, float(h), _p(Z), _p(dZ), _p(P), _stream()), 'table_hermite')
    return P


@_timed
def table_hermite_adj(G, h):
    """Adjoint of table_hermite: G[n_grid, 4, wn] (float32) -> (gZ, gdZ) [n_grid + 1, wn] (float64)."""
    n, _, wn = G.shape
    gZ = torch.empty(n + 1, wn, dtype=torch.float64, device=G.device)
    gdZ = torch.empty_like(gZ)
    check(lib().b200_table_hermite_adj(n, wn, float(h), _p(G), _p(gZ), _p(gdZ), _stream()), 'table_hermite_adj')
    return gZ, gdZ


@_timed
def table_interp(m, P, inv_h, r, order=None, coef=None):
    """out[e, :] = coef[e] * d^m w / d r^m (r[e]) from per-interval cubic coefficients P[n_grid, 4, wn] on the grid r_i = i / inv_h
    (csrc/table_train.cu)."""
    n_grid, _, wn = P.shape
    out = torch.empty(r.shape[0], wn, dtype=torch.float32, device=r.device)
    check(lib().b200_table_interp(m, wn, r.shape[0], inv_h, n_grid, _p(P), _p(r), _p(order), _p(coef), _p(out), _stream()),
          'table_interp')
    return out


@_timed
def table_dot(m, P, inv_h, r, g, order=None):
    """out[e] = sum_k g[e, k] * d^m w_k / d r^m (r[e])."""
    n_grid, _, wn = P.shape
    out = torch.empty(r.shape[0], dtype=torch.float32, device=r.device)
    check(lib().b200_table_dot(m, wn, r.shape[0], inv_h, n_grid, _p(P), _p(r), _p(order), _p(g), _p(out), _stream()), 'table_dot')
    return out


@_timed
def table_scatter(m, n_grid, inv_h, g, r, order=None, coef=None):
    """G[n_grid, 4, wn]: gradient with respect to P of sum_e coef[e] sum_k g[e, k] d^m w_k / d r^m (r[e])."""
    wn = g.shape[1]
    G = torch.empty(n_grid, 4, wn, dtype=torch.float32, device=r.device)
    check(lib().b200_table_scatter(m, wn, r.shape[0], inv_h, n_grid, _p(g), _p(coef), _p(r), _p(order), _p(G), _stream()),
          'table_scatter')
    return G


@_timed
def edge_embed_dk(params, vec, coeffs, demb):
    """d bessel_coeffs of the first reverse pass (deterministic two-stage sum over the edges)."""
    dk = torch.empty(params.num_basis, dtype=torch.float32, device=vec.device)
    check(lib().b200_edge_embed_dk(C.byref(params), vec.shape[0], _p(vec), _p(coeffs), _p(demb), _p(dk), _stream()),
          'edge_embed_dk')
    return dk


@_timed
def edge_embed_bwd2(params, vec, coeffs, gY, gemb, a, want_dk=True):
    """Second-order pass of the edge embedding: a = dL/d(dvec) -> (dL/dvec, dL/d gY, dL/d gemb, dL/d coeffs)."""
    E = vec.shape[0]
    dev = vec.device
    dvec2 = torch.empty(E, 3, dtype=torch.float32, device=dev)
    dgY = torch.empty(E, 9, dtype=torch.float32, device=dev)
    dgemb = torch.empty(E, params.num_basis, dtype=torch.float32, device=dev)
    dk = torch.empty(params.num_basis, dtype=torch.float32, device=dev) if want_dk else None
    check(lib().b200_edge_embed_bwd2(C.byref(params), E, _p(vec), _p(coeffs), _p(gY), _p(gemb), _p(a), _p(dvec2), _p(dgY),
                                     _p(dgemb), _p(dk), _stream()), 'edge_embed_bwd2')
    return dvec2, dgY, dgemb, dk


# ---- tabulated radial weights (MD path) -----------------------------------------------------------------------------------
def radial_table_build(params, coeffs, W0, W1, W2):
    d = MlpDesc(W0.shape[0], W0.shape[1], W1.shape[1], W2.shape[1])
    n = lib().b200_radial_table_floats(C.byref(params), C.byref(d))
    T = torch.empty(n, dtype=torch.float32, device=W0.device)
    check(lib().b200_radial_table_build(C.byref(params), C.byref(d), _p(coeffs), _p(W0), _p(W1), _p(W2), _p(T), _stream()),
          'radial_table_build')
    return T


@_timed
def radial_table_fwd(params, wn, nb, h1, h2, length, T, order=None):
    d = MlpDesc(nb, h1, h2, wn)
    w = torch.empty(length.shape[0], wn, dtype=torch.float32, device=length.device)
    check(lib().b200_radial_table_fwd(C.byref(params), C.byref(d), length.shape[0], _p(length), _p(order), _p(T), _p(w),
                                      _stream()), 'radial_table_fwd')
    return w


@_timed
def radial_table_bwd(params, wn, nb, h1, h2, length, T, dw, dr=None, order=None):
    d = MlpDesc(nb, h1, h2, wn)
    acc = dr is not None
    if dr is None:
        dr = torch.empty(length.shape[0], dtype=torch.float32, device=length.device)
    check(lib().b200_radial_table_bwd(C.byref(params), C.byref(d), length.shape[0], _p(length), _p(order), _p(T), _p(dw), _p(dr),
                                      int(acc), _stream()), 'radial_table_bwd')
    return dr
