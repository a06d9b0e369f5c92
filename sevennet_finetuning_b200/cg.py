"""Coupling tables of the channel-wise (uvu) tensor product, derived on the host.

The reference gets these from e3nn (``o3.TensorProduct`` built in
``sevenn/nn/convolution.py:88-95``; un-vendored dependency).  Here they are derived from the
definition: ``C[m1,m2,m3]`` is the (unique) tensor invariant under simultaneous rotation of
three real spherical-harmonic multiplets, in the component order the reference's frozen graph
uses (l=1: x,y,z; l=2: sqrt3 xz, sqrt3 xy, y^2-(x^2+z^2)/2, sqrt3 yz, sqrt3/2 (z^2-x^2)).
Invariance is imposed through the three infinitesimal rotation generators (null space of the
summed Lie-algebra action), normalised to Frobenius norm 1 and multiplied by the 'element'
path weight sqrt(2 l3 + 1).  The overall sign of each table is a convention; it is anchored
on one entry per table (``_SIGN_ANCHOR``) and verified against the constants of the reference
graph by ``tests/test_cg_tables.py``.

``python -m sevennet_finetuning_b200.cg`` regenerates ``csrc/cg_generated.cuh``.
"""
import functools
import os

import numpy as np


def _polys(l):
    """Real SH polynomials as coefficient dicts {(i,j,k): c} for x^i y^j z^k."""
    s3 = np.sqrt(3.0)
    if l == 0:
        return [{(0, 0, 0): 1.0}]
    if l == 1:
        return [{(1, 0, 0): 1.0}, {(0, 1, 0): 1.0}, {(0, 0, 1): 1.0}]
    if l == 2:
        return [{(1, 0, 1): s3}, {(1, 1, 0): s3},
                {(0, 2, 0): 1.0, (2, 0, 0): -0.5, (0, 0, 2): -0.5},
                {(0, 1, 1): s3}, {(0, 0, 2): 0.5 * s3, (2, 0, 0): -0.5 * s3}]
    raise NotImplementedError('lmax <= 2')


def _apply_generator(p, axis):
    """(G_a p)(v) = grad p . (e_a x v) as a coefficient dict."""
    out = {}
    # e_a x v components: a=0: (0,-z,y); a=1: (z,0,-x); a=2: (-y,x,0)
    cross = {0: [None, (2, -1.0), (1, 1.0)], 1: [(2, 1.0), None, (0, -1.0)],
             2: [(1, -1.0), (0, 1.0), None]}[axis]
    for mono, c in p.items():
        for d in range(3):
            if mono[d] == 0 or cross[d] is None:
                continue
            var, sgn = cross[d]
            m = list(mono)
            coef = c * mono[d] * sgn
            m[d] -= 1
            m[var] += 1
            out[tuple(m)] = out.get(tuple(m), 0.0) + coef
    return out


def _generator_matrix(l, axis):
    polys = _polys(l)
    monos = sorted({m for p in polys for m in p} |
                   {m for p in polys for m in _apply_generator(p, axis)})
    idx = {m: i for i, m in enumerate(monos)}
    B = np.zeros((len(polys), len(monos)))
    G = np.zeros((len(polys), len(monos)))
    for r, p in enumerate(polys):
        for m, c in p.items():
            B[r, idx[m]] = c
        for m, c in _apply_generator(p, axis).items():
            G[r, idx[m]] = c
    J, res, *_ = np.linalg.lstsq(B.T, G.T, rcond=None)   # G = J^T-ish: G_r = sum_s J[s,r] B_s
    assert np.abs(B.T @ J - G.T).max() < 1e-12
    return J.T


# one anchored entry per table: (m1, m2, m3) -> sign of C there (reference convention)
_SIGN_ANCHOR = {
    (1, 1, 1): ((0, 1, 2), +1), (1, 1, 2): ((0, 0, 2), -1), (1, 2, 1): ((0, 0, 2), +1),
    (1, 2, 2): ((0, 0, 1), +1), (2, 1, 1): ((0, 0, 2), +1), (2, 1, 2): ((0, 0, 1), -1),
    (2, 2, 1): ((0, 1, 0), +1), (2, 2, 2): ((0, 0, 2), -1),
}


@functools.lru_cache(maxsize=None)
def coupling(l1, l2, l3):
    """C[m1,m2,m3] including the path weight sqrt(2 l3 + 1)."""
    d = [2 * l + 1 for l in (l1, l2, l3)]
    rows = []
    for a in range(3):
        J1, J2, J3 = (_generator_matrix(l, a) for l in (l1, l2, l3))
        K = (np.kron(np.kron(J1, np.eye(d[1])), np.eye(d[2]))
             + np.kron(np.kron(np.eye(d[0]), J2), np.eye(d[2]))
             + np.kron(np.kron(np.eye(d[0]), np.eye(d[1])), J3))
        rows.append(K)
    _, s, vt = np.linalg.svd(np.concatenate(rows, 0))
    assert s[-1] < 1e-10 and (len(s) < 2 or s[-2] > 1e-3), (l1, l2, l3)
    C = vt[-1].reshape(d)
    C = C / np.linalg.norm(C)
    C[np.abs(C) < 1e-12] = 0.0
    if (l1, l2, l3) in _SIGN_ANCHOR:
        pos, sgn = _SIGN_ANCHOR[(l1, l2, l3)]
        assert abs(C[pos]) > 1e-6
        if np.sign(C[pos]) != sgn:
            C = -C
    else:  # a factor of the triple is l=0: the table is +identity / sqrt(dim)
        flat = C.reshape(-1)
        if flat[np.argmax(np.abs(flat) > 1e-9)] < 0:
            C = -C
    return C * np.sqrt(2 * l3 + 1)


def tp_paths(m0, m1, m2, lout):
    """Paths of x = m0x0e+m1x1e+m2x2e (x) 0e+1e+2e -> l3 <= lout in the reference's
    instruction order (``sevenn/nn/convolution.py:72-86``), with the weight offset of each
    path and its slot offset in the l3-sorted output.  Returns (paths, weight_numel, out_dim);
    path = dict(l1,l2,l3,mul,woff,ooff,xoff)."""
    muls = (m0, m1, m2)
    xoffs = (0, m0, m0 + 3 * m1)
    paths = []
    woff = 0
    for l1 in range(3):
        if muls[l1] == 0:
            continue
        for l2 in range(3):
            for l3 in range(abs(l1 - l2), min(l1 + l2, lout) + 1):
                paths.append(dict(l1=l1, l2=l2, l3=l3, mul=muls[l1], woff=woff, xoff=xoffs[l1]))
                woff += muls[l1]
    ooff = 0
    for l3 in range(lout + 1):           # stable sort by l3
        for p in paths:
            if p['l3'] == l3:
                p['ooff'] = ooff
                ooff += p['mul'] * (2 * l3 + 1)
    return paths, woff, ooff


ALL_PATHS = [(l1, l2, l3) for l1 in range(3) for l2 in range(3)
             for l3 in range(abs(l1 - l2), min(l1 + l2, 2) + 1)]


def _fmt(v):
    return f'{float(np.float32(v))!r}f'.replace('e-0', 'e-').replace('e+0', 'e+')


def generate_header():
    """Straight-line device code for M_p = C_p . Y (forward/backward-x) and dY += C_p : G
    (backward wrt the spherical harmonics)."""
    out = ['// GENERATED by `python -m sevennet_finetuning_b200.cg` - do not edit.',
           '// Coupling tables C[m1,m2,m3]*sqrt(2 l3+1) of the uvu tensor product, unrolled.',
           '// Y is the 9-vector (l=0 | l=1 | l=2) of an edge; M is row-major [2l1+1][2l3+1].',
           '#pragma once', '',
           '__device__ __forceinline__ float cg_fma(float c, float g, float acc) { return fmaf(c, g, acc); }',
           '// a channel pair packed in one 64-bit register: both halves with one FFMA2 (fma.rn.f32x2)',
           '__device__ __forceinline__ unsigned long long cg_fma(float c, unsigned long long g, unsigned long long acc) {',
           '  unsigned long long r, cc;',
           '  asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));',
           '  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(cc), "l"(g), "l"(acc));',
           '  return r;',
           '}', '']
    yoff = (0, 1, 4)
    for (l1, l2, l3) in ALL_PATHS:
        C = coupling(l1, l2, l3)
        d1, d2, d3 = C.shape
        tag = f'{l1}{l2}{l3}'
        out.append(f'// path ({l1},{l2},{l3}): nnz = {int((np.abs(C) > 1e-9).sum())}')
        out.append(f'__device__ __forceinline__ void cg_M_{tag}(const float* __restrict__ Y, float* M) {{')
        for a in range(d1):
            for c in range(d3):
                terms = [f'{_fmt(C[a, b, c])} * Y[{yoff[l2] + b}]' for b in range(d2)
                         if abs(C[a, b, c]) > 1e-9]
                out.append(f'  M[{a * d3 + c}] = {" + ".join(terms) if terms else "0.0f"};')
        out.append('}')
        # T = float (one channel) or float2 (a channel pair, packed FFMA2 arithmetic of sm_100a)
        out.append(f'template <class T> __device__ __forceinline__ void cg_dY_{tag}(const T* G, T* dY) {{')
        for b in range(d2):
            for a in range(d1):
                for c in range(d3):
                    if abs(C[a, b, c]) > 1e-9:
                        out.append(f'  dY[{yoff[l2] + b}] = cg_fma({_fmt(C[a, b, c])}, G[{a * d3 + c}], dY[{yoff[l2] + b}]);')
        out.append('}')
        out.append('')
    return '\n'.join(out)


if __name__ == '__main__':
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'csrc')
    for name, gen in (('cg_generated.cuh', generate_header),):
        with open(os.path.join(here, name), 'w') as f:
            f.write(gen())
        print('wrote', os.path.join(here, name))
