"""TEST INFRASTRUCTURE - float64 numpy restatement of the reference radius graph.

Reference: ``sevenn/train/dataload.py:31-68`` (``unlabeled_atoms_to_graph``) and ``:71-154``
call ``ase.neighborlist.primitive_neighbor_list('ijDS', pbc, cell, pos, cutoff,
self_interaction=True)`` then drop the zero-length self pair (``:40-47``).  ase is an
un-vendored, unpinned dependency (``pyproject.toml:21``) and absent here, so the published
contract is restated: all ordered (i, j, S) with ``|pos[j] - pos[i] + S @ cell| < cutoff``,
S integer (0 along non-periodic axes), excluding (i == j, S == 0).  Parity unpinned by any
reference test; the set definition is the contract.

Canonical arithmetic (the CUDA builder must reproduce it bit for bit, no FMA contraction):
    D_k = (pos[j,k] - pos[i,k]) + ((S0*cell[0,k] + S1*cell[1,k]) + S2*cell[2,k])
    d2  = (D0*D0 + D1*D1) + D2*D2 ;  edge  <=>  d2 < cutoff*cutoff
Canonical order of the returned list: lexicographic in (i, j, S0, S1, S2).
"""
import numpy as np


def _heights(cell):
    vol = abs(np.dot(cell[0], np.cross(cell[1], cell[2])))
    h = []
    for k in range(3):
        a, b = cell[(k + 1) % 3], cell[(k + 2) % 3]
        h.append(vol / np.linalg.norm(np.cross(a, b)))
    return np.array(h), vol


def _exact_filter(pos, cell, i, j, S, cutoff):
    Sf = S.astype(np.float64)
    shift = (Sf[:, 0:1] * cell[0][None, :] + Sf[:, 1:2] * cell[1][None, :]) + Sf[:, 2:3] * cell[2][None, :]
    D = (pos[j] - pos[i]) + shift
    d2 = (D[:, 0] * D[:, 0] + D[:, 1] * D[:, 1]) + D[:, 2] * D[:, 2]
    keep = d2 < cutoff * cutoff
    keep &= ~((i == j) & (S == 0).all(1))
    return keep, D


def radius_graph(pos, cell, pbc, cutoff):
    """Returns edge_index[2,E] int64 (row0 = centre i, row1 = neighbour j), shift[E,3] int64,
    edge_vec[E,3] float64, in canonical order."""
    pos = np.asarray(pos, np.float64)
    cell = np.asarray(cell, np.float64).reshape(3, 3)
    pbc = np.asarray(pbc, bool).reshape(3)
    N = len(pos)
    if N == 0:
        return np.zeros((2, 0), np.int64), np.zeros((0, 3), np.int64), np.zeros((0, 3))
    if not pbc.any() or abs(np.linalg.det(cell)) < 1e-12:
        # non-periodic: bounding box as a fake cell
        lo, hi = pos.min(0), pos.max(0)
        cell_eff = np.diag(np.maximum(hi - lo, 1.0) + 2 * cutoff + 1.0)
        origin = lo - cutoff - 0.5
        pbc = np.zeros(3, bool)
    else:
        cell_eff, origin = cell.copy(), np.zeros(3)
        for k in range(3):  # degenerate non-periodic axis -> keep given vector
            pass
    h, _ = _heights(cell_eff)
    inv = np.linalg.inv(cell_eff)
    frac = (pos - origin) @ inv
    off = np.floor(frac).astype(np.int64)
    off[:, ~pbc] = 0
    w = frac - off
    nimg = np.where(pbc, np.ceil(cutoff / h).astype(np.int64), 0)
    margin = cutoff / h + 1e-9
    # image atoms near the cell
    imgs_idx, imgs_S, imgs_pos = [], [], []
    for s0 in range(-nimg[0], nimg[0] + 1):
        for s1 in range(-nimg[1], nimg[1] + 1):
            for s2 in range(-nimg[2], nimg[2] + 1):
                S = np.array([s0, s1, s2])
                f = w + S
                m = np.ones(N, bool)
                for k in range(3):
                    if pbc[k]:
                        m &= (f[:, k] > -margin[k]) & (f[:, k] < 1 + margin[k])
                if m.any():
                    idx = np.nonzero(m)[0]
                    imgs_idx.append(idx)
                    imgs_S.append(np.broadcast_to(S, (len(idx), 3)))
                    imgs_pos.append(f[idx] @ cell_eff)
    imgs_idx = np.concatenate(imgs_idx)
    imgs_S = np.concatenate(imgs_S)
    imgs_pos = np.concatenate(imgs_pos)
    wpos = w @ cell_eff
    from scipy.spatial import cKDTree
    tree = cKDTree(imgs_pos)
    ctr = cKDTree(wpos)
    pairs = ctr.query_ball_tree(tree, cutoff * (1 + 1e-9) + 1e-9)
    cnt = np.fromiter((len(p) for p in pairs), np.int64, N)
    i = np.repeat(np.arange(N), cnt)
    k = np.fromiter((q for p in pairs for q in p), np.int64, cnt.sum())
    j = imgs_idx[k]
    S = imgs_S[k] - off[j] + off[i]
    keep, D = _exact_filter(pos, cell if pbc.any() else np.zeros((3, 3)), i, j, S, cutoff)
    i, j, S, D = i[keep], j[keep], S[keep], D[keep]
    order = np.lexsort((S[:, 2], S[:, 1], S[:, 0], j, i))
    return np.stack([i[order], j[order]]), S[order], D[order]


def radius_graph_bruteforce(pos, cell, pbc, cutoff):
    """O(N^2 * images) loop form of the same definition, for small cases."""
    pos = np.asarray(pos, np.float64)
    cell = np.asarray(cell, np.float64).reshape(3, 3)
    pbc = np.asarray(pbc, bool).reshape(3)
    N = len(pos)
    if pbc.any():
        h, _ = _heights(cell)
        # +1 image: positions need not be wrapped into the cell
        span = (pos @ np.linalg.inv(cell))
        ext = np.ceil(span.max(0) - span.min(0)).astype(np.int64) if N else np.zeros(3, np.int64)
        nimg = np.where(pbc, np.ceil(cutoff / h).astype(np.int64) + ext, 0)
    else:
        nimg = np.zeros(3, np.int64)
    I, J, SS = [], [], []
    ii, jj = np.meshgrid(np.arange(N), np.arange(N), indexing='ij')
    ii, jj = ii.ravel(), jj.ravel()
    for s0 in range(-nimg[0], nimg[0] + 1):
        for s1 in range(-nimg[1], nimg[1] + 1):
            for s2 in range(-nimg[2], nimg[2] + 1):
                S = np.broadcast_to(np.array([s0, s1, s2]), (len(ii), 3))
                keep, _ = _exact_filter(pos, cell, ii, jj, S, cutoff)
                I.append(ii[keep]); J.append(jj[keep]); SS.append(S[keep])
    i, j, S = np.concatenate(I), np.concatenate(J), np.concatenate(SS)
    order = np.lexsort((S[:, 2], S[:, 1], S[:, 0], j, i))
    i, j, S = i[order], j[order], S[order]
    _, D = _exact_filter(pos, cell, i, j, S, cutoff)
    return np.stack([i, j]), S, D
