"""TEST INFRASTRUCTURE - real-basis Wigner-3j tables from first principles (numpy, float64).

e3nn (un-vendored third-party dependency of the reference, version unpinned,
``pyproject.toml:24``) supplies the coupling tables used by
``sevenn/nn/convolution.py:88-95`` (``o3.TensorProduct``).  They are restated here as
the unique rotation-invariant 3-tensor in the real spherical-harmonic basis that the
frozen reference graph uses (``deployed_serial.pt`` dump lines 41-66: l=1 -> (x,y,z),
l=2 -> (sqrt3 xz, sqrt3 xy, y^2-(x^2+z^2)/2, sqrt3 yz, sqrt3/2 (z^2-x^2))).

The invariant tensor is unique up to sign; the sign is pinned by the golden tables
extracted from the frozen graph (constants c22..c29), see ``tests/test_oracle_golden.py``.
"""
import functools

import numpy as np


def real_sh(l: int, v: np.ndarray) -> np.ndarray:
    """Un-normalised real SH polynomials of unit vectors v[...,3] in the e3nn order."""
    x, y, z = v[..., 0], v[..., 1], v[..., 2]
    if l == 0:
        return np.ones(v.shape[:-1] + (1,))
    if l == 1:
        return np.stack([x, y, z], -1)
    if l == 2:
        s3 = np.sqrt(3.0)
        return np.stack(
            [s3 * x * z, s3 * x * y, y * y - 0.5 * (x * x + z * z), s3 * y * z,
             0.5 * s3 * (z * z - x * x)], -1)
    raise NotImplementedError('oracle supports l <= 2')


def _rand_rot(rng):
    q, r = np.linalg.qr(rng.standard_normal((3, 3)))
    q = q * np.sign(np.diag(r))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    return q


def wigner_D(l: int, R: np.ndarray) -> np.ndarray:
    """D with real_sh(l, R v) = D @ real_sh(l, v)."""
    rng = np.random.default_rng(12345 + l)
    v = rng.standard_normal((64, 3))
    v /= np.linalg.norm(v, axis=-1, keepdims=True)
    A = real_sh(l, v)                 # [K, d]
    B = real_sh(l, v @ R.T)           # [K, d]
    Dt, *_ = np.linalg.lstsq(A, B, rcond=None)
    return Dt.T


@functools.lru_cache(maxsize=None)
def w3j_unsigned(l1: int, l2: int, l3: int) -> np.ndarray:
    """Frobenius-normalised invariant tensor C[m1,m2,m3]; sign convention:
    the entry of largest magnitude with lowest flat index is positive."""
    d1, d2, d3 = 2 * l1 + 1, 2 * l2 + 1, 2 * l3 + 1
    rng = np.random.default_rng(777)
    rows = []
    for _ in range(6):
        R = _rand_rot(rng)
        K = np.kron(np.kron(wigner_D(l1, R), wigner_D(l2, R)), wigner_D(l3, R))
        rows.append(K - np.eye(d1 * d2 * d3))
    M = np.concatenate(rows, 0)
    _, s, vt = np.linalg.svd(M)
    assert s[-1] < 1e-10 and (len(s) < 2 or s[-2] > 1e-3), (l1, l2, l3, s[-3:])
    C = vt[-1].reshape(d1, d2, d3)
    C = C / np.linalg.norm(C)
    flat = C.reshape(-1)
    k = int(np.argmax(np.abs(flat) > 0.999 * np.abs(flat).max()))
    if flat[k] < 0:
        C = -C
    C[np.abs(C) < 1e-12] = 0.0
    return C


# Sign of the e3nn table relative to ``w3j_unsigned``; pinned against constants c22..c29 of
# the frozen reference graph (all other l<=2 triples carry +1).
_E3NN_SIGN = {(2, 1, 2): -1.0, (2, 2, 2): -1.0}


def w3j(l1: int, l2: int, l3: int) -> np.ndarray:
    """e3nn-convention real Wigner-3j (Frobenius norm 1)."""
    return _E3NN_SIGN.get((l1, l2, l3), 1.0) * w3j_unsigned(l1, l2, l3)
