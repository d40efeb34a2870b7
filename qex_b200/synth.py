"""Synthetic inputs for benchmarks and large-lattice property tests, generated with numpy only
(the product may not use the oracle): Haar-like random SU(3) links and gaussian colour vectors in
the V=1 QEX host layout."""
import numpy as np


def random_su3_links(vol, n=4, seed=0, scale=None):
    """n link fields float64[vol,3,3,2]; Gram-Schmidt of gaussian 3x3 complex matrices, det fixed to 1.
    scale: optional per-field factor (e.g. 0.1 for synthetic long links, benchStagProp.nim:84-92)."""
    rs = np.random.default_rng(seed)
    out = []
    for k in range(n):
        a = rs.standard_normal((vol, 3, 3)) + 1j * rs.standard_normal((vol, 3, 3))
        v0 = a[:, 0] / np.linalg.norm(a[:, 0], axis=1, keepdims=True)
        v1 = a[:, 1] - v0 * np.sum(v0.conj() * a[:, 1], axis=1, keepdims=True)
        v1 /= np.linalg.norm(v1, axis=1, keepdims=True)
        v2 = np.cross(v0.conj(), v1.conj())
        u = np.stack([v0, v1, v2], axis=1)
        if scale is not None and scale[k] != 1.0:
            u = u * scale[k]
        out.append(np.ascontiguousarray(np.stack([u.real, u.imag], axis=-1)))
    return out


def gaussian_vector(vol, seed=0):
    rs = np.random.default_rng(1000 + seed)
    return rs.standard_normal((vol, 3, 2))
