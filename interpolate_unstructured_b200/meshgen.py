"""Synthetic unstructured meshes of the shapes BASELINE.json names, plus a numpy face-neighbour
builder following the reference converter's conventions (convert_to_binary.py:118-162):
face ``i`` of a cell is made of vertices ``cell[(i + k) % n_vertices]``, ``k = 0..n_points_face-1``;
``neighbors[cell, i]`` is the other cell sharing face ``i`` (only if exactly two cells share it),
else -1.  Coincident points are merged before matching (``np.unique(points, axis=0)``, :130).

For meshes beyond a few million cells use ``api.build_cell_neighbors`` (C++ sort-based builder in
libiu_b200.so); this numpy version is the small-scale cross-check.
"""
from __future__ import annotations

import itertools

import numpy as np

N_POINTS_FACE = {"triangle": 2, "quad": 2, "tetra": 3}


def cell_neighbors(cells, points, cell_type):
    cells = np.asarray(cells)
    nc, nv = cells.shape
    npf = N_POINTS_FACE[cell_type]
    _, inv = np.unique(np.asarray(points), axis=0, return_inverse=True)
    cu = inv.reshape(-1)[cells.reshape(-1)].reshape(cells.shape).astype(np.int64)
    faces = np.stack([cu[:, [(i + k) % nv for k in range(npf)]] for i in range(nv)], axis=1)  # (nc, nv, npf)
    f = np.sort(faces.reshape(-1, npf), axis=1)
    order = np.lexsort(f.T[::-1])
    fs = f[order]
    new = np.ones(len(fs), dtype=bool)
    new[1:] = np.any(fs[1:] != fs[:-1], axis=1)
    gid = np.cumsum(new) - 1
    cnt = np.bincount(gid)
    neigh = np.full(nc * nv, -1, dtype=np.int32)
    start = np.flatnonzero(new)
    two = start[cnt == 2]
    a, b = order[two], order[two + 1]
    neigh[a] = b // nv
    neigh[b] = a // nv
    return neigh.reshape(nc, nv)


def _jittered_lattice(shape, jitter, rng):
    """Points of a regular lattice on the unit square/cube; interior points moved by +-jitter*h."""
    dims = len(shape)
    axes = [np.arange(n + 1, dtype=np.float64) / n for n in shape]
    grid = np.stack(np.meshgrid(*axes, indexing="ij"), axis=-1)  # (n0+1, n1+1[, n2+1], dims)
    if jitter > 0:
        interior = np.ones(grid.shape[:-1], dtype=bool)
        for d, n in enumerate(shape):
            idx = np.arange(n + 1)
            sh = [1] * dims
            sh[d] = n + 1
            interior &= ((idx > 0) & (idx < n)).reshape(sh)
        for d, n in enumerate(shape):
            grid[..., d] += interior * rng.uniform(-jitter / n, jitter / n, size=grid.shape[:-1])
    return grid


def kuhn_tet_mesh(n, jitter=0.10, seed=1234, shuffle=False):
    """Kuhn 6-tet split of n^3 cubes on [0,1]^3: 6 n^3 tets,
    (n+1)^3 points, interior vertices jittered by +-jitter*h (valid up to 0.2 h).
    Returns points (Np,3) f64, cells (Nc,4) int32 0-based."""
    rng = np.random.default_rng(seed)
    g = _jittered_lattice((n, n, n), jitter, rng)
    points = g.reshape(-1, 3)
    m = n + 1

    def vid(i, j, k):
        return (i * m + j) * m + k

    i, j, k = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
    base = np.stack([i, j, k], axis=-1).reshape(-1, 3)
    tets = []
    for perm in itertools.permutations(range(3)):
        cur = base.copy()
        vs = [vid(cur[:, 0], cur[:, 1], cur[:, 2])]
        for ax in perm:
            cur[:, ax] += 1
            vs.append(vid(cur[:, 0], cur[:, 1], cur[:, 2]))
        tets.append(np.stack(vs, axis=1))
    cells = np.stack(tets, axis=1).reshape(-1, 4).astype(np.int32)   # cube-major, 6 tets each
    if shuffle:
        cells = cells[rng.permutation(len(cells))]
    return np.ascontiguousarray(points), np.ascontiguousarray(cells)


def quad_mesh(n, jitter=0.2, seed=1234, clockwise=True):
    """n x n quads on [0,1]^2, z = 0 (config C2).  ``clockwise`` mimics quad.vtu's (0,3,4,1) order."""
    rng = np.random.default_rng(seed)
    g = _jittered_lattice((n, n), jitter, rng)
    points = np.concatenate([g.reshape(-1, 2), np.zeros(((n + 1) ** 2, 1))], axis=1)
    m = n + 1
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    i, j = i.reshape(-1), j.reshape(-1)
    v00, v10, v11, v01 = i * m + j, (i + 1) * m + j, (i + 1) * m + j + 1, i * m + j + 1
    order = [v00, v01, v11, v10] if clockwise else [v00, v10, v11, v01]
    cells = np.stack(order, axis=1).astype(np.int32)
    return np.ascontiguousarray(points), np.ascontiguousarray(cells)


def triangle_mesh(n, jitter=0.2, seed=1234):
    """n x n squares on [0,1]^2 each split into two triangles, z = 0."""
    rng = np.random.default_rng(seed)
    g = _jittered_lattice((n, n), jitter, rng)
    points = np.concatenate([g.reshape(-1, 2), np.zeros(((n + 1) ** 2, 1))], axis=1)
    m = n + 1
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    i, j = i.reshape(-1), j.reshape(-1)
    v00, v10, v11, v01 = i * m + j, (i + 1) * m + j, (i + 1) * m + j + 1, i * m + j + 1
    t1 = np.stack([v00, v10, v01], axis=1)
    t2 = np.stack([v10, v11, v01], axis=1)
    cells = np.stack([t1, t2], axis=1).reshape(-1, 3).astype(np.int32)
    return np.ascontiguousarray(points), np.ascontiguousarray(cells)


def smooth_fields(points, n_vars):
    """n_vars smooth non-linear point fields (Np, n_vars), column v = sin/cos mixtures; column 0 is
    the linear field x+y+z+1 used by the reference tests as analytic self-check."""
    x, y, z = points[:, 0], points[:, 1], points[:, 2]
    cols = [x + y + z + 1.0]
    for v in range(1, n_vars):
        cols.append(np.sin((1 + v) * x + 0.3 * v) * np.cos((2 + v) * y - 0.1 * v) + 0.5 * z * (v + 1) + v)
    return np.ascontiguousarray(np.stack(cols[:n_vars], axis=1))


def splitmix64(idx, seed):
    """Counter-based RNG (SplitMix64 finaliser) -> uint64; same sequence as the device generator."""
    z = (idx.astype(np.uint64) + np.uint64(seed)) * np.uint64(0x9E3779B97F4A7C15)
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def uniform_queries(n, lo, hi, seed, start=0):
    """n query points uniform in the box [lo, hi] (benchmark.f90:28-38: bbox shrunk by 10 % per
    side).  Point i, component k uses counter 3*(start+i)+k."""
    with np.errstate(over="ignore"):
        idx = (np.arange(3 * n, dtype=np.uint64) + np.uint64(3 * start))
        u = (splitmix64(idx, seed) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    u = u.reshape(n, 3)
    lo = np.asarray(lo, dtype=np.float64)
    hi = np.asarray(hi, dtype=np.float64)
    return np.ascontiguousarray(lo + u * (hi - lo))
