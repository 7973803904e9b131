"""CPU tests pinning the oracle (oracle/iu_oracle.cpp) against everything the reference's own
tests hold for the hot path: the analytic linear-field criterion of test_triangle.f90 /
test_quad.f90 / test_tetra.f90 (|res - (x+y+z+1)| <= 1e-14 on the three fixtures), plus
independent cross-checks (containment, brute-force nearest centre)."""
import os

import numpy as np
import pytest

from interpolate_unstructured_b200 import meshgen
from oracle.oracle import OracleGrid

THRESHOLD = 1e-14  # test_triangle.f90:23


def _bbox_queries(g, n, seed, shrink=0.0):
    p = g.points
    lo, hi = p.min(0), p.max(0)
    d = hi - lo
    return meshgen.uniform_queries(n, lo + shrink * d, hi - shrink * d, seed)


@pytest.mark.parametrize("name", ["triangle", "quad", "tetra"])
def test_reference_linear_field_criterion(golden_dir, name):
    g = OracleGrid.from_binda(os.path.join(golden_dir, name + ".binda"))
    ivar = g.get_point_data_index("Polynomial")
    assert ivar != -1
    if name == "tetra":  # test_tetra.f90:37-40 overwrites the (Float32) data with the exact solution
        g.set_point_data(ivar, g.points.sum(1) + 1)
    r = _bbox_queries(g, 1000, seed=7)
    for nn in ("kdtree2", "brute"):
        res, ic = g.interpolate_at(r, [ivar], nn=nn)
        assert (ic >= 1).all()
        assert np.abs(res[:, 0] - (r.sum(1) + 1)).max() <= THRESHOLD
        assert g.point_is_inside_cell(r, ic).all()


def test_fixture_shapes(golden_dir):
    g = OracleGrid.from_binda(os.path.join(golden_dir, "tetra.binda"))
    assert (g.n_points, g.n_cells, g.ppc, g.cell_type) == (35, 96, 4, 3)
    np.testing.assert_allclose(g.cell_volume, 1 / 12, rtol=1e-15)
    assert int((g.neighbors < 1).sum()) == 48
    g = OracleGrid.from_binda(os.path.join(golden_dir, "triangle.binda"))
    assert (g.n_points, g.n_cells, g.ppc, g.cell_type) == (9, 8, 3, 1)
    assert g.get_point_data_index("DistanceToCenter") == 1 and g.get_point_data_index("Polynomial") == 2
    assert g.get_point_data_index("nope") == -1
    g = OracleGrid.from_binda(os.path.join(golden_dir, "quad.binda"))
    np.testing.assert_allclose(g.cell_volume, 20.0, rtol=1e-15)


def test_hand_checkable_tets(golden_dir):
    g = OracleGrid.from_binda(os.path.join(golden_dir, "tetra_cell.binda"))
    # vertices reproduce their own data 1 2 2 3
    res, ic = g.interpolate_at(g.points * 0.999 + 0.001 * g.points.mean(0), [1])
    np.testing.assert_allclose(res[:, 0], [1, 2, 2, 3], atol=5e-3)
    assert (ic == 1).all()
    # outside the single tet -> not found, value untouched
    out, ic = g.interpolate_at(np.array([[2.0, 2.0, 2.0]]), [1], var_at_r=np.array([[-7.0]]))
    assert ic[0] <= 0 and out[0, 0] == -7.0
    g2 = OracleGrid.from_binda(os.path.join(golden_dir, "tetra_two_cell.binda"))
    _, ic = g2.interpolate_at(np.array([[0.2, 0.2, 0.1], [0.8, 0.8, 0.1]]), [1])
    assert list(ic) == [1, 2]


def test_kdtree2_restatement_equals_brute_force():
    pts, cells = meshgen.kuhn_tet_mesh(8, jitter=0.1, seed=3)
    nb = meshgen.cell_neighbors(cells, pts, "tetra")
    g = OracleGrid.from_arrays("tetra", pts, cells, nb, meshgen.smooth_fields(pts, 2))
    r = meshgen.uniform_queries(20000, [0.0] * 3, [1.0] * 3, seed=11)
    idx_b, best, second = g.nearest(r, "brute")
    idx_k = g.nearest(r, "kdtree2")
    ties = best == second
    assert (idx_b[~ties] == idx_k[~ties]).all()
    assert ties.sum() < 5


@pytest.mark.parametrize("kind", ["tetra", "quad", "triangle"])
def test_walk_finds_containing_cell_on_jittered_meshes(kind):
    if kind == "tetra":
        pts, cells = meshgen.kuhn_tet_mesh(6, jitter=0.1, seed=5)
    elif kind == "quad":
        pts, cells = meshgen.quad_mesh(24, jitter=0.2, seed=5)
    else:
        pts, cells = meshgen.triangle_mesh(24, jitter=0.2, seed=5)
    nb = meshgen.cell_neighbors(cells, pts, kind)
    g = OracleGrid.from_arrays(kind, pts, cells, nb, meshgen.smooth_fields(pts, 3))
    lo = [0.0, 0.0, 0.0]
    hi = [1.0, 1.0, 1.0 if kind == "tetra" else 0.0]
    r = meshgen.uniform_queries(5000, lo, hi, seed=13)
    res, ic = g.interpolate_at(r, [1, 2, 3])
    assert (ic >= 1).all()
    assert g.point_is_inside_cell(r, ic).all()
    # linear field exact to rounding; the reference's quad solver (:612-622) cancels when A is
    # small but above its 1e-20 switch, so jittered quads only reach ~1e-10
    assert np.abs(res[:, 0] - (r.sum(1) + 1)).max() < (1e-8 if kind == "quad" else 1e-12)
    # warm start from the found cell: one step, same cell (benchmark.f90:57-69 second pass)
    ic2, ns, _ = g.get_cell(r, ic, diagnostics=True)
    assert (ic2 == ic).all() and (ns == 1).all()
    # outside points leave the domain
    far = r + np.array([2.0, 0.0, 0.0])
    assert (g.get_cell(far) <= 0).all()


def test_neighbor_builder_matches_dict_restatement():
    """meshgen.cell_neighbors vs a literal dict-of-faces restatement of convert_to_binary.py:118-162."""
    pts, cells = meshgen.kuhn_tet_mesh(3, jitter=0.1, seed=1)
    nb = meshgen.cell_neighbors(cells, pts, "tetra")
    from collections import defaultdict
    f2c = defaultdict(list)
    for cid, cell in enumerate(cells):
        for i in range(4):
            f2c[tuple(sorted(cell[(i + k) % 4] for k in range(3)))].append(cid)
    ref = np.full_like(cells, -1)
    for cid, cell in enumerate(cells):
        for i in range(4):
            lst = f2c[tuple(sorted(cell[(i + k) % 4] for k in range(3)))]
            if len(lst) == 2:
                ref[cid, i] = lst[1] if lst[0] == cid else lst[0]
    assert (ref == nb).all()


def test_walk_outputs_status_and_mask():
    pts, cells = meshgen.quad_mesh(8, jitter=0.0, seed=1)
    nb = meshgen.cell_neighbors(cells, pts, "quad")
    # material mask: left half 0, right half 1 (cells are i-major, i = x index)
    mat = (np.arange(64) // 8 >= 4).astype(np.int32)
    g = OracleGrid.from_arrays("quad", pts, cells, nb, meshgen.smooth_fields(pts, 1), icell_data=mat)
    c = g.cell_centers
    r0 = c[[0, 0, 0]]
    r1 = np.array([c[63], c[0] + [0.0, 0.01, 0.0], c[0] + [-1.0, 0.0, 0.0]])
    ic1, r_p, ns, st = g.get_cell_through_neighbors(1000, r0, r1, [1, 1, 1])
    assert list(st) == [0, 0, -1] and ic1[0] == 64 and ic1[1] == 1 and ic1[2] <= 0
    ic1, r_p, ns, st = g.get_cell_through_neighbors(1000, r0, r1, [1, 1, 1], i_icell_mask=1)
    assert st[0] == 1 and mat[ic1[0] - 1] == 1 and abs(r_p[0, 0] - 0.5) < 1e-12
    # zero-length path: early return, r_p / n_steps untouched (m_interp_unstructured.f90:689-692)
    ic1, r_p, ns, st = g.get_cell_through_neighbors(1000, r0[:1], r0[:1], [1])
    assert ic1[0] == 1 and st[0] == 0 and ns[0] == -1 and np.isnan(r_p).all()
    # max_steps exhausted: Fortran do-loop leaves n_steps = max_steps + 1
    ic1, r_p, ns, st = g.get_cell_through_neighbors(2, r0[:1], r1[:1], [1])
    assert ns[0] == 3 and st[0] == 0


def _triangle_rotation_grid(golden_dir):
    """test_trace_field.f90:16-23: triangle.binda plus point data vx = -y, vy = x."""
    from interpolate_unstructured_b200 import binda
    ent = {e.name: e.data for e in binda.read_binda(os.path.join(golden_dir, "triangle.binda")) if e.name != "point_data"}
    pts = ent["points"].astype(np.float64)
    pd = np.stack([-pts[:, 1], pts[:, 0]], axis=1)
    return pts, ent["cells"], ent["cell_neighbors"], pd


def test_trace_field_reference_criterion(golden_dir):
    """test_trace_field.f90:25-53: tracing the rotation field from (1.5, 0) with the arc length
    starting at -0.75 pi must reach the boundary within 100 steps (the only check the reference
    makes); by construction the end point is about (0, 1.5) with length about 0 (comment at :43)."""
    pts, cells, nb, pd = _triangle_rotation_grid(golden_dir)
    o = OracleGrid.from_arrays("triangle", pts, cells, nb, pd)
    y0 = np.array([[1.5, 0.0, -0.75 * np.pi]])
    y, yf, ns, bm = o.integrate_along_field(y0, [1, 2], min_dx=1e-5, max_dx=1e-1, max_steps=100, rtol=1e-3, atol=1e-3,
                                            nvar=1, integrand=1)
    assert 1 < ns[0] <= 100                               # "Boundary not reached" otherwise (:50)
    end = y[0, ns[0] - 1]
    assert abs(end[0]) < 2e-2 and abs(end[1] - 1.5) < 2e-2 and abs(end[2]) < 2e-2
    assert bm[0] == -1                                    # physical boundary
    assert np.isnan(y[0, ns[0]:]).all()                   # later columns untouched
    # stays on the circle of radius 1.5 within the tolerances, field = (-y, x) along the curve
    rad = np.hypot(y[0, :ns[0], 0], y[0, :ns[0], 1])
    assert np.abs(rad - 1.5).max() < 1e-2
    assert np.abs(yf[0, :ns[0], 0] + y[0, :ns[0], 1]).max() < 1e-12
    # reverse direction from the end comes back towards the start; a seed outside is reported at once
    y2, _, ns2, bm2 = o.integrate_along_field(np.array([[5.0, 5.0, 0.0]]), [1, 2], 1e-5, 1e-1, 100, 1e-3, 1e-3, nvar=1, integrand=1)
    assert ns2[0] == 1 and bm2[0] == -1
    # max_steps too small: n_steps = max_steps + 1 and nothing written beyond the array (:1167-1168)
    y3, _, ns3, _ = o.integrate_along_field(y0, [1, 2], 1e-5, 1e-1, 5, 1e-3, 1e-3, nvar=1, integrand=1)
    assert ns3[0] == 6 and (y3[0, :5] == y[0, :5]).all()
