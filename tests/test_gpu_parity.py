"""Parity of the CUDA hot path (through the C ABI, via api.Grid) against the CPU oracle.

Bar (BASELINE.json north_star): cell indices bit-exact, interpolated values within 1e-12 relative
in fp64.  Because the device code is compiled with --fmad=false and the oracle with
-ffp-contract=off, both evaluate the reference's expressions in source order and the values are
in fact expected to be bit-identical; RTOL below is the contractual bound.

Nearest-centre ties (two centres at exactly the same fp64 squared distance) are the one place the
reference's kdtree2 order cannot be reproduced: the device returns the lowest cell
index, like the oracle's brute-force mode, and comparisons with the kdtree2 restatement exclude
those queries and count them.
"""
import os

import numpy as np
import pytest

from interpolate_unstructured_b200 import api, meshgen
from oracle.oracle import OracleGrid

pytestmark = pytest.mark.gpu
RTOL = 1e-12


def _rel_ok(a, b):
    return np.abs(a - b) <= RTOL * np.maximum(1.0, np.abs(b))


def _both_from_binda(path):
    return api.Grid.iu_read_grid(path), OracleGrid.from_binda(path)


def _mesh(kind, n, seed=5, **kw):
    if kind == "tetra":
        pts, cells = meshgen.kuhn_tet_mesh(n, jitter=0.1, seed=seed, **kw)
    elif kind == "quad":
        pts, cells = meshgen.quad_mesh(n, jitter=0.2, seed=seed)
    else:
        pts, cells = meshgen.triangle_mesh(n, jitter=0.2, seed=seed)
    nb = api.build_cell_neighbors(cells, pts, kind)
    return pts, cells, nb


def _both_from_mesh(kind, n, n_vars=4, seed=5, cell_data=False, **kw):
    pts, cells, nb = _mesh(kind, n, seed, **kw)
    pd = meshgen.smooth_fields(pts, n_vars)
    cd = icd = None
    if cell_data:
        rng = np.random.default_rng(seed)
        cd = rng.normal(size=(len(cells), 2))
        icd = rng.integers(0, 3, size=(len(cells), 2)).astype(np.int32)
    g = api.Grid.from_arrays(kind, pts, cells, nb, pd, cd, icd)
    o = OracleGrid.from_arrays(kind, pts, cells, nb, pd, cd, icd)
    return g, o, pts


def _queries(pts, n, seed, shrink=0.1, planar=False):
    lo, hi = pts.min(0), pts.max(0)
    d = hi - lo
    return meshgen.uniform_queries(n, lo + shrink * d, hi - shrink * d, seed)


# ---------------------------------------------------------------------------------------------
# derived geometry (set_cell_points, set_face_normal_vectors, set_cell_volumes, centres)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["triangle", "quad", "tetra", "tetra_cell", "tetra_two_cell"])
def test_geometry_bit_exact_on_fixtures(golden_dir, name):
    g, o = _both_from_binda(os.path.join(golden_dir, name + ".binda"))
    cp, fn, vol, ce = g.export_geometry()
    assert (cp == o.cell_points).all()
    assert (fn == o.cell_face_normals).all()
    assert (vol == o.cell_volume).all()
    assert (ce == o.cell_centers).all()
    assert (g.n_points, g.n_cells, g.n_points_per_cell, g.cell_type) == (o.n_points, o.n_cells, o.ppc, o.cell_type)
    for nm in ("Polynomial", "DistanceToCenter", "nope"):
        assert g.iu_get_point_data_index(nm) == o.get_point_data_index(nm)


@pytest.mark.parametrize("kind,n", [("tetra", 10), ("quad", 50), ("triangle", 50)])
def test_geometry_bit_exact_on_synthetic(kind, n):
    g, o, _ = _both_from_mesh(kind, n)
    cp, fn, vol, ce = g.export_geometry()
    assert (cp == o.cell_points).all() and (fn == o.cell_face_normals).all()
    assert (vol == o.cell_volume).all() and (ce == o.cell_centers).all()


def test_coord_scale_factor(golden_dir):
    path = os.path.join(golden_dir, "quad.binda")      # coordinates up to 10 -> scaled norm2 branch
    g, o = api.Grid.iu_read_grid(path, 2.5), OracleGrid.from_binda(path, 2.5)
    cp, fn, vol, ce = g.export_geometry()
    assert (cp == o.cell_points).all() and (fn == o.cell_face_normals).all() and (vol == o.cell_volume).all()
    r = _queries(o.points, 500, 3, shrink=0.0)
    res, ic = g.iu_interpolate_scalar_at(r, 1)
    ores, oic = o.interpolate_at(r, [1], nn="brute")
    assert (ic == oic).all() and (res == ores[:, 0]).all()


# ---------------------------------------------------------------------------------------------
# the reference's own acceptance criterion (test_triangle/quad/tetra.f90): linear field, 1e-14
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["triangle", "quad", "tetra"])
def test_reference_linear_field_criterion(golden_dir, name):
    g, o = _both_from_binda(os.path.join(golden_dir, name + ".binda"))
    ivar = g.iu_get_point_data_index("Polynomial")
    assert ivar != -1
    if name == "tetra":                                  # test_tetra.f90:37-40
        exact = o.points.sum(1) + 1
        g.set_point_data(ivar, exact)
        o.set_point_data(ivar, exact)
    r = _queries(o.points, 1000, 7, shrink=0.0)
    res, ic = g.iu_interpolate_scalar_at(r, ivar)
    assert (ic >= 1).all()
    assert np.abs(res - (r.sum(1) + 1)).max() <= 1e-14   # test_triangle.f90:23
    assert o.point_is_inside_cell(r, ic).all()
    ores, oic = o.interpolate_at(r, [ivar], nn="brute")
    assert (ic == oic).all()
    assert (res == ores[:, 0]).all()


# ---------------------------------------------------------------------------------------------
# nearest centre (find_nearby_cell_kdtree)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind,n", [("tetra", 12), ("quad", 64), ("triangle", 64)])
def test_nearest_centre_exact(kind, n):
    g, o, pts = _both_from_mesh(kind, n, n_vars=1)
    r = _queries(pts, 100_000, 11, shrink=-0.05)         # slightly beyond the bbox too
    idx = g.find_nearby_cell_kdtree(r)
    oidx, best, second = o.nearest(r, "brute", nthreads=0)
    ties = best == second
    assert (idx[~ties] == oidx[~ties]).all()
    assert (idx == oidx).all()                            # both resolve ties to the lowest index
    kidx = o.nearest(r, "kdtree2")
    assert (idx[~ties] == kidx[~ties]).all()
    assert ties.mean() < 1e-3


def test_nearest_centre_regular_lattice_ties():
    """Unjittered quads: queries on cell-centre bisectors are exact ties -> lowest index, like brute force."""
    pts, cells = meshgen.quad_mesh(16, jitter=0.0)
    nb = api.build_cell_neighbors(cells, pts, "quad")
    g = api.Grid.from_arrays("quad", pts, cells, nb, meshgen.smooth_fields(pts, 1))
    o = OracleGrid.from_arrays("quad", pts, cells, nb, meshgen.smooth_fields(pts, 1))
    k = np.arange(1, 16) / 16.0
    X, Y = np.meshgrid(k, k, indexing="ij")               # lattice vertices: 4-way ties
    r = np.stack([X.ravel(), Y.ravel(), np.zeros(X.size)], axis=1)
    idx = g.find_nearby_cell_kdtree(r)
    oidx, best, second = o.nearest(r, "brute")
    assert (best == second).mean() > 0.5
    assert (idx == oidx).all()


# ---------------------------------------------------------------------------------------------
# iu_get_cell / iu_interpolate_at, cold and warm, sorted and unsorted, chunked host pipeline
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind,n,nq", [("tetra", 12, 200_000), ("quad", 64, 200_000), ("triangle", 64, 200_000)])
@pytest.mark.parametrize("sort", [0, 1])
def test_cold_and_warm_parity(kind, n, nq, sort):
    g, o, pts = _both_from_mesh(kind, n, n_vars=4)
    g.set_option(api.OPT_SORT_QUERIES, sort)
    g.set_option(api.OPT_HOST_CHUNK, 30_000)             # several pipeline chunks, last one ragged
    r = _queries(pts, nq, 21, shrink=-0.02)              # a few queries fall outside -> i_cell <= 0
    i_vars = [4, 1, 3]
    sentinel = -123.25
    res = np.full((nq, 3), sentinel)
    res, ic = g.iu_interpolate_at(r, i_vars, var_at_r=res)
    ores, oic = o.interpolate_at(r, i_vars, var_at_r=np.full((nq, 3), sentinel), nn="brute")
    assert (ic == oic).all()
    assert (res == ores).all()
    assert _rel_ok(res, ores).all()
    nf = ic < 1
    assert 0 < nf.sum() < nq // 5
    assert (res[nf] == sentinel).all()                    # untouched rows (:492)
    assert o.point_is_inside_cell(r[~nf], ic[~nf]).all()
    # get_cell alone
    ic2 = g.iu_get_cell(r)
    assert (ic2 == oic).all()
    # warm pass with hints = previous cells, positions moved by ~half a cell
    h = 1.0 / n
    rng = np.random.default_rng(5)
    r2 = r[~nf] + rng.normal(scale=0.5 * h, size=(int((~nf).sum()), 3)) * ([1, 1, 1] if kind == "tetra" else [1, 1, 0])
    hint = ic[~nf].copy()
    wres, wic = g.iu_interpolate_at(r2, i_vars, i_cell=hint.copy())
    owres, owic = o.interpolate_at(r2, i_vars, i_cell=hint.copy())
    assert (wic == owic).all() and (wres == owres).all()
    # n_steps parity through the explicit walk API, starting at the hinted cell's centre (:429)
    _, _, _, ce = g.export_geometry()
    ic1, r_p, ns, st = g.iu_get_cell_through_neighbors(2**31 - 1, ce[hint - 1], r2, hint)
    oic1, or_p, ons, ost = o.get_cell_through_neighbors(2**31 - 1, ce[hint - 1], r2, hint)
    assert (ic1 == oic1).all() and (ns == ons).all() and (st == ost).all()
    assert (r_p == or_p).all()
    assert (ic1 == owic).all()


def test_kdtree2_start_gives_same_cells_except_ties():
    g, o, pts = _both_from_mesh("tetra", 12, n_vars=2)
    r = _queries(pts, 100_000, 33)
    ic = g.iu_get_cell(r)
    oic, ons, _ = o.get_cell(r, nn="kdtree2", diagnostics=True)
    _, best, second = o.nearest(r, "brute")
    ties = best == second
    assert (ic[~ties] == oic[~ties]).all()
    assert ties.sum() < 100


def test_scalar_and_cell_data_parity():
    g, o, pts = _both_from_mesh("tetra", 10, n_vars=2, cell_data=True)
    g.set_option(api.OPT_HOST_CHUNK, 7_000)
    nq = 50_000
    r = _queries(pts, nq, 41, shrink=-0.03)
    res, ic = g.iu_interpolate_scalar_at(r, 2, var_at_r=np.full(nq, 9.5))
    ores, oic = o.interpolate_at(r, [2], var_at_r=np.full((nq, 1), 9.5), nn="brute")
    assert (ic == oic).all() and (res == ores[:, 0]).all()
    for iv in (1, 2):
        c, ic = g.iu_get_cell_scalar_at(r, iv, res=np.full(nq, -1.5))
        oc, oic = o.get_cell_scalar_at(r, iv, res=np.full(nq, -1.5), nn="brute")
        assert (ic == oic).all() and (c == oc).all()
        k, ic = g.iu_get_icell_scalar_at(r, iv, res=np.full(nq, -7, dtype=np.int32))
        ok, oic = o.get_icell_scalar_at(r, iv, res=np.full(nq, -7, dtype=np.int32), nn="brute")
        assert (ic == oic).all() and (k == ok).all()
        assert (k[ic < 1] == -7).all()
    # interpolate_at_icell: cells given, entries < 1 skipped
    v = g.iu_interpolate_at_icell(r, [1, 2], oic, var_at_r=np.full((nq, 2), 3.0))
    found = oic >= 1
    ov = o.interpolate_at_icell(r[found], [1, 2], oic[found])
    assert (v[found] == ov).all() and (v[~found] == 3.0).all()


def test_point_data_update_and_add():
    """Callers mutate ug%point_data after iu_read_grid (test_tetra.f90:38-40) and append columns
    with iu_add_point_data (test_trace_field.f90:17-23)."""
    g, o, pts = _both_from_mesh("triangle", 20, n_vars=1)
    r = _queries(pts, 5_000, 43)
    assert g.n_point_data == 1
    i_vx = g.iu_add_point_data("vx")
    i_vy = g.iu_add_point_data("vy")
    assert (i_vx, i_vy) == (2, 3) and g.iu_get_point_data_index("vy") == 3
    z, _ = g.iu_interpolate_at(r, [i_vx, i_vy])
    assert (z == 0).all()                                  # new columns are zero-filled (:134)
    g.set_point_data(i_vx, -pts[:, 1])
    g.set_point_data(i_vy, pts[:, 0])
    o.set_point_data(2, -pts[:, 1])
    o.set_point_data(3, pts[:, 0])
    res, ic = g.iu_interpolate_at(r, [1, 2, 3])
    ores, oic = o.interpolate_at(r, [1, 2, 3], nn="brute")
    assert (ic == oic).all() and (res == ores).all()
    for k in range(4, 7):                                  # grow beyond the padded row stride
        assert g.iu_add_point_data(f"extra{k}") == k
    g.set_point_data(6, pts[:, 0] * 2)
    res2, _ = g.iu_interpolate_at(r, [1, 2, 3, 6])
    assert (res2[:, :3] == ores).all()
    assert np.abs(res2[:, 3] - 2 * r[:, 0]).max() < 1e-12


def test_walk_status_mask_early_and_step_cap():
    pts, cells = meshgen.quad_mesh(8, jitter=0.0, seed=1)
    nb = api.build_cell_neighbors(cells, pts, "quad")
    mat = (np.arange(64) // 8 >= 4).astype(np.int32)
    pd = meshgen.smooth_fields(pts, 1)
    g = api.Grid.from_arrays("quad", pts, cells, nb, pd, icell_data=mat)
    o = OracleGrid.from_arrays("quad", pts, cells, nb, pd, icell_data=mat)
    c = o.cell_centers
    r0 = c[[0, 0, 0, 0]]
    r1 = np.array([c[63], c[0] + [0.0, 0.01, 0.0], c[0] + [-1.0, 0.0, 0.0], c[0]])
    for mask in (0, 1):
        for cap in (1000, 2):
            got = g.iu_get_cell_through_neighbors(cap, r0, r1, [1, 1, 1, 1], i_icell_mask=mask)
            exp = o.get_cell_through_neighbors(cap, r0, r1, [1, 1, 1, 1], i_icell_mask=mask)
            for a, b in zip(got, exp):
                assert np.array_equal(a, b, equal_nan=True), (mask, cap)
    ic1, r_p, ns, st = g.iu_get_cell_through_neighbors(1000, r0, r1, [1, 1, 1, 1])
    assert list(st) == [0, 0, -1, 0] and ic1[0] == 64 and ic1[2] <= 0
    assert ns[3] == -1 and np.isnan(r_p[3]).all()          # zero-length ray: untouched (:689-692)
    ic1, r_p, ns, st = g.iu_get_cell_through_neighbors(2, r0[:1], r1[:1], [1])
    assert ns[0] == 3 and st[0] == 0                       # Fortran do-loop: max_steps + 1


def test_walk_batch_sorted_and_unsorted_agree_with_oracle():
    """iu_get_cell_through_neighbors on a batch, processed in input order and (IU_B200_OPT_SORT_QUERIES = 1) in Morton
    order of the start positions: same bits, equal to the oracle, material mask included."""
    g, o, pts = _both_from_mesh("tetra", 10, n_vars=1, cell_data=True)
    n = 120_000
    r0 = _queries(pts, n, 31)
    ic0 = g.iu_get_cell(r0)
    assert (ic0 >= 1).all()
    rng = np.random.default_rng(3)
    r1 = r0 + rng.normal(scale=0.08, size=(n, 3))                # several cells away, some leave the domain
    exp = o.get_cell_through_neighbors(50, r0, r1, ic0, i_icell_mask=2)
    for sort in (0, 1):
        g.set_option(api.OPT_SORT_QUERIES, sort)
        g.set_option(api.OPT_HOST_CHUNK, 50_000)
        got = g.iu_get_cell_through_neighbors(50, r0, r1, ic0, i_icell_mask=2)
        for a, b in zip(got, exp):
            assert np.array_equal(a, b, equal_nan=True), sort
    assert (exp[3] == -1).sum() > 10 and (exp[3] == 1).sum() > 100 and (exp[3] == 0).sum() > 100


def test_walk_batch_auto_sort_decision():
    """IU_B200_OPT_SORT_QUERIES = 2 (default): large batches of rays are sorted only when the latest sample of ray lengths is above
    about half a mean cell size; either way the results are the bits of the unsorted walk."""
    g, o, pts = _both_from_mesh("tetra", 10, n_vars=1, cell_data=True)
    n = (1 << 20) + 1000
    r0 = _queries(pts, n, 77)
    ic0 = g.iu_get_cell(r0)
    assert (ic0 >= 1).all()
    h_cell = (np.prod(pts.max(0) - pts.min(0)) / g.n_cells) ** (1 / 3)
    rng = np.random.default_rng(5)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    g.set_option(api.OPT_HOST_CHUNK, 1 << 21)
    g.set_option(api.OPT_TIME_KERNELS, 1)
    for length, sorted_expected in ((0.1 * h_cell, False), (2.0 * h_cell, True)):
        r1 = r0 + length * d
        g.set_option(api.OPT_SORT_QUERIES, 0)
        exp = g.iu_get_cell_through_neighbors(50, r0, r1, ic0)
        g.set_option(api.OPT_SORT_QUERIES, 2)
        g.iu_get_cell_through_neighbors(50, r0, r1, ic0)       # its sample decides for the calls that follow
        g.reset_stats()
        got = g.iu_get_cell_through_neighbors(50, r0, r1, ic0)
        assert (g.stats()["sort_ns"] > 0) == sorted_expected
        for a, b in zip(got, exp):
            assert np.array_equal(a, b, equal_nan=True)


def test_error_behaviour():
    g, o, pts = _both_from_mesh("tetra", 4, n_vars=2)
    r = _queries(pts, 100, 1)
    with pytest.raises(api.IUError) as e:                  # error stop at :490
        g.iu_interpolate_at(r, [1], i_cell=np.full(100, g.n_cells + 1, dtype=np.int32))
    assert e.value.code == api.E_HINT
    with pytest.raises(api.IUError) as e:
        g.iu_interpolate_at(r, [3])
    assert e.value.code == api.E_INVAL
    with pytest.raises(api.IUError) as e:
        g.iu_get_cell_scalar_at(r, 1)
    assert e.value.code == api.E_INVAL
    # the grid is still usable afterwards
    res, ic = g.iu_interpolate_at(r, [1])
    assert (ic >= 1).all()
    # empty batch is a no-op
    res, ic = g.iu_interpolate_at(np.zeros((0, 3)), [1])
    assert res.shape == (0, 1) and ic.shape == (0,)


def test_device_buffer_api_matches_host_api():
    import torch
    g, o, pts = _both_from_mesh("tetra", 12, n_vars=4, cell_data=True)
    nq = 300_000
    lo, hi = pts.min(0) - 0.02, pts.max(0) + 0.02
    dev = torch.device("cuda", g.device)
    r_d = torch.empty((nq, 3), dtype=torch.float64, device=dev)
    g.generate_queries_dev(r_d, first=17, seed=99, lo=lo, hi=hi)
    r = r_d.cpu().numpy()
    assert (r == meshgen.uniform_queries(nq, lo, hi, 99, start=17)).all()   # same counter-based stream
    ic_d = torch.zeros(nq, dtype=torch.int32, device=dev)
    out_d = torch.full((nq, 4), -5.0, dtype=torch.float64, device=dev)
    g.iu_interpolate_at_dev(r_d, [1, 2, 3, 4], out_d, ic_d)
    g.check_dev()
    ores, oic = o.interpolate_at(r, [1, 2, 3, 4], var_at_r=np.full((nq, 4), -5.0), nn="brute")
    assert (ic_d.cpu().numpy() == oic).all()
    assert (out_d.cpu().numpy() == ores).all()
    idx_d = torch.zeros(nq, dtype=torch.int32, device=dev)
    g.find_nearby_cell_kdtree_dev(r_d, idx_d)
    oidx, _, _ = o.nearest(r, "brute")
    assert (idx_d.cpu().numpy() == oidx).all()
    cd_d = torch.full((nq,), 2.0, dtype=torch.float64, device=dev)
    ic2_d = torch.zeros(nq, dtype=torch.int32, device=dev)
    g.iu_get_cell_scalar_at_dev(r_d, 2, ic2_d, cd_d)
    oc, _ = o.get_cell_scalar_at(r, 2, res=np.full(nq, 2.0), nn="brute")
    assert (cd_d.cpu().numpy() == oc).all()
    # hint error is reported by check_dev
    bad = torch.full((nq,), g.n_cells + 5, dtype=torch.int32, device=dev)
    g.iu_get_cell_dev(r_d, bad)
    with pytest.raises(api.IUError) as e:
        g.check_dev()
    assert e.value.code == api.E_HINT
    g.check_dev()                                          # cleared


def test_shuffled_cell_order_and_mid_size_mesh():
    """384k tets in random file order, 1M queries: sorted and unsorted batches agree with the oracle
    (OpenMP over queries) bit for bit."""
    g, o, pts = _both_from_mesh("tetra", 40, n_vars=4, shuffle=True)
    nq = 1_000_000
    r = _queries(pts, nq, 77)
    ores, oic = o.interpolate_at(r, [1, 2, 3, 4], nn="kdtree2", nthreads=0)
    _, best, second = o.nearest(r[:50_000], "brute")
    for sort in (0, 1):
        g.set_option(api.OPT_SORT_QUERIES, sort)
        res, ic = g.iu_interpolate_at(r, [1, 2, 3, 4])
        assert (ic == oic).all()
        assert (res == ores).all()
    assert (ic >= 1).all()
    assert np.abs(res[:, 0] - (r.sum(1) + 1)).max() < 1e-12
    st = g.stats()
    assert st["kernel_launches"] > 0 and st["walk_limit_hits"] == 0 and st["queries"] >= 2 * nq


@pytest.mark.parametrize("kind,n", [("tetra", 24), ("quad", 200), ("triangle", 150)])
def test_kernel_forms_agree_with_oracle(kind, n):
    """IU_B200_OPT_DYNAMIC_TILES: static grid-stride and ticket-tile scheduling (dense, sparse and
    tiny sorted batches; mixed cold / warm lanes) return the brute-force nearest centres and the
    oracle's cells and values."""
    g, o, pts = _both_from_mesh(kind, n, n_vars=4)
    g.set_option(api.OPT_SORT_QUERIES, 1)
    for nq, seed in ((200_000, 5), (3_000, 6), (33, 7)):
        r = _queries(pts, nq, seed, shrink=-0.03)
        oidx, best, second = o.nearest(r, "brute", nthreads=0)
        ores, oic = o.interpolate_at(r, [1, 2, 3, 4], var_at_r=np.full((nq, 4), -7.5), nn="brute", nthreads=0)
        hint = np.where((np.arange(nq) % 3 == 0) & (oic >= 1), oic, 0).astype(np.int32)    # every third query warm
        ohres, ohic = o.interpolate_at(r, [1, 2, 3, 4], var_at_r=np.full((nq, 4), -7.5), i_cell=hint.copy(), nn="brute", nthreads=0)
        for form in (0, 1):
            g.set_option(api.OPT_DYNAMIC_TILES, form)
            assert (g.find_nearby_cell_kdtree(r) == oidx).all(), (form, nq)
            res, ic = g.iu_interpolate_at(r, [1, 2, 3, 4], var_at_r=np.full((nq, 4), -7.5))
            assert (ic == oic).all() and (res == ores).all(), (form, nq)
            res, ic = g.iu_interpolate_at(r, [1, 2, 3, 4], var_at_r=np.full((nq, 4), -7.5), i_cell=hint.copy())
            assert (ic == ohic).all() and (res == ohres).all(), (form, nq)


# ---------------------------------------------------------------------------------------------
# where the start cell matters: non-convex domains, queries outside, far away, NaN
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind,n", [("quad", 48), ("tetra", 10)])
def test_nonconvex_domain_outside_and_nan_queries(kind, n):
    """In a domain with a hole the ray from the nearest centre may leave the mesh (i_cell = 0) although
    the query lies inside another cell (:427-432): the result then depends on the EXACT nearest centre,
    so this pins the kd-tree search (boxes, float bounds, leaf blocks) against brute force far beyond
    what convex meshes exercise.  Also: queries outside the bounding box, far away, and NaN."""
    pts, cells, _ = _mesh(kind, n, seed=9)
    if kind == "quad":
        c = pts[cells].mean(1)
        keep = ~((np.abs(c[:, 0] - 0.5) < 0.2) & (np.abs(c[:, 1] - 0.45) < 0.25)) & ~((c[:, 0] > 0.7) & (c[:, 1] > 0.7))
    else:
        c = pts[cells].mean(1)
        keep = ~((np.abs(c - 0.5) < 0.22).all(1)) & ~((c[:, 0] > 0.75) & (c[:, 2] < 0.3))
    cells = np.ascontiguousarray(cells[keep])
    nb = api.build_cell_neighbors(cells, pts, kind)
    pd = meshgen.smooth_fields(pts, 2)
    g = api.Grid.from_arrays(kind, pts, cells, nb, pd)
    o = OracleGrid.from_arrays(kind, pts, cells, nb, pd)
    nq = 150_000
    r = meshgen.uniform_queries(nq, [-0.3, -0.3, -0.3 if kind == "tetra" else 0.0], [1.3, 1.3, 1.3 if kind == "tetra" else 0.0], 77)
    r[:50] *= 1e6                                                 # far away
    r[50:60] = np.nan
    r_oracle = r.copy()                                           # the reference (and the oracle) index out of bounds on an
    r[60, 0] = np.inf                                             # infinite coordinate; the library reports "not found"
    for sort in (0, 1):
        g.set_option(api.OPT_SORT_QUERIES, sort)
        sentinel = 7.5
        res, ic = g.iu_interpolate_at(r, [1, 2], var_at_r=np.full((nq, 2), sentinel))
        ores, oic = o.interpolate_at(r_oracle, [1, 2], var_at_r=np.full((nq, 2), sentinel), nn="brute")
        finite = np.isfinite(r).all(1)
        assert (ic[finite] == oic[finite]).all()
        assert (res[finite] == ores[finite]).all()
        assert (ic[~finite] < 1).all() and (res[~finite] == sentinel).all()
        nn = g.find_nearby_cell_kdtree(r[finite])
        onn, _, _ = o.nearest(r[finite], "brute")
        assert (nn == onn).all()
    inside_hole_or_out = (ic < 1) & finite
    found = ic >= 1
    assert inside_hole_or_out.sum() > nq // 10 and found.sum() > nq // 10
    assert o.point_is_inside_cell(r[found], ic[found]).all()
