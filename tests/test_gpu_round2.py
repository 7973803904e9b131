"""Round-2 parity cases: BASELINE-size meshes cell for cell against the oracle (kdtree2-style start, OpenMP), the
8-variable and generic-variable-count kernel forms, the fused iu_interpolate_at + iu_get_cell_scalar_at pass, the
pre-sorted kernel form (IU_B200_OPT_SORT_QUERIES = 3) and two streams on one grid.

Bar: cell indices bit-exact, values <= 1e-12 relative (in practice bit-identical, asserted where stated).  Queries whose
nearest centre is an exact fp64 distance tie are the one documented exclusion; on the jittered meshes used
here there are none, which the tests assert instead of assuming."""
import os

import numpy as np
import pytest

from interpolate_unstructured_b200 import api, meshgen
from oracle import oracle
from oracle.oracle import OracleGrid

pytestmark = pytest.mark.gpu
RTOL = 1e-12
NT = max(1, len(os.sched_getaffinity(0)))


def _rel_ok(a, b):
    return np.abs(a - b) <= RTOL * np.maximum(1.0, np.abs(b))


def _full_size_parity(kind, pts, cells, nv, n_cold, n_warm, alpha, lo, hi):
    import torch
    nb = oracle.cell_neighbors(cells, pts, kind)
    assert np.array_equal(nb, api.build_cell_neighbors(cells, pts, kind))          # N1 at full size, both builders
    pd = meshgen.smooth_fields(pts, nv)
    g = api.Grid.from_arrays(kind, pts, cells, nb, pd, device=0)
    o = OracleGrid.from_arrays(kind, pts, cells, nb, pd)
    iv = list(range(1, nv + 1))
    # cold: default sort mode (2 = auto -> sorted, ticket tiles), device API, generated on the device
    dev = torch.device("cuda", 0)
    r_t = torch.empty((n_cold, 3), dtype=torch.float64, device=dev)
    g.generate_queries_dev(r_t, first=0, seed=4711, lo=lo, hi=hi)
    r = meshgen.uniform_queries(n_cold, lo, hi, seed=4711)
    assert np.array_equal(r_t.cpu().numpy(), r)                                    # host and device generators agree
    ic_t = torch.zeros(n_cold, dtype=torch.int32, device=dev)
    out_t = torch.full((n_cold, nv), -3.0, dtype=torch.float64, device=dev)
    g.iu_interpolate_at_dev(r_t, iv, out_t, ic_t)
    g.check_dev()
    res, ic = out_t.cpu().numpy(), ic_t.cpu().numpy()
    ores, oic = o.interpolate_at(r, iv, var_at_r=np.full((n_cold, nv), -3.0), nn="kdtree2", nthreads=NT)
    assert (oic >= 1).all()
    assert int((ic != oic).sum()) == 0, f"{int((ic != oic).sum())} of {n_cold} cells differ from the oracle"
    assert _rel_ok(res, ores).all()
    assert np.array_equal(res, ores)                                               # bit-identical in practice
    # warm (C4): displaced by alpha mean cell sizes, previous cells as hints, host-buffer API
    rw = r[:n_warm]
    d = meshgen.uniform_queries(n_warm, [-1.0] * 3, [1.0] * 3, seed=4712)
    if kind != "tetra":
        d[:, 2] = 0.0
    d /= np.sqrt((d * d).sum(1))[:, None]
    ext = pts.max(0) - pts.min(0)
    h = float(np.prod(ext[ext > 0]) / len(cells)) ** (1.0 / int((ext > 0).sum()))
    r1 = np.ascontiguousarray(rw + alpha * h * d)
    hint = ic[:n_warm].copy()
    wres, wic = g.iu_interpolate_at(r1, iv, var_at_r=np.full((n_warm, nv), -3.0), i_cell=hint.copy())
    owres, owic = o.interpolate_at(r1, iv, var_at_r=np.full((n_warm, nv), -3.0), i_cell=hint.copy(), nthreads=NT)
    assert int((wic != owic).sum()) == 0
    assert np.array_equal(wres, owres)
    assert g.stats()["walk_limit_hits"] == 0
    g.close()


def test_c3_full_size_cell_for_cell_vs_oracle():
    """BASELINE config C3 / C4 mesh (10 110 954 tets): 6M cold queries through the default sorted ticket-tile kernel and
    2M hinted queries displaced by 0.5 h, against the oracle started from its kdtree2-style tree."""
    pts, cells = meshgen.kuhn_tet_mesh(119, jitter=0.10, seed=1234)
    assert len(cells) == 10_110_954
    _full_size_parity("tetra", pts, cells, 4, 6_000_000, 2_000_000, 0.5, [0.1] * 3, [0.9] * 3)


def test_c2_full_size_cell_for_cell_vs_oracle():
    """BASELINE config C2 mesh (4.0M jittered quads): 6M cold + 2M hinted queries against the oracle."""
    pts, cells = meshgen.quad_mesh(2000, jitter=0.2, seed=1234)
    assert len(cells) == 4_000_000
    _full_size_parity("quad", pts, cells, 2, 6_000_000, 2_000_000, 0.5, [0.1, 0.1, 0.0], [0.9, 0.9, 0.0])


@pytest.mark.parametrize("kind,n", [("tetra", 14), ("quad", 90), ("triangle", 90)])
def test_variable_counts_8_aligned_8_permuted_5_generic(kind, n):
    """k_query<.., NV = 8> with adjacent aligned columns (32-byte gathers, pairs == 2), the same instantiation with a
    permuted column list (scalar gathers), n_vars = 5 (generic NV = 0 loop), 6 (NV = 0, even) and a repeated column."""
    pts, cells = {"tetra": meshgen.kuhn_tet_mesh, "quad": meshgen.quad_mesh, "triangle": meshgen.triangle_mesh}[kind](n, seed=3)
    nb = api.build_cell_neighbors(cells, pts, kind)
    pd = meshgen.smooth_fields(pts, 10)
    g = api.Grid.from_arrays(kind, pts, cells, nb, pd)
    o = OracleGrid.from_arrays(kind, pts, cells, nb, pd)
    lo, hi = pts.min(0), pts.max(0)
    r = meshgen.uniform_queries(150_000, lo - 0.02 * (hi - lo), hi + 0.02 * (hi - lo), seed=8)   # some outside: rows untouched
    for iv in ([1, 2, 3, 4, 5, 6, 7, 8], [8, 3, 1, 6, 2, 7, 5, 4], [3, 4, 5, 6, 7, 8, 9, 10], [2, 4, 6, 8, 10], [1, 2, 5, 6, 9, 10],
               [5, 5, 5, 5], [9, 10], [7]):
        for sort in (0, 1):
            g.set_option(api.OPT_SORT_QUERIES, sort)
            res, ic = g.iu_interpolate_at(r, iv, var_at_r=np.full((len(r), len(iv)), -9.25))
            ores, oic = o.interpolate_at(r, iv, var_at_r=np.full((len(r), len(iv)), -9.25), nn="brute", nthreads=NT)
            assert (ic == oic).all(), (iv, sort)
            assert np.array_equal(res, ores), (iv, sort)
            assert (oic < 1).any() and (res[oic < 1] == -9.25).all()
        # the cells are known: iu_interpolate_at_icell
        res2 = g.iu_interpolate_at_icell(r, iv, oic, var_at_r=np.full((len(r), len(iv)), -9.25))
        assert np.array_equal(res2, ores), iv


@pytest.mark.parametrize("kind,n,nv", [("tetra", 14, 8), ("tetra", 14, 4), ("quad", 90, 3), ("triangle", 90, 1)])
def test_fused_interpolate_and_cell_scalar(kind, n, nv):
    """iu_b200_interpolate_and_cell_scalar_at_batch = iu_interpolate_at followed by iu_get_cell_scalar_at with the found
    cell as i_guess (m_interp_unstructured.f90:480-495, then :437-448): same results as the library's own two calls and
    as the oracle's two calls; rows of queries outside the mesh stay untouched; cold and hinted inputs."""
    import torch
    pts, cells = {"tetra": meshgen.kuhn_tet_mesh, "quad": meshgen.quad_mesh, "triangle": meshgen.triangle_mesh}[kind](n, seed=11)
    nb = api.build_cell_neighbors(cells, pts, kind)
    pd = meshgen.smooth_fields(pts, nv)
    rng = np.random.default_rng(2)
    cd = rng.normal(size=(len(cells), 3))
    g = api.Grid.from_arrays(kind, pts, cells, nb, pd, cell_data=cd)
    o = OracleGrid.from_arrays(kind, pts, cells, nb, pd, cell_data=cd)
    lo, hi = pts.min(0), pts.max(0)
    nq = 120_000
    r = meshgen.uniform_queries(nq, lo - 0.02 * (hi - lo), hi + 0.02 * (hi - lo), seed=21)
    iv = list(range(1, nv + 1))
    for hinted in (False, True):
        for sort in (0, 1):
            g.set_option(api.OPT_SORT_QUERIES, sort)
            hint = np.zeros(nq, dtype=np.int32)
            if hinted:
                hint = np.where(np.arange(nq) % 2 == 0, rng.integers(1, len(cells) + 1, size=nq), 0).astype(np.int32)
            # oracle: the two reference calls in sequence
            ores, oic = o.interpolate_at(r, iv, var_at_r=np.full((nq, nv), -1.5), i_cell=hint.copy(), nn="brute", nthreads=NT)
            ocres, oic2 = o.get_cell_scalar_at(r, 2, i_guess=oic.copy(), res=np.full(nq, -2.5), nn="brute", nthreads=NT)
            # library: two calls
            res_a, ic_a = g.iu_interpolate_at(r, iv, var_at_r=np.full((nq, nv), -1.5), i_cell=hint.copy())
            cres_a, ic_a2 = g.iu_get_cell_scalar_at(r, 2, i_guess=ic_a.copy(), res=np.full(nq, -2.5))
            # library: fused, host buffers
            res_f, cres_f, ic_f = g.iu_interpolate_and_cell_scalar_at(r, iv, 2, var_at_r=np.full((nq, nv), -1.5),
                                                                      cell_res=np.full(nq, -2.5), i_cell=hint.copy())
            for res, cres, ic in ((res_a, cres_a, ic_a2), (res_f, cres_f, ic_f)):
                assert (ic == oic2).all(), (hinted, sort)
                assert np.array_equal(res, ores) and np.array_equal(cres, ocres), (hinted, sort)
            assert (oic2 < 1).any() and (cres_f[oic2 < 1] == -2.5).all() and (res_f[oic2 < 1] == -1.5).all()
            # fused, device buffers
            dev = torch.device("cuda", 0)
            r_t = torch.from_numpy(r).to(dev)
            ic_t = torch.from_numpy(hint.copy()).to(dev)
            out_t = torch.full((nq, nv), -1.5, dtype=torch.float64, device=dev)
            cr_t = torch.full((nq,), -2.5, dtype=torch.float64, device=dev)
            g.iu_interpolate_and_cell_scalar_at_dev(r_t, iv, 2, out_t, cr_t, ic_t)
            g.check_dev()
            assert np.array_equal(ic_t.cpu().numpy(), oic2) and np.array_equal(out_t.cpu().numpy(), ores)
            assert np.array_equal(cr_t.cpu().numpy(), ocres)
    with pytest.raises(api.IUError):
        g.iu_interpolate_and_cell_scalar_at(r[:10], iv, 4)          # cell_data has 3 columns


@pytest.mark.parametrize("kind,n", [("tetra", 20), ("quad", 150)])
def test_presorted_kernel_form_sort3(kind, n):
    """IU_B200_OPT_SORT_QUERIES = 3 (ticket tiles without a permutation; the caller's queries are taken as spatially
    ordered): results equal the oracle's whether the queries really are ordered or not."""
    pts, cells = {"tetra": meshgen.kuhn_tet_mesh, "quad": meshgen.quad_mesh}[kind](n, seed=5)
    nb = api.build_cell_neighbors(cells, pts, kind)
    pd = meshgen.smooth_fields(pts, 4)
    g = api.Grid.from_arrays(kind, pts, cells, nb, pd)
    o = OracleGrid.from_arrays(kind, pts, cells, nb, pd)
    lo, hi = pts.min(0), pts.max(0)
    r = meshgen.uniform_queries(100_000, lo - 0.03 * (hi - lo), hi + 0.03 * (hi - lo), seed=6)
    key = np.lexsort((r[:, 0], r[:, 1], r[:, 2]))
    for rr in (r, np.ascontiguousarray(r[key])):
        ores, oic = o.interpolate_at(rr, [1, 2, 3, 4], var_at_r=np.full((len(rr), 4), -7.5), nn="brute", nthreads=NT)
        hint = np.where((np.arange(len(rr)) % 3 == 0) & (oic >= 1), oic, 0).astype(np.int32)
        ohres, ohic = o.interpolate_at(rr, [1, 2, 3, 4], var_at_r=np.full((len(rr), 4), -7.5), i_cell=hint.copy(), nn="brute", nthreads=NT)
        g.set_option(api.OPT_SORT_QUERIES, 3)
        for dyn in (0, 1):
            g.set_option(api.OPT_DYNAMIC_TILES, dyn)
            res, ic = g.iu_interpolate_at(rr, [1, 2, 3, 4], var_at_r=np.full((len(rr), 4), -7.5))
            assert (ic == oic).all() and np.array_equal(res, ores), dyn
            res, ic = g.iu_interpolate_at(rr, [1, 2, 3, 4], var_at_r=np.full((len(rr), 4), -7.5), i_cell=hint.copy())
            assert (ic == ohic).all() and np.array_equal(res, ohres), dyn
            assert (g.find_nearby_cell_kdtree(rr) == o.nearest(rr, "brute", nthreads=NT)[0]).all()


def test_two_streams_on_one_grid():
    """*_dev calls on different streams share the grid's sort workspace and ticket counter; the library orders
    them on the device (event), so interleaved calls on two streams return what serial calls return."""
    import torch
    pts, cells = meshgen.kuhn_tet_mesh(24, seed=5)
    nb = api.build_cell_neighbors(cells, pts, "tetra")
    pd = meshgen.smooth_fields(pts, 4)
    g = api.Grid.from_arrays("tetra", pts, cells, nb, pd)
    g.set_option(api.OPT_SORT_QUERIES, 1)
    dev = torch.device("cuda", 0)
    nq = 400_000
    rs = [torch.from_numpy(meshgen.uniform_queries(nq, [0.05] * 3, [0.95] * 3, seed=40 + k)).to(dev) for k in range(2)]
    ref = []
    for r in rs:
        ic = torch.zeros(nq, dtype=torch.int32, device=dev)
        out = torch.zeros((nq, 4), dtype=torch.float64, device=dev)
        g.iu_interpolate_at_dev(r, [1, 2, 3, 4], out, ic)
        g.check_dev()
        ref.append((out.clone(), ic.clone()))
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    torch.cuda.synchronize()
    outs = [torch.zeros((nq, 4), dtype=torch.float64, device=dev) for _ in range(2)]
    ics = [torch.zeros(nq, dtype=torch.int32, device=dev) for _ in range(2)]
    for rep in range(6):
        for k in range(2):
            with torch.cuda.stream(streams[k]):
                ics[k].zero_()
                g.iu_interpolate_at_dev(rs[k], [1, 2, 3, 4], outs[k], ics[k])
    torch.cuda.synchronize()
    g.check_dev()
    for k in range(2):
        assert torch.equal(ics[k], ref[k][1]) and torch.equal(outs[k], ref[k][0])


def _multi_checks(g, o, pts, n_rep):
    """host-buffer batch calls of a handle with n_rep replicas against the oracle and the single-GPU results"""
    assert g.device_count == n_rep and g.stats()["devices"] == n_rep
    lo, hi = pts.min(0), pts.max(0)
    nq = 300_001                                           # not a multiple of the slab alignment
    r = meshgen.uniform_queries(nq, lo - 0.02 * (hi - lo), hi + 0.02 * (hi - lo), seed=31)
    ores, oic = o.interpolate_at(r, [1, 2, 3, 4], var_at_r=np.full((nq, 4), -4.5), nn="brute", nthreads=NT)
    res, ic = g.iu_interpolate_at(r, [1, 2, 3, 4], var_at_r=np.full((nq, 4), -4.5))
    assert (ic == oic).all() and np.array_equal(res, ores)
    st = g.stats()
    assert st["queries"] >= nq and st["not_found"] == int((oic < 1).sum())
    # every replica worked on its slab
    per = [g.replica(i).stats()["queries"] for i in range(n_rep)]
    assert all(q > 0 for q in per[1:]) and per[0] > sum(per[1:]), per   # replica 0 = the handle: reports the sum over all
    # fused call, cell scalars, nearest centre, walk
    cres, ic2 = g.iu_get_cell_scalar_at(r, 1, res=np.full(nq, -2.5))
    ocres, oic2 = o.get_cell_scalar_at(r, 1, res=np.full(nq, -2.5), nn="brute", nthreads=NT)
    assert (ic2 == oic2).all() and np.array_equal(cres, ocres)
    f_res, f_cres, f_ic = g.iu_interpolate_and_cell_scalar_at(r, [1, 2, 3, 4], 1, var_at_r=np.full((nq, 4), -4.5), cell_res=np.full(nq, -2.5))
    assert (f_ic == oic2).all() and np.array_equal(f_res, ores) and np.array_equal(f_cres, ocres)
    assert (g.find_nearby_cell_kdtree(r) == o.nearest(r, "brute", nthreads=NT)[0]).all()
    ok = oic >= 1
    r0 = o.cell_centers[oic[ok] - 1]
    ic1, r_p, ns, stt = g.iu_get_cell_through_neighbors(2 ** 20, r0, r[ok] + 0.01, oic[ok])
    oic1, or_p, ons, ostt = o.get_cell_through_neighbors(2 ** 20, r0, r[ok] + 0.01, oic[ok])
    assert (ic1 == oic1).all() and (ns == ons).all() and (stt == ostt).all() and np.array_equal(r_p, or_p)
    # a column update reaches every replica
    col = np.cos(7 * pts[:, 0]) + pts[:, 1]
    g.set_point_data(2, col)
    o.set_point_data(2, col)
    ores, oic = o.interpolate_at(r, [2, 1], var_at_r=np.full((nq, 2), -4.5), nn="brute", nthreads=NT)
    res, ic = g.iu_interpolate_at(r, [2, 1], var_at_r=np.full((nq, 2), -4.5))
    assert (ic == oic).all() and np.array_equal(res, ores)
    # an added column too
    iv = g.iu_add_point_data("extra")
    g.set_point_data(iv, 2.0 * col)
    res, ic = g.iu_interpolate_at(r, [iv], var_at_r=np.full((nq, 1), -4.5))
    assert np.array_equal(res[oic >= 1, 0], 2.0 * ores[oic >= 1, 0])
    # a hint above n_cells is an error on whichever device sees it (:490)
    bad = np.zeros(nq, dtype=np.int32)
    bad[-5] = g.n_cells + 1
    with pytest.raises(api.IUError):
        g.iu_interpolate_at(r, [1], i_cell=bad)


def _multi_fixture():
    pts, cells = meshgen.kuhn_tet_mesh(16, seed=9)
    nb = api.build_cell_neighbors(cells, pts, "tetra")
    pd = meshgen.smooth_fields(pts, 4)
    cd = np.arange(1, len(cells) + 1, dtype=np.float64)
    return pts, cells, nb, pd, cd


def test_multi_replica_handle_on_one_gpu(monkeypatch):
    """iu_b200_grid_replicate + the slab threads of the host-buffer calls, exercised on a single GPU: the test hook
    IU_B200_ALLOW_SAME_DEVICE_REPLICAS puts three replicas (device-to-device clones) on device 0."""
    monkeypatch.setenv("IU_B200_ALLOW_SAME_DEVICE_REPLICAS", "1")
    pts, cells, nb, pd, cd = _multi_fixture()
    g = api.Grid.from_arrays("tetra", pts, cells, nb, pd, cell_data=cd, device=0)
    o = OracleGrid.from_arrays("tetra", pts, cells, nb, pd, cell_data=cd)
    g.replicate([0, 0])
    _multi_checks(g, o, pts, 3)
    g.close()


def test_multi_gpu_handle_two_devices():
    """one handle, two GPUs (skipped on a one-GPU box): results identical to the oracle's and to the single-GPU handle's"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    pts, cells, nb, pd, cd = _multi_fixture()
    g = api.Grid.from_arrays("tetra", pts, cells, nb, pd, cell_data=cd, devices=[0, 1])
    o = OracleGrid.from_arrays("tetra", pts, cells, nb, pd, cell_data=cd)
    assert g.replica(1).device == 1
    _multi_checks(g, o, pts, 2)
    # the device-buffer calls of replica 1 with tensors of GPU 1
    dev1 = torch.device("cuda", 1)
    r = meshgen.uniform_queries(200_000, [0.05] * 3, [0.95] * 3, seed=77)
    with torch.cuda.device(1):
        r_t = torch.from_numpy(r).to(dev1)
        ic_t = torch.zeros(len(r), dtype=torch.int32, device=dev1)
        out_t = torch.zeros((len(r), 4), dtype=torch.float64, device=dev1)
        g1 = g.replica(1)
        g1.iu_interpolate_at_dev(r_t, [1, 2, 3, 4], out_t, ic_t)
        g1.check_dev()
    res, ic = g.replica(0).iu_interpolate_at(r, [1, 2, 3, 4])
    assert np.array_equal(ic_t.cpu().numpy(), ic) and np.array_equal(out_t.cpu().numpy(), res)
    g.close()


def test_plain_c_driver_runs_the_reference_test_programs():
    """tests/c_driver/iu_driver.c: test_triangle / test_quad / test_tetra (1e-14 on the linear field, one point per call
    like the reference's loops) and benchmark.f90's two passes, through the C ABI from plain C."""
    import subprocess
    import __graft_entry__ as ge
    exe = ge.build_c_driver()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([exe, os.path.join(root, "tests", "golden")], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.count("PASSED") == 5


@pytest.mark.parametrize("kind,n", [("tetra", 16), ("quad", 120)])
def test_sort_of_next_sub_batch_overlaps_the_query_kernel(kind, n):
    """run_query_overlapped: several sorted sub-batches per call, the sort of sub-batch k+1 on a second stream beside
    k_query of sub-batch k (short-lived CTAs, two permutation buffers).  Small IU_B200_OPT_DEV_BATCH values force many
    sub-batches; results must equal the oracle's and the non-overlapped path's for cold, hinted and fused calls."""
    import torch
    pts, cells = {"tetra": meshgen.kuhn_tet_mesh, "quad": meshgen.quad_mesh}[kind](n, seed=5)
    nb = api.build_cell_neighbors(cells, pts, kind)
    pd = meshgen.smooth_fields(pts, 4)
    cd = np.arange(1, len(cells) + 1, dtype=np.float64)
    g = api.Grid.from_arrays(kind, pts, cells, nb, pd, cell_data=cd)
    o = OracleGrid.from_arrays(kind, pts, cells, nb, pd, cell_data=cd)
    lo, hi = pts.min(0), pts.max(0)
    nq = 300_001
    r = meshgen.uniform_queries(nq, lo - 0.02 * (hi - lo), hi + 0.02 * (hi - lo), seed=61)
    ores, oic = o.interpolate_at(r, [1, 2, 3, 4], var_at_r=np.full((nq, 4), -6.5), nn="brute", nthreads=NT)
    hint = np.where((np.arange(nq) % 3 == 0) & (oic >= 1), oic, 0).astype(np.int32)
    ohres, ohic = o.interpolate_at(r, [1, 2, 3, 4], var_at_r=np.full((nq, 4), -6.5), i_cell=hint.copy(), nn="brute", nthreads=NT)
    dev = torch.device("cuda", 0)
    r_t = torch.from_numpy(r).to(dev)
    g.set_option(api.OPT_SORT_QUERIES, 1)
    for batch, tpw, overlap in ((40_000, 1, 1), (37_123, 8, 1), (150_000, 3, 1), (40_000, 8, 0)):
        g.set_option(api.OPT_DEV_BATCH, batch)
        g.set_option(api.OPT_TILES_PER_WARP, tpw)
        g.set_option(api.OPT_OVERLAP_SORT, overlap)
        for rep in range(2):                               # the second call reuses buffers, stream and events
            ic_t = torch.zeros(nq, dtype=torch.int32, device=dev)
            out_t = torch.full((nq, 4), -6.5, dtype=torch.float64, device=dev)
            g.iu_interpolate_at_dev(r_t, [1, 2, 3, 4], out_t, ic_t)
            g.check_dev()
            assert np.array_equal(ic_t.cpu().numpy(), oic) and np.array_equal(out_t.cpu().numpy(), ores), (batch, tpw, overlap)
        ic_t = torch.from_numpy(hint.copy()).to(dev)
        out_t = torch.full((nq, 4), -6.5, dtype=torch.float64, device=dev)
        g.iu_interpolate_at_dev(r_t, [1, 2, 3, 4], out_t, ic_t)
        g.check_dev()
        assert np.array_equal(ic_t.cpu().numpy(), ohic) and np.array_equal(out_t.cpu().numpy(), ohres), (batch, tpw, overlap)
        # fused call and nearest centre through the same path
        ic_t = torch.zeros(nq, dtype=torch.int32, device=dev)
        cr_t = torch.full((nq,), -2.5, dtype=torch.float64, device=dev)
        g.iu_interpolate_and_cell_scalar_at_dev(r_t, [1, 2, 3, 4], 1, out_t, cr_t, ic_t)
        idx_t = torch.zeros(nq, dtype=torch.int32, device=dev)
        g.find_nearby_cell_kdtree_dev(r_t, idx_t)
        g.check_dev()
        assert np.array_equal(ic_t.cpu().numpy(), oic)
        assert np.array_equal(cr_t.cpu().numpy()[oic >= 1], oic[oic >= 1].astype(np.float64))
        assert np.array_equal(idx_t.cpu().numpy(), o.nearest(r, "brute", nthreads=NT)[0])
    st = g.stats()
    assert st["walk_limit_hits"] == 0


@pytest.mark.parametrize("kind,n,jitter", [("tetra", 8, 0.0), ("tetra", 8, 0.1), ("quad", 24, 0.0), ("quad", 24, 0.2), ("triangle", 24, 0.0)])
def test_queries_on_vertices_edges_faces_and_centres(kind, n, jitter):
    """north_star: 'query points lying exactly on shared faces are ... resolved by the reference's own comparison order'.
    Queries AT mesh vertices, edge midpoints, face centroids and cell centres (zero-length rays, :689-692) of lattice and
    jittered meshes: every face-distance tie (strict '<', first face wins, :755), every `distance_left > 0` decision on
    a face (:704) and every nearest-centre tie (lowest cell index, as the oracle's exact search) must come out as in
    the oracle, bit for bit, cold and with hints; walks with explicit start positions too."""
    gen = {"tetra": meshgen.kuhn_tet_mesh, "quad": meshgen.quad_mesh, "triangle": meshgen.triangle_mesh}[kind]
    pts, cells = gen(n, jitter=jitter, seed=4)
    nb = api.build_cell_neighbors(cells, pts, kind)
    pd = meshgen.smooth_fields(pts, 3)
    g = api.Grid.from_arrays(kind, pts, cells, nb, pd)
    o = OracleGrid.from_arrays(kind, pts, cells, nb, pd)
    P = pts[cells]                                                   # (Nc, ppc, 3)
    ppc = cells.shape[1]
    q = [pts, P.mean(1)]                                             # vertices, (approximate) cell centres
    q.append(o.cell_centers)                                         # the exact centres: zero-length rays
    for a in range(ppc):
        for b in range(a + 1, ppc):
            q.append(0.5 * (P[:, a] + P[:, b]))                      # edge midpoints
            q.append(P[:, a] + 0.25 * (P[:, b] - P[:, a]))
    if ppc == 4 and kind == "tetra":
        for a in range(4):
            q.append(P[:, [k for k in range(4) if k != a]].mean(1))  # face centroids
    r = np.ascontiguousarray(np.concatenate(q))
    nq = len(r)
    for sort in (0, 1):
        g.set_option(api.OPT_SORT_QUERIES, sort)
        oidx = o.nearest(r, "brute", nthreads=NT)[0]
        assert (g.find_nearby_cell_kdtree(r) == oidx).all(), sort
        ores, oic = o.interpolate_at(r, [1, 2, 3], var_at_r=np.full((nq, 3), -8.5), nn="brute", nthreads=NT)
        res, ic = g.iu_interpolate_at(r, [1, 2, 3], var_at_r=np.full((nq, 3), -8.5))
        assert (ic == oic).all(), (sort, int((ic != oic).sum()))
        assert np.array_equal(res, ores), sort
        # hints: the cell found, and a neighbour of it (the walk then crosses exactly the face / vertex the query sits on)
        nbr = np.where(oic >= 1, nb[np.maximum(oic, 1) - 1, 0] + 1, 0)
        for hint in (np.maximum(oic, 0).astype(np.int32), np.where(nbr >= 1, nbr, np.maximum(oic, 0)).astype(np.int32)):
            ohres, ohic = o.interpolate_at(r, [1, 2, 3], var_at_r=np.full((nq, 3), -8.5), i_cell=hint.copy(), nn="brute", nthreads=NT)
            hres, hic = g.iu_interpolate_at(r, [1, 2, 3], var_at_r=np.full((nq, 3), -8.5), i_cell=hint.copy())
            assert (hic == ohic).all() and np.array_equal(hres, ohres), sort
    # explicit rays between special points: from every cell's first vertex to its last vertex's neighbour cell centre
    r0 = np.ascontiguousarray(P[:, 0])
    r1 = np.ascontiguousarray(o.cell_centers[np.maximum(nb[:, -1], 0)])
    ic0 = np.arange(1, len(cells) + 1, dtype=np.int32)
    want = o.get_cell_through_neighbors(10 ** 6, r0, r1, ic0)
    got = g.iu_get_cell_through_neighbors(10 ** 6, r0, r1, ic0)
    for a, b in zip(got, want):
        assert np.array_equal(a, b, equal_nan=True)


def test_device_built_kdtree_exact_search_all_sizes():
    """tree_build.cu: the kd-tree is built on the device (level-synchronous: boxes, radix sort by (node, coordinate),
    middle elements).  With the voxel table switched off every query runs the tree search, which must return the exact
    nearest centre (ties -> lowest cell index) for leaf counts of every shape: 1 cell, fewer cells than a leaf, one
    more than a leaf, non-powers of two, lattices full of ties, planar meshes; inside, outside and far-away queries."""
    cases = [("quad", 1, 0.0), ("quad", 2, 0.0), ("quad", 3, 0.2), ("triangle", 2, 0.0), ("triangle", 3, 0.2), ("tetra", 1, 0.0),
             ("tetra", 2, 0.1), ("tetra", 3, 0.0), ("tetra", 5, 0.1), ("quad", 37, 0.0), ("quad", 64, 0.2), ("triangle", 45, 0.2),
             ("tetra", 11, 0.0), ("tetra", 13, 0.1)]
    for kind, n, jitter in cases:
        gen = {"tetra": meshgen.kuhn_tet_mesh, "quad": meshgen.quad_mesh, "triangle": meshgen.triangle_mesh}[kind]
        pts, cells = gen(n, jitter=jitter, seed=7)
        nb = api.build_cell_neighbors(cells, pts, kind)
        pd = meshgen.smooth_fields(pts, 2)
        g = api.Grid.from_arrays(kind, pts, cells, nb, pd)
        o = OracleGrid.from_arrays(kind, pts, cells, nb, pd)
        g.set_option(api.OPT_NN_VOXEL_DENSITY, 0)                      # kd-tree search only
        assert g.stats()["nn_voxels"] == 0 and g.stats()["tree_build_us"] > 0
        lo, hi = pts.min(0), pts.max(0)
        ext = np.where(hi > lo, hi - lo, 0.0)
        r = np.concatenate([meshgen.uniform_queries(20_000, lo - 0.3 * ext, hi + 0.3 * ext, seed=n),
                            meshgen.uniform_queries(500, lo - 50 * ext - 1, hi + 50 * ext + 1, seed=n + 1),
                            o.cell_centers, pts])
        if kind != "tetra":
            r[:20_500, 2] = 0.0
        for sort in (0, 1):
            g.set_option(api.OPT_SORT_QUERIES, sort)
            idx = g.find_nearby_cell_kdtree(r)
            oidx, best, second = o.nearest(r, "brute", nthreads=NT)
            assert (idx == oidx).all(), (kind, n, jitter, sort, int((idx != oidx).sum()))
        res, ic = g.iu_interpolate_at(r, [1, 2], var_at_r=np.full((len(r), 2), -1.0))
        ores, oic = o.interpolate_at(r, [1, 2], var_at_r=np.full((len(r), 2), -1.0), nn="brute", nthreads=NT)
        assert (ic == oic).all() and np.array_equal(res, ores), (kind, n, jitter)
        # the voxel table built from this tree gives the same answers
        g.set_option(api.OPT_NN_VOXEL_DENSITY, 400)
        assert (g.find_nearby_cell_kdtree(r) == oidx).all(), (kind, n, jitter, "voxels")
        g.close()
