#!/usr/bin/env python3
"""Small end-to-end exercise of every kernel of libiu_b200.so for compute-sanitizer (memcheck / racecheck / initcheck):
three cell types, cold / hinted / pre-located interpolation in every scheduling form, nearest centre (voxel table and
kd-tree fallback), cell scalars, the fused call, ray batches with a mask, the field-line tracer, column updates,
geometry / mesh export and a two-replica handle.  Results are checked against the oracle so that a silent corruption
fails the run as well.

    compute-sanitizer --tool memcheck python tools/sanitize_smoke.py
    IU_B200_LIB=$PWD/interpolate_unstructured_b200/lib_checked/libiu_b200.so python tools/sanitize_smoke.py

The second form is what this pool allows (compute-sanitizer is closed on it): the library built with -DIU_CHECKED=1
range-checks every index its kernels derive from mesh or query data and counts violations (IU_B200_STAT_BOUNDS_ERRORS)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["IU_B200_ALLOW_SAME_DEVICE_REPLICAS"] = "1"
from interpolate_unstructured_b200 import api, meshgen  # noqa: E402
from oracle.oracle import OracleGrid  # noqa: E402


def main():
    nq = int(os.environ.get("IU_SANITIZE_QUERIES", "20000"))
    for kind, n in (("tetra", 7), ("quad", 30), ("triangle", 30)):
        pts, cells = {"tetra": meshgen.kuhn_tet_mesh, "quad": meshgen.quad_mesh, "triangle": meshgen.triangle_mesh}[kind](n, seed=3)
        nb = api.build_cell_neighbors(cells, pts, kind)
        pd = meshgen.smooth_fields(pts, 8)
        rng = np.random.default_rng(1)
        cd = rng.normal(size=(len(cells), 2))
        icd = rng.integers(0, 2, size=(len(cells), 1)).astype(np.int32)
        g = api.Grid.from_arrays(kind, pts, cells, nb, pd, cd, icd)
        o = OracleGrid.from_arrays(kind, pts, cells, nb, pd, cd, icd)
        lo, hi = pts.min(0), pts.max(0)
        r = meshgen.uniform_queries(nq, lo - 0.05 * (hi - lo), hi + 0.05 * (hi - lo), seed=5)
        for density in (400, 0):                                   # voxel table, then kd-tree only
            g.set_option(api.OPT_NN_VOXEL_DENSITY, density)
            assert (g.find_nearby_cell_kdtree(r) == o.nearest(r, "brute")[0]).all()
        g.set_option(api.OPT_NN_VOXEL_DENSITY, 400)
        for iv in ([1, 2, 3, 4], [1, 2, 3, 4, 5, 6, 7, 8], [3], [2, 5, 7]):
            ores, oic = o.interpolate_at(r, iv, var_at_r=np.full((nq, len(iv)), -1.0), nn="brute")
            hint = np.where(np.arange(nq) % 2 == 0, np.maximum(oic, 0), 0).astype(np.int32)
            ohres, ohic = o.interpolate_at(r, iv, var_at_r=np.full((nq, len(iv)), -1.0), i_cell=hint.copy(), nn="brute")
            for sort, dyn in ((0, 0), (1, 0), (1, 1), (3, 1)):
                g.set_option(api.OPT_SORT_QUERIES, sort); g.set_option(api.OPT_DYNAMIC_TILES, dyn)
                res, ic = g.iu_interpolate_at(r, iv, var_at_r=np.full((nq, len(iv)), -1.0))
                assert (ic == oic).all() and np.array_equal(res, ores), (kind, iv, sort, dyn)
                res, ic = g.iu_interpolate_at(r, iv, var_at_r=np.full((nq, len(iv)), -1.0), i_cell=hint.copy())
                assert (ic == ohic).all() and np.array_equal(res, ohres), (kind, iv, sort, dyn)
            assert np.array_equal(g.iu_interpolate_at_icell(r, iv, oic, var_at_r=np.full((nq, len(iv)), -1.0)), ores)
        g.set_option(api.OPT_SORT_QUERIES, 1)
        cres, ic = g.iu_get_cell_scalar_at(r, 2, res=np.full(nq, -2.0))
        ocres, oic = o.get_cell_scalar_at(r, 2, res=np.full(nq, -2.0), nn="brute")
        assert (ic == oic).all() and np.array_equal(cres, ocres)
        ires, ic = g.iu_get_icell_scalar_at(r, 1, res=np.full(nq, -3, dtype=np.int32))
        oires, _ = o.get_icell_scalar_at(r, 1, res=np.full(nq, -3, dtype=np.int32), nn="brute")
        assert np.array_equal(ires, oires)
        f_res, f_cres, f_ic = g.iu_interpolate_and_cell_scalar_at(r, [1, 2, 3, 4], 2, var_at_r=np.full((nq, 4), -1.0), cell_res=np.full(nq, -2.0))
        assert (f_ic == oic).all() and np.array_equal(f_cres, ocres)
        # rays with a material mask, sorted and unsorted
        ok = oic >= 1
        r0 = o.cell_centers[oic[ok] - 1]
        r1 = r[ok] + 0.3 * (hi - lo) * rng.normal(size=(int(ok.sum()), 3)) * (np.array([1, 1, 1 if kind == "tetra" else 0]))
        want = o.get_cell_through_neighbors(10 ** 6, r0, r1, oic[ok], i_icell_mask=1)
        for sort in (0, 1):
            g.set_option(api.OPT_SORT_QUERIES, sort)
            got = g.iu_get_cell_through_neighbors(10 ** 6, r0, r1, oic[ok], i_icell_mask=1)
            assert all(np.array_equal(a, b) for a, b in zip(got, want)), (kind, sort)
        # tracer
        nl, ms = 300, 40
        nd = 3 if kind == "tetra" else 2
        c = pts - 0.5
        field = np.stack([-c[:, 1] + 0.3 * c[:, 0], c[:, 0] + 0.3 * c[:, 1], 0.2 + 0 * c[:, 2]], axis=1)[:, :nd]
        for k in range(nd):
            g.set_point_data(1 + k, field[:, k]); o.set_point_data(1 + k, field[:, k])
        y = np.full((nl, ms, nd + 1), np.nan)
        y[:, 0, :nd] = rng.uniform(0.2, 0.8, size=(nl, nd)); y[:, 0, nd] = 0
        h = 1.0 / n
        yy, yf, ns, bm = g.iu_integrate_along_field(y.copy(), list(range(1, nd + 1)), 1e-3 * h, 2 * h, 1e-4, 1e-6, nvar=1, integrand=1)
        oy, oyf, ons, obm = o.integrate_along_field(y[:, 0, :], list(range(1, nd + 1)), 1e-3 * h, 2 * h, ms, 1e-4, 1e-6, nvar=1, integrand=1, nn="brute")
        assert (ns == ons).all()
        # exports, replicas
        geo = g.export_geometry()
        assert np.array_equal(geo["cell_volume"] if isinstance(geo, dict) else geo[2], o.cell_volume)
        g.export_mesh()
        g.replicate([0])
        res, ic = g.iu_interpolate_at(np.tile(r, (5, 1)), [1, 2])
        assert g.device_count == 2 and (ic[:nq] == ic[nq:2 * nq]).all()
        be = g.stats()["bounds_errors"]
        assert be <= 0, f"{be} index-range violations counted by the checked build"
        g.close()
        print(kind, "ok", "(checked build: 0 index-range violations)" if be == 0 else "(unchecked build)", flush=True)
    print("sanitize smoke ok")


if __name__ == "__main__":
    main()
