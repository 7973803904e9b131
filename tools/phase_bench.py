#!/usr/bin/env python3
"""Phase timing of the cold path on the C3 mesh (diagnostic, not the bench contract).

Times the query kernel of each device entry point on the same 67M Morton-sorted queries:
  nearest   find_nearby_cell_kdtree      (kd-tree search only)
  cell      iu_get_cell                  (search + walk)
  interp    iu_interpolate_at            (search + walk + interpolation)        = the bench kernel
  warm      iu_interpolate_at with hints (walk from the hinted cell's centre + interpolation)
  icell     iu_interpolate_at_icell      (interpolation only)
usage: python tools/phase_bench.py [--mesh-n 119] [--queries 67108864] [--reps 3]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from interpolate_unstructured_b200 import api  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mesh-n", type=int, default=119)
    ap.add_argument("--queries", type=float, default=1 << 26)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    wl = bench.Workload("C3", a.mesh_n).mesh()
    pts, cells, nb, pd, qlo, qhi = wl.pts, wl.cells, wl.nb, wl.pd, wl.qlo, wl.qhi
    N_VARS, I_VARS = wl.n_vars, wl.i_vars
    g = api.Grid.from_arrays("tetra", pts, cells, nb, pd, device=0)
    g.set_option(api.OPT_TIME_KERNELS, 1)
    nq = int(a.queries)
    r = torch.empty((nq, 3), dtype=torch.float64, device=dev)
    g.generate_queries_dev(r, first=0, seed=20261104, lo=qlo, hi=qhi)
    ic = torch.zeros(nq, dtype=torch.int32, device=dev)
    idx = torch.zeros(nq, dtype=torch.int32, device=dev)
    out = torch.empty((nq, N_VARS), dtype=torch.float64, device=dev)
    hint = torch.zeros(nq, dtype=torch.int32, device=dev)
    g.iu_get_cell_dev(r, hint)

    def nearest():
        g.find_nearby_cell_kdtree_dev(r, idx)

    def cell():
        ic.zero_()
        g.iu_get_cell_dev(r, ic)

    def interp():
        ic.zero_()
        g.iu_interpolate_at_dev(r, I_VARS, out, ic)

    def warm():
        ic.copy_(hint)
        g.iu_interpolate_at_dev(r, I_VARS, out, ic)

    def icell():
        g.iu_interpolate_at_icell_dev(r, I_VARS, out, hint)

    # the same queries physically in Morton order (24-bit key like the library's), processed without a permutation
    def spread(v):
        v = v & 0x3ff
        v = (v | (v << 16)) & 0x030000ff
        v = (v | (v << 8)) & 0x0300f00f
        v = (v | (v << 4)) & 0x030c30c3
        v = (v | (v << 2)) & 0x09249249
        return v
    lo = torch.tensor(pts.min(0), device=dev)
    sc = 256.0 / torch.tensor(pts.max(0) - pts.min(0), device=dev)
    q = ((r - lo) * sc).clamp_(0, 255).to(torch.int64)
    key = spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)
    order = torch.argsort(key)
    del q, key
    rs = r[order].contiguous()
    hs = hint[order].contiguous()
    del order

    def presorted(fn2):
        def f():
            g.set_option(api.OPT_SORT_QUERIES, 3)
            fn2()
            g.set_option(api.OPT_SORT_QUERIES, 2)
        return f

    def nearest_ps():
        g.find_nearby_cell_kdtree_dev(rs, idx)

    def interp_ps():
        ic.zero_()
        g.iu_interpolate_at_dev(rs, I_VARS, out, ic)

    def warm_ps():
        ic.copy_(hs)
        g.iu_interpolate_at_dev(rs, I_VARS, out, ic)

    def icell_ps():
        g.iu_interpolate_at_icell_dev(rs, I_VARS, out, hs)

    # upper bound of what grouping the queries of a warp by CELL could give: queries physically ordered by the cell that
    # contains them (original cell id: cube-major for the Kuhn mesh), processed without a permutation
    order_c = torch.argsort(hint.long() * 64 + torch.randint(0, 64, (nq,), device=dev))
    rc = r[order_c].contiguous()
    del order_c

    def interp_bycell():
        ic.zero_()
        g.iu_interpolate_at_dev(rc, I_VARS, out, ic)

    res = {}
    for name, fn in [("interp_grouped_by_cell", presorted(interp_bycell)), ("nearest_presorted", presorted(nearest_ps)), ("interp_presorted", presorted(interp_ps)),
                     ("warm_presorted", presorted(warm_ps)), ("icell_presorted", presorted(icell_ps)),("nearest", nearest), ("cell", cell), ("interp", interp), ("warm", warm), ("icell", icell)]:
        fn()
        g.check_dev()
        torch.cuda.synchronize()
        g.reset_stats()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        st = g.stats()
        n_k = max(1, st["query_kernel_launches"])
        res[name] = {"kernel_ms": st["query_kernel_ns"] / n_k * 1e-6, "sort_ms": st["sort_ns"] / n_k * 1e-6,
                     "wall_ms": e0.elapsed_time(e1) / a.reps, "launches": n_k}
        print(f"{name:18s} kernel {res[name]['kernel_ms']:8.3f} ms  sort {res[name]['sort_ms']:7.3f} ms  wall {res[name]['wall_ms']:8.3f} ms", flush=True)
    print(json.dumps({"queries": nq, "phases": res}))


if __name__ == "__main__":
    main()
