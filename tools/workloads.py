#!/usr/bin/env python3
"""The other BASELINE.json configurations, measured like bench.py measures C3 (one JSON line each;
bench.py stays the contract line for C3, these are reported in profiles/README.md).

  C2  synthetic 2-D quad mesh, 2000 x 2000 cells, 100M uniform cold queries, one scalar
  C4  warm start on the C3 mesh: 100M tracers located cold, displaced by alpha*h in a random
      direction (alpha = 0.1, 0.5, 1.0), then (a) iu_interpolate_at with the previous cells as hints
      (walk from the old cell centre, m_interp_unstructured.f90:424-432) and
      (b) iu_get_cell_through_neighbors from the old position (:667-725), 3 variables
  C5  Kuhn tet mesh 203^3 (50.2M tets), 8 point variables + one cell-data variable
      (iu_interpolate_at followed by iu_get_cell_scalar_at with the found cells as hints),
      --queries per GPU (default 5e8)

usage: python tools/workloads.py C2|C4|C5 [--queries Q] [--steps K] [--warmup W]
Times on the device with CUDA events; queries are generated on the device and are larger than L2."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from interpolate_unstructured_b200 import api, meshgen  # noqa: E402


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
    except (OSError, KeyError, ValueError):
        return 6650.0


def line(workload, nq, ms, per_query, m_raw, extra=None):
    alg = nq * per_query + m_raw
    d = {"workload": workload, "queries": nq, "ms_per_step": ms, "value": nq / (ms * 1e-3), "unit": "queries/s",
         "algorithmic_bytes_per_query": per_query, "mesh_raw_bytes": m_raw,
         "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "peak": peak_gbs(), "unit": "GB/s",
                      "frac": alg / (ms * 1e-3) / 1e9 / peak_gbs(), "timed": "whole step (sort + k_query)"}}
    if extra:
        d.update(extra)
    print(json.dumps(d), flush=True)


def run_c2(a):
    dev = torch.device("cuda", 0)
    n = 2000
    pts, cells = meshgen.quad_mesh(n, jitter=0.2, seed=1234)
    nb = api.build_cell_neighbors(cells, pts, "quad")
    pd = meshgen.smooth_fields(pts, 2)            # column 1 linear (self-check), column 2 smooth non-linear
    g = api.Grid.from_arrays("quad", pts, cells, nb, pd, device=0)
    nq = int(a.queries or 1e8)
    r = torch.empty((nq, 3), dtype=torch.float64, device=dev)
    g.generate_queries_dev(r, first=0, seed=2, lo=[0.1, 0.1, 0.0], hi=[0.9, 0.9, 0.0])
    ic = torch.zeros(nq, dtype=torch.int32, device=dev)
    out = torch.empty((nq, 1), dtype=torch.float64, device=dev)

    def step(var=2):
        ic.zero_()
        g.iu_interpolate_at_dev(r, [var], out, ic)

    step(1)
    g.check_dev()
    assert int((ic >= 1).sum()) == nq
    # the reference's inverse bilinear map (:591-641) loses digits by cancellation on nearly
    # parallelogram-shaped quads: errors up to ~1e-6 are the REFERENCE arithmetic (oracle agrees bit for bit)
    e = (out[:, 0] - (r.sum(1) + 1)).abs()
    err = float(e.max())
    assert err < 1e-5 and float(e.median()) < 1e-13, err
    ms = timed(step, a.steps, a.warmup)
    m_raw = 24 * len(pts) + 8 * 4 * len(cells) + 8 * len(pts)
    line(f"C2: {n}x{n} jittered quads ({len(cells)} cells / {len(pts)} points), {nq} cold queries, iu_interpolate_scalar_at semantics (1 variable)",
         nq, ms, 24 + 4 + 8, m_raw, {"checks": {"located": nq, "linear_field_max_abs_err": err}})


def run_c4(a):
    dev = torch.device("cuda", 0)
    n = 119
    pts, cells = meshgen.kuhn_tet_mesh(n, jitter=0.10, seed=1234)
    nb = api.build_cell_neighbors(cells, pts, "tetra")
    pd = meshgen.smooth_fields(pts, 3)
    g = api.Grid.from_arrays("tetra", pts, cells, nb, pd, device=0)
    if os.environ.get("IU_WL_SORT"):            # A/B of IU_B200_OPT_SORT_QUERIES (default 2 = auto)
        g.set_option(api.OPT_SORT_QUERIES, int(os.environ["IU_WL_SORT"]))
    nq = int(a.queries or 1e8)
    h = 1.0 / n
    r0 = torch.empty((nq, 3), dtype=torch.float64, device=dev)
    g.generate_queries_dev(r0, first=0, seed=4, lo=[0.1] * 3, hi=[0.9] * 3)
    ic0 = torch.zeros(nq, dtype=torch.int32, device=dev)
    g.iu_get_cell_dev(r0, ic0)
    g.check_dev()
    assert int((ic0 >= 1).sum()) == nq
    gen = torch.Generator(device=dev)
    gen.manual_seed(44)
    d = torch.randn((nq, 3), dtype=torch.float64, device=dev, generator=gen)
    d /= d.norm(dim=1, keepdim=True)
    ic = torch.empty_like(ic0)
    out = torch.empty((nq, 3), dtype=torch.float64, device=dev)
    ic1 = torch.empty_like(ic0)
    r_p = torch.empty((nq, 3), dtype=torch.float64, device=dev)
    n_steps = torch.empty_like(ic0)
    status = torch.empty_like(ic0)
    m_raw = 24 * len(pts) + 8 * 4 * len(cells) + 8 * 3 * len(pts)
    for alpha in (0.1, 0.5, 1.0):
        r1 = r0 + alpha * h * d

        def hinted():
            ic.copy_(ic0)
            g.iu_interpolate_at_dev(r1, [1, 2, 3], out, ic)

        def through():
            g.iu_get_cell_through_neighbors_dev(2 ** 20, r0, r1, ic0, ic1, r_p, n_steps, status)

        hinted(); through()
        g.check_dev()
        # both routes end in the same cell (queries on shared faces aside: none expected for random fp64 points)
        same = int((ic == ic1).sum())
        err = float((out[:, 0] - (r1.sum(1) + 1)).abs()[ic >= 1].max())
        assert err < 1e-12, err
        ms_h = timed(hinted, a.steps, a.warmup)
        ms_t = timed(through, a.steps, a.warmup)
        extra = {"alpha": alpha, "mean_walk_steps_from_old_position": float(n_steps.double().mean()),
                 "checks": {"found": int((ic >= 1).sum()), "same_cell_both_routes": same, "linear_field_max_abs_err": err}}
        line(f"C4: warm iu_interpolate_at (hint = previous cell, walk from its centre), 3 vars, displacement {alpha} h, C3 mesh", nq, ms_h, 24 + 8 + 24, m_raw, extra)
        line(f"C4: iu_get_cell_through_neighbors from the previous position, displacement {alpha} h, C3 mesh", nq, ms_t, 48 + 4 + 4 + 24 + 4 + 4,
             24 * len(pts) + 8 * 4 * len(cells), extra)


def run_c5(a):
    dev = torch.device("cuda", 0)
    n = a.mesh_n or 203
    t0 = time.time()
    pts, cells = meshgen.kuhn_tet_mesh(n, jitter=0.10, seed=1234)
    nb = api.build_cell_neighbors(cells, pts, "tetra")
    pd = meshgen.smooth_fields(pts, 8)
    cd = np.arange(1, len(cells) + 1, dtype=np.float64)
    g = api.Grid.from_arrays("tetra", pts, cells, nb, pd, cell_data=cd, device=0)
    setup = time.time() - t0
    nq = int(a.queries or 5e8)
    r = torch.empty((nq, 3), dtype=torch.float64, device=dev)
    g.generate_queries_dev(r, first=0, seed=5, lo=[0.1] * 3, hi=[0.9] * 3)
    ic = torch.zeros(nq, dtype=torch.int32, device=dev)
    out = torch.empty((nq, 8), dtype=torch.float64, device=dev)
    res = torch.empty(nq, dtype=torch.float64, device=dev)

    def step():
        ic.zero_()
        g.iu_interpolate_at_dev(r, list(range(1, 9)), out, ic)
        g.iu_get_cell_scalar_at_dev(r, 1, ic, res)      # second call of the config, the found cells as hints

    step()
    g.check_dev()
    assert int((ic >= 1).sum()) == nq
    err = float((out[:, 0] - (r.sum(1) + 1)).abs().max())
    assert err < 1e-12, err
    assert torch.equal(res.to(torch.int32), ic)
    ms = timed(step, a.steps, a.warmup)
    m_raw = 24 * len(pts) + 8 * 4 * len(cells) + 8 * 8 * len(pts) + 8 * len(cells)
    line(f"C5 (one GPU's share): Kuhn {n}^3 = {len(cells)} tets / {len(pts)} points, 8 point vars + 1 cell var, {nq} cold queries per GPU",
         nq, ms, 24 + 4 + 64 + 8, m_raw,
         {"setup_seconds": setup, "mesh_device_bytes": g.stats()["mesh_bytes"], "checks": {"located": nq, "linear_field_max_abs_err": err}})


def run_trace(a):
    """Batched field-line tracer (iu_integrate_along_field, test_trace_field style) on the C3 mesh: n lines from random
    seeds through a swirl + drift field, arc-length integrand, host-buffer API (y crosses PCIe both ways)."""
    n = 119
    pts, cells = meshgen.kuhn_tet_mesh(n, jitter=0.10, seed=1234)
    nb = api.build_cell_neighbors(cells, pts, "tetra")
    c = pts - 0.5
    field = np.stack([-c[:, 1] + 0.3 * c[:, 0], c[:, 0] + 0.3 * c[:, 1], 0.2 + 0.0 * c[:, 2]], axis=1)
    g = api.Grid.from_arrays("tetra", pts, cells, nb, np.ascontiguousarray(field), device=0)
    nl = int(a.queries or 5e5)
    max_steps = 64
    rng = np.random.default_rng(11)
    seeds = rng.uniform(0.2, 0.8, size=(nl, 3))
    h = 1.0 / n

    def run():
        y = np.full((nl, max_steps, 4), np.nan)
        y[:, 0, :3] = seeds
        y[:, 0, 3] = 0.0
        t0 = time.perf_counter()
        out = g.iu_integrate_along_field(y, [1, 2, 3], 1e-3 * h, 2.0 * h, 1e-4, 1e-6, nvar=1, integrand=1)
        return time.perf_counter() - t0, out

    run()
    ts = []
    for _ in range(max(1, a.steps)):
        t, (y, yf, ns, bm) = run()
        ts.append(t)
    t = float(np.mean(ts))
    steps = int(np.minimum(ns, max_steps).sum())
    # the same lines with y resident in HBM (device API): seeds cold-located + k_trace
    dev = torch.device("cuda", 0)
    y_d = torch.full((nl, max_steps, 4), float("nan"), dtype=torch.float64, device=dev)
    ns_d = torch.zeros(nl, dtype=torch.int32, device=dev)
    seeds_d = torch.from_numpy(seeds).to(dev)

    def run_dev():
        y_d[:, 0, :3] = seeds_d
        y_d[:, 0, 3] = 0.0
        g.iu_integrate_along_field_dev(y_d, [1, 2, 3], 1e-3 * h, 2.0 * h, 1e-4, 1e-6, ns_d, nvar=1, integrand=1)

    ms_dev = timed(run_dev, max(1, a.steps), 1)
    g.check_dev()
    assert torch.equal(ns_d.cpu(), torch.from_numpy(ns))
    print(json.dumps({"workload": f"trace: {nl} field lines on the C3 mesh ({len(cells)} tets), RK3(2), max_steps {max_steps}, max_dx 2 h, rtol 1e-4, "
                                  "arc-length integrand, host-buffer API", "lines": nl, "seconds": t, "lines_per_s": nl / t,
                      "stored_steps": steps, "stored_steps_per_s": steps / t,
                      "device_ms": ms_dev, "device_lines_per_s": nl / (ms_dev * 1e-3), "device_stored_steps_per_s": steps / (ms_dev * 1e-3), "mean_steps_per_line": steps / nl,
                      "lines_that_left_the_domain": int((bm == -1).sum())}), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("workload", choices=["C2", "C4", "C5", "trace"])
    ap.add_argument("--queries", type=float, default=0)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--mesh-n", type=int, default=0)
    a = ap.parse_args()
    {"C2": run_c2, "C4": run_c4, "C5": run_c5, "trace": run_trace}[a.workload](a)
