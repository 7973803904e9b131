#!/usr/bin/env python3
"""One process, one handle, several GPUs (iu_b200_grid_replicate): set-up time, device-resident throughput through the
replicas' *_dev calls, and the host-buffer batch call that the library itself splits into one slab per GPU.

    python tools/multi_bench.py [--workload C3|C5] [--gpus N] [--queries Q per GPU] [--e2e-queries Q total] [--steps K]

Compare `value` with bench.py's torchrun number at the same N (within 2 %)."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from interpolate_unstructured_b200 import api  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C3", choices=["C3", "C5"])
    ap.add_argument("--gpus", type=int, default=0)
    ap.add_argument("--queries", type=float, default=0)
    ap.add_argument("--e2e-queries", type=float, default=0)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--mesh-n", type=int, default=0)
    a = ap.parse_args()
    n_gpu = a.gpus or torch.cuda.device_count()
    wl = bench.Workload(a.workload, a.mesh_n).mesh()
    t0 = time.time()
    g = api.Grid.from_arrays(wl.cell_type, wl.pts, wl.cells, wl.nb, wl.pd, cell_data=wl.cd, device=0)
    t_build = time.time() - t0
    t0 = time.time()
    g.replicate(list(range(n_gpu)))
    t_rep = time.time() - t0
    assert g.device_count == n_gpu
    nq = int(a.queries or (wl.queries if a.workload == "C3" else 250_000_000))
    nv = wl.n_vars
    reps, bufs = [], []
    for d in range(n_gpu):
        with torch.cuda.device(d):
            dev = torch.device("cuda", d)
            gd = g.replica(d)
            r = torch.empty((nq, 3), dtype=torch.float64, device=dev)
            gd.generate_queries_dev(r, first=d * nq, seed=bench.QUERY_SEED, lo=wl.qlo, hi=wl.qhi)
            ic = torch.zeros(nq, dtype=torch.int32, device=dev)
            out = torch.empty((nq, nv), dtype=torch.float64, device=dev)
            cres = torch.empty(nq, dtype=torch.float64, device=dev) if wl.cell_var else None
            reps.append(gd); bufs.append((r, ic, out, cres))

    def step():
        for d in range(n_gpu):
            with torch.cuda.device(d):
                r, ic, out, cres = bufs[d]
                ic.zero_()
                if wl.cell_var:
                    reps[d].iu_interpolate_and_cell_scalar_at_dev(r, wl.i_vars, wl.cell_var, out, cres, ic)
                else:
                    reps[d].iu_interpolate_at_dev(r, wl.i_vars, out, ic)

    def sync():
        for d in range(n_gpu):
            torch.cuda.synchronize(d)

    for _ in range(a.warmup):
        step()
    sync()
    ev = []
    for d in range(n_gpu):
        with torch.cuda.device(d):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ev.append((e0, e1))
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step()
    for d in range(n_gpu):
        with torch.cuda.device(d):
            ev[d][1].record()
    sync()
    wall = time.perf_counter() - t0
    ms = max(e0.elapsed_time(e1) for e0, e1 in ev)          # device time, max over the GPUs
    for d in range(n_gpu):
        with torch.cuda.device(d):
            reps[d].check_dev()
            r, ic, out, cres = bufs[d]
            assert int((ic >= 1).sum()) == nq
            assert float((out[:, 0] - (r.sum(1) + 1)).abs().max()) < 1e-12
    value = n_gpu * nq * a.steps / (ms * 1e-3)
    line = {"what": f"one process, one handle, {n_gpu} GPUs: {wl.text()}", "n_gpus": n_gpu, "queries_per_step_per_gpu": nq, "steps": a.steps,
            "value": value, "unit": "queries/s", "ms_per_step_device_max": ms / a.steps, "ms_per_step_wall": wall / a.steps * 1e3,
            "setup": {"build_on_gpu0_s": t_build, "replicate_to_the_others_s": t_rep, "replicate_us_library": g.stats()["replicate_us"],
                      "mesh_device_bytes_per_gpu": reps[0].stats()["mesh_bytes"] if n_gpu == 1 else reps[1].stats()["mesh_bytes"]}}
    # host-buffer leg: ONE call on the handle, the library splits it over the GPUs
    del bufs
    ne = int(a.e2e_queries or 100_000_000 * n_gpu // (1 if a.workload == "C3" else 2))
    r_h = api.pinned_empty((ne, 3), np.float64)
    ic_h = api.pinned_empty((ne,), np.int32)
    out_h = api.pinned_empty((ne, nv), np.float64)
    cres_h = api.pinned_empty((ne,), np.float64) if wl.cell_var else None
    r_h[:] = wl.host_queries(min(ne, 20_000_000))[np.arange(ne) % min(ne, 20_000_000)]
    ts = []
    for k in range(a.steps + 1):
        ic_h[:] = 0
        t0 = time.perf_counter()
        if wl.cell_var:
            g.iu_interpolate_and_cell_scalar_at(r_h, wl.i_vars, wl.cell_var, var_at_r=out_h, cell_res=cres_h, i_cell=ic_h)
        else:
            g.iu_interpolate_at(r_h, wl.i_vars, var_at_r=out_h, i_cell=ic_h)
        ts.append(time.perf_counter() - t0)
    assert (ic_h >= 1).all() and np.abs(out_h[:, 0] - (r_h.sum(1) + 1)).max() < 1e-12
    t = float(np.mean(ts[1:]))
    line["e2e"] = {"value": ne / t, "unit": "queries/s", "queries_per_call": ne, "ms_per_call": t * 1e3,
                   "h2d_bytes_per_call": ne * 28, "d2h_bytes_per_call": ne * (4 + 8 * nv + (8 if wl.cell_var else 0))}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
