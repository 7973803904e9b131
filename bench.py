#!/usr/bin/env python3
"""Headline benchmark: cold point location + interpolation on a synthetic 3-D tetrahedral mesh.

Workload (BASELINE.json configs[2], SURVEY.md 8d "C3"): Kuhn 6-tet split of 119^3 cubes on the unit
cube (10 110 954 tets, 1 728 000 points, interior vertices jittered by 0.1 h), 4 point-data
variables, queries uniform in the bbox shrunk by 10 % per side (benchmark.f90:28-38), every query
cold (i_cell = 0: kd-tree nearest centre -> ray walk -> interpolation = iu_interpolate_at).
One step = one batch of --queries queries per GPU (default 1e9), resident in HBM before the timed
region.  The mesh is replicated per GPU and queries are sharded by rank ("weak" scaling, no
collective on the data path).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--queries Q] [--impl reference]

Prints ONE JSON line (rank 0).  `value` = device-resident throughput over all ranks; `e2e` = the
same metric through the host-buffer C ABI call (H2D and D2H inside the timed region); `roofline`
= the hot kernel (k_query) against the measured HBM copy bandwidth; `cpu_baseline` = the CPU
restatement of the reference path (oracle/, OpenMP over queries) on a bounded sample.
`--impl reference` times only that CPU restatement (the reference itself is Fortran + a missing
kdtree2 submodule and cannot be built on this image).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "queries/s (kd-tree+walk+interp), 3D tet mesh"
N_VARS = 4
I_VARS = [1, 2, 3, 4]


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--mesh-n", type=int, default=119, help="cubes per edge of the Kuhn mesh (119 -> 10.1M tets)")
    p.add_argument("--queries", type=float, default=1e9, help="queries per step per GPU (device-resident leg)")
    p.add_argument("--e2e-queries", type=float, default=1e8, help="queries per step per GPU (host-buffer leg)")
    p.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    p.add_argument("--sort", type=int, default=-1, help="override IU_B200_OPT_SORT_QUERIES (0 never, 1 always, 2 auto)")
    p.add_argument("--dynamic", type=int, default=-1, help="override IU_B200_OPT_DYNAMIC_TILES (0 static grid-stride, 1 ticket tiles)")
    p.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    p.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg")
    return p.parse_args()


def make_mesh(n):
    from interpolate_unstructured_b200 import api, meshgen
    cache = f"/tmp/iu_b200_bench_mesh_{n}_{N_VARS}.npz"      # same process tree re-runs (profiling) skip the generation
    if os.path.exists(cache):
        z = np.load(cache)
        pts, cells, nb, pd = z["pts"], z["cells"], z["nb"], z["pd"]
    else:
        pts, cells = meshgen.kuhn_tet_mesh(n, jitter=0.10, seed=1234)
        nb = api.build_cell_neighbors(cells, pts, "tetra")
        pd = meshgen.smooth_fields(pts, N_VARS)
        if int(os.environ.get("RANK", "0")) == 0:
            try:
                np.savez(cache + ".tmp.npz", pts=pts, cells=cells, nb=nb, pd=pd)
                os.replace(cache + ".tmp.npz", cache)
            except OSError:
                pass
    lo, hi = pts.min(0), pts.max(0)
    d = hi - lo
    return pts, cells, nb, pd, lo + 0.1 * d, hi - 0.1 * d


def algorithmic_bytes(nq, n_points, n_cells, ppc=4, nv=N_VARS):
    """queries in, cells + values out, raw mesh once per launch."""
    per_query = 24 + 4 + 8 * nv
    m_raw = 24 * n_points + 4 * ppc * n_cells + 4 * ppc * n_cells + 8 * nv * n_points
    return nq * per_query + m_raw, per_query, m_raw


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            self.thread.join(timeout=5)

    def summary(self):
        sm, mx, reasons, power = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power)}


def cpu_reference_leg(mesh, target_seconds, steps=1, warmup=0):
    """CPU restatement of benchmark.f90's cold pass (oracle/iu_oracle.cpp, kdtree2-style tree, OpenMP
    over queries) on a bounded sample of the same workload.  Returns (queries/s, cores, sample text, seconds/step)."""
    from interpolate_unstructured_b200 import meshgen
    from oracle import oracle
    from oracle.oracle import OracleGrid
    pts, cells, nb, pd, qlo, qhi = mesh
    cores = int(oracle.lib().iu_oracle_max_threads())
    g = OracleGrid.from_arrays("tetra", pts, cells, nb, pd)
    # calibrate on a small sample, then size the timed sample for ~target_seconds of wall time
    r = meshgen.uniform_queries(200_000, qlo, qhi, seed=4242)
    ic = np.zeros(len(r), dtype=np.int32)
    t, _ = g.benchmark_pass(r, I_VARS, ic, nthreads=cores)
    rate = len(r) / max(t, 1e-9)
    n = int(min(max(rate * target_seconds, 200_000), 50_000_000))
    r = meshgen.uniform_queries(n, qlo, qhi, seed=777)
    times = []
    for s in range(warmup + steps):
        ic = np.zeros(n, dtype=np.int32)
        t, res = g.benchmark_pass(r, I_VARS, ic, nthreads=cores)
        if s >= warmup:
            times.append(t)
    assert (ic >= 1).all()
    assert np.abs(res[:, 0] - (r.sum(1) + 1)).max() < 1e-12
    tm = float(np.mean(times))
    sample = (f"{n} cold queries of the same workload ({len(cells)} tets, {N_VARS} vars), C++ restatement of "
              f"benchmark.f90:46-55 with OpenMP over queries, {cores} threads, {tm:.2f} s per pass")
    return n / tm, cores, sample, tm


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    mesh = make_mesh(args.mesh_n)
    value, cores, sample, tm = cpu_reference_leg(mesh, args.cpu_seconds, steps=max(1, args.steps), warmup=min(args.warmup, 1))
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": tm * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"C3: Kuhn tet mesh {len(mesh[1])} tets / {len(mesh[0])} points, cold iu_interpolate_at, {N_VARS} vars",
                   "note": "reference (Fortran + missing kdtree2 submodule) is unbuildable on this image; this is its C++ restatement on host cores"},
        "cpu_baseline": {"value": value, "unit": "queries/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


def bind_to_gpu_numa_node(local):
    """Pin this rank (and therefore its first-touch pinned host buffers) to the CPUs of the NUMA node its GPU hangs
    off: the host-buffer leg moves 64 B per query over PCIe and should not cross the socket interconnect."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return node
    except (OSError, ValueError, AttributeError):
        return None


def run_b200(args):
    import torch
    import torch.distributed as dist
    from interpolate_unstructured_b200 import api

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libiu_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_node = bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    t0 = time.time()
    mesh = make_mesh(args.mesh_n)
    pts, cells, nb, pd, qlo, qhi = mesh
    g = api.Grid.from_arrays("tetra", pts, cells, nb, pd, device=local)
    setup_s = time.time() - t0
    if args.sort >= 0:
        g.set_option(api.OPT_SORT_QUERIES, args.sort)
    if args.dynamic >= 0:
        g.set_option(api.OPT_DYNAMIC_TILES, args.dynamic)
    g.set_option(api.OPT_TIME_KERNELS, 1)

    nq = int(args.queries)
    r = torch.empty((nq, 3), dtype=torch.float64, device=dev)
    g.generate_queries_dev(r, first=rank * nq, seed=20261104, lo=qlo, hi=qhi)
    ic = torch.zeros(nq, dtype=torch.int32, device=dev)
    out = torch.empty((nq, N_VARS), dtype=torch.float64, device=dev)

    def step():
        ic.zero_()                                   # cold: i_cell = 0 for every query (benchmark.f90:41)
        g.iu_interpolate_at_dev(r, I_VARS, out, ic)

    for _ in range(args.warmup):
        step()
    g.check_dev()
    barrier()
    g.reset_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    st = g.stats()
    value = world * nq * args.steps / (ms * 1e-3)

    # correctness of what was timed (size-independent properties; the oracle parity lives in tests/)
    found = ic >= 1
    n_found = int(found.sum().item())
    lin_err = float((out[:, 0] - (r.sum(1) + 1.0)).abs()[found].max().item())
    assert n_found == nq, f"{nq - n_found} queries not located"
    assert lin_err < 1e-12, f"linear field not reproduced: {lin_err}"
    assert st["walk_limit_hits"] == 0

    # roofline of the dominant kernel: algorithmic bytes per launch / mean launch duration
    n_k = max(1, st["query_kernel_launches"])
    k_ms = st["query_kernel_ns"] / n_k * 1e-6
    alg_bytes, per_query, m_raw = algorithmic_bytes(nq * args.steps / n_k, len(pts), len(cells))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    traffic, ncu_counters = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj.get("k_query_dram_bytes_per_launch")
        ncu_counters = dict(tj.get("counters") or {}, source="profiles/" + tj.get("source", ""))   # what actually limits k_query (ncu, last capture)
    except (OSError, ValueError):
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "kernel": "k_query<TETRA, INTERP, 4>", "kernel_ms_per_launch": k_ms, "launches": n_k,
                "algorithmic_bytes_per_launch": alg_bytes, "bytes_per_query": per_query, "mesh_raw_bytes": m_raw,
                "peak_source": "MEASURED_PEAKS.json (measured)" if peaks else "fallback 6650 GB/s",
                "kernel_share_of_step": st["query_kernel_ns"] * 1e-6 / (e0.elapsed_time(e1)),
                "sort_share_of_step": st["sort_ns"] * 1e-6 / (e0.elapsed_time(e1)), "ncu_counters": ncu_counters}
    gpu_launches = st["kernel_launches"] + args.steps      # + the i_cell memset kernel of each step

    # end to end through the host-buffer C ABI (pinned host memory; H2D + kernels + D2H per step)
    e2e = None
    if not args.no_e2e:
        del out, ic
        ne = int(args.e2e_queries)
        r_h = api.pinned_empty((ne, 3), np.float64)
        ic_h = api.pinned_empty((ne,), np.int32)
        out_h = api.pinned_empty((ne, N_VARS), np.float64)
        torch.cuda.synchronize()
        r_h[:] = r[:ne].cpu().numpy() if ne <= nq else np.resize(r.cpu().numpy(), (ne, 3))
        for _ in range(max(1, min(args.warmup, 2))):
            ic_h[:] = 0
            g.iu_interpolate_at(r_h, I_VARS, var_at_r=out_h, i_cell=ic_h)
        barrier()
        tt = 0.0
        for _ in range(args.steps):
            ic_h[:] = 0
            barrier()
            t1 = time.perf_counter()
            g.iu_interpolate_at(r_h, I_VARS, var_at_r=out_h, i_cell=ic_h)   # synchronous: returns after the last D2H
            tt += time.perf_counter() - t1
        tt = max_over_ranks(tt)
        assert (ic_h >= 1).all() and np.abs(out_h[:, 0] - (r_h.sum(1) + 1)).max() < 1e-12
        e2e = {"value": world * ne * args.steps / tt, "unit": "queries/s", "h2d_bytes_per_step": ne * 28, "d2h_bytes_per_step": ne * (4 + 8 * N_VARS),
               "queries_per_step": ne, "ms_per_step": tt / args.steps * 1e3, "host_memory": "pinned (iu_b200_host_alloc)",
               "numa_node_rank0": numa_node}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, cores, sample, _ = cpu_reference_leg(mesh, args.cpu_seconds)
        cpu = {"value": v, "unit": "queries/s", "cores": cores, "kind": "port", "sample": sample}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "queries/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"C3: Kuhn tet mesh {len(cells)} tets / {len(pts)} points (jitter 0.1 h), cold iu_interpolate_at, "
                                   f"{N_VARS} vars, {nq} uniform queries per GPU per step, mesh replicated per GPU",
                       "queries_per_step_per_gpu": nq, "l2_policy": "inputs larger than L2 (24 B x queries per step >> 126 MB)",
                       "sort_queries": args.sort if args.sort >= 0 else 2, "dynamic_tiles": args.dynamic if args.dynamic >= 0 else 1, "setup_seconds": setup_s,
                       "mesh_device_bytes": st["mesh_bytes"], "tree_depth": st["tree_depth"],
                       "nn_voxel_table": {k: st[k] for k in ("nn_voxels", "nn_voxels_tree", "nn_voxel_entries", "nn_tree_queries", "nn_voxel_build_us")}},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": gpu_launches, "clocks": clocks.summary(),
            "checks": {"located": n_found, "linear_field_max_abs_err": lin_err},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
