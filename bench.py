#!/usr/bin/env python3
"""Benchmark of the hot path (nearest cell centre -> ray walk -> interpolation) on the BASELINE.json configurations.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C1|C2|C3|C4|C5] [--queries Q] [--impl reference]

Default workload C3 = BASELINE.json configs[2]: Kuhn 6-tet split of 119^3 cubes on the unit cube
(10 110 954 tets, 1 728 000 points, interior vertices jittered by 0.1 h), 4 point-data variables, queries uniform in
the bbox shrunk by 10 % per side (benchmark.f90:28-38), every query cold (i_cell = 0: nearest centre -> ray walk ->
interpolation = iu_interpolate_at).  One step = one batch of --queries queries per GPU (default 1e9), resident in HBM
before the timed region.  The mesh is replicated per GPU and queries are sharded by rank ("weak" scaling, no collective
on the data path).  The other workloads: C1 benchmark.f90 itself (triangle fixture, 1M queries, one scalar), C2 4M
jittered quads / 1e8 queries / one scalar, C4 1e8 tracers on the C3 mesh displaced by --alpha cell sizes with the
previous cell as hint (3 variables), C5 50.2M tets / 8 variables + one cell-data scalar in one fused pass / 5e8 queries
per GPU.

Prints ONE JSON line (rank 0).  `value` = device-resident throughput over all ranks; `e2e` = the same metric through the
host-buffer C ABI call (H2D and D2H inside the timed region); `roofline` = the hot kernel (k_query) against the measured
HBM copy bandwidth; `cpu_baseline` = the CPU restatement of the reference path (oracle/) on a bounded sample of the same
query stream: both passes of benchmark.f90:46-69 (cold, then warm with the cells just found), with ONE thread (the
reference is serial) and with all host cores (OpenMP over queries); `checks.oracle_*` compare the GPU results of the
timed stream with that oracle pass query for query.  `--impl reference` times only the CPU restatement (the reference
itself is Fortran + a missing kdtree2 submodule and cannot be built on this image) and never loads libiu_b200.so.
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

QUERY_SEED = 20261104


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--workload", default="C3", choices=["C1", "C2", "C3", "C4", "C5"])
    p.add_argument("--mesh-n", type=int, default=0, help="override the mesh size (cubes / squares per edge)")
    p.add_argument("--queries", type=float, default=0, help="queries per step per GPU (device-resident leg); 0 = the workload's own")
    p.add_argument("--e2e-queries", type=float, default=0, help="queries per step per GPU (host-buffer leg); 0 = the workload's own")
    p.add_argument("--alpha", type=float, default=0.5, help="C4: displacement of the tracers in mean cell sizes")
    p.add_argument("--cpu-seconds", type=float, default=8.0, help="target wall time of one all-cores CPU pass (cpu_baseline)")
    p.add_argument("--sort", type=int, default=-1, help="override IU_B200_OPT_SORT_QUERIES (0 never, 1 always, 2 auto)")
    p.add_argument("--dynamic", type=int, default=-1, help="override IU_B200_OPT_DYNAMIC_TILES (0 static grid-stride, 1 ticket tiles)")
    p.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (and the oracle check)")
    p.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg")
    return p.parse_args()


# ------------------------------------------------------------------------------------------------
# workloads
# ------------------------------------------------------------------------------------------------
class Workload:
    """Mesh, query box, requested variables and entry point of one BASELINE.json configuration.  Everything here is
    numpy + the oracle's neighbour builder: the reference arm builds its inputs without the product library."""

    def __init__(self, name, mesh_n=0, alpha=0.5):
        self.name = name
        self.alpha = alpha
        self.cell_var = 0          # > 0: fused iu_interpolate_at + iu_get_cell_scalar_at (C5)
        self.warm = False          # True: the timed call gets the previous cells as hints (C4)
        if name == "C1":
            self.cell_type, self.n, self.i_vars = "triangle", 0, [1]
            self.queries, self.e2e_queries = 1_000_000, 1_000_000
        elif name == "C2":
            self.cell_type, self.n, self.i_vars = "quad", mesh_n or 2000, [2]
            self.queries, self.e2e_queries = 100_000_000, 100_000_000
        elif name == "C3":
            self.cell_type, self.n, self.i_vars = "tetra", mesh_n or 119, [1, 2, 3, 4]
            self.queries, self.e2e_queries = 1_000_000_000, 100_000_000
        elif name == "C4":
            self.cell_type, self.n, self.i_vars = "tetra", mesh_n or 119, [1, 2, 3]
            self.queries, self.e2e_queries = 100_000_000, 100_000_000
            self.warm = True
        else:
            self.cell_type, self.n, self.i_vars = "tetra", mesh_n or 203, [1, 2, 3, 4, 5, 6, 7, 8]
            self.queries, self.e2e_queries = 500_000_000, 50_000_000
            self.cell_var = 1
        self.n_vars = len(self.i_vars)
        self.ppc = 3 if self.cell_type == "triangle" else 4
        self.n_pd = {"C1": 2, "C2": 2, "C3": 4, "C4": 3, "C5": 8}[name]

    # ---- mesh (cached in /tmp for the sibling arm / profiling re-runs of the same command) ----
    def mesh(self):
        from interpolate_unstructured_b200 import meshgen          # pure numpy generators
        from oracle import oracle
        if self.name == "C1":
            from interpolate_unstructured_b200 import binda       # pure Python reader of the fixture
            ents = {}
            for e in binda.read_binda(os.path.join(ROOT, "tests", "golden", "triangle.binda")):
                ents.setdefault(e.name, []).append(e)
            pts = np.ascontiguousarray(ents["points"][0].data, dtype=np.float64)
            cells = np.ascontiguousarray(ents["cells"][0].data, dtype=np.int32)
            nb = np.ascontiguousarray(ents["cell_neighbors"][0].data, dtype=np.int32)
            pd = np.ascontiguousarray(np.stack([np.asarray(e.data, dtype=np.float64) for e in ents["point_data"]], axis=1))
            cd = None
        else:
            cache = f"/tmp/iu_b200_bench_{self.cell_type}_{self.n}_{self.n_pd}.npz"
            if os.path.exists(cache):
                z = np.load(cache)
                pts, cells, nb, pd = z["pts"], z["cells"], z["nb"], z["pd"]
            else:
                if self.cell_type == "quad":
                    pts, cells = meshgen.quad_mesh(self.n, jitter=0.2, seed=1234)
                else:
                    pts, cells = meshgen.kuhn_tet_mesh(self.n, jitter=0.10, seed=1234)
                nb = oracle.cell_neighbors(cells, pts, self.cell_type)     # convert_to_binary.py:118-162 restated
                pd = meshgen.smooth_fields(pts, self.n_pd)                 # column 1 linear (self-check), others smooth
                if int(os.environ.get("RANK", "0")) == 0:
                    try:
                        np.savez(cache + ".tmp.npz", pts=pts, cells=cells, nb=nb, pd=pd)
                        os.replace(cache + ".tmp.npz", cache)
                    except OSError:
                        pass
            cd = np.arange(1, len(cells) + 1, dtype=np.float64) if self.cell_var else None
        self.pts, self.cells, self.nb, self.pd, self.cd = pts, cells, nb, pd, cd
        lo, hi = pts.min(0), pts.max(0)
        d = hi - lo
        self.qlo, self.qhi = lo + 0.1 * d, hi - 0.1 * d               # benchmark.f90:28-38
        nd = int((d > 0).sum())
        self.h = float(np.prod(d[d > 0]) / len(cells)) ** (1.0 / nd)  # mean cell size
        return self

    def text(self):
        """The `config.workload` string: the same in both arms."""
        nc, npt = len(self.cells), len(self.pts)
        if self.name == "C1":
            return f"C1: benchmark.f90 on test_data/triangle ({nc} triangles / {npt} points), cold iu_interpolate_scalar_at, ivar 1, uniform queries in the 10 %-shrunk bbox"
        if self.name == "C2":
            return f"C2: {self.n}x{self.n} jittered quads ({nc} cells / {npt} points), cold iu_interpolate_scalar_at (1 variable), uniform queries in the 10 %-shrunk bbox"
        if self.name == "C3":
            return f"C3: Kuhn tet mesh {nc} tets / {npt} points (jitter 0.1 h), cold iu_interpolate_at, {self.n_vars} vars, uniform queries in the 10 %-shrunk bbox, mesh replicated per GPU"
        if self.name == "C4":
            return (f"C4: warm iu_interpolate_at on the C3 mesh ({nc} tets), {self.n_vars} vars, tracers displaced by {self.alpha} mean cell sizes in a "
                    "random direction, previous cell as hint (walk from its centre, m_interp_unstructured.f90:424-432)")
        return (f"C5: Kuhn tet mesh {nc} tets / {npt} points, cold iu_interpolate_at with {self.n_vars} vars + iu_get_cell_scalar_at "
                "(1 cell-data var, found cell as hint) per query, mesh replicated per GPU")

    def algorithmic_bytes(self, nq):
        """queries in, cells (+ hints) and values out, raw mesh once per batch."""
        per_query = 24 + 4 + (4 if self.warm else 0) + 8 * self.n_vars + (8 if self.cell_var else 0)
        m_raw = 24 * len(self.pts) + 8 * self.ppc * len(self.cells) + 8 * self.n_vars * len(self.pts) + (8 * len(self.cells) if self.cell_var else 0)
        return nq * per_query + m_raw, per_query, m_raw

    def host_queries(self, n, start=0):
        """Queries start .. start+n of rank 0's stream: the numbers iu_b200_generate_queries_dev produces on the device."""
        from interpolate_unstructured_b200 import meshgen
        return meshgen.uniform_queries(n, self.qlo, self.qhi, seed=QUERY_SEED, start=start)

    def displacement(self, n, start=0):
        """C4: unit vectors (counter-based, so that host and device agree) scaled by alpha * h."""
        from interpolate_unstructured_b200 import meshgen
        u = meshgen.uniform_queries(n, [-1.0] * 3, [1.0] * 3, seed=QUERY_SEED + 1, start=start)
        # elementwise IEEE operations in this order on host and device: the displaced positions agree bit for bit
        nrm = np.sqrt(u[:, 0] * u[:, 0] + u[:, 1] * u[:, 1] + u[:, 2] * u[:, 2])
        return u / nrm[:, None] * (self.alpha * self.h)


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


# ------------------------------------------------------------------------------------------------
# CPU leg: the oracle (C++ restatement of the reference path) on the host cores
# ------------------------------------------------------------------------------------------------
def host_cores():
    """Threads the CPU leg may use: the CPUs this process is allowed on, whatever OMP_NUM_THREADS says
    (torchrun exports OMP_NUM_THREADS=1)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


class CpuLeg:
    """benchmark.f90:14-71 restated: build the grid, pass 1 cold, pass 2 warm, wall clock per pass."""

    def __init__(self, wl):
        from oracle.oracle import OracleGrid
        self.wl = wl
        self.g = OracleGrid.from_arrays(wl.cell_type, wl.pts, wl.cells, wl.nb, wl.pd, cell_data=wl.cd)
        self.cores = host_cores()

    def one_pass(self, r, ic, nthreads):
        """One timed pass of the workload's call over the queries r; ic is updated in place.  Returns (seconds, res, cres)."""
        wl = self.wl
        t, res = self.g.benchmark_pass(r, wl.i_vars, ic, nthreads=nthreads)       # iu_interpolate_at, kdtree2 start
        cres = None
        if wl.cell_var:                                                           # second call of C5, found cells as hints
            t0 = time.perf_counter()
            cres, ic2 = self.g.get_cell_scalar_at(r, wl.cell_var, i_guess=ic, nthreads=nthreads)
            t += time.perf_counter() - t0
            ic[:] = ic2
        return t, res, cres

    def queries(self, n):
        """The first n queries of rank 0's device stream (C4: their displaced positions + the cells of the undisplaced ones)."""
        wl = self.wl
        r = wl.host_queries(n)
        if not wl.warm:
            return r, np.zeros(n, dtype=np.int32)
        ic0 = np.zeros(n, dtype=np.int32)
        self.g.benchmark_pass(r, wl.i_vars, ic0, nthreads=self.cores)             # untimed: where the tracers were
        return np.ascontiguousarray(r + wl.displacement(n)), ic0

    def calibrate(self, seconds, nthreads, n_max):
        n0 = int(min(n_max, 20_000 if nthreads == 1 else 200_000))
        r, ic = self.queries(n0)
        t, _, _ = self.one_pass(r, ic.copy(), nthreads)
        return int(min(max(n0 / max(t, 1e-9) * seconds, min(n0, n_max)), n_max))

    def timed(self, n, nthreads, steps=1, warmup=0):
        """-> dict(first = queries/s of the workload's pass, second = queries/s of benchmark.f90's second pass (the same
        points with the cells of the first pass as hints; dt = 0 as on planar meshes), r, ic, res, cres of the first pass)."""
        r, ic_in = self.queries(n)
        t1s, t2s = [], []
        for s in range(warmup + steps):
            ic = ic_in.copy()
            t1, res, cres = self.one_pass(r, ic, nthreads)
            ic_first = ic.copy()
            t2, _, _ = self.one_pass(r, ic, nthreads)
            if s >= warmup:
                t1s.append(t1); t2s.append(t2)
        return {"n": n, "first": n / float(np.mean(t1s)), "second": n / float(np.mean(t2s)), "seconds": float(np.mean(t1s)),
                "r": r, "ic_in": ic_in, "ic": ic_first, "res": res, "cres": cres}


def cpu_baseline(wl, seconds, n_max, steps=1, warmup=0):
    """All-cores and one-thread timings of the oracle on a bounded sample of the workload's own query stream."""
    leg = CpuLeg(wl)
    n_all = leg.calibrate(seconds, leg.cores, n_max)
    all_cores = leg.timed(n_all, leg.cores, steps, warmup)
    n_one = leg.calibrate(max(2.0, seconds / 2), 1, n_all)
    one = leg.timed(n_one, 1)
    what = "warm (hinted) pass" if wl.warm else "cold pass"
    sample = (f"first {n_all} queries of rank 0's stream ({what} {all_cores['seconds']:.2f} s on {leg.cores} threads; one thread: first {n_one} queries), "
              f"C++ restatement of benchmark.f90:46-69 (oracle/iu_oracle.cpp, kdtree2-style tree, OpenMP over queries); the reference is serial Fortran")
    cb = {"value": all_cores["first"], "unit": "queries/s", "cores": leg.cores, "kind": "port", "sample": sample,
          "serial": {"value": one["first"], "cores": 1, "queries": n_one},
          "second_pass_with_hints": {"value": all_cores["second"], "serial_value": one["second"],
                                     "what": "benchmark.f90:57-69: the same points again with the cells of the first pass as hints"}}
    return cb, all_cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = Workload(args.workload, args.mesh_n, args.alpha).mesh()
    k, w = max(1, args.steps), min(max(0, args.warmup), 1)
    # the whole --steps K --warmup W run (two passes per step, plus the one-thread sample) stays within ~2 minutes
    per_pass = min(args.cpu_seconds, max(0.5, 90.0 / (2 * (k + w) + 2)))
    n_max = int(args.queries) if args.queries else (wl.queries if wl.name in ("C1",) else 50_000_000)
    cb, ac = cpu_baseline(wl, per_pass, n_max, steps=k, warmup=w)
    out = {
        "impl": "reference", "metric": metric_name(wl), "value": cb["value"], "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ac["seconds"] * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": config_of(wl),
        "note": "the reference (Fortran + missing kdtree2 submodule) is unbuildable on this image; this is its C++ restatement on all host cores",
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


def metric_name(wl):
    return {"C1": "queries/s (kd-tree+walk+interp), benchmark.f90 triangle fixture",
            "C2": "queries/s (kd-tree+walk+interp), 2D quad mesh",
            "C3": "queries/s (kd-tree+walk+interp), 3D tet mesh",
            "C4": "queries/s (hinted walk+interp), 3D tet mesh",
            "C5": "queries/s (kd-tree+walk+interp+cell scalar), 3D tet mesh"}[wl.name]


def config_of(wl):
    return {"workload": wl.text(), "l2_policy": "inputs larger than L2 (24 B x queries per step >> 126 MB)" if wl.name != "C1" else
            "C1 is the reference's own CPU-runnable case (1M queries = 24 MB of inputs, an 8-cell mesh): smaller than L2 by definition, not flushed"}


def bind_to_gpu_numa_node(local):
    """Pin this rank (and therefore its first-touch pinned host buffers) to the CPUs of the NUMA node its GPU hangs
    off: the host-buffer leg moves 64 B per query over PCIe and should not cross the socket interconnect.
    Returns the node, or "single-node host" when the machine exposes no NUMA placement."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return "single-node host (numa_node = -1)"
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return node
    except (OSError, ValueError, AttributeError):
        return "unknown"


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
    affinity0 = os.sched_getaffinity(0)
    numa_node = bind_to_gpu_numa_node(local)
    if world == 1:
        os.sched_setaffinity(0, affinity0)          # one rank: keep every core for the CPU leg
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
    wl = Workload(args.workload, args.mesh_n, args.alpha).mesh()
    t_mesh = time.time() - t0
    g = api.Grid.from_arrays(wl.cell_type, wl.pts, wl.cells, wl.nb, wl.pd, cell_data=wl.cd, device=local)
    setup_s = time.time() - t0 - t_mesh
    if args.sort >= 0:
        g.set_option(api.OPT_SORT_QUERIES, args.sort)
    if args.dynamic >= 0:
        g.set_option(api.OPT_DYNAMIC_TILES, args.dynamic)
    g.set_option(api.OPT_TIME_KERNELS, 1)

    nq = int(args.queries or wl.queries)
    nv = wl.n_vars
    r = torch.empty((nq, 3), dtype=torch.float64, device=dev)
    g.generate_queries_dev(r, first=rank * nq, seed=QUERY_SEED, lo=wl.qlo, hi=wl.qhi)
    ic = torch.zeros(nq, dtype=torch.int32, device=dev)
    out = torch.empty((nq, nv), dtype=torch.float64, device=dev)
    cres = torch.empty(nq, dtype=torch.float64, device=dev) if wl.cell_var else None
    ic0 = None
    if wl.warm:                                      # C4: locate the tracers, then move them
        ic0 = torch.zeros(nq, dtype=torch.int32, device=dev)
        g.iu_get_cell_dev(r, ic0)
        g.check_dev()
        d = torch.empty((nq, 3), dtype=torch.float64, device=dev)
        g.generate_queries_dev(d, first=rank * nq, seed=QUERY_SEED + 1, lo=[-1.0] * 3, hi=[1.0] * 3)
        nrm = torch.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])   # as Workload.displacement, bit for bit
        r.copy_(r + d / nrm[:, None] * (wl.alpha * wl.h))
        del d, nrm

    def step():
        if wl.warm:
            ic.copy_(ic0)                            # hints = the cells the tracers were in
        else:
            ic.zero_()                               # cold: i_cell = 0 for every query (benchmark.f90:41)
        if wl.cell_var:
            g.iu_interpolate_and_cell_scalar_at_dev(r, wl.i_vars, wl.cell_var, out, cres, ic)
        else:
            g.iu_interpolate_at_dev(r, wl.i_vars, out, ic)

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
    ms_local = e0.elapsed_time(e1)
    ms = max_over_ranks(ms_local)
    st = g.stats()
    value = world * nq * args.steps / (ms * 1e-3)

    # correctness of what was timed: size-independent properties here, the oracle comparison below
    found = ic >= 1
    n_found = int(found.sum().item())
    checks = {"located": n_found, "queries": nq}
    if wl.name in ("C3", "C4", "C5"):                # their first variable is the linear field x + y + z + 1
        lin = (out[:, 0] - (r.sum(1) + 1.0)).abs()[found]
        checks["linear_field_max_abs_err"] = float(lin.max().item()) if n_found else None
        assert n_found == 0 or checks["linear_field_max_abs_err"] < 1e-12, f"linear field not reproduced: {checks}"
    if not wl.warm:
        assert n_found == nq, f"{nq - n_found} queries not located"
    if wl.cell_var:
        assert torch.equal(cres[found].to(torch.int32), ic[found]), "cell-data scalar is not the found cell's"
    assert st["walk_limit_hits"] == 0

    # roofline of the dominant kernel: algorithmic bytes (queries x bytes per query + the raw mesh ONCE per batch)
    # over the k_query time of one batch
    n_k = max(1, st["query_kernel_launches"])
    k_ms = st["query_kernel_ns"] / n_k * 1e-6
    k_ms_per_batch = st["query_kernel_ns"] * 1e-6 / args.steps
    alg_bytes, per_query, m_raw = wl.algorithmic_bytes(nq)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg_bytes / (k_ms_per_batch * 1e-3) / 1e9
    traffic, traffic_src, ncu_counters = None, None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if wl.name == "C3":
            traffic = tj.get("k_query_dram_bytes_per_launch")
            traffic_src = f"profiles/{tj.get('source', '')} (ncu --set full capture of the same kernel, {tj.get('queries_per_launch')} queries per launch; not measured in this run)"
            ncu_counters = dict(tj.get("counters") or {}, source="profiles/" + tj.get("source", ""))
    except (OSError, ValueError):
        pass
    kname = f"k_query<{wl.cell_type.upper()}, {'INTERP_CELL' if wl.cell_var else 'INTERP'}, {nv}>"
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": traffic_src, "kernel": kname, "kernel_ms_per_launch": k_ms, "launches": n_k,
                "launches_per_batch": n_k / args.steps, "kernel_ms_per_batch": k_ms_per_batch,
                "algorithmic_bytes_per_batch": alg_bytes, "bytes_per_query": per_query, "mesh_raw_bytes": m_raw,
                "peak_source": "MEASURED_PEAKS.json (measured)" if peaks else "fallback 6650 GB/s",
                "kernel_share_of_step": st["query_kernel_ns"] * 1e-6 / ms_local,
                "sort_share_of_step": st["sort_ns"] * 1e-6 / ms_local, "ncu_counters": ncu_counters}
    gpu_launches = st["kernel_launches"] + args.steps      # + the i_cell memset / copy kernel of each step

    # what the oracle comparison needs from the device-resident run (rank 0's first queries)
    n_keep = min(nq, 50_000_000)
    keep = None
    if rank == 0 and not args.no_cpu:
        keep = (out[:n_keep].cpu().numpy(), ic[:n_keep].cpu().numpy(), cres[:n_keep].cpu().numpy() if wl.cell_var else None,
                r[:min(n_keep, 1000)].cpu().numpy())

    # end to end through the host-buffer C ABI (pinned host memory; H2D + kernels + D2H per step)
    e2e = None
    if not args.no_e2e:
        ne = int(args.e2e_queries or min(wl.e2e_queries, nq))
        r_h = api.pinned_empty((ne, 3), np.float64)
        ic_h = api.pinned_empty((ne,), np.int32)
        out_h = api.pinned_empty((ne, nv), np.float64)
        cres_h = api.pinned_empty((ne,), np.float64) if wl.cell_var else None
        hint_h = ic0[:ne].cpu().numpy() if wl.warm else None
        torch.cuda.synchronize()
        r_h[:] = r[:ne].cpu().numpy()
        del out, ic, cres
        r_sum = r_h.sum(1)

        def e2e_call():
            if wl.cell_var:
                g.iu_interpolate_and_cell_scalar_at(r_h, wl.i_vars, wl.cell_var, var_at_r=out_h, cell_res=cres_h, i_cell=ic_h)
            else:
                g.iu_interpolate_at(r_h, wl.i_vars, var_at_r=out_h, i_cell=ic_h)   # synchronous: returns after the last D2H

        def e2e_reset():
            if wl.warm:
                ic_h[:] = hint_h
            else:
                ic_h[:] = 0

        for _ in range(max(1, min(args.warmup, 2))):
            e2e_reset()
            e2e_call()
        barrier()
        g.reset_stats()
        tt = 0.0
        for _ in range(args.steps):
            e2e_reset()
            barrier()
            t1 = time.perf_counter()
            e2e_call()
            tt += time.perf_counter() - t1
        tt = max_over_ranks(tt)
        skipped = g.stats()["h2d_skipped_bytes"] // args.steps     # i_cell of all-cold chunks is zeroed on the device, not uploaded
        ok = ic_h >= 1
        assert wl.warm or ok.all()
        if wl.name in ("C3", "C4", "C5"):
            assert np.abs(out_h[:, 0] - (r_sum + 1))[ok].max() < 1e-12
        h2d = ne * 28 - skipped
        d2h = ne * (4 + 8 * nv + (8 if wl.cell_var else 0))
        e2e = {"value": world * ne * args.steps / tt, "unit": "queries/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "queries_per_step": ne, "ms_per_step": tt / args.steps * 1e3, "host_memory": "pinned (iu_b200_host_alloc)",
               "h2d_note": "24 B of coordinates per query + 4 B of i_cell only for pipeline chunks that carry a hint" if skipped else "24 B of coordinates + 4 B of i_cell per query",
               "numa_node_rank0": numa_node}

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()                 # the CPU leg below runs on rank 0 alone, the other ranks are done
    if rank != 0:
        return
    g.close()
    torch.cuda.empty_cache()

    cpu = None
    if not args.no_cpu:
        os.sched_setaffinity(0, affinity0)
        cpu, ac = cpu_baseline(wl, args.cpu_seconds, n_keep)
        # the oracle's answers for the first queries of the timed device stream, query for query
        g_out, g_ic, g_cres, g_r = keep
        n = ac["n"]
        assert np.array_equal(ac["r"][:len(g_r)], g_r), "host and device query generators disagree"
        mism = g_ic[:n] != ac["ic"]
        fnd = ac["ic"] >= 1
        rel = np.abs(g_out[:n] - ac["res"]) / np.maximum(1.0, np.abs(ac["res"]))
        checks.update({"oracle_sample": int(n), "oracle_cell_mismatches": int(mism.sum()),
                       "oracle_value_max_rel": float(rel[fnd & ~mism].max()) if (fnd & ~mism).any() else 0.0,
                       "oracle_values_bit_identical": bool(np.array_equal(g_out[:n][fnd & ~mism], ac["res"][fnd & ~mism])),
                       "oracle_start": "kdtree2-style tree (ties aside, the start cell does not change the result inside a convex domain)"})
        if wl.cell_var:
            checks["oracle_cell_scalar_mismatches"] = int((g_cres[:n][fnd] != ac["cres"][fnd]).sum())

    line = {
        "metric": metric_name(wl), "value": value, "unit": "queries/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": config_of(wl),
        "run": {"queries_per_step_per_gpu": nq, "sort_queries": args.sort if args.sort >= 0 else 2,
                "dynamic_tiles": args.dynamic if args.dynamic >= 0 else 1, "mesh_seconds": t_mesh, "setup_seconds": setup_s,
                "mesh_device_bytes": st["mesh_bytes"], "tree_depth": st["tree_depth"],
                "nn_voxel_table": {k: st[k] for k in ("nn_voxels", "nn_voxels_tree", "nn_voxel_entries", "nn_tree_queries", "nn_voxel_build_us")}},
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": gpu_launches, "clocks": clocks.summary(),
        "checks": checks,
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
