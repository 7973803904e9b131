import sys, time, numpy as np
sys.path.insert(0, "/root/repo")
import torch, bench
from interpolate_unstructured_b200 import api
wl = bench.Workload("C3").mesh(); pts, cells, nb, pd, qlo, qhi = wl.pts, wl.cells, wl.nb, wl.pd, wl.qlo, wl.qhi
g = api.Grid.from_arrays("tetra", pts, cells, nb, pd, device=0)
ne = 100_000_000
r = torch.empty((ne, 3), dtype=torch.float64, device="cuda")
g.generate_queries_dev(r, first=0, seed=1, lo=qlo, hi=qhi)
r_h = api.pinned_empty((ne, 3), np.float64); ic_h = api.pinned_empty((ne,), np.int32); out_h = api.pinned_empty((ne, 4), np.float64)
r_h[:] = r.cpu().numpy(); del r
for chunk in (1 << 20, 1 << 21, 1 << 22, 1 << 23, 1 << 24):
    g.set_option(api.OPT_HOST_CHUNK, chunk)
    ts = []
    for k in range(4):
        ic_h[:] = 0
        t = time.perf_counter(); g.iu_interpolate_at(r_h, [1, 2, 3, 4], var_at_r=out_h, i_cell=ic_h); ts.append(time.perf_counter() - t)
    print(f"host chunk {chunk:9d}: {ne / min(ts[1:]):.3e} q/s  ({min(ts[1:]) * 1e3:.1f} ms)", flush=True)
