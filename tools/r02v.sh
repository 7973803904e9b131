#!/bin/bash
# r02v: voxel-table build with the children's boxes tested at push time: parity of everything that uses the table, set-up times
mkdir -p gpurun_out
{
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python - <<'PY'
import os, sys, time
sys.path.insert(0, os.getcwd())
import bench
from interpolate_unstructured_b200 import api
for name in ("C3", "C2", "C5"):
    wl = bench.Workload(name).mesh()
    for rep in range(2):
        t0 = time.time()
        g = api.Grid.from_arrays(wl.cell_type, wl.pts, wl.cells, wl.nb, wl.pd, cell_data=wl.cd, device=0)
        t = time.time() - t0
        st = g.stats()
        print(f"{name}: grid creation {t:.2f} s, kd-tree build {st['tree_build_us'] / 1e6:.3f} s, voxel table {st['nn_voxel_build_us'] / 1e6:.3f} s ({st['nn_voxels']} voxels, {st['nn_voxels_tree']} flagged, {st['nn_voxel_entries']} entries)", flush=True)
        g.close()
PY
} 2>&1 | tee gpurun_out/r02v.txt
