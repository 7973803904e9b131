#!/usr/bin/env python3
"""Creates the grid of a bench workload and prints the set-up statistics (profiling target for the set-up kernels)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from interpolate_unstructured_b200 import api  # noqa: E402

wl = bench.Workload(sys.argv[1] if len(sys.argv) > 1 else "C3", int(sys.argv[2]) if len(sys.argv) > 2 else 0).mesh()
t0 = time.time()
g = api.Grid.from_arrays(wl.cell_type, wl.pts, wl.cells, wl.nb, wl.pd, cell_data=wl.cd, device=0)
st = g.stats()
print(f"grid creation {time.time() - t0:.2f} s, kd-tree {st['tree_build_us'] / 1e6:.3f} s, voxel table {st['nn_voxel_build_us'] / 1e6:.3f} s")
