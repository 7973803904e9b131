#!/bin/bash
# r02u: kd-tree built on the device: parity (all GPU tests), set-up time against the host builder, bench
mkdir -p gpurun_out
{
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for host in "" 1; do
  echo "== IU_B200_HOST_TREE=$host"
  IU_B200_HOST_TREE=$host python - <<'PY'
import os, sys, time
sys.path.insert(0, os.getcwd())
import bench
from interpolate_unstructured_b200 import api
if os.environ.get("IU_B200_HOST_TREE") == "": del os.environ["IU_B200_HOST_TREE"]
for name, n in (("C3", 0), ("C5", 0)):
    wl = bench.Workload(name, n).mesh()
    for rep in range(2):
        t0 = time.time()
        g = api.Grid.from_arrays(wl.cell_type, wl.pts, wl.cells, wl.nb, wl.pd, cell_data=wl.cd, device=0)
        t = time.time() - t0
        st = g.stats()
        print(f"{name}: grid creation {t:.2f} s, kd-tree build {st['tree_build_us'] / 1e6:.3f} s, voxel table {st['nn_voxel_build_us'] / 1e6:.3f} s, tree depth {st['tree_depth']}", flush=True)
        g.close()
PY
done
python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value %.3e  setup %.2f s  voxels %s'%(d['value'], d['run']['setup_seconds'], d['run']['nn_voxel_table']))"
} 2>&1 | tee gpurun_out/r02u.txt
