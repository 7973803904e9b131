#!/bin/bash
# r02a: first run of the voxel-table nearest-centre search: parity tests, then the bench at several table densities
tag=r02a
mkdir -p gpurun_out
nproc > gpurun_out/${tag}_nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/${tag}_pytest_gpu.log
for d in 400 0 200 800 100; do
  IU_B200_NN_VOXEL_DENSITY=$d timeout 600 python bench.py --queries 2e8 --steps 3 --warmup 2 --no-cpu --no-e2e > gpurun_out/${tag}_bench_d$d.json 2> gpurun_out/${tag}_bench_d$d.err; echo "bench d=$d rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${tag}_bench_d$d.json"))
    print($d, "value %.3e"%d["value"], "k_ms %.3f"%d["roofline"]["kernel_ms_per_launch"], "sort_share %.3f"%d["roofline"]["sort_share_of_step"], d["config"]["nn_voxel_table"], "setup", d["config"]["setup_seconds"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/${tag}_bench_d$d.err").read()[-2000:])
PY
done
