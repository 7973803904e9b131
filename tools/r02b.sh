#!/bin/bash
# r02b: phase timings and one full ncu capture of k_query with the voxel table
tag=${1:-r02b}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${tag}_pytest_gpu.log
timeout 600 python tools/phase_bench.py > gpurun_out/${tag}_phases.txt 2>&1; echo "phases rc=$?"; cat gpurun_out/${tag}_phases.txt | tail -20
CMD="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_query -s 3 -c 1 -f -o gpurun_out/${tag}_k_query $CMD > gpurun_out/${tag}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out/ | grep ${tag}
