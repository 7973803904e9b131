#!/bin/bash
# Round profile recipe (run under gpurun): GPU parity tests, the default bench line, then the two
# ncu passes of /opt/skills/guides/B200_PROFILING.md on a shortened bench command.
# usage: tools/profile.sh [tag]   -> everything lands in gpurun_out/<tag>_*
tag=${1:-r01b}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${tag}_smi.csv 2>&1
nproc > gpurun_out/${tag}_nproc.txt
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/${tag}_pytest_gpu.log
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
cat gpurun_out/${tag}_bench.json
CMD="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
$CMD > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_query -s 3 -c 2 -f -o gpurun_out/${tag}_k_query $CMD > gpurun_out/${tag}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out/
