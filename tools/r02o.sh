#!/bin/bash
# r02o: final validation of HEAD: all GPU tests, smoke, reference arm + bench line (driver-style flags), launch list + full capture
tag=r02o
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; cat gpurun_out/${tag}_bench.json; tail -3 gpurun_out/${tag}_bench.err
CMD="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_query -s 3 -c 1 -f -o gpurun_out/${tag}_k_query $CMD > gpurun_out/${tag}_ncu_full.log 2>&1
echo "ncu full rc=$?"
python tools/workloads.py trace --steps 2 > gpurun_out/${tag}_trace.json 2>&1; tail -c 600 gpurun_out/${tag}_trace.json
