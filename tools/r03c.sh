#!/bin/bash
# r03c: validation of HEAD with k_query at 3 CTAs per SM / 64-query tiles: GPU tests, smoke, bench + reference arm, ncu
tag=r03c
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
IU_B200_LIB=$PWD/interpolate_unstructured_b200/lib_checked/libiu_b200.so timeout 600 python tools/sanitize_smoke.py 2>&1 | tail -1
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/${tag}_bench.json')); r=json.load(open('gpurun_out/${tag}_ref.json'))
print('value %.4e e2e %.4e ms/step %.2f k_query %.3f ms frac %.4f setup %.2f s checks %s'%(d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['roofline']['frac'], d['run']['setup_seconds'], d['checks']))
print('reference arm %.4e (%d threads), serial %.4e; same config: %s'%(r['value'], r['cpu_baseline']['cores'], r['cpu_baseline']['serial']['value'], r['config'] == d['config']))"
CMD="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_query -s 3 -c 1 -f -o gpurun_out/${tag}_k_query $CMD > gpurun_out/${tag}_ncu_full.log 2>&1
echo "ncu full rc=$?"
