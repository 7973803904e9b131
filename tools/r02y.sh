#!/bin/bash
# r02y: validation of the final HEAD: all GPU tests, smoke, reference arm + bench line with the driver's flags
tag=r02y
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/${tag}_bench.json')); r=json.load(open('gpurun_out/${tag}_ref.json'))
print('value %.4e e2e %.4e ms/step %.2f setup %.2f s checks %s'%(d['value'], d['e2e']['value'], d['ms_per_step'], d['run']['setup_seconds'], d['checks']))
print('reference arm %.4e (%d threads), serial %.4e; same config: %s'%(r['value'], r['cpu_baseline']['cores'], r['cpu_baseline']['serial']['value'], r['config'] == d['config']))"
IU_B200_LIB=$PWD/interpolate_unstructured_b200/lib_checked/libiu_b200.so timeout 900 python tools/sanitize_smoke.py 2>&1 | tail -2
