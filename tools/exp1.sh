#!/bin/bash
# experiment: scheduling (static / ticket tiles) x resident CTAs per SM (4 / 3)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="python bench.py --queries 2e8 --steps 3 --warmup 2 --no-cpu --no-e2e"
for lib in lib lib3; do for dyn in 1 0; do
  echo "== $lib dynamic=$dyn"
  IU_B200_LIB=$PWD/interpolate_unstructured_b200/$lib/libiu_b200.so $B --dynamic $dyn | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.3e q/s  kernel %.2f ms/launch  share %.2f sort %.2f'%(d['value'], r['kernel_ms_per_launch'], r['kernel_share_of_step'], r['sort_share_of_step']))"
done; done 2>&1 | tee gpurun_out/exp1.txt
