#!/bin/bash
# r03f: ticket tiles of 32 queries (one warp-load per ticket) against 64 at 3 CTAs per SM
mkdir -p gpurun_out
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('   %-3s value %.3e q/s  kernel %.2f ms/launch  ms/step %.2f'%(d['config']['workload'][:2], d['value'], r['kernel_ms_per_launch'], d['ms_per_step']))"; }
{
for lib in lib lib_t32; do
  echo "== $lib"; export IU_B200_LIB=$PWD/interpolate_unstructured_b200/$lib/libiu_b200.so
  python bench.py --queries 2e8 --steps 3 --warmup 3 --no-cpu --no-e2e | show
  python bench.py --workload C5 --queries 1e8 --steps 2 --warmup 1 --no-cpu --no-e2e | show
  python bench.py --workload C2 --steps 2 --warmup 1 --no-cpu --no-e2e | show
done
} 2>&1 | tee gpurun_out/r03f.txt
