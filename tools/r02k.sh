#!/bin/bash
# r02k: sort of sub-batch k+1 beside k_query of sub-batch k: parity, then A/B over tiles per warp
mkdir -p gpurun_out
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.3e q/s  kernel %.2f ms/launch  share %.2f sort %.2f  ms/step %.2f'%(d['value'], r['kernel_ms_per_launch'], r['kernel_share_of_step'], r['sort_share_of_step'], d['ms_per_step']), d['checks'])"; }
{
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -k "overlaps or two_streams or full_size" 2>&1 | tail -3
B="python bench.py --queries 1e9 --steps 3 --warmup 2 --no-cpu --no-e2e"
echo "== overlap off"; IU_B200_OVERLAP_SORT=0 $B | show
for t in 1 2 4 8 16 64; do echo "== overlap on, tiles per warp $t"; IU_B200_TILES_PER_WARP=$t $B | show; done
echo "== overlap off, non-persistent effect alone is not separable (tiles_per_warp only applies beside a sort)"
} 2>&1 | tee gpurun_out/r02k.txt
