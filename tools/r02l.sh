#!/bin/bash
# r02l: radix-sort variants (micro), then the e2e leg with the i_cell upload skipped for all-cold chunks
mkdir -p gpurun_out
{
./tools/micro/sort_variants
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r02l_bench.json; python -c "
import json
d=json.load(open('gpurun_out/r02l_bench.json')); print('value %.3e  e2e %.3e  h2d %d d2h %d ms %.2f'%(d['value'], d['e2e']['value'], d['e2e']['h2d_bytes_per_step'], d['e2e']['d2h_bytes_per_step'], d['e2e']['ms_per_step']))"
} 2>&1 | tee gpurun_out/r02l.txt
