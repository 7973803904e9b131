#!/bin/bash
# r02e: hint flag in the permutation entry (cold queries skip the scattered hint read), L2 fetch granularity A/B, new tests
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02e_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02e_pytest_gpu.log
B="python bench.py --queries 2e8 --steps 3 --warmup 3 --no-cpu --no-e2e"
P="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,l1tex__throughput.avg.pct_of_peak_sustained_elapsed
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.3e q/s  kernel %.2f ms/launch  share %.2f sort %.2f  ms/step %.2f'%(d['value'], r['kernel_ms_per_launch'], r['kernel_share_of_step'], r['sort_share_of_step'], d['ms_per_step']))"; }
ncu1() { $P > /dev/null 2>&1 && ncu --metrics $M --clock-control none -k regex:k_query -s 3 -c 1 --csv --log-file gpurun_out/r02e_$1.csv $P > /dev/null 2>&1
  grep -E "dram__|hit_rate|time_duration|inst_executed|throughput" gpurun_out/r02e_$1.csv | awk -F'","' '{print "   ", $(NF-2), $(NF-1), $NF}'; }
{
echo "== hint flag, default L2 fetch granularity"; $B | show; ncu1 base
for f in 32 128; do echo "== IU_B200_L2_FETCH=$f"; export IU_B200_L2_FETCH=$f; $B | show; ncu1 l2f$f; unset IU_B200_L2_FETCH; done
echo "== C4 (hinted) 1e8"; python bench.py --workload C4 --steps 3 --warmup 2 --no-cpu --no-e2e | show
} 2>&1 | tee gpurun_out/r02e.txt
