#!/bin/bash
# r02j: precomputed first walk step (64-byte start entries): parity, timing, counters, phases
mkdir -p gpurun_out
B="python bench.py --queries 2e8 --steps 3 --warmup 3 --no-cpu --no-e2e"
P="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts.sum
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.3e q/s  kernel %.2f ms/launch  share %.2f sort %.2f  ms/step %.2f'%(d['value'], r['kernel_ms_per_launch'], r['kernel_share_of_step'], r['sort_share_of_step'], d['ms_per_step']), d['checks'])"; }
ncu1() { $P $2 > /dev/null 2>&1 && ncu --metrics $M --clock-control none -k regex:k_query -s 3 -c 1 --csv --log-file gpurun_out/r02j_$1.csv $P $2 > /dev/null 2>&1
  grep -E "dram__|hit_rate|time_duration|inst_executed|throughput|wavefronts" gpurun_out/r02j_$1.csv | awk -F'","' '{print "   ", $(NF-2), $(NF-1), $NF}'; }
{
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "== first-step build"; $B | show; ncu1 base
timeout 600 python tools/phase_bench.py 2>&1 | grep -v "^{"
for wl in C2 C4 C5; do echo "== $wl"; python bench.py --workload $wl --steps 3 --warmup 2 --no-cpu --no-e2e | show; done
} 2>&1 | tee gpurun_out/r02j.txt
