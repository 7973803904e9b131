#!/bin/bash
# r02m: 32-byte gathers as two 16-byte loads (lib_split); k_trace compiled for 1 / 4 / 6 / 8 resident CTAs per SM
mkdir -p gpurun_out
B="python bench.py --queries 2e8 --steps 3 --warmup 3 --no-cpu --no-e2e"
P="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
M=dram__bytes_read.sum,gpu__time_duration.sum,smsp__inst_executed.sum,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts.sum,l1tex__t_sector_hit_rate.pct
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.3e q/s  kernel %.2f ms/launch  share %.2f sort %.2f  ms/step %.2f'%(d['value'], r['kernel_ms_per_launch'], r['kernel_share_of_step'], r['sort_share_of_step'], d['ms_per_step']))"; }
ncu1() { $P > /dev/null 2>&1 && ncu --metrics $M --clock-control none -k regex:k_query -s 3 -c 1 --csv --log-file gpurun_out/r02m_$1.csv $P > /dev/null 2>&1
  grep -E "dram__|hit_rate|time_duration|inst_executed|throughput|wavefronts" gpurun_out/r02m_$1.csv | awk -F'","' '{print "   ", $(NF-2), $(NF-1), $NF}'; }
{
for lib in lib lib_split; do
  echo "== $lib"; export IU_B200_LIB=$PWD/interpolate_unstructured_b200/$lib/libiu_b200.so
  $B | show; ncu1 $lib
done
for lib in lib lib_tr4 lib_tr6 lib_tr8; do
  echo "== trace $lib"; export IU_B200_LIB=$PWD/interpolate_unstructured_b200/$lib/libiu_b200.so
  python tools/workloads.py trace --steps 2 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('device %.2f ms  %.3e lines/s  %.3e steps/s   host call %.3f s'%(d['device_ms'], d['device_lines_per_s'], d['device_stored_steps_per_s'], d['seconds']))"
done
} 2>&1 | tee gpurun_out/r02m.txt
