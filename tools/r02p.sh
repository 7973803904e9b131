#!/bin/bash
# r02p: padded duplicate candidate loads predicated off + streaming loads in k_walk (lib_dup) against HEAD (lib)
mkdir -p gpurun_out
B="python bench.py --queries 2e8 --steps 3 --warmup 3 --no-cpu --no-e2e"
P="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
M=dram__bytes_read.sum,gpu__time_duration.sum,smsp__inst_executed.sum,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts.sum,smsp__thread_inst_executed_per_inst_executed.ratio
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.3e q/s  kernel %.2f ms/launch  share %.2f sort %.2f  ms/step %.2f'%(d['value'], r['kernel_ms_per_launch'], r['kernel_share_of_step'], r['sort_share_of_step'], d['ms_per_step']))"; }
ncu1() { $P > /dev/null 2>&1 && ncu --metrics $M --clock-control none -k regex:k_query -s 3 -c 1 --csv --log-file gpurun_out/r02p_$1.csv $P > /dev/null 2>&1
  grep -E "dram__|time_duration|inst_executed|throughput|wavefronts" gpurun_out/r02p_$1.csv | awk -F'","' '{print "   ", $(NF-2), $(NF-1), $NF}'; }
{
for lib in lib lib_dup; do
  echo "== $lib"; export IU_B200_LIB=$PWD/interpolate_unstructured_b200/$lib/libiu_b200.so
  $B | show; ncu1 $lib
  python bench.py --workload C2 --steps 3 --warmup 2 --no-cpu --no-e2e | show
  python tools/workloads.py C4 --steps 3 --warmup 2 | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('   %-100s %.3e q/s'%(d['workload'][:100], d['value']))"
done
export IU_B200_LIB=$PWD/interpolate_unstructured_b200/lib_dup/libiu_b200.so
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_vtk.py tests/test_gpu_trace.py -m gpu -x -q 2>&1 | tail -3
} 2>&1 | tee gpurun_out/r02p.txt
