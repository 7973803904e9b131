#!/bin/bash
# r02f: .L2::64B on the scattered query IO (lib_ltcq) and also on the mesh gathers (lib)
mkdir -p gpurun_out
B="python bench.py --queries 2e8 --steps 3 --warmup 3 --no-cpu --no-e2e"
P="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,l1tex__throughput.avg.pct_of_peak_sustained_elapsed
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.3e q/s  kernel %.2f ms/launch  share %.2f sort %.2f  ms/step %.2f'%(d['value'], r['kernel_ms_per_launch'], r['kernel_share_of_step'], r['sort_share_of_step'], d['ms_per_step']))"; }
ncu1() { $P > /dev/null 2>&1 && ncu --metrics $M --clock-control none -k regex:k_query -s 3 -c 1 --csv --log-file gpurun_out/r02f_$1.csv $P > /dev/null 2>&1
  grep -E "dram__|hit_rate|time_duration|inst_executed|throughput" gpurun_out/r02f_$1.csv | awk -F'","' '{print "   ", $(NF-2), $(NF-1), $NF}'; }
{
for lib in lib_ltcq lib; do
  echo "== $lib"; export IU_B200_LIB=$PWD/interpolate_unstructured_b200/$lib/libiu_b200.so
  $B | show; ncu1 $lib
done
export IU_B200_LIB=$PWD/interpolate_unstructured_b200/lib/libiu_b200.so
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python tools/phase_bench.py 2>&1 | grep -v "^{" 
echo "== copy ceiling"; timeout 300 python tools/copy_ceiling.py
} 2>&1 | tee gpurun_out/r02f.txt
