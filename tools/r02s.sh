#!/bin/bash
# r02s: C5 on one GPU: counters of k_query<TETRA, INTERP_CELL, 8>, and 64-byte DRAM fetches for the mesh gathers (lib_ltcm)
mkdir -p gpurun_out
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,dram__throughput.avg.pct_of_peak_sustained_elapsed
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.3e q/s  kernel %.2f ms/launch  share %.2f sort %.2f  ms/step %.2f'%(d['value'], r['kernel_ms_per_launch'], r['kernel_share_of_step'], r['sort_share_of_step'], d['ms_per_step']))"; }
{
for lib in lib lib_ltcm; do
  echo "== $lib"; export IU_B200_LIB=$PWD/interpolate_unstructured_b200/$lib/libiu_b200.so
  python bench.py --workload C5 --queries 2e8 --steps 3 --warmup 2 --no-cpu --no-e2e | show
done
export IU_B200_LIB=$PWD/interpolate_unstructured_b200/lib/libiu_b200.so
P="python bench.py --workload C5 --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
ncu --metrics $M --clock-control none -k regex:k_query -s 3 -c 1 --csv --log-file gpurun_out/r02s_c5.csv $P > /dev/null 2>&1
grep -E "dram__|hit_rate|time_duration|throughput" gpurun_out/r02s_c5.csv | awk -F'","' '{print "   ", $(NF-2), $(NF-1), $NF}'
} 2>&1 | tee gpurun_out/r02s.txt
