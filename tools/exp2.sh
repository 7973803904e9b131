#!/bin/bash
# experiment: streaming cache hints (lib = on, lib3 = off) x scheduling; timing + DRAM bytes of k_query
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
B="python bench.py --queries 2e8 --steps 3 --warmup 2 --no-cpu --no-e2e"
P="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,gpu__time_duration.sum
for lib in lib lib3; do for dyn in 1 0; do
  echo "== $lib dynamic=$dyn"
  export IU_B200_LIB=$PWD/interpolate_unstructured_b200/$lib/libiu_b200.so
  $B --dynamic $dyn | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.3e q/s  kernel %.2f ms/launch  share %.2f sort %.2f'%(d['value'], r['kernel_ms_per_launch'], r['kernel_share_of_step'], r['sort_share_of_step']))"
  $P --dynamic $dyn > /dev/null 2>&1 && ncu --metrics $M --clock-control none -k regex:k_query -s 3 -c 1 --csv --log-file gpurun_out/exp2_${lib}_${dyn}.csv $P --dynamic $dyn > /dev/null 2>&1
  grep -E "dram__|hit_rate|time_duration" gpurun_out/exp2_${lib}_${dyn}.csv | awk -F'","' '{print "   ", $(NF-2), $(NF-1), $NF}'
done; done 2>&1 | tee gpurun_out/exp2.txt
