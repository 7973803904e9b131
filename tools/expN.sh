#!/bin/bash
# A/B harness: tools/expN.sh <tag> <libdir> [<libdir> ...]   (lib dirs under interpolate_unstructured_b200/)
# GPU parity tests with the first lib, then per lib: bench line (2e8 queries) + a few ncu counters of k_query.
tag=$1; shift
mkdir -p gpurun_out
export IU_B200_LIB=$PWD/interpolate_unstructured_b200/$1/libiu_b200.so
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
B="python bench.py --queries 2e8 --steps 3 --warmup 3 --no-cpu --no-e2e"
P="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__data_pipe_lsu_wavefronts.sum,smsp__issue_active.avg.pct
for lib in "$@"; do
  echo "== $lib"
  export IU_B200_LIB=$PWD/interpolate_unstructured_b200/$lib/libiu_b200.so
  $B $BENCH_EXTRA | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('value %.3e q/s  kernel %.2f ms/launch  share %.2f sort %.2f'%(d['value'], r['kernel_ms_per_launch'], r['kernel_share_of_step'], r['sort_share_of_step']))"
  if [ -z "$NO_NCU" ]; then
  $P $BENCH_EXTRA > /dev/null 2>&1 && ncu --metrics $M --clock-control none -k regex:k_query -s 3 -c 1 --csv --log-file gpurun_out/${tag}_${lib}.csv $P $BENCH_EXTRA > /dev/null 2>&1
  grep -E "dram__|hit_rate|time_duration|inst_executed|throughput|warps_active|l1tex__|issue_active" gpurun_out/${tag}_${lib}.csv | awk -F'","' '{print "   ", $(NF-2), $(NF-1), $NF}'
  fi
done 2>&1 | tee gpurun_out/${tag}.txt
