#!/bin/bash
# r02g: bench lines of the other BASELINE configurations on one GPU, then compute-sanitizer over tools/sanitize_smoke.py
mkdir -p gpurun_out
for wl in C1 C2 C4 C5; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 > gpurun_out/r02g_${wl}_bench.json 2> gpurun_out/r02g_${wl}_bench.err; echo "$wl rc=$?"
  python -c "
import json
d=json.load(open('gpurun_out/r02g_${wl}_bench.json')); r=d['roofline']; c=d['cpu_baseline']; e=d['e2e']
print('$wl value %.3e q/s  k_query %.2f ms/launch  frac %.4f  e2e %.3e  cpu all-cores %.3e (%d) serial %.3e  checks %s'%(d['value'], r['kernel_ms_per_launch'], r['frac'], e['value'], c['value'], c['cores'], c['serial']['value'], d['checks']))" || tail -5 gpurun_out/r02g_${wl}_bench.err
done
python tools/sanitize_smoke.py > gpurun_out/r02g_smoke_plain.log 2>&1; echo "smoke plain rc=$?"; tail -3 gpurun_out/r02g_smoke_plain.log
for tool in memcheck racecheck initcheck synccheck; do
  IU_SANITIZE_QUERIES=4000 timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 python tools/sanitize_smoke.py > gpurun_out/r02g_sanitizer_$tool.log 2>&1; echo "$tool rc=$?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize smoke ok|Error|hazard" gpurun_out/r02g_sanitizer_$tool.log | head -8
done
CMD="python bench.py --queries 2e8 --steps 1 --warmup 1 --no-cpu --no-e2e"
$CMD > gpurun_out/r02g_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02g_launches.csv $CMD > gpurun_out/r02g_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_query -s 3 -c 1 -f -o gpurun_out/r02g_k_query $CMD > gpurun_out/r02g_ncu_full.log 2>&1
echo "ncu full rc=$?"
