#!/bin/bash
# r03g: bench lines of the other BASELINE configurations on one GPU with HEAD (3 CTAs per SM, 64-query tiles, device tree build)
mkdir -p gpurun_out
for wl in C1 C2 C4 C5; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 > gpurun_out/r03g_${wl}_bench.json 2> gpurun_out/r03g_${wl}_bench.err; echo "$wl rc=$?"
  python -c "
import json
d=json.load(open('gpurun_out/r03g_${wl}_bench.json')); r=d['roofline']; c=d['cpu_baseline']; e=d['e2e']; k=d['checks']
print('$wl value %.3e q/s  k_query %.2f ms/launch  frac %.4f  e2e %.3e  cpu all-cores %.3e (%d) serial %.3e  setup %.2f s  oracle sample %d mismatches %d bit-identical %s'%(d['value'], r['kernel_ms_per_launch'], r['frac'], e['value'], c['value'], c['cores'], c['serial']['value'], d['run']['setup_seconds'], k['oracle_sample'], k['oracle_cell_mismatches'], k['oracle_values_bit_identical']))" || tail -5 gpurun_out/r03g_${wl}_bench.err
done
