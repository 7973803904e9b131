#!/bin/bash
# r02d: GPU tests (with the round-2 cases), the new bench line (C3 + reference arm), DRAM bytes of every k_query form
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02d_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02d_pytest_gpu.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02d_ref.json 2> gpurun_out/r02d_ref.err; echo "ref rc=$?"; tail -c 1500 gpurun_out/r02d_ref.json; tail -5 gpurun_out/r02d_ref.err
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r02d_bench.json 2> gpurun_out/r02d_bench.err; echo "bench rc=$?"; cat gpurun_out/r02d_bench.json; tail -5 gpurun_out/r02d_bench.err
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct
timeout 900 ncu --metrics $M --clock-control none -k regex:k_query --csv --log-file gpurun_out/r02d_phases_ncu.csv python tools/phase_bench.py --reps 1 > gpurun_out/r02d_phases_ncu.log 2>&1; echo "ncu phases rc=$?"
python - <<'PY'
import csv, io
rows = list(csv.DictReader(io.StringIO("".join(l for l in open("gpurun_out/r02d_phases_ncu.csv") if l.startswith('"')))))
by = {}
for r in rows:
    by.setdefault(r["ID"], {"k": r["Kernel Name"][:60]})[r["Metric Name"]] = r["Metric Value"]
for i, d in by.items():
    print(i, d)
PY
