#!/bin/bash
mkdir -p gpurun_out
P="python bench.py --queries 7e7 --steps 1 --warmup 1 --no-cpu --no-e2e --dynamic 2"
$P > gpurun_out/exp6_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_query -s 2 -c 1 -f -o gpurun_out/exp6_packet $P > gpurun_out/exp6_ncu.log 2>&1
echo rc=$?
