#!/bin/bash
# r02h (8 GPUs of one box): C5 as BASELINE.json states it (torchrun, 8 ranks x 5e8 queries), the single-process multi-GPU
# handle, the plain-copy ceiling at N = 1/2/4/8, the two-GPU parity test, and the checked build's smoke run
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8; nproc
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
echo "== two-GPU handle test"; timeout 600 python -m pytest tests/test_gpu_round2.py -m gpu -q -k "multi" 2>&1 | tail -3
echo "== checked build smoke"; IU_B200_LIB=$PWD/interpolate_unstructured_b200/lib_checked/libiu_b200.so timeout 900 python tools/sanitize_smoke.py > gpurun_out/r02h_checked_smoke.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r02h_checked_smoke.log
echo "== C5 reference arm"; timeout 900 $TR --nproc-per-node 8 --master-port 29501 bench.py --gpus 8 --workload C5 --impl reference --steps 2 --warmup 1 > gpurun_out/r02h_c5_n8_ref.json 2> gpurun_out/r02h_c5_n8_ref.err; echo "rc=$?"; tail -c 600 gpurun_out/r02h_c5_n8_ref.json
echo "== C5 N=8"; timeout 1500 $TR --nproc-per-node 8 --master-port 29502 bench.py --gpus 8 --workload C5 --steps 3 --warmup 3 > gpurun_out/r02h_c5_n8_bench.json 2> gpurun_out/r02h_c5_n8_bench.err; echo "rc=$?"; cat gpurun_out/r02h_c5_n8_bench.json; tail -3 gpurun_out/r02h_c5_n8_bench.err
echo "== C3 N=8 (short)"; timeout 900 $TR --nproc-per-node 8 --master-port 29503 bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu > gpurun_out/r02h_c3_n8_bench.json 2> gpurun_out/r02h_c3_n8_bench.err; echo "rc=$?"; cat gpurun_out/r02h_c3_n8_bench.json; tail -3 gpurun_out/r02h_c3_n8_bench.err
echo "== one process, one handle, 8 GPUs: C3"; timeout 900 python tools/multi_bench.py --workload C3 --gpus 8 > gpurun_out/r02h_multi_c3_n8.json 2> gpurun_out/r02h_multi_c3_n8.err; echo "rc=$?"; cat gpurun_out/r02h_multi_c3_n8.json; tail -3 gpurun_out/r02h_multi_c3_n8.err
echo "== one process, one handle, 8 GPUs: C5 (set-up time)"; timeout 900 python tools/multi_bench.py --workload C5 --gpus 8 --queries 1e8 --e2e-queries 2e8 --steps 2 --warmup 1 > gpurun_out/r02h_multi_c5_n8.json 2> gpurun_out/r02h_multi_c5_n8.err; echo "rc=$?"; cat gpurun_out/r02h_multi_c5_n8.json; tail -3 gpurun_out/r02h_multi_c5_n8.err
echo "== copy ceiling"; for n in 1 2 4 8; do
  if [ $n = 1 ]; then timeout 300 python tools/copy_ceiling.py; else timeout 300 $TR --nproc-per-node $n --master-port 2951$n tools/copy_ceiling.py 2>/dev/null; fi
done | tee gpurun_out/r02h_copy_ceiling.jsonl
