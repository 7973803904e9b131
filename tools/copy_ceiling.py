#!/usr/bin/env python3
"""Plain-copy ceiling of the host-buffer leg: the bytes bench.py's e2e call moves per query
(28 B host -> device, 36 B device -> host for C3), as concurrent cudaMemcpyAsync of pinned buffers on two streams in
chunks of 4M queries, NO kernels.  One process per GPU:

    python tools/copy_ceiling.py                                   (1 GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/copy_ceiling.py

Prints one JSON line: query-equivalents/s summed over the ranks (max-over-ranks time) and GB/s per rank and direction;
also each direction alone.  bench.py's e2e value divided by `both.queries_per_s` is the fraction of the ceiling reached."""
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    nq, chunk = 100_000_000, 1 << 22
    h_in = torch.empty(nq * 28, dtype=torch.uint8, pin_memory=True); h_in.zero_()
    h_out = torch.empty(nq * 36, dtype=torch.uint8, pin_memory=True); h_out.zero_()
    d_in = [torch.empty(chunk * 28, dtype=torch.uint8, device=dev) for _ in range(3)]
    d_out = [torch.zeros(chunk * 36, dtype=torch.uint8, device=dev) for _ in range(3)]
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

    def run(do_in, do_out):
        k = 0
        for off in range(0, nq, chunk):
            m = min(chunk, nq - off)
            if do_in:
                with torch.cuda.stream(s_in):
                    d_in[k % 3][:m * 28].copy_(h_in[off * 28:(off + m) * 28], non_blocking=True)
            if do_out:
                with torch.cuda.stream(s_out):
                    h_out[off * 36:(off + m) * 36].copy_(d_out[k % 3][:m * 36], non_blocking=True)
            k += 1
        torch.cuda.synchronize()

    def timed(do_in, do_out, reps=3):
        run(do_in, do_out)
        best = 1e9
        for _ in range(reps):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            run(do_in, do_out)
            t = time.perf_counter() - t0
            if world > 1:
                tt = torch.tensor([t], dtype=torch.float64, device=dev)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                t = float(tt.item())
            best = min(best, t)
        return best

    res = {}
    for name, (a, b) in {"both": (True, True), "h2d_only": (True, False), "d2h_only": (False, True)}.items():
        t = timed(a, b)
        res[name] = {"seconds": t, "queries_per_s": world * nq / t, "h2d_gbs_per_rank": nq * 28 / t / 1e9 if a else 0.0,
                     "d2h_gbs_per_rank": nq * 36 / t / 1e9 if b else 0.0}
    if rank == 0:
        print(json.dumps({"what": "plain pinned cudaMemcpyAsync ceiling, 28 B in + 36 B out per query-equivalent, 4M-query chunks, no kernels",
                          "n_gpus": world, "queries_per_rank": nq, **res}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
