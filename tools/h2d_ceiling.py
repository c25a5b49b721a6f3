#!/usr/bin/env python3
"""Host -> device ceiling of the box, measured the way the strip pipelines load it: raw pinned cudaMemcpyAsync on
1 .. N GPUs at once, one process per GPU (the e2e 1 -> 8 curve is bounded by this, not by a kernel).

    python tools/h2d_ceiling.py                                   # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        tools/h2d_ceiling.py --out profiles/r2_h2d_ceiling_nN.json

Per configuration every participating rank copies the same 1 GiB pinned buffer `--reps` times, bracketed by a barrier
on both sides; the rate is bytes of all participants / the slowest participant's device time (CUDA events).  Runs:
all ranks together, every rank alone (its own root port / NUMA distance), and every PAIR (a pair that shares a PCIe
switch or root port shows less than twice the single rate).  Rank 0 prints / writes one JSON object.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=float, default=1.0)
    ap.add_argument("--reps", type=int, default=8)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-pairs", action="store_true")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    try:
        from openeo_processes_dask_b200 import sharding

        sharding.bind_rank_to_gpu_numa(local)  # the binding the bench uses
    except Exception:
        pass
    n = int(args.gib * 2**30) // 4
    host = torch.empty(n, dtype=torch.float32).pin_memory()
    host.fill_(1.0)
    dev = torch.empty(n, dtype=torch.float32, device="cuda")
    flag = torch.zeros(1, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.all_reduce(flag)
            torch.cuda.synchronize()

    def run(participants):
        mine = rank in participants
        if mine:
            for _ in range(2):
                dev.copy_(host, non_blocking=True)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        if mine:
            for _ in range(args.reps):
                dev.copy_(host, non_blocking=True)
        b.record()
        barrier()
        ms = torch.tensor([a.elapsed_time(b) if mine else 0.0], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = float(ms.item())
        return len(participants) * args.reps * n * 4 / (ms * 1e-3) / 1e9

    out = {"world": world, "gib_per_copy": args.gib, "reps": args.reps, "unit": "GB/s",
           "all_ranks_together": run(list(range(world)))}
    out["per_rank_alone"] = [run([r]) for r in range(world)]
    if world > 1 and not args.no_pairs:
        out["pairs"] = {f"{i}-{j}": run([i, j]) for i in range(world) for j in range(i + 1, world)}
    if world >= 4:
        out["first_half_together"] = run(list(range(world // 2)))
        out["second_half_together"] = run(list(range(world // 2, world)))
    if rank == 0:
        txt = json.dumps(out)
        print(txt)
        if args.out:
            with open(args.out, "w") as f:
                f.write(txt + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
