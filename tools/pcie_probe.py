#!/usr/bin/env python
"""Aggregate host<->device copy throughput with one process per GPU (torchrun): what the box gives the end-to-end path.
    python -m torch.distributed.run --nproc-per-node N tools/pcie_probe.py"""
import os, time, torch, torch.distributed as dist
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 500 << 20
h_out, h_in = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n // 4, dtype=torch.uint8).pin_memory()
d_out, d_in = torch.empty(n, dtype=torch.uint8, device="cuda"), torch.empty(n // 4, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for mode in ("d2h", "h2d", "both"):
    for it in range(2):
        if world > 1: dist.barrier()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(6):
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s1): h_out.copy_(d_out, non_blocking=True)
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s2): d_in.copy_(h_in, non_blocking=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    gb = 6 * ((n if mode != "h2d" else 0) + (n // 4 if mode != "d2h" else 0)) / 1e9
    t = torch.tensor([gb / dt], device="cuda")
    if world > 1: dist.all_reduce(t)
    if rank == 0: print(f"{mode}: {float(t):.1f} GB/s aggregate over {world} GPU(s), {float(t) / world:.1f} per GPU")
# the same traffic over a larger pinned footprint (three rotating buffer sets, as the end-to-end loop of bench.py uses)
outs = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(3)]
ins = [torch.empty(n // 4, dtype=torch.uint8).pin_memory() for _ in range(3)]
for it in range(2):
    if world > 1: dist.barrier()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for k in range(6):
        with torch.cuda.stream(s1): outs[k % 3].copy_(d_out, non_blocking=True)
        with torch.cuda.stream(s2): d_in.copy_(ins[k % 3], non_blocking=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
t = torch.tensor([6 * (n + n // 4) / 1e9 / dt], device="cuda")
per = [torch.zeros_like(t) for _ in range(world)]
if world > 1: dist.all_gather(per, t)
if world > 1: dist.all_reduce(t)
if rank == 0: print(f"both, 3 rotating sets: {float(t):.1f} GB/s aggregate over {world} GPU(s); per rank: " + " ".join(f"{float(x):.1f}" for x in per))
# ... and with a bandwidth-bound kernel running beside the copies
big = torch.empty(2 << 30, dtype=torch.uint8, device="cuda")
s3 = torch.cuda.Stream()
if world > 1: dist.barrier()
torch.cuda.synchronize(); t0 = time.perf_counter()
with torch.cuda.stream(s3):
    for _ in range(40): big.add_(1)
for k in range(6):
    with torch.cuda.stream(s1): outs[k % 3].copy_(d_out, non_blocking=True)
    with torch.cuda.stream(s2): d_in.copy_(ins[k % 3], non_blocking=True)
s1.synchronize(); s2.synchronize(); dt = time.perf_counter() - t0
t = torch.tensor([6 * (n + n // 4) / 1e9 / dt], device="cuda")
if world > 1: dist.all_reduce(t)
if rank == 0: print(f"both, 3 sets, HBM-bound kernel beside: {float(t):.1f} GB/s aggregate over {world} GPU(s)")
torch.cuda.synchronize()
if world > 1: dist.destroy_process_group()
