"""Host<->device copy bandwidth per rank, alone and all ranks at once, with
the pinned buffers allocated before / after binding the rank to its GPU's
local CPUs (distributed.bind_to_gpu_numa).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N \
        --master-addr 127.0.0.1 --master-port 29533 tools/diag_pcie.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ocsmesh_b200 import distributed as D

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
dev = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{dev}"))
NB = 1 << 30


def bw(host, devbuf, both):
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    out = torch.empty_like(host).pin_memory() if both else None
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_stream(torch.cuda.current_stream())
    s2.wait_stream(torch.cuda.current_stream())
    for _ in range(3):
        with torch.cuda.stream(s1):
            devbuf.copy_(host, non_blocking=True)
        if both:
            with torch.cuda.stream(s2):
                out.copy_(devbuf2, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)
    e1.record()
    torch.cuda.synchronize()
    return 3 * NB / (e0.elapsed_time(e1) * 1e-3) / 1e9


devbuf = torch.empty(NB, dtype=torch.uint8, device="cuda")
devbuf2 = torch.zeros(NB, dtype=torch.uint8, device="cuda")
res = {"rank": rank, "affinity_before": len(os.sched_getaffinity(0))}
for phase in ("unbound", "bound"):
    if phase == "bound":
        res["numa"] = D.bind_to_gpu_numa(dev)
    host = torch.empty(NB, dtype=torch.uint8).pin_memory()
    host.fill_(1)
    res[f"{phase}_all_h2d_GBs"] = round(bw(host, devbuf, False), 1)
    res[f"{phase}_all_duplex_GBs_each_way"] = round(bw(host, devbuf, True), 1)
    del host
# one rank at a time (bound buffers)
host = torch.empty(NB, dtype=torch.uint8).pin_memory()
for r in range(world):
    if world > 1:
        dist.barrier()
    if r == rank:
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            devbuf.copy_(host, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        res["alone_h2d_GBs"] = round(3 * NB / (e0.elapsed_time(e1) * 1e-3) / 1e9, 1)
    if world > 1:
        dist.barrier()
print(json.dumps(res), flush=True)
if world > 1:
    dist.destroy_process_group()
