#!/usr/bin/env python
"""Timeline of one data-parallel C3 train step on rank 0 (CUPTI through torch.profiler): which kernels the NCCL
collectives overlap and what is left exposed at the end of the backward pass.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29530 \
        profiles/dp_timeline.py > profiles/<round>_dp_timeline_nN.txt
"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.distributed as dist
from torch.profiler import ProfilerActivity, profile

import bench
from bsarec_b200 import BSARecModel, FusedAdam
from helpers import make_args

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
out = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
desc, cfg, B, _ = bench.WORKLOADS["c3"]
torch.manual_seed(42)
model = BSARecModel(make_args(cfg, bench.P_DROP, bench.P_DROP, batch_size=B)).to(dev)
model.enable_data_parallel()
optim = FusedAdam(model.parameters(), lr=1e-3)
batches = [{k: v.to(dev) for k, v in b.items()} for b in bench.synth_batches(cfg, B, 1234 + rank, 2)]


def step(b):
    loss = model.calculate_loss(b["ids"], b["answers"], None, None, None)
    optim.zero_grad()
    loss.backward()
    optim.step()


model.train()
for i in range(4):
    step(batches[i % 2])
torch.cuda.synchronize()
dist.barrier()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step(batches[0])
    torch.cuda.synchronize()
if rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    t0 = evs[0].time_range.start
    end = max(e.time_range.end for e in evs)
    print(f"world {world}: one train step on rank 0, {(end - t0) / 1e3:.3f} ms from first kernel start to last kernel end", file=out)
    print("  start_ms    dur_ms  kernel (>= 40 us, and every nccl / memcpy)", file=out)
    for e in evs:
        name = re.sub(r"\(.*$", "", e.name.replace("(anonymous namespace)::", "")).replace("bsa::umma::", "").replace("bsa::", "")
        dur = e.time_range.end - e.time_range.start
        if dur >= 40 or "nccl" in name.lower() or "memcpy" in name.lower():
            print(f"{(e.time_range.start - t0) / 1e3:10.3f} {dur / 1e3:9.3f}  {name[:90]}", file=out)
dist.destroy_process_group()
