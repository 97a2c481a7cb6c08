#!/usr/bin/env python
"""Warm per-kernel breakdown of one C3 train step and one eval step (CUPTI activity trace through torch.profiler:
kernels run back to back with warm caches, unlike the serialised cold-cache ncu launch list).

    python profiles/step_breakdown.py [workload] [steps] > profiles/<round>_step_breakdown.txt
"""
import collections
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from torch.profiler import ProfilerActivity, profile

import bench
from bsarec_b200 import BSARecModel, FusedAdam
from helpers import make_args

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
desc, cfg, B, _ = bench.WORKLOADS[wl]
if len(sys.argv) > 3:
    B = int(sys.argv[3])
dev = torch.device("cuda:0")
torch.manual_seed(42)
model = BSARecModel(make_args(cfg, bench.P_DROP, bench.P_DROP, batch_size=B)).to(dev)
optim = (torch.optim.Adam if os.environ.get('OPT') == 'torch' else FusedAdam)(model.parameters(), lr=1e-3)
batches = [{k: v.to(dev) for k, v in b.items()} for b in bench.synth_batches(cfg, B, 1234, 2)]


def train_step(b):
    loss = model.calculate_loss(b["ids"], b["answers"], None, None, None)
    optim.zero_grad()
    loss.backward()
    optim.step()


def eval_step(b):
    return model.full_sort_topk(b["ids"], None, b["indptr"], b["indices"], k=20)


def short(name):
    name = name.replace("(anonymous namespace)::", "")
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("bsa::umma::", "").replace("bsa::", "")
    name = name.replace("void ", "")
    return name[:100]


def run(label, fn, train):
    model.train(train)
    for i in range(3):
        fn(batches[i % 2])
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for i in range(steps):
            fn(batches[i % 2])
        torch.cuda.synchronize()
    agg = collections.OrderedDict()
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            a = agg.setdefault(short(ev.name), [0, 0.0])
            a[0] += 1
            a[1] += ev.device_time
    tot = sum(v[1] for v in agg.values())
    print(f"== {label}: {desc}, B={B}; {steps} steps; sum of kernel time {tot / steps / 1e3:.3f} ms/step")
    for n, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{us / steps / 1e3:9.3f} ms {100 * us / tot:5.1f}%  x{c / steps:<6.1f} {n}")


run("train step (fwd+bwd+Adam)", train_step, True)
run("eval step (encoder + score + top-20)", eval_step, False)
