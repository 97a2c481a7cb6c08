#!/usr/bin/env python
"""Per-kernel device time INSIDE graph replays of the train step (CUPTI through torch.profiler): which kernels a
launch-bound workload's step is made of once launches are free.  usage: graph_kernels.py c1|c2 [steps]"""
import collections, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
from helpers import make_args
from bsarec_b200 import BSARecModel, FusedAdam, GraphedTrainStep

name = sys.argv[1] if len(sys.argv) > 1 else "c1"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
desc, cfg, B, _ = bench.WORKLOADS[name]
dev = torch.device("cuda", 0)
torch.manual_seed(42)
model = BSARecModel(make_args(cfg, 0.5, 0.5, batch_size=B)).to(dev).train()
opt = FusedAdam(model.parameters(), lr=1e-3)
batches = [{k: v.to(dev) for k, v in b.items()} for b in bench.synth_batches(cfg, B, 1234, 4)]
step = GraphedTrainStep(model, opt)
scratch = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for i in range(5):
    step(batches[i % 4]["ids"], batches[i % 4]["answers"])
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(steps):
        scratch.fill_(1)
        step(batches[i % 4]["ids"], batches[i % 4]["answers"])
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
agg = collections.OrderedDict()
def short(n):
    n = n.replace("(anonymous namespace)::", "").replace("void ", "")
    n = re.sub(r"\(.*$", "", n).replace("bsa::umma::", "").replace("bsa::", "")
    return n[:100]
tot = 0.0
for e in ev:
    n = short(e.name)
    if "FillFunctor<unsigned char>" in e.name or ("Memset" in n and e.device_time > 20):
        continue
    a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += e.device_time; tot += e.device_time
print(f"{name}: {steps} replays, kernel time {tot / steps / 1e3:.3f} ms/step, {sum(a[0] for a in agg.values()) / steps:.0f} kernels/step")
for n, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{us / steps:8.1f} us {100 * us / tot:5.1f}%  x{c / steps:<5.1f} {us / c:6.1f} us each  {n}")
