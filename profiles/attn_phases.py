#!/usr/bin/env python
"""Where one CTA of the tcgen05 attention kernels spends its time: globaltimer stamps at the phase boundaries of one
mid-grid CTA per q-tile, taken during a warm C3 train step.  Needs the instrumented build:

    BSA_EXTRA_NVCC_FLAGS=-DBSA_ATTN_STAMPS python -m bsarec_b200.build --force
    BSA_EXTRA_NVCC_FLAGS=-DBSA_ATTN_STAMPS python profiles/attn_phases.py > profiles/<round>_attn_phases.txt
"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

import bench
from bsarec_b200 import BSARecModel, FusedAdam
from bsarec_b200.build import LIB
from helpers import make_args

desc, cfg, B, _ = bench.WORKLOADS["c3"]
dev = torch.device("cuda:0")
torch.manual_seed(42)
model = BSARecModel(make_args(cfg, bench.P_DROP, bench.P_DROP, batch_size=B)).to(dev)
optim = FusedAdam(model.parameters(), lr=1e-3)
batches = [{k: v.to(dev) for k, v in b.items()} for b in bench.synth_batches(cfg, B, 1234, 2)]
model.train(True)
for i in range(4):
    b = batches[i % 2]
    loss = model.calculate_loss(b["ids"], b["answers"], None, None, None)
    optim.zero_grad()
    loss.backward()
    optim.step()
torch.cuda.synchronize()
lib = ctypes.CDLL(LIB)
buf = (ctypes.c_ulonglong * 64)()
rc = lib.bsa_debug_attn_stamps(buf)
assert rc == 0, rc
FWD = [(7, "kernel entry"), (0, "ids -> valid[], TMEM alloc, barrier init + TMA issue"), (1, "wait Q, K tiles"), (2, "split Q, K -> (hi, lo)"),
       (3, "MMA 1 S = Q K^T (+ wait V)"), (4, "softmax + dropout, P pair -> TMEM, V split"), (5, "MMA 2 O = P V"), (6, "epilogue ctx, lse")]
BWD = [(31, "kernel entry"), (16, "ids -> valid[], TMEM alloc, barrier init + TMA issue"), (17, "wait Q, K tiles"), (18, "split Q, K"),
       (19, "MMA 1 S = Q K^T (+ delta row sums)"), (20, "reload dO, V by TMA"), (21, "split dO, V"), (22, "MMA 2 dPd = dO V^T (+ wait K rows)"),
       (23, "split K rows, dS pass -> TMEM"), (24, "MMA 3 dQ = dS K"), (25, "epilogue dQ")]
print(f"{desc}: phase times of one mid-grid CTA (us, from %globaltimer; ~32 ns resolution)")
for name, table in (("attn_fwd_umma_kernel", FWD), ("attn_bwd_q_umma_kernel", BWD)):
    for qt in range(2):
        t = [buf[qt * 32 + i] for i, _ in table]
        print(f"-- {name}, q-tile {qt}: {(t[-1] - t[0]) / 1e3:.2f} us entry -> last barrier")
        for k in range(1, len(table)):
            print(f"   {(t[k] - t[k - 1]) / 1e3:7.2f}  {table[k][1]}")
