#!/usr/bin/env python
"""bench.py — BSARec hot-path throughput on B200 (train seqs/s fwd+bwd+Adam and full-ranking eval users/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c3|c2|c1|c5]

N>1 is launched by torchrun (one rank per GPU, NCCL).  One JSON line is printed by rank 0; its keys follow
the driver's contract.  A "step" is one pass of the hot path over one synthetic batch:
  train: calculate_loss -> zero_grad -> backward -> [gradient all-reduce] -> Adam.step   (trainers.py:103-107)
  eval : ids -> encoder -> full-catalog scores -> history mask -> top-20                  (trainers.py:126-149)
`value` times the train step with the batches already resident in HBM; `e2e` times the same step through the
public API from pinned HOST batches (H2D of ids/answers inside the timed region, D2H read of the loss).
`--impl reference` times the CPU restatement of the reference (oracle/, kind "port") on the host cores.
"""
import argparse
import json
import os
import sys
import threading
import time

_json_out = sys.stdout

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (description, cfg kwargs, per-GPU batch, cpu sample batch)
    "c3": ("C3 synthetic 1M-item catalog, L=200, d=256, 4 heads, 2 layers, c=5, batch 512 per GPU (4096 on 8)",
           dict(item_size=1_000_000, hidden_size=256, max_seq_length=200, num_attention_heads=4, num_hidden_layers=2,
                c=5, alpha=0.7), 512, 32),
    "c2": ("C2 ML-1M shape: 3417 items, L=200, d=64, 2 heads, 2 layers, c=5, batch 256 per GPU",
           dict(item_size=3417, hidden_size=64, max_seq_length=200, num_attention_heads=2, num_hidden_layers=2,
                c=5, alpha=0.7), 256, 64),
    "c1": ("C1 Beauty shape: 12102 items, L=50, d=64, 1 head, 2 layers, c=5, alpha=0.7, batch 256 per GPU",
           dict(item_size=12102, hidden_size=64, max_seq_length=50, num_attention_heads=1, num_hidden_layers=2,
                c=5, alpha=0.7), 256, 256),
    "c5": ("C5 long sequence: 12102 items, L=1024, d=128, 2 heads, 2 layers, c=32, batch 128 per GPU",
           dict(item_size=12102, hidden_size=128, max_seq_length=1024, num_attention_heads=2, num_hidden_layers=2,
                c=32, alpha=0.7), 128, 8),
}
P_DROP = 0.5        # both dropout probabilities, the reference's defaults (utils.py:86-87)
TOPK = 20
NBATCH = 4          # distinct synthetic batches cycled through


def flops_train_per_seq(c):
    L, d, P, N = c["max_seq_length"], c["hidden_size"], c["num_hidden_layers"], c["item_size"]
    return 3.0 * (P * (24.0 * L * d * d + 6.0 * L * L * d) + 2.0 * N * d)      # SURVEY.md §8d canonical count


def flops_eval_per_user(c):
    L, d, P, N = c["max_seq_length"], c["hidden_size"], c["num_hidden_layers"], c["item_size"]
    return P * (24.0 * L * d * d + 6.0 * L * L * d) + 2.0 * N * d


def synth_batches(cfg, B, seed, n):
    """Seeded synthetic batches: lengths U{1..L} left-padded, ids U{1..N-1}, one all-padding
    row per 8 (Beauty has 15 % such rows), answers U{1..N-1}; history = the user's own non-zero ids."""
    import numpy as np
    import torch
    g = torch.Generator().manual_seed(seed)
    L, N = cfg["max_seq_length"], cfg["item_size"]
    out = []
    for _ in range(n):
        lens = torch.randint(1, L + 1, (B,), generator=g)
        lens[1::8] = 0
        ids = torch.randint(1, N, (B, L), generator=g)
        pos = torch.arange(L).unsqueeze(0)
        ids = torch.where(pos >= (L - lens).unsqueeze(1), ids, torch.zeros_like(ids))
        answers = torch.randint(1, N, (B,), generator=g)
        srt = torch.sort(ids, dim=1).values
        keep = (srt > 0) & torch.cat([torch.ones(B, 1, dtype=torch.bool), srt[:, 1:] != srt[:, :-1]], dim=1)
        cnt = keep.sum(1)
        indptr = torch.zeros(B + 1, dtype=torch.int64)
        indptr[1:] = torch.cumsum(cnt, 0)
        indices = srt[keep]
        if indices.numel() == 0:
            indices = torch.zeros(1, dtype=torch.int64)
        out.append(dict(ids=ids, answers=answers, indptr=indptr, indices=indices))
    return out


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons every 200 ms while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                     getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                     getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                     getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap"}
            while not self.stop_flag:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
                time.sleep(0.2)
        except Exception as e:   # never let monitoring break the bench
            self.reasons.add(f"sampler_error:{type(e).__name__}")

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def run_reference(args, desc, cfg, cpu_B):
    """The reference's algorithm on the host cores: oracle port, all threads, bounded sample per step."""
    import torch
    import oracle
    torch.set_num_threads(os.cpu_count())
    ocfg = oracle.OracleConfig(hidden_dropout_prob=P_DROP, attention_probs_dropout_prob=P_DROP, **cfg)
    P = oracle.init_params(ocfg, seed=42)
    step = oracle.adam_train_step(P, ocfg)
    batches = synth_batches(cfg, cpu_B, 1234, max(2, min(NBATCH, args.steps)))
    for i in range(args.warmup):
        step(batches[i % len(batches)]["ids"], batches[i % len(batches)]["answers"])
    t0 = time.perf_counter()
    for i in range(args.steps):
        b = batches[i % len(batches)]
        step(b["ids"], b["answers"])
    dt = time.perf_counter() - t0
    import numpy as np
    b = batches[0]
    Pd = {k: v.detach() for k, v in step.P.items()}
    oracle.eval_step(Pd, b["ids"], ocfg, b["indptr"].numpy(), b["indices"].numpy(), np.arange(cpu_B), TOPK)
    t1 = time.perf_counter()
    neval = max(1, min(3, args.steps))
    for _ in range(neval):
        oracle.eval_step(Pd, b["ids"], ocfg, b["indptr"].numpy(), b["indices"].numpy(), np.arange(cpu_B), TOPK)
    dte = time.perf_counter() - t1
    val = args.steps * cpu_B / dt
    sample = f"{args.steps} train steps of {cpu_B} sequences + {neval} eval batches of {cpu_B} users, torch CPU, {os.cpu_count()} threads"
    line = {
        "impl": "reference", "metric": "train seqs/sec (fwd+bwd+Adam)", "value": val, "unit": "seqs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "cpu_batch": cpu_B},
        "eval": {"value": neval * cpu_B / dte, "unit": "users/s"},
        "cpu_baseline": {"value": val, "unit": "seqs/s", "cores": os.cpu_count(), "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "seqs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=_json_out, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ce-impl", type=int, default=0)
    ap.add_argument("--dp", default="overlap", choices=["overlap", "bucket"],
                    help="N>1 gradient exchange: overlap = GradExchange inside backward (dense CE part all-reduced under the "
                         "encoder backward, sparse rows all-gathered); bucket = one flat all-reduce after backward")
    ap.add_argument("--optimizer", default="fused", choices=["fused", "torch"],
                    help="fused: bsarec_b200.FusedAdam (one kernel pass per tensor); torch: torch.optim.Adam (foreach)")
    args = ap.parse_args()
    # stdout carries the one JSON line and nothing else: native libraries that write to fd 1 (NCCL prints its version banner
    # there when NCCL_DEBUG is set) are pointed at stderr, the JSON line goes to the saved descriptor
    global _json_out
    sys.stdout.flush()
    _json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    desc, cfg, B, cpu_B = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank == 0:
            run_reference(args, desc, cfg, cpu_B)
        return

    import torch
    import torch.distributed as dist
    from bsarec_b200 import BSARecModel, FusedAdam, _cabi
    from bsarec_b200.parallel import allreduce_gradients
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import make_args

    if args.warmup < 3:
        args.warmup = 3
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(42)
    model = BSARecModel(make_args(cfg, P_DROP, P_DROP, batch_size=B)).to(dev)
    model.ce_impl = args.ce_impl
    if world > 1 and args.dp == "overlap":
        model.enable_data_parallel()
    opt_cls = FusedAdam if args.optimizer == "fused" else torch.optim.Adam
    optim = opt_cls(model.parameters(), lr=1e-3, betas=(0.9, 0.999), weight_decay=0.0)
    params = list(model.parameters())
    host = synth_batches(cfg, B, 1234 + rank, NBATCH)
    pinned = [{k: v.pin_memory() for k, v in b.items()} for b in host]
    resident = [{k: v.to(dev) for k, v in b.items()} for b in host]

    def train_step(b):
        loss = model.calculate_loss(b["ids"], b["answers"], None, None, None)
        optim.zero_grad()
        loss.backward()
        if world > 1 and args.dp == "bucket":
            allreduce_gradients(params)
        optim.step()
        return loss

    def e2e_step(hb):
        ids = hb["ids"].to(dev, non_blocking=True)
        answers = hb["answers"].to(dev, non_blocking=True)
        return train_step({"ids": ids, "answers": answers}).item()

    if world > 1:
        # item-sharded evaluation: users split across ranks for the encoder, every rank scores ALL users of
        # the global batch against its row shard of the item table, candidate lists are all-gathered and merged.
        from bsarec_b200 import functional as Fn
        from bsarec_b200.parallel import shard_range, sharded_full_sort_topk
        lo, hi = shard_range(cfg["item_size"], world, rank)
        all_host = [synth_batches(cfg, B, 1234 + r, NBATCH) for r in range(world)]
        for i in range(NBATCH):          # global history CSR, rank-major, same on every rank
            ptrs, idxs, base = [torch.zeros(1, dtype=torch.int64)], [], 0
            for r in range(world):
                hb = all_host[r][i]
                n = int(hb["indptr"][-1])
                ptrs.append(hb["indptr"][1:] + base)
                idxs.append(hb["indices"][:n])
                base += n
            gi = torch.cat(idxs) if base > 0 else torch.zeros(1, dtype=torch.int64)
            resident[i]["g_indptr"], resident[i]["g_indices"] = torch.cat(ptrs).to(dev), gi.to(dev)
            pinned[i]["g_indptr"], pinned[i]["g_indices"] = torch.cat(ptrs).pin_memory(), gi.pin_memory()
        del all_host

        def eval_step(b):
            s_loc = model.last_hidden(b["ids"])
            shard = model.item_embeddings.weight.detach()[lo:hi]
            return sharded_full_sort_topk(s_loc, shard, lo, b["g_indptr"], b["g_indices"], TOPK,
                                          lambda s_, e_, ip, ix, k, off: Fn.score_topk(s_, e_, ip, ix, k, off, 0, cache=model._topk_cache), Fn.topk_merge)
    else:
        def eval_step(b):
            return model.full_sort_topk(b["ids"], None, b["indptr"], b["indices"], k=TOPK)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        """n calls of fn(i) bracketed by barrier+sync; device time by CUDA events, max over ranks."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        barrier()
        return float(ms.item())

    model.train()
    for i in range(args.warmup):
        train_step(resident[i % NBATCH])
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = _cabi.launch_count()
    ms_train = timed(lambda i: train_step(resident[i % NBATCH]), args.steps)
    launches = _cabi.launch_count() - l0
    sampler.stop_flag = True
    # same region again with per-kernel event timing switched on (roofline of the dominant kernel)
    _cabi.profile(True)
    timed(lambda i: train_step(resident[i % NBATCH]), args.steps)
    prof = _cabi.profile_results()
    _cabi.profile(False)
    # end to end: pinned host batches in, loss out
    for i in range(2):
        e2e_step(pinned[i % NBATCH])
    ms_e2e = timed(lambda i: e2e_step(pinned[i % NBATCH]), args.steps)
    # evaluation
    model.eval()
    for i in range(args.warmup):
        eval_step(resident[i % NBATCH])
    ms_eval = timed(lambda i: eval_step(resident[i % NBATCH]), args.steps)

    def eval_e2e(i):
        hb = pinned[i % NBATCH]
        out = eval_step({k: v.to(dev, non_blocking=True) for k, v in hb.items()})
        return out.cpu()
    eval_e2e(0)
    ms_eval_e2e = timed(eval_e2e, args.steps)
    sampler.join(timeout=2)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    bf16_peak = peaks.get("bf16_tflops_sustained", 1400.0)
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md: 6.65 TB/s, 1.4 PF/s sustained bf16)"
    tensor_peak = bf16_peak / 2.0 / 3.0          # fp32-parity bar => 3xTF32: effective tensor peak
    hbm_tags = {"adam", "embed_fwd", "embed_bwd", "scatter_add", "last_row_fwd", "last_row_bwd", "filter_fwd", "filter_bwd",
                "residual_ln_fwd", "ln_bwd"}       # work counted in algorithmic bytes

    def kernel_line(name, rec):
        sec = rec["ms"] * 1e-3
        if name in hbm_tags:
            ach = rec["work"] / sec / 1e9
            return {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak}
        ach = rec["work"] / sec / 1e12
        return {"bound": "tensor", "achieved": ach, "peak": tensor_peak, "unit": "TFLOP/s", "frac": ach / tensor_peak}

    prof = {k: v for k, v in prof.items() if v["launches"] > 0 and v["ms"] > 0}
    traffic = {}
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f)
    except Exception:
        pass
    dom = max(prof.items(), key=lambda kv: kv[1]["ms"]) if prof else (None, None)
    roofline = None
    if dom[0]:
        roofline = kernel_line(dom[0], dom[1])
        roofline.update({
            "kernel": dom[0], "traffic": traffic.get(dom[0]), "traffic_note": traffic.get("_note") if dom[0] in traffic else None,
            "peak_source": src + ("; tensor peak = bf16 sustained / 2 (TF32 operands) / 3 (split-TF32 passes)" if roofline["bound"] == "tensor" else ""),
            "avg_launch_ms": dom[1]["ms"] / dom[1]["launches"], "launches_per_step": dom[1]["launches"] / args.steps,
            "share_of_step": dom[1]["ms"] / (ms_train if ms_train > 0 else 1.0),
            "all_kernels": {k: dict(kernel_line(k, v), ms_per_step=v["ms"] / args.steps) for k, v in prof.items()}})
    seqs = args.steps * B * world
    value = seqs / (ms_train * 1e-3)
    h2d = B * cfg["max_seq_length"] * 8 + B * 8
    line = {
        "metric": "train seqs/sec (fwd+bwd+Adam)", "value": value, "unit": "seqs/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_train / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "per_gpu_batch": B, "global_batch": B * world, "dropout": P_DROP,
                   "parallelism": f"dp{world}" + (f" (train: replica + NCCL gradient exchange [{args.dp}]; eval: item table row-sharded {world} ways, top-k merge)" if world > 1 else ""), "l2": "inputs_exceed_l2 (item table + activations >> 126 MB)",
                   "optimizer": ("bsarec_b200.FusedAdam" if args.optimizer == "fused" else "torch.optim.Adam") + " inside the step",
                   "ce_impl": args.ce_impl},
        "model_tflops": value * flops_train_per_seq(cfg) / 1e12,
        "eval": {"metric": "full-ranking eval users/sec", "value": seqs / (ms_eval * 1e-3), "unit": "users/s",
                 "ms_per_step": ms_eval / args.steps,
                 "e2e": {"value": seqs / (ms_eval_e2e * 1e-3), "unit": "users/s",
                         "h2d_bytes_per_step": int(sum(v.numel() * v.element_size() for v in pinned[0].values())), "d2h_bytes_per_step": B * TOPK * 8},
                 "model_tflops": seqs / (ms_eval * 1e-3) * flops_eval_per_user(cfg) / 1e12},
        "e2e": {"value": seqs / (ms_e2e * 1e-3), "unit": "seqs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
        "gpu_launches": int(launches),
        "clocks": sampler.summary(),
        "roofline": roofline,
    }
    if not args.no_cpu_baseline and world == 1:
        import oracle
        torch.set_num_threads(os.cpu_count())
        ocfg = oracle.OracleConfig(hidden_dropout_prob=P_DROP, attention_probs_dropout_prob=P_DROP, **cfg)
        P = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        step = oracle.adam_train_step(P, ocfg)
        cb = synth_batches(cfg, cpu_B, 99, 2)
        step(cb[0]["ids"], cb[0]["answers"])
        t0 = time.perf_counter()
        n = 0
        while n < 3 or (time.perf_counter() - t0 < 10.0 and n < 50):
            step(cb[n % 2]["ids"], cb[n % 2]["answers"])
            n += 1
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": n * cpu_B / dt, "unit": "seqs/s", "cores": os.cpu_count(), "kind": "port",
                                "sample": f"{n} train steps (fwd+bwd+Adam) of {cpu_B} sequences of the same workload, torch CPU, {os.cpu_count()} threads"}
    print(json.dumps(line), file=_json_out, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
