"""Multi-GPU self-check (run under torchrun on the GPU box, not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dp_selfcheck.py

Every rank runs its shard of one global batch through the CUDA path with the overlapped GradExchange; rank 0 checks the
resulting gradients against (a) the flat-bucket all-reduce of the same local gradients and (b) the CPU oracle on the
GLOBAL batch (<= 1e-4 relative), and all ranks check that replicas are bit-identical."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle                                                        # noqa: E402
from helpers import build_model, oracle_cfg, synth_batch, perturbed_params   # noqa: E402
from bsarec_b200.parallel import allreduce_gradients                # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    cfgkw = dict(item_size=5003, hidden_size=64, max_seq_length=50, num_attention_heads=2, num_hidden_layers=2, c=5, alpha=0.7)
    cfg = oracle_cfg(cfgkw)
    P = perturbed_params(cfg, 7)
    Bl = 24
    ids, answers = synth_batch(17, Bl * world, 50, 5003)
    mine = slice(rank * Bl, (rank + 1) * Bl)
    res = {}
    for mode in ("overlap", "bucket"):
        model = build_model(cfgkw, P, dev).train()
        if mode == "overlap":
            model.enable_data_parallel()
        model.calculate_loss(ids[mine].to(dev), answers[mine].to(dev), None, None, None).backward()
        if mode == "bucket":
            allreduce_gradients(list(model.parameters()))
        res[mode] = {n: p.grad.detach().clone() for n, p in model.named_parameters()}
    worst = 0.0
    for n in res["overlap"]:
        a, b = res["overlap"][n], res["bucket"][n]
        den = b.abs().max().item() or 1.0
        worst = max(worst, (a - b).abs().max().item() / den)
        gathered = [torch.empty_like(a) for _ in range(world)]
        dist.all_gather(gathered, a.contiguous())
        for g in gathered:
            assert torch.equal(g, gathered[0]), f"replicas differ on {n}"
    assert worst < 2e-6, worst
    if rank == 0:
        _, g64 = oracle.loss_and_grads({k: v.double() for k, v in P.items()}, ids, answers, cfg, use_fft=False)
        bad = 0.0
        for n, g in res["overlap"].items():
            if n.endswith("key.bias"):
                continue
            den = g64[n].abs().max().item() or 1.0
            bad = max(bad, (g.cpu().double() - g64[n]).abs().max().item() / den)
        assert bad < 1e-4, bad
        print(f"dp_selfcheck ok: world {world}, overlap vs bucket {worst:.2e}, vs oracle(global batch) {bad:.2e}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
