#!/usr/bin/env python
"""Time bsa_gemm (tcgen05 3xTF32 engine) on the GEMM shapes of one BSARecBlock at C3 (M = 512*200 tokens, d = 256).
Reports whole-call time (operand split kernels included) and the GEMM kernel alone (library event tags)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bsarec_b200 import _cabi

lib = _cabi.load()
dev = torch.device("cuda:0")
Mtok, d = int(os.environ.get("MTOK", 102400)), 256
shapes = [  # name, M, N, K, a_mn, b_mn
    ("fwd d->d", Mtok, d, d, 0, 0), ("fwd d->4d", Mtok, 4 * d, d, 0, 0), ("fwd 4d->d", Mtok, d, 4 * d, 0, 0),
    ("dX d<-d", Mtok, d, d, 0, 1), ("dX d<-4d", Mtok, d, 4 * d, 0, 1), ("dX 4d<-d", Mtok, 4 * d, d, 0, 1),
    ("dW dxd", d, d, Mtok, 1, 1), ("dW dx4d", d, 4 * d, Mtok, 1, 1), ("dW 4dxd", 4 * d, d, Mtok, 1, 1),
]
only = os.environ.get("SHAPES")
if only:
    shapes = [s for s in shapes if s[0] in only.split(",")]
st = torch.cuda.current_stream().cuda_stream
for name, M, N, K, a_mn, b_mn in shapes:
    A = torch.randn((K, M) if a_mn else (M, K), device=dev)
    B = torch.randn((K, N) if b_mn else (N, K), device=dev)
    Cc = torch.empty(M, N, device=dev)
    lda, ldb = A.stride(0), B.stride(0)
    wsb = lib.bsa_gemm_workspace_bytes(M, N, K, lda, ldb, a_mn, b_mn)
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    def run():
        _cabi.check(lib.bsa_gemm(2, M, N, K, A.data_ptr(), lda, a_mn, B.data_ptr(), ldb, b_mn, Cc.data_ptr(), N, None, 1.0, 0,
                                 ws.data_ptr(), wsb, st))
    for _ in range(3): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record()
    for _ in range(n): run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    _cabi.profile(True)
    for _ in range(n): run()
    torch.cuda.synchronize()
    pr = _cabi.profile_results(); _cabi.profile(False)
    kms = pr["linear_umma"]["ms"] / n
    fl = 2.0 * M * N * K
    print(json.dumps({"shape": name, "M": M, "N": N, "K": K, "a_mn": a_mn, "b_mn": b_mn, "call_ms": round(ms, 4), "kernel_ms": round(kms, 4),
                      "kernel_tflops": round(fl / kms / 1e9, 1), "call_tflops": round(fl / ms / 1e9, 1)}))
    del A, B, Cc, ws
