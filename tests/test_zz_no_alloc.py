"""bsa_info_nce_fwd_static — the no-allocation variant of bsa_info_nce_fwd (scratch in a file-scope device array of the
library, include/bsarec_b200.h).  It has no counterpart in the reference; it was added
for callers that cannot allocate device memory.  Called through the C ABI and compared with the existing entry
point on two calls in a row (the second with a smaller batch, so stale scratch contents from the first call would show),
with the tolerance of the other parity tests.  Collected last on purpose (file name).
"""
import ctypes as C

import pytest
import torch

from bsarec_b200 import _cabi

TOL = 1e-4          # the parity bar of tests/test_gpu_parity.py


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _reference_loss(lib, z, B, d, tau):
    wsb = lib.bsa_info_nce_workspace_bytes(B)
    ws = torch.empty(wsb, dtype=torch.uint8, device=z.device)
    loss = torch.empty((), dtype=torch.float32, device=z.device)
    _cabi.check(lib.bsa_info_nce_fwd(B, d, C.c_float(tau), _ptr(z), _ptr(loss), _ptr(ws), wsb, _stream()), "bsa_info_nce_fwd")
    return loss


@pytest.mark.gpu
def test_info_nce_static_scratch_matches_workspace_entry_point_twice_in_a_row(cuda_device):
    lib = _cabi.load()
    max_b = lib.bsa_info_nce_static_max_b()
    assert max_b == 4096
    g = torch.Generator().manual_seed(2024)
    with torch.cuda.device(cuda_device):
        for call, (B, d, tau) in enumerate([(96, 64, 1.0), (17, 32, 0.5), (96, 64, 1.0)]):
            z = (0.3 * torch.randn(2 * B, d, generator=g)).to(cuda_device)
            want = _reference_loss(lib, z, B, d, tau)
            got = torch.full((), float("nan"), dtype=torch.float32, device=cuda_device)
            torch.cuda.synchronize()
            free0, alloc0 = torch.cuda.mem_get_info(cuda_device)[0], torch.cuda.memory_allocated(cuda_device)
            _cabi.check(lib.bsa_info_nce_fwd_static(B, d, C.c_float(tau), _ptr(z), _ptr(got), _stream()),
                        "bsa_info_nce_fwd_static")
            torch.cuda.synchronize()
            # nothing was allocated by the call: neither through torch nor behind its back
            assert torch.cuda.memory_allocated(cuda_device) == alloc0
            assert torch.cuda.mem_get_info(cuda_device)[0] >= free0
            w, o = float(want.item()), float(got.item())
            assert o == o and abs(o - w) <= TOL * abs(w), (call, B, d, o, w)
            # the oracle's closed form on the host as a second opinion (same bar)
            zd = z.double().cpu()
            sim = zd @ zd.T / tau
            sim.fill_diagonal_(float("-inf"))
            idx = (torch.arange(2 * B) + B) % (2 * B)
            ref = float((torch.logsumexp(sim, dim=1) - sim[torch.arange(2 * B), idx]).mean())
            assert abs(o - ref) <= TOL * abs(ref), (call, o, ref)


@pytest.mark.gpu
def test_info_nce_static_rejects_sizes_beyond_its_scratch(cuda_device):
    lib = _cabi.load()
    max_b = lib.bsa_info_nce_static_max_b()
    with torch.cuda.device(cuda_device):
        z = torch.zeros(8, 16, device=cuda_device)
        loss = torch.zeros((), device=cuda_device)
        rc = lib.bsa_info_nce_fwd_static(max_b + 1, 16, C.c_float(1.0), _ptr(z), _ptr(loss), _stream())
        assert rc != 0
        msg = lib.bsa_last_error_string().decode()
        assert "outside" in msg and str(max_b) in msg, msg
        assert lib.bsa_info_nce_fwd_static(0, 16, C.c_float(1.0), _ptr(z), _ptr(loss), _stream()) != 0
        # inside the array but beyond the kernel's shared-memory plan (2B + d floats <= 48 KB): also a clean error
        assert lib.bsa_info_nce_fwd_static(max_b, 8192, C.c_float(1.0), _ptr(z), _ptr(loss), _stream()) != 0
        assert "shared-memory" in lib.bsa_last_error_string().decode()
        assert lib.bsa_info_nce_fwd_static(4, 16, C.c_float(1.0), None, _ptr(loss), _stream()) != 0
        torch.cuda.synchronize()
        # and the entry point still works afterwards
        z = torch.randn(8, 16, device=cuda_device)
        _cabi.check(lib.bsa_info_nce_fwd_static(4, 16, C.c_float(1.0), _ptr(z), _ptr(loss), _stream()), "bsa_info_nce_fwd_static")
        want = _reference_loss(lib, z, 4, 16, 1.0)
        torch.cuda.synchronize()
        assert abs(float(loss) - float(want)) <= TOL * abs(float(want))
