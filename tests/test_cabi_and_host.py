"""CPU-side checks of the drop-in boundary: the C-ABI library exports exactly what include/bsarec_b200.h
declares, argument checking fails loudly without a GPU, and the host-side mirror of the reference class
contract (state_dict keys, seeded init, ctor errors) holds.  No compute call is made here."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest
import torch

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "bsarec_b200.h")


def _header_functions():
    """Names of every function prototype in the header."""
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    src = re.sub(r"typedef\s+struct\s*\{.*?\}\s*\w+\s*;", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bsa_\w+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from bsarec_b200 import build, _cabi
    build.build(force=False, verbose=False)
    return _cabi.load()


def test_header_and_binding_agree(lib):
    from bsarec_b200 import _cabi
    declared = _header_functions()
    assert declared, "no prototypes parsed from the header"
    assert sorted(_cabi.SIGNATURES) == declared


def test_library_exports_every_declared_symbol_and_nothing_else(lib):
    from bsarec_b200 import _cabi
    out = subprocess.run(["nm", "-D", "--defined-only", _cabi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l and l.split()[-1].startswith("bsa_"))
    assert exported == _header_functions()
    # built with -fvisibility=hidden: no C++ internals leak through the boundary
    leaked = [l.split()[-1] for l in out.splitlines() if " T " in l and not l.split()[-1].startswith(("bsa_", "_init", "_fini"))]
    assert leaked == []


def test_library_is_sm100a_with_tcgen05_and_tma(lib):
    from bsarec_b200 import _cabi
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("no cuobjdump")
    elf = subprocess.run([cuobjdump, "-lelf", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in elf and "sm_90" not in elf
    sass = subprocess.run([cuobjdump, "-sass", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    assert re.search(r"UTC\w*MMA", sass), "no tcgen05.mma in SASS"
    assert "UTMALDG" in sass, "no TMA loads in SASS"
    assert "LDTM" in sass, "no tcgen05.ld in SASS"


def test_abi_version_and_pure_host_queries(lib):
    from bsarec_b200._cabi import BlockDesc, CeDesc, EmbedDesc, TopkDesc
    from bsarec_b200 import _cabi
    assert lib.bsa_abi_version() == _cabi.ABI_VERSION == 4
    # filter rank = 2*kept-1, capped by L (exact low-pass operator, bsarec.py:87,93-97)
    assert lib.bsa_filter_rank(50, 3) == 5
    assert lib.bsa_filter_rank(200, 3) == 5
    assert lib.bsa_filter_rank(1024, 17) == 33
    bd = BlockDesc(B=4, L=50, d=64, H=2, kept_bins=3, alpha=0.7, p_hidden=0.5, p_attn=0.5, training=1, seed=1, layer=0)
    assert lib.bsa_block_saved_bytes(C.byref(bd)) > 4 * 50 * 64 * 4
    assert lib.bsa_block_workspace_bytes(C.byref(bd)) > 0
    assert lib.bsa_ce_workspace_bytes(C.byref(CeDesc(B=4, d=64, item_size=1000, impl=0))) > 0
    assert lib.bsa_score_topk_workspace_bytes(C.byref(TopkDesc(B=4, d=64, k=20, shard_rows=1000, shard_offset=0, impl=0))) > 0
    assert lib.bsa_embed_bwd_workspace_bytes(C.byref(EmbedDesc(B=4, L=50, d=64, item_size=1000, p_hidden=0.5, training=1, seed=1))) > 0
    assert lib.bsa_sort_scatter_plan_bytes(200) > 0


def test_null_arguments_are_rejected_with_a_message(lib):
    """Every export returns non-zero and sets the error string instead of dereferencing a null pointer."""
    from bsarec_b200._cabi import CeDesc, check
    cd = CeDesc(B=4, d=64, item_size=1000, impl=0)
    rc = lib.bsa_ce_fwd(C.byref(cd), None, None, None, None, None, None, 0, None)
    assert rc != 0
    assert b"null argument" in lib.bsa_last_error_string()
    with pytest.raises(RuntimeError, match="null argument"):
        check(rc, "bsa_ce_fwd")
    assert lib.bsa_gemm(2, 8, 8, 8, None, 8, 0, None, 8, 0, None, 8, None, C.c_float(1.0), 0, None, 0, None) != 0


def test_product_path_has_no_cpu_fallback():
    """CPU tensors must raise, not silently compute somewhere else; the package never imports the oracle."""
    from bsarec_b200 import BSARecModel
    from helpers import make_args
    cfg = dict(item_size=101, hidden_size=32, max_seq_length=10, num_attention_heads=2, num_hidden_layers=1, c=3, alpha=0.5)
    m = BSARecModel(make_args(cfg))
    ids = torch.randint(1, 101, (3, 10))
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU fallback|CUDA"):
            m.predict(ids, None)
    pkg = os.path.join(ROOT, "bsarec_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(import|from)\s+oracle", src, flags=re.M), fn


def test_state_dict_contract_matches_appendix_c():
    """42 keys in the reference order with the reference shapes."""
    import oracle
    from bsarec_b200 import BSARecModel, MODEL_DICT
    from helpers import make_args, oracle_cfg
    cfg = dict(item_size=301, hidden_size=64, max_seq_length=50, num_attention_heads=2, num_hidden_layers=2, c=5, alpha=0.7)
    m = MODEL_DICT["bsarec"](args=make_args(cfg))
    assert isinstance(m, BSARecModel)
    sd = m.state_dict()
    assert list(sd.keys()) == oracle.param_names(2)
    assert len(sd) == 42
    N, L, d, P = 301, 50, 64, 2
    assert sum(p.numel() for p in m.parameters()) == N * d + L * d + 2 * d + P * (12 * d * d + 16 * d)
    ref = oracle.init_params(oracle_cfg(cfg), seed=42)
    for k in sd:
        assert sd[k].shape == ref[k].shape, k
    assert list(dict(m.named_buffers())) == []


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/model"), reason="live reference not mounted")
def test_seeded_init_is_bit_identical_to_live_reference():
    """Same torch.manual_seed -> same 42 initial tensors as the unmodified reference class (bsarec.py:8-14,
    _abstract_model.py:26-39): the module tree is constructed in the reference's order."""
    import sys
    from bsarec_b200 import BSARecModel
    from helpers import make_args
    cfg = dict(item_size=301, hidden_size=64, max_seq_length=50, num_attention_heads=2, num_hidden_layers=2, c=5, alpha=0.7)
    sys.path.insert(0, "/root/reference/src")
    try:
        from model.bsarec import BSARecModel as RefModel
    finally:
        sys.path.remove("/root/reference/src")
    args = make_args(cfg, 0.5, 0.5)
    torch.manual_seed(42)
    ours = BSARecModel(args).state_dict()
    torch.manual_seed(42)
    ref = RefModel(args=args).state_dict()
    assert list(ours.keys()) == list(ref.keys())
    for k in ref:
        assert torch.equal(ours[k], ref[k]), k


def test_ctor_errors_match_reference():
    from bsarec_b200 import BSARecModel
    from helpers import make_args
    cfg = dict(item_size=101, hidden_size=30, max_seq_length=10, num_attention_heads=4, num_hidden_layers=1, c=3, alpha=0.5)
    with pytest.raises(ValueError, match="not a multiple of the number of attention"):   # _modules.py:79-82
        BSARecModel(make_args(cfg))


def test_metrics_vectorised_equal_reference_loops():
    """trainer.recall_at_k / ndcg_k are vectorised restatements of metrics.py:3-23; check against the oracle's."""
    import oracle
    from bsarec_b200 import trainer
    rng = np.random.default_rng(0)
    pred = np.stack([rng.permutation(200)[:20] for _ in range(300)])
    ans = rng.integers(0, 200, 300)
    for k in (5, 10, 15, 20):
        assert abs(trainer.recall_at_k(ans, pred, k) - oracle.recall_at_k(ans.tolist(), pred, k)) < 1e-12
        assert abs(trainer.ndcg_k(ans, pred, k) - oracle.ndcg_k(ans.tolist(), pred, k)) < 1e-12
