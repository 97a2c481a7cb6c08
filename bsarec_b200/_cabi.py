"""ctypes binding of include/bsarec_b200.h (libbsarec_sm100a.so).

There is no CPU fallback: if the library cannot be loaded the import of the package's compute path
raises, and every call checks the C return code and raises ``RuntimeError`` with the library's message.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbsarec_sm100a.so")

ABI_VERSION = 4          # == BSA_ABI_VERSION of include/bsarec_b200.h (checked at load)

c_f32p = C.c_void_p
c_i64p = C.c_void_p


class EmbedDesc(C.Structure):
    _fields_ = [("B", C.c_int), ("L", C.c_int), ("d", C.c_int), ("item_size", C.c_int64),
                ("p_hidden", C.c_float), ("training", C.c_int), ("seed", C.c_uint64)]


class BlockDesc(C.Structure):
    _fields_ = [("B", C.c_int), ("L", C.c_int), ("d", C.c_int), ("H", C.c_int), ("kept_bins", C.c_int),
                ("alpha", C.c_float), ("p_hidden", C.c_float), ("p_attn", C.c_float), ("training", C.c_int),
                ("seed", C.c_uint64), ("layer", C.c_int), ("inference", C.c_int), ("kind", C.c_int)]


BLOCK_FIELDS = ["sqrt_beta", "f_ln_w", "f_ln_b", "wq", "bq", "wk", "bk", "wv", "bv", "wo", "bo",
                "a_ln_w", "a_ln_b", "w1", "b1", "w2", "b2", "o_ln_w", "o_ln_b"]

# member of bsa_block_weights -> state_dict key suffix under item_encoder.blocks.<l>.
BLOCK_PARAM_KEYS = {
    "sqrt_beta": "layer.filter_layer.sqrt_beta",
    "f_ln_w": "layer.filter_layer.LayerNorm.weight", "f_ln_b": "layer.filter_layer.LayerNorm.bias",
    "wq": "layer.attention_layer.query.weight", "bq": "layer.attention_layer.query.bias",
    "wk": "layer.attention_layer.key.weight", "bk": "layer.attention_layer.key.bias",
    "wv": "layer.attention_layer.value.weight", "bv": "layer.attention_layer.value.bias",
    "wo": "layer.attention_layer.dense.weight", "bo": "layer.attention_layer.dense.bias",
    "a_ln_w": "layer.attention_layer.LayerNorm.weight", "a_ln_b": "layer.attention_layer.LayerNorm.bias",
    "w1": "feed_forward.dense_1.weight", "b1": "feed_forward.dense_1.bias",
    "w2": "feed_forward.dense_2.weight", "b2": "feed_forward.dense_2.bias",
    "o_ln_w": "feed_forward.LayerNorm.weight", "o_ln_b": "feed_forward.LayerNorm.bias",
}


class BlockPtrs(C.Structure):           # bsa_block_weights and bsa_block_grads share this layout
    _fields_ = [(n, C.c_void_p) for n in BLOCK_FIELDS]


class CeDesc(C.Structure):
    _fields_ = [("B", C.c_int), ("d", C.c_int), ("item_size", C.c_int64), ("impl", C.c_int), ("reuse_split", C.c_int)]


class TopkDesc(C.Structure):
    _fields_ = [("B", C.c_int), ("d", C.c_int), ("k", C.c_int), ("shard_rows", C.c_int64),
                ("shard_offset", C.c_int64), ("impl", C.c_int), ("reuse_split", C.c_int)]


# name -> (restype, argtypes); mirrors include/bsarec_b200.h one to one
_VP, _SZ, _I, _I64 = C.c_void_p, C.c_size_t, C.c_int, C.c_int64
SIGNATURES = {
    "bsa_abi_version": (_I, []),
    "bsa_last_error_string": (C.c_char_p, []),
    "bsa_launch_count": (C.c_ulonglong, []),
    "bsa_set_replay_state": (None, [_VP]),
    "bsa_get_replay_state": (_VP, []),
    "bsa_profile_enable": (None, [_I]),
    "bsa_profile_reset": (None, []),
    "bsa_profile_num_tags": (_I, []),
    "bsa_profile_get": (_I, [_I, C.POINTER(C.c_char_p), C.POINTER(C.c_double), C.POINTER(C.c_double),
                            C.POINTER(C.c_longlong)]),
    "bsa_embed_ln_dropout_fwd": (_I, [C.POINTER(EmbedDesc)] + [_VP] * 9),
    "bsa_embed_bwd_workspace_bytes": (_SZ, [C.POINTER(EmbedDesc)]),
    "bsa_embed_ln_dropout_bwd": (_I, [C.POINTER(EmbedDesc)] + [_VP] * 12 + [_SZ, _VP]),
    "bsa_embed_ln_dropout_bwd_rows": (_I, [C.POINTER(EmbedDesc)] + [_VP] * 12 + [_SZ, _VP]),
    "bsa_scatter_add_workspace_bytes": (_SZ, [_I64]),
    "bsa_scatter_add_rows": (_I, [_VP, _I64, _I64, _VP, _I, _VP, _VP, _SZ, _VP]),
    "bsa_sort_scatter_plan_bytes": (_SZ, [_I64]),
    "bsa_sort_scatter_plan": (_I, [_VP, _I64, _I64, _VP, _SZ, _VP, _VP, _VP]),
    "bsa_filter_rank": (_I, [_I, _I]),
    "bsa_build_filter_basis": (_I, [_VP, _I, _I, _VP]),
    "bsa_set_ce_mode": (None, [_I]),
    "bsa_get_ce_mode": (_I, []),
    "bsa_set_linear_impl": (None, [_I]),
    "bsa_get_linear_impl": (_I, []),
    "bsa_set_attn_impl": (None, [_I]),
    "bsa_get_attn_impl": (_I, []),
    "bsa_block_saved_bytes": (_SZ, [C.POINTER(BlockDesc)]),
    "bsa_block_workspace_bytes": (_SZ, [C.POINTER(BlockDesc)]),
    "bsa_block_fwd": (_I, [C.POINTER(BlockDesc), C.POINTER(BlockPtrs), _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "bsa_block_bwd": (_I, [C.POINTER(BlockDesc), C.POINTER(BlockPtrs), _VP, _VP, _VP, _VP, _VP, _VP,
                           C.POINTER(BlockPtrs), _VP, _SZ, _VP]),
    "bsa_block_last_supported": (_I, [C.POINTER(BlockDesc)]),
    "bsa_block_last_saved_bytes": (_SZ, [C.POINTER(BlockDesc)]),
    "bsa_block_last_workspace_bytes": (_SZ, [C.POINTER(BlockDesc)]),
    "bsa_block_last_fwd": (_I, [C.POINTER(BlockDesc), C.POINTER(BlockPtrs), _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "bsa_block_last_bwd": (_I, [C.POINTER(BlockDesc), C.POINTER(BlockPtrs), _VP, _VP, _VP, _VP, _VP, _VP,
                                C.POINTER(BlockPtrs), _VP, _SZ, _VP]),
    "bsa_ce_workspace_bytes": (_SZ, [C.POINTER(CeDesc)]),
    "bsa_ce_fwd": (_I, [C.POINTER(CeDesc), _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "bsa_ce_fwd_partial": (_I, [C.POINTER(CeDesc), _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "bsa_ce_bwd": (_I, [C.POINTER(CeDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "bsa_score_topk_workspace_bytes": (_SZ, [C.POINTER(TopkDesc)]),
    "bsa_score_topk": (_I, [C.POINTER(TopkDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "bsa_topk_merge": (_I, [_VP, _VP, _I, _I, _I, _VP, _VP, _VP]),
    "bsa_gemm_workspace_bytes": (_SZ, [_I, _I, _I, _I64, _I64, _I, _I]),
    "bsa_gemm": (_I, [_I, _I, _I, _I, _VP, _I64, _I, _VP, _I64, _I, _VP, _I64, _VP, C.c_float, _I, _VP, _SZ, _VP]),
    "bsa_adam_step": (_I, [_VP, _VP, _VP, _VP, _I64] + [C.c_float] * 5 + [_I64, _VP]),
    "bsa_adam_step_multi": (_I, [_I, _VP, _VP, _VP, _VP, _VP] + [C.c_float] * 5 + [_I64, _VP]),
    "bsa_build_batch": (_I, [_I, _VP, _VP, _VP, _I64, _VP, _I, _I, _VP, _VP, _VP, _VP]),
    "bsa_hit_rank_hist": (_I, [_VP, _VP, _I64, _I, _VP, _VP]),
    "bsa_gather_last": (_I, [_VP, _VP, _I, _I, _I, _VP]),
    "bsa_scatter_last": (_I, [_VP, _VP, _I, _I, _I, _VP]),
    "bsa_gather_rows": (_I, [_VP, _I64, _I64, _VP, _I, _VP, _VP]),
    "bsa_scatter_rows": (_I, [_VP, _I64, _I64, _VP, _I, _VP, _VP]),
    "bsa_fmlp_supported": (_I, [_I, _I]),
    "bsa_fmlp_tables": (_I, [_VP, _VP, _I, _VP]),
    "bsa_fmlp_kernel_fwd": (_I, [_VP, _VP, _VP, _VP, _I, _I, _VP]),
    "bsa_fmlp_kernel_bwd": (_I, [_VP, _VP, _VP, _VP, _I, _I, _VP]),
    "bsa_info_nce_workspace_bytes": (_SZ, [_I]),
    "bsa_info_nce_fwd": (_I, [_I, _I, C.c_float, _VP, _VP, _VP, _SZ, _VP]),
    "bsa_info_nce_bwd": (_I, [_I, _I, C.c_float, _VP, _VP, C.c_float, _VP, _VP, _SZ, _VP]),
    "bsa_info_nce_static_max_b": (_I, []),
    "bsa_info_nce_fwd_static": (_I, [_I, _I, C.c_float, _VP, _VP, _VP]),
    "bsa_bce_pair_workspace_bytes": (_SZ, [_I, _I]),
    "bsa_bce_pair_fwd": (_I, [_I, _I, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _I, _VP]),
    "bsa_bce_pair_bwd": (_I, [_I, _I, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _I, _VP]),
}

_lib = None


def load():
    """Load the shared library (once).  Raises RuntimeError when it is missing: no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"bsarec_b200: {LIB_PATH} is not built. Run `python -m bsarec_b200.build` (needs nvcc) — "
            "there is no CPU or PyTorch fallback for the BSARec hot path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here == header/library mismatch
        fn.restype = res
        fn.argtypes = args
    if lib.bsa_abi_version() != ABI_VERSION:
        raise RuntimeError("bsarec_b200: ABI version mismatch between _cabi.py and the shared library")
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().bsa_last_error_string().decode("utf-8", "replace")
        raise RuntimeError(f"bsarec_b200 {what} failed (code {rc}): {msg}")


def launch_count() -> int:
    return int(load().bsa_launch_count())


def profile(on: bool):
    lib = load()
    lib.bsa_profile_reset()
    lib.bsa_profile_enable(int(on))


def profile_results():
    """{tag: {"ms": total device ms, "work": algorithmic FLOPs or bytes, "launches": n}} for tags that ran."""
    lib = load()
    out = {}
    for t in range(lib.bsa_profile_num_tags()):
        name, ms, work, n = C.c_char_p(), C.c_double(), C.c_double(), C.c_longlong()
        check(lib.bsa_profile_get(t, C.byref(name), C.byref(ms), C.byref(work), C.byref(n)), "bsa_profile_get")
        if n.value:
            out[name.value.decode()] = {"ms": ms.value, "work": work.value, "launches": n.value}
    return out
