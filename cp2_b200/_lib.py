"""ctypes binding of the C ABI in include/cp2_b200.h (libcp2_b200.so, sm_100a).

There is NO fallback: if the shared library is missing or does not export every
symbol the header declares, loading raises.  Build it with
`python -c "import __graft_entry__ as g; g.build()"` or `make -C cp2_b200/csrc`.
"""
import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libcp2_b200.so")

# name -> (restype, argtypes); mirrors include/cp2_b200.h one to one
SIGNATURES = {
    "cp2_abi_version": (c_int, []),
    "cp2_last_error": (c_char_p, []),
    "cp2_composite": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p]),
    "cp2_composite2": (c_int, [c_void_p] * 10 + [c_int] * 4 + [c_void_p]),
    "cp2_prep_tiles": (c_int, [c_int]),
    "cp2_prep_features": (c_int, [c_void_p] * 13 + [c_int] * 3 + [c_void_p]),
    "cp2_prep_features_static": (c_int, [c_void_p] * 11 + [c_int] * 3 + [c_void_p]),
    "cp2_prep_features2": (c_int, [c_void_p] * 23 + [c_int] * 3 + [c_void_p]),
    "cp2_dense_lse": (c_int, [c_void_p] * 4 + [c_float] + [c_int] * 3 + [c_void_p]),
    "cp2_dense_pad": (c_int, [c_int]),
    "cp2_dense_grad": (c_int, [c_void_p] * 13 + [c_float] + [c_int] * 3 + [c_void_p]),
    "cp2_dense_small_ok": (c_int, [c_int, c_int]),
    "cp2_dense_small": (c_int, [c_void_p] * 14 + [c_float] + [c_int] * 3 + [c_void_p]),
    "cp2_dense_small_bwd": (c_int, [c_void_p] * 10 + [c_int] * 3 + [c_void_p]),
    "cp2_loss_combine": (c_int, [c_void_p] * 21 + [c_float] * 3 + [c_int] * 3 + [c_void_p]),
    "cp2_global_scratch_bytes": (c_int64, [c_int, c_int]),
    "cp2_queue_shadow": (c_int, [c_void_p] * 2 + [c_int] * 2 + [c_void_p]),
    "cp2_global_nce": (c_int, [c_void_p] * 11 + [c_float] + [c_int] * 3 + [c_void_p]),
    "cp2_ema": (c_int, [c_void_p] * 5 + [c_int, c_float, c_float, c_void_p]),
    "cp2_enqueue": (c_int, [c_void_p] * 5 + [c_int] * 3 + [c_void_p]),
    "cp2_id_iou": (c_int, [c_void_p] * 6 + [c_int] * 4 + [c_void_p]),
    "cp2_dense_weighted_tiles": (c_int, [c_int]),
    "cp2_dense_weighted_lse": (c_int, [c_void_p] * 18 + [c_float] * 3 + [c_int, c_float, c_float] + [c_int] * 6 + [c_void_p]),
    "cp2_dense_weighted_grad": (c_int, [c_void_p] * 16 + [c_float] * 3 + [c_int, c_float, c_float] + [c_int] * 6 + [c_void_p]),
    "cp2_dense_weighted_small": (c_int, [c_void_p] * 19 + [c_float] * 3 + [c_int, c_float, c_float] + [c_int] * 6 + [c_void_p]),
    "cp2_dense_weighted_stream_ids_bytes": (c_int64, [c_int, c_int]),
    "cp2_dense_weighted_stream": (c_int, [c_void_p] * 20 + [c_float] * 3 + [c_int, c_float, c_float] + [c_int] * 6 + [c_void_p]),
    "cp2_dense_weighted_stream_grad": (c_int, [c_void_p] * 15 + [c_float] * 3 + [c_int, c_float, c_float] + [c_int] * 3 + [c_void_p]),
    "cp2_dense_top1_scratch_floats": (c_int64, [c_int, c_int]),
    "cp2_dense_top1": (c_int, [c_void_p] * 7 + [c_int] * 3 + [c_void_p]),
    "cp2_rows_nce": (c_int, [c_void_p] * 7 + [c_float, c_int, c_int, c_int, c_int64, c_int, c_int64, c_void_p]),
    "cp2_queue_sums_split": (c_int, [c_int, c_int]),
    "cp2_queue_sums": (c_int, [c_void_p] * 5 + [c_float] + [c_int] * 3 + [c_void_p]),
    "cp2_positive_keys": (c_int, [c_void_p] * 5 + [c_float] + [c_int] * 6 + [c_void_p]),
    "cp2_densecl_match_scratch_bytes": (c_int64, [c_int, c_int]),
    "cp2_densecl_match": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_void_p]),
    "cp2_mirror_partials": (c_int, [c_int, c_int64]),
    "cp2_mirror_loss": (c_int, [c_void_p] * 8 + [c_float, c_float, c_int64, c_int, c_int, c_int64, c_void_p]),
    "cp2_mirror_up_partials": (c_int, [c_int] * 3),
    "cp2_mirror_loss_upsampled": (c_int, [c_void_p] * 7 + [c_float, c_float, c_int64] + [c_int] * 6 + [c_void_p]),
}
# include/cp2_b200_debug.h — test-only CUDA-core reference kernels, in their own library
# (libcp2_b200_debug.so, `load_debug()`); the product library does not contain them
DEBUG_LIB_PATH = os.path.join(_HERE, "csrc", "libcp2_b200_debug.so")
DEBUG_SIGNATURES = {
    "cp2_dbg_dense_lse_simt": (c_int, [c_void_p] * 3 + [c_float] + [c_int] * 3 + [c_void_p]),
    "cp2_dbg_dense_grad_simt": (c_int, [c_void_p] * 10 + [c_float] + [c_int] * 3 + [c_void_p]),
    "cp2_dbg_dense_trace": (c_int, [c_void_p]),
    "cp2_dbg_global_nce_simt": (c_int, [c_void_p] * 4 + [c_float] + [c_int] * 3 + [c_void_p]),
}

EMA_CHUNK = 4096  # CP2_EMA_CHUNK
ABI_VERSION = 10  # CP2_ABI_VERSION

_lib = None
_dbg = None


def _bind(lib, table, path):
    for name, (res, args) in table.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise RuntimeError(f"cp2_b200: {path} does not export {name}") from e
        fn.restype = res
        fn.argtypes = args


def load_debug():
    """TEST-ONLY: libcp2_b200_debug.so (SIMT cross-check kernels, traced / tunable copy of the dense kernels).
    cp2_b200/ never calls this."""
    global _dbg
    if _dbg is None:
        if not os.path.isfile(DEBUG_LIB_PATH):
            raise RuntimeError(f"cp2_b200: {DEBUG_LIB_PATH} not found (make -C cp2_b200/csrc)")
        lib = ctypes.CDLL(DEBUG_LIB_PATH)
        _bind(lib, DEBUG_SIGNATURES, DEBUG_LIB_PATH)
        for name in ("cp2_dense_lse", "cp2_dense_grad", "cp2_dense_pad", "cp2_last_error"):
            getattr(lib, name).restype, getattr(lib, name).argtypes = SIGNATURES[name]
        _dbg = lib
    return _dbg


def load():
    """Loads libcp2_b200.so once; raises if absent or incomplete."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise RuntimeError(
            f"cp2_b200: {LIB_PATH} not found. The CUDA extension is mandatory (no CPU or "
            "PyTorch fallback exists): run `make -C cp2_b200/csrc` or `__graft_entry__.build()`.")
    lib = ctypes.CDLL(LIB_PATH)
    _bind(lib, SIGNATURES, LIB_PATH)
    if lib.cp2_abi_version() != ABI_VERSION:
        raise RuntimeError(f"cp2_b200: ABI version {lib.cp2_abi_version()} != {ABI_VERSION}")
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().cp2_last_error()
        raise RuntimeError(f"cp2_b200: {what} failed (code {rc}): {msg.decode() if msg else ''}")
