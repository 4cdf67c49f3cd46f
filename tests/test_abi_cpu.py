"""CPU: the C-ABI library builds, loads and exports every symbol include/cp2_b200.h declares; argument
validation returns CP2_E_INVALID without touching a device; the host ops refuse CPU tensors (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    from cp2_b200 import _lib
    if not os.path.isfile(_lib.LIB_PATH):
        g.build()
    return _lib.load()


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cp2_\w+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    from cp2_b200 import _lib
    names = _declared("cp2_b200.h")
    assert len(names) >= 12
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} declared in include/ but not exported"
    # and the ctypes table mirrors the header one to one
    assert set(names) == set(_lib.SIGNATURES)
    # the test-only kernels live in their own library; the product library carries none of them
    dbg = ctypes.CDLL(_lib.DEBUG_LIB_PATH)
    for n in _declared("cp2_b200_debug.h"):
        assert hasattr(dbg, n), f"{n} declared in cp2_b200_debug.h but not exported by the debug library"
        assert not hasattr(raw, n), f"test-only symbol {n} in the product library"
    assert set(_declared("cp2_b200_debug.h")) == set(_lib.DEBUG_SIGNATURES)


def test_abi_version_and_host_helpers(lib):
    from cp2_b200 import _lib
    assert lib.cp2_abi_version() == _lib.ABI_VERSION == 10
    assert lib.cp2_mirror_up_partials(2, 512, 512) == 2 * 32 * 4 and lib.cp2_mirror_partials(0, 16) == 1
    assert lib.cp2_prep_tiles(196) == 7 and lib.cp2_prep_tiles(32) == 1
    assert lib.cp2_global_scratch_bytes(32, 128) == 32 * 8 + 32 * 128 * 8 + 32 * 12 + 16
    assert lib.cp2_dense_pad(196) == 384 and lib.cp2_dense_pad(3136) == 3456 and lib.cp2_dense_pad(384) == 384


def test_invalid_arguments_are_rejected_before_any_launch(lib):
    # bad shapes / null pointers -> CP2_E_INVALID (-1) and a message; no device is needed for these paths
    assert lib.cp2_composite(None, None, None, None, None, 1, 0, 4, 16, None) == -1
    assert b"cp2_composite" in lib.cp2_last_error()
    assert lib.cp2_dense_lse(None, None, None, None, 1.0, 1, 128, 196, None) == -1
    assert lib.cp2_dense_lse(1, 1, None, 1, 1.0, 1, 96, 196, None) == -1     # unsupported C
    assert b"C=96" in lib.cp2_last_error()
    assert lib.cp2_dense_lse(1, 1, None, 1, 50.0, 1, 128, 196, None) == -1    # 1/tau outside (0, 16]
    assert b"1/tau" in lib.cp2_last_error()
    assert lib.cp2_global_nce(8, 16, 8, None, None, None, 8, 8, None, None, 8, 5.0, 4, 64, 4096, None) == -1   # C
    assert b"C=64" in lib.cp2_last_error()
    assert lib.cp2_global_nce(8, 16, None, 8, None, None, 8, 8, 8, 8, 8, 5.0, 4, 128, 4096, None) == -1   # meter w/o queue
    assert lib.cp2_enqueue(1, 1, None, 1, 1, 10, 128, 5, None) == -1          # n > K
    assert lib.cp2_ema(None, None, None, None, None, 3, 0.999, 0.001, None) == -1
    assert lib.cp2_mirror_loss(1, 1, 1, 1, 1, None, None, None, 0.01, 0.5, 0, 2, 9, 64, None) == -1   # Cls > 8
    assert b"Cls=9" in lib.cp2_last_error()
    assert lib.cp2_mirror_loss(1, 1, 1, 1, 1, None, None, None, 0.01, 0.5, 0, 0, 2, 64, None) == -1   # empty mean
    assert lib.cp2_mirror_loss_upsampled(1, 1, None, 1, 1, None, None, 0.01, 0.5, 0, 2, 2, 4, 4, 64, 64, None) == -1
    assert lib.cp2_id_iou(1, 1, 1, 1, 1, 1, 2, 8, 8, 16, None) == -1          # empty sub-sampled map
    # empty work is a no-op success, as the reference's ops are on empty tensors
    assert lib.cp2_composite(None, None, None, None, None, 0, 8, 8, 16, None) == 0
    assert lib.cp2_enqueue(None, None, None, None, None, 0, 128, 16, None) == 0
    assert lib.cp2_ema(None, None, None, None, None, 0, 0.999, 0.001, None) == 0


def test_ops_refuse_cpu_tensors(lib):
    from cp2_b200 import ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.composite(torch.zeros(1, 3, 16, 16), torch.zeros(1, 3, 16, 16), 16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.enqueue(torch.zeros(2, 128), torch.zeros(128, 16), torch.zeros(1, dtype=torch.int64))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.prep_features(torch.zeros(1, 128, 4), torch.zeros(1, 4))


def test_no_getenv_on_the_product_launch_path():
    """Tuning knobs exist only in the -DCP2_DEBUG_BUILD copy of the dense kernels (libcp2_b200_debug.so)."""
    src = open(os.path.join(ROOT, "cp2_b200", "csrc", "dense_umma.cu")).read()
    src = re.sub(r"//[^\n]*", "", src)
    product = re.sub(r"#ifdef CP2_DEBUG_BUILD.*?#(else|endif)", "", src, flags=re.S)
    assert "getenv" not in product


def test_product_path_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "cp2_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in re.sub(r'""".*?"""', "", src, flags=re.S).replace("# ", ""), fn
