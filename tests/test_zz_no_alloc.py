"""GPU: cp2_prep_features_static — the allocation-free variant of cp2_prep_features (work space in a file-scope
device array of the library; no counterpart in the reference).
Called through the C-ABI on two calls in a row and compared with the existing entry point, which runs the same
kernel on caller-owned scratch: the outputs must be bit-identical.  Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_BYTES = 8 << 20          # CP2_PREP_STATIC_BYTES
STATIC_MAX_B = 4096             # CP2_PREP_STATIC_MAX_B


def _outputs(B, C, P, dev, compact):
    f32 = dict(device=dev, dtype=torch.float32)
    out = {
        "fhat_t": torch.zeros((B, P, C), device=dev, dtype=torch.bfloat16),
        "inv_norm": torch.zeros((B, P), **f32), "pos_sum": torch.zeros((B, C), **f32),
        "neg_sum": torch.zeros((B, C), **f32), "pos_unit": torch.zeros((B, C), **f32),
        "neg_unit": torch.zeros((B, C), **f32), "norms": torch.zeros((B, 2), **f32),
        "key_count": torch.zeros((B,), device=dev, dtype=torch.int32) if compact else None,
        "perm": torch.zeros((B, P), device=dev, dtype=torch.int32) if compact else None,
    }
    return out


def _p(t):
    return None if t is None else t.data_ptr()


def _run_pair(lib, _lib, B, C, P, seed, compact):
    """One call of each entry point on the same inputs -> (static outputs, reference-entry outputs)."""
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(seed)
    feat = torch.randn(B, C, P, device=dev, generator=g)
    mask = (torch.rand(B, P, device=dev, generator=g) < 0.6).float()
    st = torch.cuda.current_stream().cuda_stream
    # everything either call needs is allocated HERE, by the caller
    a, b = _outputs(B, C, P, dev, compact), _outputs(B, C, P, dev, compact)
    scratch = torch.empty((B, lib.cp2_prep_tiles(P), 2, C), device=dev, dtype=torch.float32)
    counters = torch.zeros((B,), device=dev, dtype=torch.int32)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = lib.cp2_prep_features_static(_p(feat), _p(mask), _p(a["fhat_t"]), _p(a["inv_norm"]), _p(a["pos_sum"]),
                                      _p(a["neg_sum"]), _p(a["key_count"]), _p(a["perm"]), _p(a["pos_unit"]),
                                      _p(a["neg_unit"]), _p(a["norms"]), B, C, P, st)
    _lib.check(rc, "cp2_prep_features_static")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    rc = lib.cp2_prep_features(_p(feat), _p(mask), _p(b["fhat_t"]), _p(b["inv_norm"]), _p(b["pos_sum"]),
                               _p(b["neg_sum"]), _p(scratch), _p(b["key_count"]), _p(b["perm"]), _p(b["pos_unit"]),
                               _p(b["neg_unit"]), _p(b["norms"]), _p(counters), B, C, P, st)
    _lib.check(rc, "cp2_prep_features")
    torch.cuda.synchronize()
    return a, b, free0 - free1


def _same(a, b):
    for k in a:
        if a[k] is None:
            assert b[k] is None
            continue
        assert torch.equal(a[k], b[k]), f"{k}: static-scratch entry differs from cp2_prep_features"
    assert torch.isfinite(a["pos_unit"]).all() and float(a["inv_norm"].min()) > 0


def test_prep_features_static_matches_prep_features_twice():
    from cp2_b200 import _lib
    lib = _lib.load()
    # call 1: the pre-training shape (configs[1]) with key compaction; call 2 right behind it: another shape
    # and other data on the SAME static array (stale partial sums / counters of call 1 must not leak in)
    a, b, _ = _run_pair(lib, _lib, B=32, C=128, P=196, seed=11, compact=True)
    _same(a, b)
    a2, b2, _ = _run_pair(lib, _lib, B=24, C=256, P=784, seed=12, compact=False)
    _same(a2, b2)
    # and the first shape again: same answer as the first time (nothing of call 2 survives)
    a3, b3, grew = _run_pair(lib, _lib, B=32, C=128, P=196, seed=11, compact=True)
    _same(a3, b3)
    _same(a3, a)
    # the kernel of this shape is loaded by now (lazy module loading happens at a first launch): the device's free
    # memory must not shrink across the allocation-free call
    assert grew <= 0, f"device free memory shrank by {grew} bytes across an allocation-free call"


def test_prep_features_static_rejects_sizes_beyond_the_array():
    from cp2_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    x = torch.zeros(16, device=dev)            # never dereferenced: the call must fail before any launch
    p = x.data_ptr()
    # B = 256, P = 3136, C = 128 needs 25.7 MB of partial sums
    rc = lib.cp2_prep_features_static(p, p, p, p, p, p, None, None, None, None, None, 256, 128, 3136, None)
    assert rc != 0
    msg = lib.cp2_last_error()                 # c_char_p restype -> bytes
    assert b"static" in msg and b"bytes" in msg, msg
    rc = lib.cp2_prep_features_static(p, p, p, p, p, p, None, None, None, None, None, STATIC_MAX_B + 1, 64, 1, None)
    assert rc != 0
    # the largest documented shape that fits does fit (B = 256, 28 x 28, C = 128: 6 553 600 bytes)
    assert 256 * lib.cp2_prep_tiles(784) * 2 * 128 * 4 <= STATIC_BYTES - 4 * STATIC_MAX_B
    torch.cuda.synchronize()
