"""GPU parity: every kernel of libcp2_b200.so, called through the C ABI (ctypes), against the
CPU oracle (oracle/cp2_oracle.py) on the same seeded inputs and against the golden vectors
recorded from the unmodified reference.

Tolerances (BASELINE.json north_star): composite / EMA / enqueue bit-exact; losses within 1e-4
relative; gradients of the bf16-operand tensor-core path within rel-L2 2e-2 of the fp32
reference (measured: ~3e-3), TF32 global path likewise.
"""
import numpy as np
import pytest
import torch

from conftest import T, load_golden
from oracle import cp2_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from cp2_b200 import _lib, ops as _ops
    _lib.load()
    return _ops


def dev(t):
    return t.cuda()


def rel_l2(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


# --------------------------------------------------------------------------- K1
@pytest.mark.parametrize("B,H,W,os_", [(4, 64, 64, 16), (3, 224, 224, 16), (2, 30, 45, 8), (1, 17, 19, 4),
                                        (2, 224, 224, 32), (5, 96, 100, 16)])
def test_composite_bit_exact(ops, B, H, W, os_):
    g = torch.Generator().manual_seed(B * 1000 + H)
    img = torch.rand(B, 3, H, W, generator=g) - 0.3
    bg = torch.rand(B, 3, H, W, generator=g)
    bg[:, :, H // 4: H // 2 + 3, W // 5: W // 2 + 1] = 0.0
    bg[0, 0, 0, :] = 0.0                 # channel 0 only -> still fg, bg added on channels 1,2
    bg[-1, 1:, -1, :] = 0.0              # channels 1,2 only -> not fg
    bg[0, 0, 1, 0] = -0.0                # -0.0 == 0 is foreground
    ref, mask = O.composite(img, bg)
    out, mds, mfull = ops.composite(dev(img), dev(bg), os_, want_full_mask=True)
    assert torch.equal(out.cpu().view(torch.int32), ref.view(torch.int32))
    assert torch.equal(mfull.cpu(), mask)
    assert torch.equal(mds.cpu(), O.stride_subsample(mask, os_))
    out2, mds2, none = ops.composite(dev(img), dev(bg), os_)
    assert none is None and torch.equal(out2, out) and torch.equal(mds2, mds)


def test_composite_golden_and_empty(ops):
    g = load_golden("cp2_step_small.npz")
    for s in (0, 1):
        out, mds, _ = ops.composite(dev(T(g[f"s{s}.img_a"])), dev(T(g[f"s{s}.bg0"])), 16)
        assert torch.equal(out.cpu(), T(g[f"s{s}.img_a_comp"]))
    out, mds, _ = ops.composite(torch.empty(0, 3, 32, 32).cuda(), torch.empty(0, 3, 32, 32).cuda(), 16)
    assert out.shape == (0, 3, 32, 32) and mds.shape == (0, 2, 2)


def test_composite_unaligned_views(ops):
    # a non-16-byte-aligned base pointer must take the scalar path and still be exact
    g = torch.Generator().manual_seed(9)
    base_i = torch.rand(2 * 3 * 16 * 16 + 1, generator=g).cuda()
    base_b = torch.rand(2 * 3 * 16 * 16 + 1, generator=g).cuda()
    img = base_i[1:].view(2, 3, 16, 16)
    bg = base_b[1:].view(2, 3, 16, 16)
    bg[:, :, 4:12, 4:12] = 0
    ref, mask = O.composite(img.cpu(), bg.cpu())
    out, mds, _ = ops.composite(img, bg, 8)
    assert torch.equal(out.cpu(), ref) and torch.equal(mds.cpu(), O.stride_subsample(mask, 8))


# --------------------------------------------------------------------------- K4
def test_ema_golden_bit_exact(ops):
    g = load_golden("ema.npz")
    n = int(g["n"])
    q = [dev(T(g[f"q{i}"])) for i in range(n)]
    k = [dev(T(g[f"k{i}"])) for i in range(n)]
    plan = ops.EmaPlan(q, k)
    plan.apply(0.999)
    for i in range(n):
        assert torch.equal(k[i].cpu().view(torch.int32), T(g[f"k1_{i}"]).view(torch.int32)), i
    plan.apply(0.99)
    for i in range(n):
        assert torch.equal(k[i].cpu().view(torch.int32), T(g[f"k2_{i}"]).view(torch.int32)), i


def test_ema_many_tensors_and_unaligned(ops):
    g = torch.Generator().manual_seed(4)
    sizes = [1, 3, 4097, 4096, 8191, 64 * 3 * 7 * 7, 2048 * 512 + 5, 7, 256, 12289]
    big_q = torch.randn(sum(sizes) + 3, generator=g).cuda()
    big_k = torch.randn(sum(sizes) + 3, generator=g).cuda()
    q, k, o = [], [], 1                       # offset 1 float: most slices are not 16 B aligned
    for s in sizes:
        q.append(big_q[o:o + s]); k.append(big_k[o:o + s]); o += s
    ref = O.ema_update([t.cpu() for t in q], [t.cpu() for t in k], 0.999)
    guard = (big_k[0].item(), big_k[-2:].clone())
    plan = ops.EmaPlan(q, k)
    plan.apply(0.999)
    for a, b in zip(k, ref):
        assert torch.equal(a.cpu().view(torch.int32), b.view(torch.int32))
    assert big_k[0].item() == guard[0] and torch.equal(big_k[-2:], guard[1])
    # storage moves -> tables rebuilt
    k2 = [t.clone() for t in k]
    ref2 = O.ema_update([t.cpu() for t in q], [t.cpu() for t in k2], 0.5)
    plan.rebind(q, k2)
    plan.apply(0.5)
    for a, b in zip(k2, ref2):
        assert torch.equal(a.cpu().view(torch.int32), b.view(torch.int32))


# --------------------------------------------------------------------------- K5
def test_enqueue_golden_wraparound(ops):
    g = load_golden("enqueue.npz")
    queue = dev(T(g["queue0"]).clone())
    ptr = torch.zeros(1, dtype=torch.int64).cuda()
    for i, _n in enumerate(g["batches"]):
        ops.enqueue(dev(T(g[f"keys{i}"])), queue, ptr)
        assert int(ptr.item()) == int(g[f"ptr{i + 1}"][0])
        assert torch.equal(queue.cpu(), T(g[f"queue{i + 1}"]))


@pytest.mark.parametrize("n,C,K,ptr0", [(256, 128, 65536, 65536 - 100), (32, 128, 4096, 0), (33, 128, 100, 90),
                                         (100, 128, 100, 37), (8, 256, 72, 70), (0, 128, 64, 5)])
def test_enqueue_random(ops, n, C, K, ptr0):
    g = torch.Generator().manual_seed(n + K)
    queue = torch.randn(C, K, generator=g)
    keys = torch.randn(n, C, generator=g)
    ref = queue.clone()
    p = ptr0
    if n:
        p = O.dequeue_and_enqueue(ref, ptr0, keys)
    qd, pd = dev(queue), torch.tensor([ptr0], dtype=torch.int64).cuda()
    ops.enqueue(dev(keys), qd, pd)
    assert int(pd.item()) == p and torch.equal(qd.cpu(), ref)


# -------------------------------------------------------------------------- K2a
@pytest.mark.parametrize("B,C,P", [(2, 128, 196), (3, 256, 50), (1, 128, 784), (2, 64, 33)])
def test_prep_features(ops, B, C, P):
    g = torch.Generator().manual_seed(P)
    feat = torch.randn(B, C, P, generator=g) * 3
    feat[0, :, 1] = 0.0                                   # zero pixel: eps path of F.normalize
    mask = (torch.rand(B, P, generator=g) < 0.6).float()
    f_hat, pos, neg = O.normalize_and_pool(feat, mask)
    fh, inv, ps, ns, cnt = ops.prep_features(dev(feat), dev(mask))
    assert cnt is None
    assert fh.shape == (B, P, C) and fh.dtype == torch.bfloat16
    want = f_hat.transpose(1, 2)
    assert torch.allclose(fh.float().cpu(), want, rtol=0, atol=2 ** -8)     # bf16 rounding of |v|<=1
    assert rel_l2(fh.float(), want) < 3e-3
    n = feat.norm(dim=1).clamp_min(1e-12)
    assert torch.allclose(inv.cpu(), 1.0 / n, rtol=2e-6)
    raw_pos = torch.einsum("ncx,nx->nc", f_hat, mask)
    raw_neg = torch.einsum("ncx,nx->nc", f_hat, 1 - mask)
    assert torch.allclose(ps.cpu(), raw_pos, rtol=1e-5, atol=1e-5)
    assert torch.allclose(ns.cpu(), raw_neg, rtol=1e-5, atol=1e-5)
    assert torch.allclose(torch.nn.functional.normalize(ps.cpu(), dim=1), pos, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("B,C,P", [(2, 128, 196), (2, 256, 50), (1, 128, 3136), (3, 128, 33)])
def test_prep_features_compact_keys(ops, B, C, P):
    """compact=True: foreground rows first (pixel order), background rows after (pixel order), key_count = #fg,
    perm = the pixel of each row; everything else identical to the plain call."""
    g = torch.Generator().manual_seed(P + 1)
    feat = torch.randn(B, C, P, generator=g)
    mask = (torch.rand(B, P, generator=g) < 0.6).float()
    mask[0] = 1.0                                         # all foreground
    if B > 1:
        mask[1] = 0.0                                     # no foreground at all
    fh, inv, ps, ns, _ = ops.prep_features(dev(feat), dev(mask))
    perm = torch.full((B, P), -1, dtype=torch.int32, device="cuda")
    fc, inv2, ps2, ns2, cnt = ops.prep_features(dev(feat), dev(mask), compact=True, perm=perm)
    assert torch.equal(inv, inv2) and torch.equal(ps, ps2) and torch.equal(ns, ns2)
    assert cnt.dtype == torch.int32 and torch.equal(cnt.cpu().long(), mask.sum(1).long())
    for b in range(B):
        n = int(cnt[b])
        fg = mask[b].bool().cuda()
        assert torch.equal(fc[b, :n], fh[b][fg])
        assert torch.equal(fc[b, n:], fh[b][~fg])
        want = torch.cat([torch.nonzero(fg).flatten(), torch.nonzero(~fg).flatten()]).int()
        assert torch.equal(perm[b], want)


# ------------------------------------------------------------------- K2 forward
def _unit_bf16(B, P, C, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.nn.functional.normalize(torch.randn(B, P, C, generator=g), dim=2)
    return x.to(torch.bfloat16)


def _lse_ref(qh, kh, inv_tau):
    s = torch.einsum("nxc,nyc->nxy", qh.double(), kh.double()) * inv_tau
    return torch.logsumexp(s, dim=1)


@pytest.mark.parametrize("B,C,P,tau", [(2, 128, 196, 0.2), (1, 128, 128, 1.0), (3, 128, 50, 0.2),
                                       (2, 256, 196, 0.2), (1, 128, 784, 0.2), (1, 256, 300, 1.0),
                                       (2, 128, 1, 0.2)])
def test_dense_lse_vs_oracle(ops, B, C, P, tau):
    qh, kh = _unit_bf16(B, P, C, 1), _unit_bf16(B, P, C, 2)
    ref = _lse_ref(qh.float(), kh.float(), 1.0 / tau)
    got = ops.dense_lse(dev(qh), dev(kh), 1.0 / tau)
    torch.cuda.synchronize()
    # a quarter of the exponentials run as a degree-3 polynomial on the FMA pipe (max rel. error 7.6e-5, mean
    # bias 5e-6): |d lse| stays below 5e-5, i.e. < 1e-5 relative on a loss whose tolerance is 1e-4
    assert torch.allclose(got.double().cpu(), ref, rtol=0, atol=5e-5), float((got.double().cpu() - ref).abs().max())


def test_dense_lse_compact_count(ops):
    """key_count: rows < count match the full computation, tiles entirely past count are zeroed."""
    B, C, P = 3, 128, 700
    qh, kh = dev(_unit_bf16(B, P, C, 21)), dev(_unit_bf16(B, P, C, 22))
    cnt = torch.tensor([700, 130, 0], dtype=torch.int32).cuda()
    full = ops.dense_lse(qh, kh, 5.0)
    got = ops.dense_lse(qh, kh, 5.0, cnt)
    torch.cuda.synchronize()
    assert torch.equal(got[0], full[0])
    assert torch.equal(got[1, :256], full[1, :256]) and not got[1, 256:].any()
    assert not got[2].any()


# ------------------------------------------------------------------ K2 backward
def _grad_ref(qh, kh, lse, ma, mb, inv_norm, alpha, v1, v2, inv_tau):
    """fp64 restatement of the cp2_dense_grad contract (include/cp2_b200.h)."""
    qh, kh = qh.double(), kh.double()
    s = torch.einsum("nxc,nyc->nxy", qh, kh) * inv_tau
    p = torch.exp(s - lse.double()[:, None, :]) * mb.double()[:, None, :]
    acc = torch.einsum("nxy,nyc->nxc", p, kh)
    g = alpha.double()[:, None, None] * acc + ma.double()[:, :, None] * v1.double()[:, None, :]
    if v2 is not None:
        g = g + (1 - ma.double())[:, :, None] * v2.double()[:, None, :]
    d = (g - qh * (qh * g).sum(-1, keepdim=True)) * inv_norm.double()[:, :, None]
    return d.transpose(1, 2)                                # [B,C,P]


@pytest.mark.parametrize("B,C,P,tau,bg", [(2, 128, 196, 0.2, False), (1, 128, 128, 1.0, True),
                                          (2, 128, 50, 0.2, True), (1, 256, 196, 0.2, False),
                                          (1, 128, 784, 0.2, True), (1, 256, 300, 1.0, True)])
def test_dense_grad_vs_oracle(ops, B, C, P, tau, bg):
    g = torch.Generator().manual_seed(P + C)
    qh, kh = _unit_bf16(B, P, C, 5), _unit_bf16(B, P, C, 6)
    ma = (torch.rand(B, P, generator=g) < 0.6).float()
    mb = (torch.rand(B, P, generator=g) < 0.6).float()
    inv_norm = torch.rand(B, P, generator=g) + 0.5
    alpha = torch.rand(B, generator=g) + 0.1
    v1 = torch.randn(B, C, generator=g) * 0.1
    v2 = torch.randn(B, C, generator=g) * 0.1 if bg else None
    lse = _lse_ref(qh.float(), kh.float(), 1.0 / tau).float()
    ref = _grad_ref(qh.float(), kh.float(), lse, ma, mb, inv_norm, alpha, v1, v2, 1.0 / tau)
    got = ops.dense_grad(dev(qh), dev(kh), dev(lse), dev(ma), dev(mb), dev(inv_norm), dev(alpha), dev(v1),
                         None if v2 is None else dev(v2), 1.0 / tau)
    torch.cuda.synchronize()
    assert got.shape == (B, C, P)
    err = rel_l2(got, ref)
    assert err < 5e-3, err                                   # Ptilde is rounded to bf16 (2^-9)


@pytest.mark.parametrize("C,P", [(128, 500), (256, 333)])
def test_dense_grad_compact_equals_masked(ops, C, P):
    """Compacted keys + key_count give the gradient of the mask_b formulation (same fp64 reference)."""
    B, tau = 3, 0.2
    g = torch.Generator().manual_seed(C + P)
    feat_k = torch.randn(B, C, P, generator=g)
    qh = _unit_bf16(B, P, C, 31)
    ma = (torch.rand(B, P, generator=g) < 0.6).float()
    mb = (torch.rand(B, P, generator=g) < 0.6).float()
    mb[1] = 0.0                                           # an image without foreground keys
    mb[2, :P // 2] = 1.0
    inv_norm = torch.rand(B, P, generator=g) + 0.5
    alpha = torch.rand(B, generator=g) + 0.1
    v1 = torch.randn(B, C, generator=g) * 0.1
    kh_full, _, _, _, _ = ops.prep_features(dev(feat_k), dev(mb))
    kh_c, _, _, _, cnt = ops.prep_features(dev(feat_k), dev(mb), compact=True)
    lse_c = ops.dense_lse(dev(qh), kh_c, 1.0 / tau, cnt)
    got = ops.dense_grad(dev(qh), kh_c, lse_c, dev(ma), None, dev(inv_norm), dev(alpha), dev(v1), None,
                         1.0 / tau, key_count=cnt)
    torch.cuda.synchronize()
    lse = _lse_ref(qh.float(), kh_full.float().cpu(), 1.0 / tau).float()
    ref = _grad_ref(qh.float(), kh_full.float().cpu(), lse, ma, mb, inv_norm, alpha, v1, None, 1.0 / tau)
    assert rel_l2(got, ref) < 5e-3, rel_l2(got, ref)
    # image 1 has no fg keys: only the rank-1 terms remain
    only_rank1 = _grad_ref(qh[1:2].float(), kh_full[1:2].float().cpu(), lse[1:2], ma[1:2], mb[1:2], inv_norm[1:2],
                           alpha[1:2] * 0, v1[1:2], None, 1.0 / tau)
    assert rel_l2(got[1:2], only_rank1) < 1e-5


# ------------------------------------------------------------------ K2-small (maps that fit TMEM)
@pytest.mark.parametrize("B,C,P,tau,bg", [(2, 128, 196, 0.2, False), (3, 128, 196, 1.0, True), (1, 128, 128, 0.2, True),
                                          (2, 128, 50, 0.2, False), (2, 128, 256, 0.2, True), (1, 128, 129, 0.2, False),
                                          (2, 128, 1, 0.2, False), (2, 256, 49, 0.2, True), (1, 256, 128, 1.0, False),
                                          (33, 128, 196, 0.2, False)])
def test_dense_small_vs_oracle(ops, B, C, P, tau, bg):
    """cp2_dense_small + cp2_dense_small_bwd against the fp64 restatement of the lse / grad contracts."""
    assert ops.dense_small_ok(C, P)
    g = torch.Generator().manual_seed(P * 3 + C + B)
    qh, kh = _unit_bf16(B, P, C, 50 + P), _unit_bf16(B, P, C, 51 + P)
    ma = (torch.rand(B, P, generator=g) < 0.6).float()
    mb = (torch.rand(B, P, generator=g) < 0.6).float()
    inv_norm = torch.rand(B, P, generator=g) + 0.5
    alpha = torch.rand(B, generator=g) + 0.1
    v1 = torch.randn(B, C, generator=g) * 0.1
    v2 = torch.randn(B, C, generator=g) * 0.1 if bg else None
    lse_ref = _lse_ref(qh.float(), kh.float(), 1.0 / tau)
    lse, acc = ops.dense_small(dev(qh), dev(kh), 1.0 / tau, mask_b=dev(mb))
    gs = torch.tensor([0.75], device="cuda")
    got = ops.dense_small_bwd(acc, dev(qh), dev(ma), dev(inv_norm), dev(alpha), dev(v1),
                              None if v2 is None else dev(v2), gscale=gs)
    torch.cuda.synchronize()
    assert torch.allclose(lse.double().cpu(), lse_ref, rtol=0, atol=5e-5), float((lse.double().cpu() - lse_ref).abs().max())
    ref = 0.75 * _grad_ref(qh.float(), kh.float(), lse_ref.float(), ma, mb, inv_norm, alpha, v1, v2, 1.0 / tau)
    assert got.shape == (B, C, P)
    err = rel_l2(got, ref)
    assert err < 5e-3, err


def test_dense_small_ok_shapes(ops):
    assert ops.dense_small_ok(128, 196) and ops.dense_small_ok(128, 256) and ops.dense_small_ok(256, 128)
    assert not ops.dense_small_ok(128, 257) and not ops.dense_small_ok(256, 196) and not ops.dense_small_ok(64, 100)


@pytest.mark.parametrize("C,P", [(128, 196), (128, 250), (256, 100)])
def test_dense_small_compact_matches_streaming_kernels(ops, C, P):
    """Compacted keys + key_count: the single-kernel path equals cp2_dense_lse + cp2_dense_grad (same operands,
    same bf16 rounding of P~) and the fp64 reference of the mask_b formulation; an image without foreground
    keys keeps only the rank-1 terms."""
    B, tau = 4, 0.2
    g = torch.Generator().manual_seed(C + P)
    feat_k = torch.randn(B, C, P, generator=g)
    qh = _unit_bf16(B, P, C, 61)
    ma = (torch.rand(B, P, generator=g) < 0.6).float()
    mb = (torch.rand(B, P, generator=g) < 0.6).float()
    mb[1] = 0.0                                           # no foreground key
    mb[2] = 1.0                                           # every key
    mb[3, 5:] = 0.0                                       # a handful of keys (one 16-row slice)
    inv_norm = torch.rand(B, P, generator=g) + 0.5
    alpha = torch.rand(B, generator=g) + 0.1
    v1 = torch.randn(B, C, generator=g) * 0.1
    v2 = torch.randn(B, C, generator=g) * 0.1
    kh_full, _, _, _, _ = ops.prep_features(dev(feat_k), dev(mb))
    kh_c, _, _, _, cnt = ops.prep_features(dev(feat_k), dev(mb), compact=True)
    lse_s, acc = ops.dense_small(dev(qh), kh_c, 1.0 / tau, key_count=cnt)
    got = ops.dense_small_bwd(acc, dev(qh), dev(ma), dev(inv_norm), dev(alpha), dev(v1), dev(v2), key_count=cnt)
    lse_t = ops.dense_lse(dev(qh), kh_c, 1.0 / tau, cnt)
    ref_t = ops.dense_grad(dev(qh), kh_c, lse_t, dev(ma), None, dev(inv_norm), dev(alpha), dev(v1), dev(v2),
                           1.0 / tau, key_count=cnt)
    torch.cuda.synchronize()
    cn = cnt.cpu().tolist()
    for b in range(B):
        assert torch.allclose(lse_s[b, :cn[b]], lse_t[b, :cn[b]], rtol=0, atol=2e-5)
        assert not lse_s[b, cn[b]:].any()
    assert rel_l2(got, ref_t) < 2e-3, rel_l2(got, ref_t)
    lse = _lse_ref(qh.float(), kh_full.float().cpu(), 1.0 / tau).float()
    ref = _grad_ref(qh.float(), kh_full.float().cpu(), lse, ma, mb, inv_norm, alpha, v1, v2, 1.0 / tau)
    assert rel_l2(got, ref) < 5e-3, rel_l2(got, ref)
    only_rank1 = _grad_ref(qh[1:2].float(), kh_full[1:2].float().cpu(), lse[1:2], ma[1:2], mb[1:2], inv_norm[1:2],
                           alpha[1:2] * 0, v1[1:2], v2[1:2], 1.0 / tau)
    assert rel_l2(got[1:2], only_rank1) < 1e-5


# --------------------------------------------------------------------------- K3
@pytest.mark.parametrize("B,K,tau,C", [(32, 65536, 0.2, 128), (4, 4096, 0.2, 128), (2, 1024, 0.07, 128),
                                       (64, 8192, 0.2, 128), (100, 4096, 0.2, 128), (7, 1000, 0.2, 128),
                                       (1, 256, 1.0, 128), (32, 65536, 0.2, 256), (5, 1000, 0.2, 256),
                                       (40, 300, 0.2, 256), (3, 100, 0.2, 128)])
def test_global_nce_vs_oracle(ops, B, K, tau, C):
    g = torch.Generator().manual_seed(B + K)
    qp = torch.nn.functional.normalize(torch.randn(B, C, generator=g), dim=1)
    queue = torch.nn.functional.normalize(torch.randn(C, K, generator=g), dim=0)
    e = torch.exp((qp.double() @ queue.double() - 1.0) / tau)
    z_ref, g_ref = e.sum(dim=1), e @ queue.double().T
    z, gg = ops.global_nce(dev(qp), dev(queue), 1.0 / tau)[:2]
    torch.cuda.synchronize()
    # TF32 operands (10-bit mantissa): logits off by <~1e-3 -> e off by <~1e-3/tau relative
    assert torch.allclose(z.double().cpu(), z_ref, rtol=2e-3 / tau * 0.2 + 1e-3), (z.cpu(), z_ref)
    assert rel_l2(gg, g_ref) < 5e-3, rel_l2(gg, g_ref)


def test_global_nce_vs_simt(ops):
    from cp2_b200 import _lib
    lib = _lib.load_debug()
    B, C, K = 32, 128, 65536
    g = torch.Generator().manual_seed(0)
    qp = dev(torch.nn.functional.normalize(torch.randn(B, C, generator=g), dim=1))
    queue = dev(torch.nn.functional.normalize(torch.randn(C, K, generator=g), dim=0))
    z, gg = ops.global_nce(qp, queue, 5.0)[:2]
    z2, g2 = torch.empty_like(z), torch.empty_like(gg)
    _lib.check(lib.cp2_dbg_global_nce_simt(qp.data_ptr(), queue.data_ptr(), z2.data_ptr(), g2.data_ptr(), 5.0,
                                           B, C, K, torch.cuda.current_stream().cuda_stream), "simt")
    torch.cuda.synchronize()
    assert torch.allclose(z, z2, rtol=2e-3) and rel_l2(gg, g2) < 5e-3


# ------------------------------------------------------------------- a2 + a4
def test_id_iou_reference_cases(ops):
    """The reference's own unit-test constants (tests/test_correlation_mapping.py:73-74, 126-127) and the
    golden outputs of get_masked_correlation_map recorded by oracle/gen_golden.gen_corr; output stride 1."""
    g = load_golden("corr.npz")
    for p, iou, ioum in (("u", 12 / 38, 1 / 3), ("s", 4 / 7, 2 / 3)):
        a, b = T(g[f"{p}_map_a"]).long(), T(g[f"{p}_map_b"]).long()
        r_iou, r_ioum = ops.id_iou(dev(a), dev(b), dev(T(g[f"{p}_mask_a"])), dev(T(g[f"{p}_mask_b"])), 1)
        assert torch.equal(r_iou.cpu(), T(g[f"{p}_iou"]).reshape(-1))
        assert torch.equal(r_ioum.cpu(), T(g[f"{p}_iou_masked"]).reshape(-1))
        assert torch.all(r_iou.cpu() == torch.tensor(iou, dtype=torch.float32))
        assert torch.all(r_ioum.cpu() == torch.tensor(ioum, dtype=torch.float32))


@pytest.mark.parametrize("B,H,W,os_", [(4, 224, 224, 16), (3, 64, 48, 8), (2, 30, 45, 4), (2, 112, 112, 2)])
def test_id_iou_vs_oracle(ops, B, H, W, os_):
    """Two crops of one id canvas (pixel ids) and a coarse region-id map with repeated and zero ids,
    sub-sampled at [OS/2::OS] inside the kernel (builder.py:1165-1186); bit-exact."""
    gen = torch.Generator().manual_seed(B + H)
    canvas = torch.arange(1, (H + 9) * (W + 7) + 1).reshape(H + 9, W + 7)
    offs = torch.randint(0, 7, (B, 4), generator=gen)
    ids_a = torch.stack([canvas[int(o[0]):int(o[0]) + H, int(o[1]):int(o[1]) + W] for o in offs])
    ids_b = torch.stack([canvas[int(o[2]):int(o[2]) + H, int(o[3]):int(o[3]) + W] for o in offs])
    for kind in ("pixel", "region"):
        if kind == "region":
            ids_a, ids_b = (ids_a // 97) % 5, (ids_b // 97) % 5        # few distinct ids, zeros included
        mask_a = (torch.rand(B, H, W, generator=gen) < 0.6).float()
        mask_b = (torch.rand(B, H, W, generator=gen) < 0.6).float()
        ma, mb = O.stride_subsample(mask_a, os_), O.stride_subsample(mask_b, os_)
        ma[:, 0, 0] = 1.0
        mb[:, 0, 0] = 1.0                                             # at least one unmasked entry per map
        sa, sb = O.stride_subsample(ids_a, os_), O.stride_subsample(ids_b, os_)
        ref = O.masked_correlation(sa, sb, ma, mb)
        iou, ioum = ops.id_iou(dev(ids_a), dev(ids_b), dev(ma), dev(mb), os_)
        assert torch.equal(iou.cpu(), ref["iou"]), kind
        assert torch.equal(ioum.cpu(), ref["iou_masked"]), kind


# ------------------------------------------------------------------ f4: accuracy meters
@pytest.mark.parametrize("B,C,hw", [(4, 128, 14), (3, 256, 9), (2, 128, 28), (5, 64, 7)])
def test_dense_top1_vs_oracle(ops, B, C, hw):
    """cp2_dense_top1 against builder.py:1442-1448 as restated by the oracle, image by image."""
    g = torch.Generator().manual_seed(31 * hw + B)
    P = hw * hw
    q = torch.randn(B, C, P, generator=g)
    k = torch.randn(B, C, P, generator=g)
    ma = (torch.rand(B, P, generator=g) < 0.6).float()
    mb = (torch.rand(B, P, generator=g) < 0.6).float()
    acc, hit = ops.dense_top1(q.cuda(), k.cuda(), ma.cuda(), mb.cuda())
    qh, kh = torch.nn.functional.normalize(q, dim=1), torch.nn.functional.normalize(k, dim=1)
    want = [O.dense_accuracy(qh[i:i + 1], kh[i:i + 1], ma[i:i + 1], mb[i:i + 1]).item() / 100.0 for i in range(B)]
    assert hit.cpu().tolist() == want
    np.testing.assert_allclose(acc.item(), O.dense_accuracy(qh, kh, ma, mb).item(), rtol=1e-6)
    # a planted maximum at a known pair, on a background key: miss; then on a foreground pair: hit
    q2, k2 = q.clone(), k.clone()
    x, y = P - 1, 3
    k2[:, :, y] = q2[:, :, x] * 2.5                       # cosine 1 with query pixel x
    ma2, mb2 = ma.clone(), mb.clone()
    ma2[:, x], mb2[:, y] = 1.0, 0.0
    _, hit2 = ops.dense_top1(q2.cuda(), k2.cuda(), ma2.cuda(), mb2.cuda())
    assert hit2.cpu().tolist() == [0.0] * B
    mb2[:, y] = 1.0
    acc3, hit3 = ops.dense_top1(q2.cuda(), k2.cuda(), ma2.cuda(), mb2.cuda())
    assert hit3.cpu().tolist() == [1.0] * B and acc3.item() == 100.0


def _seg_hits(ops, q, k, ma, mb, tau=0.2):
    """acc_seg out of the dense kernel itself, the way the step calls it: compacted keys + perm."""
    B, C, P = q.shape
    qd, kd, mad, mbd = q.cuda(), k.cuda(), ma.cuda(), mb.cuda()
    (qh, q_inv, *_), (kh, k_inv, _, _, cnt, _, _, _, perm) = ops.prep_features_pair(qd, mad, kd, mbd, want_perm=True)
    _, _, hit = ops.dense_small(qh, kh, 1.0 / tau, cnt, seg=(qd, kd, q_inv, k_inv, perm, mad, mbd))
    torch.cuda.synchronize()
    return hit.cpu()


@pytest.mark.parametrize("B,C,hw", [(4, 128, 14), (6, 128, 16), (3, 128, 9), (3, 256, 9), (5, 128, 11), (40, 128, 14)])
def test_dense_small_acc_seg_vs_oracle(ops, B, C, hw):
    """The acc_seg meter found by cp2_dense_small while S is in TMEM (bf16-operand maximum, candidates within 2^-7
    re-scored in fp32) against builder.py:1442-1448 as restated by the oracle, image by image, incl. planted
    maxima and a near-tie that bf16 rounding alone would decide wrongly."""
    g = torch.Generator().manual_seed(17 * hw + B)
    P = hw * hw
    q = torch.randn(B, C, P, generator=g)
    k = torch.randn(B, C, P, generator=g)
    ma = (torch.rand(B, P, generator=g) < 0.6).float()
    mb = (torch.rand(B, P, generator=g) < 0.6).float()
    if B > 2:
        mb[1] = 0.0                                       # no foreground key at all
        mb[2] = 1.0
    qh, kh = torch.nn.functional.normalize(q, dim=1), torch.nn.functional.normalize(k, dim=1)
    want = [O.dense_accuracy(qh[i:i + 1], kh[i:i + 1], ma[i:i + 1], mb[i:i + 1]).item() / 100.0 for i in range(B)]
    assert _seg_hits(ops, q, k, ma, mb).tolist() == want
    # a planted maximum at a known pair, on a background key: miss; then on a foreground pair: hit
    q2, k2 = q.clone(), k.clone()
    x, y = P - 1, 3
    k2[:, :, y] = q2[:, :, x] * 2.5                       # cosine 1 with query pixel x
    ma2, mb2 = ma.clone(), mb.clone()
    ma2[:, x], mb2[:, y] = 1.0, 0.0
    assert _seg_hits(ops, q2, k2, ma2, mb2).tolist() == [0.0] * B
    mb2[:, y] = 1.0
    assert _seg_hits(ops, q2, k2, ma2, mb2).tolist() == [1.0] * B
    # near-tie: a second pair 2e-4 below the planted maximum in fp32 (far inside the bf16 rounding of the operands),
    # with the opposite label: the fp32 re-scoring has to pick the planted one, whatever bf16 says
    q3, k3 = q2.clone(), k2.clone()
    x2, y2 = 1, P - 2
    v = q3[:, :, x2] / q3[:, :, x2].norm(dim=1, keepdim=True)
    r = torch.randn(B, C, generator=g)
    r = r - (r * v).sum(1, keepdim=True) * v
    r = r / r.norm(dim=1, keepdim=True)
    k3[:, :, y2] = v * (1 - 2e-4) + r * (1 - (1 - 2e-4) ** 2) ** 0.5     # cosine 1 - 2e-4 with query pixel x2
    ma3, mb3 = ma2.clone(), mb2.clone()
    ma3[:, x2], mb3[:, y2] = 1.0, 0.0                     # runner-up is a non-label pair, winner a label pair
    qh3, kh3 = torch.nn.functional.normalize(q3, dim=1), torch.nn.functional.normalize(k3, dim=1)
    want3 = [O.dense_accuracy(qh3[i:i + 1], kh3[i:i + 1], ma3[i:i + 1], mb3[i:i + 1]).item() / 100.0 for i in range(B)]
    assert want3 == [1.0] * B
    assert _seg_hits(ops, q3, k3, ma3, mb3).tolist() == want3


@pytest.mark.parametrize("B,K,bg", [(32, 65536, False), (4, 1000, True), (70, 4096, False)])
def test_global_top1_vs_oracle(ops, B, K, bg):
    """cp2_global_top1 against `_accuracy(logits_moco, 0, topk=(1,))` (builder.py:1441, 1690-1706)."""
    g = torch.Generator().manual_seed(K + B)
    C = 128
    nrm = torch.nn.functional.normalize
    queue = nrm(torch.randn(C, K, generator=g), dim=0)
    q_pos = nrm(torch.randn(B, C, generator=g), dim=1)
    # positives of graded quality: some beat every negative, some do not
    k_pos = nrm(q_pos * torch.linspace(0.0, 1.2, B)[:, None] + torch.randn(B, C, generator=g) * 0.08, dim=1)
    q_neg = nrm(torch.randn(B, C, generator=g), dim=1)
    k_neg = nrm(q_pos * 0.6 + torch.randn(B, C, generator=g) * 0.08, dim=1)
    logits = O.global_logits(q_pos, k_pos, queue, q_neg, k_neg, include_background=bg)
    want = (logits.argmax(dim=1) == 0).float()
    assert 0 < want.sum() < B
    acc, hit = ops.global_top1(q_pos.cuda(), k_pos.cuda(), queue.cuda(), q_neg.cuda() if bg else None,
                               k_neg.cuda() if bg else None)
    assert torch.equal(hit.cpu(), want)
    np.testing.assert_allclose(acc.item(), O.topk_accuracy(logits, topk=(1,))[0].item(), rtol=1e-6)


def test_global_top1_near_ties_are_rescored_exactly(ops):
    """Queue entries whose logit lies within the bf16 error band of l_pos (|l_k - l_pos| ~ 1e-5 .. 1e-3, on either
    side) must be decided by their exact fp32 logits, not by the bf16 operand the tensor cores see."""
    B, C, K = 16, 128, 4096
    g = torch.Generator().manual_seed(3)
    nrm = torch.nn.functional.normalize
    queue = nrm(torch.randn(C, K, generator=g), dim=0) * 1.0
    q_pos = nrm(torch.randn(B, C, generator=g), dim=1)
    k_pos = nrm(q_pos * 2.0 + torch.randn(B, C, generator=g) * 0.1, dim=1)       # l_pos ~ 0.99 > every random negative
    l_pos = (q_pos * k_pos).sum(1)
    deltas = torch.tensor([3e-5, -3e-5, 2e-4, -2e-4, 8e-4, -8e-4, 2e-3, -2e-3] * 2)
    for b in range(B):
        # an entry u = unit vector with q_pos[b].u = l_pos[b] + delta: rotate k_pos[b] inside span(q, k)
        target = float(l_pos[b] + deltas[b])
        perp = nrm(k_pos[b] - l_pos[b] * q_pos[b], dim=0)
        u = target * q_pos[b] + (1.0 - target * target) ** 0.5 * perp
        queue[:, 37 + 101 * b] = u
    logits = O.global_logits(q_pos.double(), k_pos.double(), queue.double())
    want = (logits.argmax(dim=1) == 0).float()
    assert 0 < want.sum() < B
    margin = (logits[:, 1:].max(dim=1).values - logits[:, 0]).abs()
    assert float(margin.min()) > 1e-5                       # decidable in fp32
    acc, hit = ops.global_top1(q_pos.cuda(), k_pos.cuda(), queue.cuda())
    assert torch.equal(hit.cpu(), want), (hit.cpu(), want, margin)
