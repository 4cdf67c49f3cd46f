"""CPU: pins oracle/cp2_oracle.py against golden vectors recorded from the unmodified
reference (oracle/gen_golden.py) and against the constants in the reference's own unit
tests (tests/test_correlation_mapping.py:73-74,126-127; test_contrastive_metrics.py:53-55)."""
import numpy as np
import pytest
import torch

from conftest import T, load_golden
from oracle import cp2_oracle as O


def test_composite_and_subsample_match_reference_bit_exact():
    g = load_golden("cp2_step_small.npz")
    for s in (0, 1):
        out, mask = O.composite(T(g[f"s{s}.img_a"]), T(g[f"s{s}.bg0"]))
        assert torch.equal(out, T(g[f"s{s}.img_a_comp"]))
        assert mask.shape == (4, 64, 64)
        assert O.stride_subsample(mask, 16).shape == (4, 4, 4)


@pytest.mark.parametrize("tag", ["t1", "t02", "t02_bg"])
def test_loss_path_p196(tag):
    g = load_golden(f"cp2_loss_p196_{tag}.npz")
    q = T(g["q"]).requires_grad_(True)
    mask_a = O.stride_subsample(T(g["bg0_ch0_zero"]).float(), 16).reshape(2, -1)
    mask_b = O.stride_subsample(T(g["bg1_ch0_zero"]).float(), 16).reshape(2, -1)
    r = O.cp2_loss(q, T(g["k"]), mask_a, mask_b, T(g["queue0"]),
                   temp_global=float(g["temp_global"]), temp_local=float(g["temp_local"]),
                   lmbd_dense=float(g["lmbd"]), include_background=bool(g["include_background"]))
    r["loss"].backward()
    np.testing.assert_allclose(r["loss"].item(), g["loss"], rtol=1e-6)
    np.testing.assert_allclose(r["loss_ins"].item(), g["loss_ins"], rtol=1e-6)
    np.testing.assert_allclose(r["loss_dense"].item(), g["loss_dense"], rtol=1e-6)
    np.testing.assert_allclose(q.grad.numpy(), g["grad_q"], rtol=1e-4, atol=1e-9)
    np.testing.assert_allclose(r["acc_dense"].item(), g["acc_seg"], rtol=0, atol=1e-4)
    np.testing.assert_allclose(r["acc1"].item(), g["acc_ins"], rtol=0, atol=1e-4)
    queue = T(g["queue0"]).clone()
    ptr = O.dequeue_and_enqueue(queue, 0, r["k_pos"])
    assert ptr == int(g["queue_ptr1"][0])
    np.testing.assert_allclose(queue[:, :8].numpy(), g["queue1_head"], rtol=1e-6, atol=1e-8)
    ids = torch.arange(1, 224 * 224 + 1).reshape(1, 224, 224).repeat(2, 1, 1)
    ids = O.stride_subsample(ids, 16)
    mc = O.masked_correlation(ids, ids, mask_a.reshape(2, 14, 14), mask_b.reshape(2, 14, 14))
    np.testing.assert_allclose(mc["iou"].numpy(), g["iou"])
    np.testing.assert_allclose(mc["iou_masked"].numpy(), g["iou_masked"])
    w = O.cp2_corr_weights(ids, ids, ids, ids, mask_a, mask_b)
    assert w.dtype == torch.int64 and bool((w == 1).all())


WEIGHTED_TAGS = ["none_t1", "fixed_t05", "avg_t1_bg", "hard_t1", "median_t05"]


def weighted_case(tag):
    """Inputs of a PROPOSED-mode fixture (oracle/gen_golden.py:gen_weighted)."""
    g = load_golden(f"cp2_weighted_{tag}.npz")
    mask_a = O.stride_subsample(T(g["bg0_ch0_zero"]).float(), 16).reshape(2, -1)
    mask_b = O.stride_subsample(T(g["bg1_ch0_zero"]).float(), 16).reshape(2, -1)
    ids = {n: T(g[n + "_ds"]) for n in ("pixel_ids_a", "pixel_ids_b", "region_ids_a", "region_ids_b")}
    return g, mask_a, mask_b, ids


@pytest.mark.parametrize("tag", WEIGHTED_TAGS)
def test_weighted_loss_path(tag):
    """f3: PretrainType.PROPOSED / PIXEL_REGION_ID weights + NegativeType (builder.py:1204-1243, 1332-1392)."""
    g, mask_a, mask_b, ids = weighted_case(tag)
    wp, wr, wn = (float(x) for x in g["weights"])
    w = O.cp2_corr_weights(ids["pixel_ids_a"], ids["pixel_ids_b"], ids["region_ids_a"], ids["region_ids_b"],
                           mask_a, mask_b, wp, wr, wn)
    assert w.dtype == torch.float32 and len(torch.unique(w)) == 3     # all three kinds of pair occur
    q = T(g["q"]).requires_grad_(True)
    r = O.cp2_loss(q, T(g["k"]), mask_a, mask_b, T(g["queue0"]), temp_global=float(g["temp_global"]),
                   temp_local=float(g["temp_local"]), lmbd_dense=float(g["lmbd"]),
                   include_background=bool(g["include_background"]), weights=w,
                   negative_type=str(g["negative_type"]), negative_scale=float(g["negative_scale"]))
    r["loss"].backward()
    np.testing.assert_allclose(r["loss"].item(), g["loss"], rtol=1e-6)
    np.testing.assert_allclose(r["loss_dense"].item(), g["loss_dense"], rtol=1e-6)
    np.testing.assert_allclose(q.grad.numpy(), g["grad_q"], rtol=1e-4, atol=1e-9)
    # the fixture is not the CP2-mode loss in disguise
    q2 = T(g["q"]).requires_grad_(True)
    r2 = O.cp2_loss(q2, T(g["k"]), mask_a, mask_b, T(g["queue0"]), temp_local=float(g["temp_local"]),
                    include_background=bool(g["include_background"]))
    r2["loss"].backward()
    assert (q.grad - q2.grad).abs().max() > 2e-8 and abs(r2["loss_dense"].item() - float(g["loss_dense"])) > 1e-5
    ra, rb = ids["region_ids_a"], ids["region_ids_b"]
    mc = O.masked_correlation(ra, rb, mask_a.reshape(2, 14, 14), mask_b.reshape(2, 14, 14))
    np.testing.assert_allclose(mc["iou"].numpy(), g["iou"])
    np.testing.assert_allclose(mc["iou_masked"].numpy(), g["iou_masked"])


def test_full_step_small():
    g = load_golden("cp2_step_small.npz")
    queue = T(g["queue0"]).clone()
    ptr = 0
    for s in (0, 1):
        _, ma = O.composite(T(g[f"s{s}.img_a"]), T(g[f"s{s}.bg0"]))
        _, mb = O.composite(T(g[f"s{s}.img_b"]), T(g[f"s{s}.bg1"]))
        ma = O.stride_subsample(ma, 16).reshape(4, -1)
        mb = O.stride_subsample(mb, 16).reshape(4, -1)
        q = T(g[f"s{s}.q"]).requires_grad_(True)
        r = O.cp2_loss(q, T(g[f"s{s}.k"]), ma, mb, queue, 0.2, 0.2, 0.2)
        r["loss"].backward()
        np.testing.assert_allclose(r["loss"].item(), g[f"s{s}.loss"], rtol=1e-6)
        np.testing.assert_allclose(q.grad.numpy(), g[f"s{s}.grad_q"], rtol=1e-4, atol=1e-9)
        ptr = O.dequeue_and_enqueue(queue, ptr, r["k_pos"])
        assert ptr == int(g[f"s{s}.queue_ptr"][0])
        np.testing.assert_allclose(queue.numpy(), g[f"s{s}.queue"], rtol=1e-6, atol=1e-8)


def test_ema_bit_exact():
    g = load_golden("ema.npz")
    n = int(g["n"])
    q = [T(g[f"q{i}"]) for i in range(n)]
    k = [T(g[f"k{i}"]) for i in range(n)]
    k1 = O.ema_update(q, k, 0.999)
    for i in range(n):
        assert torch.equal(k1[i].view(torch.int32), T(g[f"k1_{i}"]).view(torch.int32)), i
    k2 = O.ema_update(q, k1, 0.99)
    for i in range(n):
        assert torch.equal(k2[i].view(torch.int32), T(g[f"k2_{i}"]).view(torch.int32)), i


def test_enqueue_bit_exact_with_wraparound():
    g = load_golden("enqueue.npz")
    queue = T(g["queue0"]).clone()
    ptr = 0
    for i, n in enumerate(g["batches"]):
        ptr = O.dequeue_and_enqueue(queue, ptr, T(g[f"keys{i}"]))
        assert ptr == int(g[f"ptr{i + 1}"][0])
        assert torch.equal(queue, T(g[f"queue{i + 1}"]))


def test_masked_iou_reference_constants():
    g = load_golden("corr.npz")
    for p, iou, ioum in (("u", 12 / 38, 1 / 3), ("s", 4 / 7, 2 / 3)):
        r = O.masked_correlation(T(g[f"{p}_map_a"]), T(g[f"{p}_map_b"]),
                                 T(g[f"{p}_mask_a"]), T(g[f"{p}_mask_b"]))
        assert torch.all(r["iou"] == torch.tensor(iou, dtype=torch.float32))
        assert torch.all(r["iou_masked"] == torch.tensor(ioum, dtype=torch.float32))
        np.testing.assert_array_equal(r["iou"].numpy(), g[f"{p}_iou"].reshape(-1))
        np.testing.assert_array_equal(r["corr_map"].numpy(), g[f"{p}_corr"])


def test_dense_stats_and_quantile_semantics():
    g = load_golden("corr.npz")
    st = O.dense_loss_stats(T(g["d_logits"]), T(g["d_labels"]))
    for side in ("positive", "negative"):
        np.testing.assert_allclose(st[side]["average"].numpy(), g[f"d_{side}_avg"], rtol=1e-6)
        np.testing.assert_allclose(torch.stack(list(st[side]["quartiles"])).numpy(),
                                   g[f"d_{side}_q"], rtol=1e-6)
    # reference tests/test_contrastive_metrics.py:16-57
    scores = torch.Tensor([[[1, 2, 3], [4, 5, 6]], [[1, 2, 3], [7, 8, 9]]])
    qt = torch.quantile(scores.flatten(1), q=torch.Tensor([0.25, 0.5, 0.75]), dim=1)
    assert torch.all(qt == torch.Tensor([[2.25, 2.25], [3.5, 5.0], [4.75, 7.75]]))


def test_shuffle_roundtrip():
    x = torch.arange(24.0).reshape(12, 2)
    perm = torch.randperm(12, generator=torch.Generator().manual_seed(0))
    world = 3
    shuf = [O.batch_shuffle(x, perm, r, world) for r in range(world)]
    gathered = O.concat_all_gather([s[0] for s in shuf])
    back = O.concat_all_gather([O.batch_unshuffle(gathered, shuf[0][1], r, world) for r in range(world)])
    assert torch.equal(back, x)


@pytest.mark.parametrize("tag", ["a", "b", "c", "d", "e"])
def test_mirror_step_matches_reference(tag):
    """oracle.mirror_step vs the unmodified MirrorModule.shared_step + SegmentationModule.forward
    (networks/mirror_network.py:40-63, segment_network.py:219-231) recorded by gen_golden.gen_mirror."""
    g = load_golden("mirror.npz")
    lmbd, temp = (float(v) for v in g[f"{tag}_cfg"])
    s_lo = T(g[f"{tag}_s_lo"]).requires_grad_(True)
    none = str(g[f"{tag}_variant"]) == "NONE"
    t_lo = None if none else T(g[f"{tag}_t_lo"]).requires_grad_(True)
    loss, cls, cmp = O.mirror_step(s_lo, t_lo, T(g[f"{tag}_masks"]), lmbd, temp)
    loss.backward()
    np.testing.assert_allclose(loss.item(), g[f"{tag}_loss"], rtol=1e-6)
    np.testing.assert_allclose(cls.item(), g[f"{tag}_class"], rtol=1e-6)
    np.testing.assert_allclose(s_lo.grad.numpy(), g[f"{tag}_ds_lo"], rtol=1e-5, atol=1e-9)
    if not none:
        np.testing.assert_allclose(cmp.item(), g[f"{tag}_compare"], rtol=1e-6)
        np.testing.assert_allclose(t_lo.grad.numpy(), g[f"{tag}_dt_lo"], rtol=1e-5, atol=1e-9)


DENSECL_TAGS = ["plain", "v2_pred"]


def densecl_case(tag):
    """Inputs of a `forward_densecl` fixture (oracle/gen_golden.py:gen_densecl), normalised as
    get_query_features / get_key_features do (builder.py:700-760)."""
    g = load_golden(f"densecl_{tag}.npz")
    nrm = torch.nn.functional.normalize
    ql = T(g["q_local_raw"]).requires_grad_(True)
    qg = T(g["q_global_raw"]).requires_grad_(True)
    feats = {
        "q_local_raw": ql, "q_global_raw": qg,
        "q_local": nrm(ql.flatten(2), dim=1), "q_global": nrm(qg, dim=1),
        "k_local": nrm(T(g["k_local_raw"]).flatten(2), dim=1), "k_global": nrm(T(g["k_global_raw"]), dim=1),
        "k_pooled": nrm(T(g["k_pooled_raw"]), dim=1),
        "embd_q": nrm(T(g["embd_q"]).flatten(2), dim=1), "embd_k": nrm(T(g["embd_k"]).flatten(2), dim=1),
    }
    return g, feats


@pytest.mark.parametrize("tag", DENSECL_TAGS)
def test_densecl_loss_path(tag):
    """f2: forward_densecl (builder.py:762-999) between the feature extractors and the loss."""
    g, f = densecl_case(tag)
    r = O.densecl_loss(f["embd_q"], f["embd_k"], f["q_local"], f["k_local"], f["q_global"], f["k_global"],
                       T(g["pix_q"]), T(g["pix_k"]), T(g["queue0"]), T(g["queue2_0"]), 0.2, 0.2, 0.5,
                       float(g["lmbd_coordinate"]))
    r["loss"].backward()
    np.testing.assert_allclose(r["loss"].item(), g["loss"], rtol=1e-6)
    np.testing.assert_allclose(r["loss_global"].item(), g["loss_global"], rtol=1e-6)
    np.testing.assert_allclose(r["loss_local"].item(), g["loss_local"], rtol=1e-6)
    np.testing.assert_allclose(f["q_local_raw"].grad.numpy(), g["grad_q_local"], rtol=1e-4, atol=1e-9)
    np.testing.assert_allclose(f["q_global_raw"].grad.numpy(), g["grad_q_global"], rtol=1e-4, atol=1e-9)
    queue, queue2 = T(g["queue0"]).clone(), T(g["queue2_0"]).clone()       # builder.py:975-976
    assert O.dequeue_and_enqueue(queue, 0, f["k_global"]) == int(g["queue_ptr1"][0])
    assert O.dequeue_and_enqueue(queue2, 0, f["k_pooled"]) == int(g["queue2_ptr1"][0])
    np.testing.assert_allclose(queue.numpy(), g["queue1"], rtol=1e-6, atol=1e-8)
    np.testing.assert_allclose(queue2.numpy(), g["queue2_1"], rtol=1e-6, atol=1e-8)


def test_densecl_coordinate_mixing_formula():
    """builder.py:839-855 on a hand-made case (the reference itself cannot run it: it raises at :861 as soon
    as one pixel id is shared): overlap pixels mix the matched score with the sum over same-id key pixels."""
    k_local = torch.nn.functional.normalize(torch.randn(1, 8, 4, generator=torch.Generator().manual_seed(0)), dim=1)
    q_local = torch.nn.functional.normalize(torch.randn(1, 8, 4, generator=torch.Generator().manual_seed(1)), dim=1)
    idx = torch.tensor([[2, 0, 1, 3]])
    pix_q, pix_k = torch.tensor([[5, 6, 7, 8]]), torch.tensor([[7, 9, 7, 10]])
    kp = O.densecl_positive_keys(k_local, idx, pix_q, pix_k, 0.25)
    sim = torch.einsum("ncx,ncy->nxy", q_local, k_local)[0]
    pos = torch.einsum("nxc,ncx->nx", kp, q_local)[0]
    want = torch.stack([sim[0, 2], sim[1, 0], 0.75 * sim[2, 1] + 0.25 * (sim[2, 0] + sim[2, 2]), sim[3, 3]])
    np.testing.assert_allclose(pos.numpy(), want.numpy(), rtol=1e-5, atol=1e-6)
