"""CPU, gloo, world_size 2: the host-side collectives of the drop-in MODEL (builder.py:609-649, 1710-1722)
against the oracle's single-process restatement.  The kernels themselves need a GPU (tests/test_*_gpu.py)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORLD = 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _Tiny(nn.Module):
    def __init__(self):
        super().__init__()
        self.proj = nn.Conv2d(3, 128, 16, 16)

    def forward(self, x):
        return self.proj(x)


def _worker(rank, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    from cp2_b200.builder import MODEL, concat_all_gather
    from oracle import cp2_oracle as O

    torch.manual_seed(100 + rank)                       # different CPU RNG state per rank (Appendix A.7)
    model = MODEL(None, rank=rank, K=64, encoder_factory=_Tiny, output_stride=16)
    res = {}
    # a10: rank-ordered concatenation
    x = torch.full((3, 5), float(rank)) + torch.arange(5.0)
    g = concat_all_gather(x)
    res["gather_ok"] = torch.equal(g, O.concat_all_gather([torch.full((3, 5), float(r)) + torch.arange(5.0)
                                                           for r in range(WORLD)]))
    # a11: shuffle -> every rank uses rank 0's permutation; unshuffle restores this rank's rows
    imgs = torch.arange(4 * 2.0).reshape(4, 2) + 100 * rank
    all_imgs = torch.cat([torch.arange(4 * 2.0).reshape(4, 2) + 100 * r for r in range(WORLD)])
    shuf, idx_unshuffle = model._batch_shuffle_ddp(imgs)
    idx_shuffle = torch.argsort(idx_unshuffle)
    ref_shuf, ref_unshuf = O.batch_shuffle(all_imgs, idx_shuffle, rank, WORLD)
    res["shuffle_ok"] = torch.equal(shuf, ref_shuf) and torch.equal(idx_unshuffle, ref_unshuf)
    perm_list = [torch.empty_like(idx_shuffle) for _ in range(WORLD)]
    dist.all_gather(perm_list, idx_shuffle)
    res["same_perm"] = all(torch.equal(p, perm_list[0]) for p in perm_list)
    back = model._batch_unshuffle_ddp(shuf * 2.0, idx_unshuffle)      # "encoder" = x2
    res["unshuffle_ok"] = torch.equal(back, imgs * 2.0)
    res["unshuffle_oracle"] = torch.equal(
        back, O.batch_unshuffle(torch.cat([O.batch_shuffle(all_imgs, idx_shuffle, r, WORLD)[0] * 2.0
                                           for r in range(WORLD)]), idx_unshuffle, rank, WORLD))
    # a non-cached idx tensor (e.g. restored by a caller) takes the device->host path and agrees
    back2 = model._batch_unshuffle_ddp(shuf * 2.0, idx_unshuffle.clone())
    res["unshuffle_uncached"] = torch.equal(back2, imgs * 2.0)
    # several more draws with 4-D payloads (ragged per-peer row counts in the all-to-all)
    ok = True
    for t in range(6):
        big = torch.randn(5, 3, 4, 4) + rank
        all_big = [torch.empty_like(big) for _ in range(WORLD)]
        dist.all_gather(all_big, big)
        s5, iu = model._batch_shuffle_ddp(big)
        ref5, _ = O.batch_shuffle(torch.cat(all_big), torch.argsort(iu), rank, WORLD)
        ok = ok and torch.equal(s5, ref5) and torch.equal(model._batch_unshuffle_ddp(s5, iu), big)
    res["shuffle_4d_ok"] = ok
    # a9: the gathered keys every rank would enqueue are identical and rank-ordered -> identical queues
    keys = torch.randn(4, 128)
    keys_all = concat_all_gather(keys)
    queue = torch.zeros(128, 64)
    ptr = O.dequeue_and_enqueue(queue, 60, keys_all)                  # wraps: 60 + 8 > 64
    res["ptr"] = ptr
    torch.save({"res": res, "queue": queue, "keys": keys}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_world2_gloo_host_logic(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(port, str(tmp_path)), nprocs=WORLD, join=True)
    outs = [torch.load(tmp_path / f"r{r}.pt") for r in range(WORLD)]
    for o in outs:
        for k in ("gather_ok", "shuffle_ok", "same_perm", "unshuffle_ok", "unshuffle_oracle", "unshuffle_uncached",
                  "shuffle_4d_ok"):
            assert o["res"][k], k
        assert o["res"]["ptr"] == 4
    assert torch.equal(outs[0]["queue"], outs[1]["queue"])
    # equal to a world-1 run fed the concatenated keys
    from oracle import cp2_oracle as O
    q1 = torch.zeros(128, 64)
    O.dequeue_and_enqueue(q1, 60, torch.cat([outs[0]["keys"], outs[1]["keys"]]))
    assert torch.equal(q1, outs[0]["queue"])


def test_world1_collectives_are_identity():
    sys.path.insert(0, ROOT)
    from cp2_b200.builder import MODEL, concat_all_gather
    x = torch.randn(3, 4)
    assert concat_all_gather(x) is x
    model = MODEL(None, rank=0, K=32, encoder_factory=_Tiny, output_stride=16)
    torch.manual_seed(5)
    y, idx = model._batch_shuffle_ddp(x)
    assert torch.equal(model._batch_unshuffle_ddp(y, idx), x)
    assert model.queue.shape == (128, 32) and model.queue_ptr.dtype == torch.int64
    assert torch.allclose(model.queue.norm(dim=0), torch.ones(32), atol=1e-6)
    assert all(not p.requires_grad for p in model.encoder_k.parameters())
    sd = model.state_dict()
    for k in ("queue", "queue_ptr", "queue2", "queue2_ptr"):            # checkpoint contract (main.py:535-542)
        assert k in sd
