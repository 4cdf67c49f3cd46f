"""GPU, NCCL, world_size 2 (skipped on a 1-GPU box): the keys all-gather feeding the enqueue kernel
(builder.py:569-587, 1710-1722), the BN batch shuffle (builder.py:609-649) and two DDP training steps.
After n enqueues the queue and pointer are bit-identical on every rank and equal to a world-1 oracle run fed
the rank-ordered concatenation of the keys."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu
WORLD = 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _Tiny(nn.Module):
    def __init__(self):
        super().__init__()
        self.proj = nn.Conv2d(3, 128, 16, 16)
        self.bn = nn.BatchNorm2d(128)

    def forward(self, x):
        return self.bn(self.proj(x))


def _worker(rank, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=WORLD, device_id=dev)
    from cp2_b200.builder import MODEL
    from oracle import cp2_oracle as O

    torch.manual_seed(0)                                   # same weights / queue on every rank
    K = 72
    model = MODEL(None, rank=rank, K=K, dense_logits_temp=0.2, device=dev, encoder_factory=_Tiny,
                  output_stride=16).to(dev)
    res = {}
    # ---- a9/a10: 5 enqueues of 16 keys per rank (32 per call) into K=72 -> wraps twice
    queue_ref = model.queue.detach().cpu().clone()
    ptr_ref = 0
    g = torch.Generator().manual_seed(7)
    for _ in range(5):
        keys_all = torch.randn(WORLD * 16, 128, generator=g)          # every rank draws the same tensor
        model._dequeue_and_enqueue(keys_all[rank * 16:(rank + 1) * 16].to(dev))
        ptr_ref = O.dequeue_and_enqueue(queue_ref, ptr_ref, O.concat_all_gather(
            [keys_all[r * 16:(r + 1) * 16] for r in range(WORLD)]))
    torch.cuda.synchronize()
    res["queue_ok"] = torch.equal(model.queue.cpu(), queue_ref)
    res["ptr_ok"] = int(model.queue_ptr.item()) == ptr_ref
    # ---- a11: shuffle / unshuffle round trip, rank 0's permutation everywhere
    torch.manual_seed(100 + rank)
    x = (torch.arange(8.0).reshape(4, 2) + 100 * rank).to(dev)
    y, idx_unshuffle = model._batch_shuffle_ddp(x)
    perms = [torch.empty_like(idx_unshuffle) for _ in range(WORLD)]
    dist.all_gather(perms, idx_unshuffle)
    res["same_perm"] = all(torch.equal(p, perms[0]) for p in perms)
    res["roundtrip"] = torch.equal(model._batch_unshuffle_ddp(y * 3.0, idx_unshuffle), x * 3.0)
    # the all-to-all row exchange lands exactly the rows the reference's all-gather + index would
    ok = True
    for t in range(4):
        big = torch.randn(6, 3, 8, 8, device=dev) + rank
        all_big = [torch.empty_like(big) for _ in range(WORLD)]
        dist.all_gather(all_big, big)
        yb, iu = model._batch_shuffle_ddp(big)
        ref, _ = O.batch_shuffle(torch.cat(all_big).cpu(), torch.argsort(iu).cpu(), rank, WORLD)
        ok = ok and torch.equal(yb.cpu(), ref) and torch.equal(model._batch_unshuffle_ddp(yb, iu), big)
    res["shuffle_vs_oracle"] = ok
    # ---- two DDP steps: finite loss, gradients all-reduced, queues stay identical across ranks
    ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[rank], broadcast_buffers=False)
    opt = torch.optim.SGD(model.encoder_q.parameters(), 0.01)
    losses = []
    for step in range(2):
        v = O.synth_views(4, 64, 64, 16, seed=10 * rank + step)
        loss = ddp(visualize=False, step=step, new_epoch=False, **{k: t.to(dev) for k, t in v.items()})
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    res["finite"] = all(torch.isfinite(torch.tensor(losses)))
    q_all = [torch.empty_like(model.queue) for _ in range(WORLD)]
    dist.all_gather(q_all, model.queue.contiguous())
    res["queues_equal"] = all(torch.equal(q, q_all[0]) for q in q_all)
    w = model.encoder_q.proj.weight.detach()
    w_all = [torch.empty_like(w) for _ in range(WORLD)]
    dist.all_gather(w_all, w.contiguous())
    res["weights_equal"] = all(torch.equal(t, w_all[0]) for t in w_all)
    res["ptr_after"] = int(model.queue_ptr.item())
    torch.save(res, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_world2_nccl_enqueue_shuffle_ddp(tmp_path):
    if torch.cuda.device_count() < WORLD:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    mp.spawn(_worker, args=(_free_port(), str(tmp_path)), nprocs=WORLD, join=True)
    for r in range(WORLD):
        res = torch.load(tmp_path / f"r{r}.pt")
        for k in ("queue_ok", "ptr_ok", "same_perm", "roundtrip", "shuffle_vs_oracle", "finite", "queues_equal", "weights_equal"):
            assert res[k], (r, k)
        assert res["ptr_after"] == (5 * 32 + 2 * 8) % 72
