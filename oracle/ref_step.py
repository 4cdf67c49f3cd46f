"""TEST / BASELINE INFRASTRUCTURE ONLY — never imported by the product path (cp2_b200/).

One pre-training iteration of the UNMODIFIED reference: `builder.MODEL(...)` driven exactly as
`train()` drives it (main.py:590-644: keyword forward with the four id maps, `optimizer.zero_grad()`,
`loss.backward()`, `optimizer.step()`, `loss.item()`), with SGD as main.py:467-474 builds it.  The
reference module comes from `oracle/ref_harness.py` (source tree here, `oracle/_ref/*.pyc` on the GPU
box).  The segmentor (`build_segmentor(cfg.model)`, mmseg — not installable) is the caller's
`encoder_factory`, the same stand-in both bench arms use.

Users: `bench.py --impl reference` (device cpu: the reference's own CPU path) and bench.py's
`gpu_eager_baseline` leg (device cuda: the reference's eager PyTorch path on the same B200).
"""
import torch

from . import ref_harness as rh


class ReferenceStep:
    def __init__(self, encoder_factory, device="cpu", dim=128, K=65536, temp_global=0.2, temp_local=1.0,
                 lmbd_dense=0.2, lr=1e-3, seed=0, channels_last=False):
        dev = torch.device(device)
        torch.manual_seed(seed)
        self.builder, model = rh.make_reference_model(
            encoder_factory, cpu=(dev.type == "cpu"), dim=dim, K=K, m=0.999, instance_logits_temp=temp_global,
            lmbd_cp2_dense_loss=lmbd_dense, dense_logits_temp=temp_local, device=dev)
        model = model.to(dev)
        if channels_last:
            model = model.to(memory_format=torch.channels_last)
        self.model, self.device = model, dev
        self.kind = "_ref" if rh.REFERENCE_KIND in ("source", "pyc") else "port"
        self.opt = torch.optim.SGD(model.parameters(), lr, momentum=0.9, weight_decay=1e-4)   # main.py:467-474
        self.n = 0

    def step(self, batch):
        """batch: dict with img_a, img_b, bg0, bg1, pixel_ids_a/b, region_ids_a/b on self.device."""
        loss = self.model(visualize=False, step=self.n, new_epoch=False, **batch)        # main.py:616-628
        self.opt.zero_grad()
        loss.backward()
        self.opt.step()
        self.n += 1
        return loss.item()                                                                 # main.py:646
