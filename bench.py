#!/usr/bin/env python
"""bench.py — CP2 pre-training throughput on B200 (BASELINE.json: "CP2 pretrain imgs/s at 1/2/4/8
B200; dense-loss fwd+bwd us vs TC roofline").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Own arm (default): one process per GPU, configs[1] of BASELINE.json per rank (ResNet-50 DeepLabV3 OS16,
224x224 views, B=32 per GPU, C=128, queue 65536, tau 0.2), weak scaling, synthetic data, random-init
weights.  One step = composite -> encoder_q -> EMA -> encoder_k -> dense + global InfoNCE -> enqueue ->
backward -> SGD step, through cp2_b200.builder.MODEL (the reference's module API).  The convolutional
backbone is PyTorch/cuDNN (out of scope); everything else on the step is a kernel of
libcp2_b200.so.  `value` keeps the inputs resident in HBM; `e2e` copies every step's inputs from pinned
host memory and reads the loss back.  Per-kernel CUDA-event timings of the timed region give `kernels`
and `roofline`; at N=1 the dense-loss microbench sweep (configs[3]) and the CPU baseline are added.

Reference arm (--impl reference): the reference is pure Python/PyTorch and cannot travel to the GPU box;
its CPU path is restated in oracle/cp2_oracle.py (CP2StepPort, kind "port") and timed on the host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "cp2_pretrain_images_per_sec"
UNIT = "images/s"
IMG, OS_, DIM, QUEUE, TAU_G, TAU_D, LMBD = 224, 16, 128, 65536, 0.2, 0.2, 0.2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=32, help="images per GPU (configs[1]: 32)")
    ap.add_argument("--nchw", action="store_true",
                    help="keep the (out-of-scope) cuDNN backbone in NCHW instead of channels_last (NHWC)")
    ap.add_argument("--no-sweep", action="store_true", help="skip the dense-loss microbench sweep")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sweep-only", action="store_true", help="run only the dense-loss sweep (profiling)")
    ap.add_argument("--sweep-points", default="196,784,3136", help="P values of the sweep")
    ap.add_argument("--sweep-dims", default="128,256")
    ap.add_argument("--sweep-batch", type=int, default=256)
    ap.add_argument("--sweep-iters", type=int, default=10)
    return ap.parse_args()


def workload_config(batch, n_gpus, nchw=False):
    return {
        "backbone_memory_format": "NCHW" if nchw else "channels_last (NHWC; fp32 values unchanged)",
        "workload": "BASELINE.json configs[1]: CP2 pretrain, ResNet-50 DeepLabV3 OS16 (torchvision stand-in "
                    "for the mmseg segmentor), 224x224 views (14x14 map), C=128, queue 65536",
        "batch_per_gpu": batch, "global_batch": batch * n_gpus, "image": IMG, "feature_map": IMG // OS_,
        "dim": DIM, "queue": QUEUE, "temp_global": TAU_G, "temp_dense": TAU_D, "lmbd_dense": LMBD,
        "optimizer": "SGD(lr=1e-3, momentum=0.9, wd=1e-4)", "precision": "fp32 encoder (as the reference); "
        "bf16 operands / fp32 accumulate in the two contraction kernels",
        "parallelism": f"dp{n_gpus}",
        "l2": "no flush: one step streams >10 GB of activations and 0.8 GB of EMA traffic through the 126 MB L2",
    }


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sustained": d["bf16_tflops_sustained"],
                "source": "MEASURED_PEAKS.json"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc = index, None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


def synth_batch(batch, seed, pinned=False):
    """Synthetic step inputs: images ~U[0,1); backgrounds ~U[0,1) with one rectangle per
    image set to exactly 0 on all channels, area 50-80 %, sides >= output stride so the rectangle covers
    a point of the [OS/2::OS] lattice.  CPU tensors (pinned if asked)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    out = {k: torch.rand(batch, 3, IMG, IMG, generator=g) for k in ("img_a", "img_b", "bg0", "bg1")}
    for name in ("bg0", "bg1"):
        frac = 0.5 + 0.3 * torch.rand(batch, generator=g)
        ar = torch.exp(torch.empty(batch).uniform_(-0.2231, 0.2231, generator=g))
        for i in range(batch):
            rh = int(min(IMG, max(OS_, round(float((frac[i] * IMG * IMG * ar[i]) ** 0.5)))))
            rw = int(min(IMG, max(OS_, round(float((frac[i] * IMG * IMG / ar[i]) ** 0.5)))))
            top = int(torch.randint(0, IMG - rh + 1, (1,), generator=g))
            left = int(torch.randint(0, IMG - rw + 1, (1,), generator=g))
            out[name][i, :, top:top + rh, left:left + rw] = 0.0
    # pixel / region id maps (main.py:616-628 passes them on every step; CP2 mapping: region ids = pixel ids,
    # loader.py:84-85): two random crops of one id canvas per image, int64 [B,H,W]
    pad = 32
    canvas = torch.arange(1, (IMG + pad) * (IMG + pad) + 1, dtype=torch.int64).reshape(IMG + pad, IMG + pad)
    off = torch.randint(0, pad + 1, (batch, 4), generator=g)
    for name, (cy, cx) in (("pixel_ids_a", (0, 1)), ("pixel_ids_b", (2, 3))):
        out[name] = torch.stack([canvas[int(o[cy]):int(o[cy]) + IMG, int(o[cx]):int(o[cx]) + IMG] for o in off])
    out["region_ids_a"], out["region_ids_b"] = out["pixel_ids_a"].clone(), out["pixel_ids_b"].clone()
    if pinned:
        out = {k: v.pin_memory() for k, v in out.items()}
    return out


# ----------------------------------------------------------------- CPU legs (oracle port)
def _cpu_port(batch, seed=0):
    """The reference's CPU path restated (oracle/cp2_oracle.py:CP2StepPort) on the same workload."""
    import torch
    from cp2_b200.encoder import DeepLabV3Encoder
    from oracle import cp2_oracle as O
    port = O.CP2StepPort(lambda: DeepLabV3Encoder(dim=DIM), dim=DIM, K=QUEUE, temp_global=TAU_G,
                         temp_local=TAU_D, lmbd_dense=LMBD, output_stride=OS_, lr=1e-3, seed=seed)
    b = synth_batch(batch, seed=1234)
    return port, b, torch.get_num_threads()


def cpu_baseline_leg(budget_s=20.0):
    """Bounded sample for the own arm's `cpu_baseline` object: batch 8, 1 warm-up + up to 5 timed steps."""
    batch = 8
    port, b, threads = _cpu_port(batch)
    t0 = time.perf_counter()
    port.step(**b)
    warm = time.perf_counter() - t0
    n = max(1, min(5, int(budget_s / max(warm, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(n):
        port.step(**b)
    dt = (time.perf_counter() - t0) / n
    return {"value": batch / dt, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{n} timed steps (after 1 warm-up) of the full pre-training step at batch {batch} "
                      f"(same model, queue {QUEUE}, 224x224), {dt * 1e3:.0f} ms/step; host has "
                      f"{os.cpu_count()} logical CPUs, torch used {threads} threads"}


def reference_arm(args):
    """--impl reference: the reference's own CPU implementation of the path (port), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    total = args.steps + args.warmup
    # calibrate at batch 2 (BatchNorm needs > 1 sample), then size each step's sample so the run ends in minutes
    port, b2, threads = _cpu_port(2)
    t0 = time.perf_counter()
    port.step(**b2)
    port.step(**b2)
    per_img = (time.perf_counter() - t0) / 4.0
    budget = 150.0
    batch = int(budget / max(total, 1) / max(per_img, 1e-4))
    batch = max(2, min(args.batch, batch))
    b = synth_batch(batch, seed=1234)
    for _ in range(args.warmup):
        port.step(**b)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        port.step(**b)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    value = batch / dt
    sample = (f"each step = the full pre-training step on {batch} of the {args.batch} images of one GPU's batch "
              f"(sized after a batch-2 calibration so that {total} steps fit in ~{budget:.0f} s); "
              f"{os.cpu_count()} logical CPUs, torch threads {threads}")
    cfg = workload_config(args.batch, args.gpus)
    cfg["sample_batch"] = batch
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------ dense-loss microbench
def dense_sweep(args, peaks):
    """configs[3]: dense pixel InfoNCE fwd+bwd at B=256 images/GPU, P in {14^2,28^2,56^2}, C in {128,256}.
    Algorithmic work 4*B*P^2*C FLOP (Q^T K forward + dQ backward; the backward's recomputation of Q^T K is
    not counted, and background key columns skipped by the compaction are still counted: `frac` is the
    dense-algorithmic figure, `executed_frac` = share of key columns actually visited).  Two mask variants
    (SURVEY.md 8d): Bernoulli(0.65) fg masks (CP2's copy-paste masks are 50-80 % foreground) and all-ones
    (nothing to skip).  `us` = cp2_dense_lse + cp2_dense_grad; `us_with_prep` adds the two
    cp2_prep_features launches.  L2 is flushed between iterations."""
    import torch
    from cp2_b200 import ops
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    B = args.sweep_batch
    rows = []
    for C in [int(x) for x in args.sweep_dims.split(",")]:
        for P in [int(x) for x in args.sweep_points.split(",")]:
            g = torch.Generator(device=dev).manual_seed(7)
            q = torch.randn(B, C, P, device=dev, generator=g)
            k = torch.randn(B, C, P, device=dev, generator=g)
            alpha = torch.full((B,), 1.0 / B, device=dev)
            v1 = torch.zeros(B, C, device=dev)
            inv_tau = 1.0 / TAU_D
            for variant in ("bernoulli(0.65)", "ones"):
                g8 = torch.Generator(device=dev).manual_seed(8)
                if variant == "ones":
                    ma = torch.ones(B, P, device=dev)
                    mb = torch.ones(B, P, device=dev)
                else:
                    ma = (torch.rand(B, P, device=dev, generator=g8) < 0.65).float()
                    mb = (torch.rand(B, P, device=dev, generator=g8) < 0.65).float()
                t_core, t_all = [], []
                for it in range(3 + args.sweep_iters):
                    flush.zero_()
                    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                    e[0].record()
                    qh, q_inv, _, _, _ = ops.prep_features(q, ma)
                    kh, _, _, _, cnt = ops.prep_features(k, mb, compact=True)
                    e[1].record()
                    lse = ops.dense_lse(qh, kh, inv_tau, cnt)
                    dq = ops.dense_grad(qh, kh, lse, ma, None, q_inv, alpha, v1, None, inv_tau, key_count=cnt)
                    e[2].record()
                    torch.cuda.synchronize()
                    if it >= 3:
                        t_core.append(e[1].elapsed_time(e[2]) * 1e3)
                        t_all.append(e[0].elapsed_time(e[2]) * 1e3)
                us, us_all = statistics.mean(t_core), statistics.mean(t_all)
                flop = 4.0 * B * P * P * C
                tf = flop / (us * 1e-6) / 1e12
                rows.append({"P": P, "C": C, "B": B, "masks": variant, "us": round(us, 2),
                             "us_with_prep": round(us_all, 2), "flop": flop, "tflops": round(tf, 2),
                             "peak": peaks["tf_burst"], "frac": round(tf / peaks["tf_burst"], 4),
                             "executed_frac": round(float(cnt.float().mean()) / P, 4),
                             "finite": bool(torch.isfinite(dq).all())})
                del qh, kh, dq, lse
            del q, k
    return rows


# ------------------------------------------------------------------ mirror-loss microbench
def mirror_sweep(peaks, iters=20):
    """configs[4] shapes (512x512 labels, Cls=2, decoder maps 32x32 = OS16 and 64x64 = OS8): the fused
    resize + loss + backward kernel (cp2_mirror_loss_upsampled) and the full-resolution kernel
    (cp2_mirror_loss) against the eager torch formulation of networks/mirror_network.py:47-63 +
    segment_network.py:219-231 on the same GPU.  CUDA events, a 1 GiB memset between iterations flushes L2
    and hides the host launch path."""
    import torch
    import torch.nn.functional as F
    from cp2_b200 import ops
    dev = torch.device("cuda")
    flush = torch.empty(1 << 30, dtype=torch.uint8, device=dev)

    def timeit(fn, n):
        ts = []
        for i in range(n + 3):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(a.elapsed_time(b) * 1e3)
        return statistics.median(ts)

    rows = []
    for B, lo in ((8, 32), (8, 64), (32, 32), (32, 64)):
        Cls, H = 2, 512
        g = torch.Generator(device=dev).manual_seed(11)
        s_lo = torch.randn(B, Cls, lo, lo, device=dev, generator=g).requires_grad_()
        t_lo = torch.randn(B, Cls, lo, lo, device=dev, generator=g).requires_grad_()
        masks = torch.randint(0, Cls, (B, H, H), device=dev, generator=g)
        s_hi = F.interpolate(s_lo.detach(), size=(H, H), mode="bilinear", align_corners=False)
        t_hi = F.interpolate(t_lo.detach(), size=(H, H), mode="bilinear", align_corners=False)

        def eager():
            s_lo.grad = t_lo.grad = None
            s = F.interpolate(s_lo, size=(H, H), mode="bilinear", align_corners=False)
            t = F.interpolate(t_lo, size=(H, H), mode="bilinear", align_corners=False)
            cmp_ = F.cross_entropy(torch.softmax(s / 2.0, 1), torch.softmax(t / 2.0, 1))
            cls_ = F.cross_entropy(torch.cat([s, t]), torch.cat([masks, masks]))
            (cls_ + 0.01 * cmp_).backward()

        up = timeit(lambda: ops.mirror_loss_upsampled_raw(s_lo.detach(), t_lo.detach(), masks, 0.01, 2.0, True), iters)
        full = timeit(lambda: ops.mirror_loss_raw(s_hi, t_hi, masks, 0.01, 2.0, True), iters)
        eag = timeit(eager, 5)
        px = B * H * H
        full_bytes = px * (8 + 16 * Cls)
        rows.append({"B": B, "Cls": Cls, "logits": lo, "labels": H, "fused_resize_loss_bwd_us": round(up, 2),
                     "label_bytes": px * 8, "full_res_loss_bwd_us": round(full, 2),
                     "full_res_GBps": round(full_bytes / full / 1e3, 1),
                     "full_res_frac_hbm": round(full_bytes / full / 1e3 / peaks["hbm_gbs"], 4),
                     "eager_torch_us": round(eag, 1), "speedup_vs_eager": round(eag / up, 1)})
    return rows


# ------------------------------------------------------------------ PROPOSED-mode dense loss microbench
def weighted_sweep(peaks, iters=20):
    """Row f3 at configs[1]'s shape (B=32, 14x14, C=128) and at 28x28: cp2_dense_weighted_lse +
    cp2_dense_weighted_grad (pair weights 10/2/0.5, NegativeType.FIXED) against the eager torch formulation
    of builder.py:1204-1243, 1289-1292, 1332-1392, 1431-1437 + autograd on the same GPU.  L2 flushed."""
    import torch
    import torch.nn.functional as F
    from cp2_b200 import ops
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timeit(fn, n):
        ts = []
        for i in range(n + 3):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(a.elapsed_time(b) * 1e3)
        return statistics.median(ts)

    rows = []
    for B, hw in ((32, 14), (32, 28)):
        C, P, os_ = 128, hw * hw, 16
        H = hw * os_
        g = torch.Generator(device=dev).manual_seed(13)
        q = torch.randn(B, C, P, device=dev, generator=g)
        k = torch.randn(B, C, P, device=dev, generator=g)
        ma = (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
        mb = (torch.rand(B, P, device=dev, generator=g) < 0.65).float()
        pa = torch.arange(1, H * H + 1, device=dev).reshape(1, H, H).repeat(B, 1, 1)
        pb = torch.roll(pa, shifts=(os_, os_), dims=(1, 2)).contiguous()
        ra = (pa - 1) // (H * 64) * 4 + ((pa - 1) % H) // 64
        rb = torch.roll(ra, shifts=(os_, os_), dims=(1, 2)).contiguous()
        wts, scale, inv_tau = (10.0, 2.0, 0.5), 2.0, 1.0 / TAU_D
        pw = ops.PairWeights(pa, pb, ra, rb, os_, *wts, "FIXED", scale)
        alpha = torch.full((B,), 1.0 / B, device=dev)
        v1 = torch.zeros(B, C, device=dev)
        qh, q_inv, sq, nsq, _ = ops.prep_features(q, ma)
        kh, _, sk, nsk, _ = ops.prep_features(k, mb)

        def ours():
            lse, dense_n, wstats, _, _ = ops.dense_weighted_lse(qh, kh, ma, mb, pw, sq, nsq, sk, nsk, inv_tau)
            return ops.dense_weighted_grad(qh, kh, ma, mb, pw, lse, wstats, q_inv, alpha, v1, None, inv_tau)

        sub = lambda t: t[:, os_ // 2::os_, os_ // 2::os_].reshape(B, -1)
        qe = q.clone().requires_grad_()

        def eager():
            qe.grad = None
            pix = sub(pa)[:, :, None] == sub(pb)[:, None, :]
            reg = (sub(ra)[:, :, None] == sub(rb)[:, None, :]) & ((sub(ra)[:, :, None] * sub(rb)[:, None, :]) != 0)
            w = wts[1] * reg
            w[pix] = wts[0]
            w = w + (w == 0) * wts[2]
            logits = torch.einsum("ncx,ncy->nxy", F.normalize(qe, dim=1), F.normalize(k, dim=1))
            labels = torch.einsum("nx,ny->nxy", ma, mb)
            neg = labels == 0
            logits = torch.where(neg, 2.0 / (1.0 + torch.exp(logits * -scale)) - 1.0, logits) * w * inv_tau
            nl = -torch.log_softmax(logits, dim=1)
            ((nl * labels).flatten(1).sum(1) / labels.flatten(1).sum(1)).mean().backward()

        t_ours, t_eager = timeit(ours, iters), timeit(eager, 5)
        rows.append({"B": B, "P": P, "C": C, "negative_type": "FIXED", "weights": list(wts),
                     "weighted_lse_grad_us": round(t_ours, 2), "launches": 4, "eager_torch_us": round(t_eager, 1),
                     "speedup_vs_eager": round(t_eager / t_ours, 1),
                     "fp32_simt_tflops": round(6.0 * B * P * P * C / t_ours / 1e6, 2)})
    return rows


# ------------------------------------------------------------------ forward_densecl loss microbench
def densecl_sweep(peaks, iters=10):
    """Row f2 at the DenseCL scripts' shape (B=32, 14x14 backbone map, 2048-d backbone features, dim 128,
    K=65536, tau 0.2): ops.densecl_loss forward + backward (the backbone match is a cuBLAS bmm in both arms)
    against the eager torch formulation of builder.py:762-931.  L2 flushed between iterations."""
    import torch
    import torch.nn.functional as F
    from cp2_b200 import ops
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    B, P, C, Cb, K, tau = 32, 196, DIM, 2048, QUEUE, 0.2
    g = torch.Generator(device=dev).manual_seed(17)
    nrm = F.normalize
    embd_q = nrm(torch.randn(B, Cb, P, device=dev, generator=g), dim=1)
    embd_k = nrm(torch.randn(B, Cb, P, device=dev, generator=g), dim=1)
    q_local = nrm(torch.randn(B, C, P, device=dev, generator=g), dim=1).requires_grad_()
    k_local = nrm(torch.randn(B, C, P, device=dev, generator=g), dim=1)
    q_global = nrm(torch.randn(B, C, device=dev, generator=g), dim=1).requires_grad_()
    k_global = nrm(torch.randn(B, C, device=dev, generator=g), dim=1)
    queue = nrm(torch.randn(C, K, device=dev, generator=g), dim=0)
    queue2 = nrm(torch.randn(C, K, device=dev, generator=g), dim=0)

    def timeit(fn, n):
        ts = []
        for i in range(n + 2):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            if i >= 2:
                ts.append(a.elapsed_time(b) * 1e3)
        return statistics.median(ts)

    def ours():
        q_local.grad = q_global.grad = None
        ops.densecl_loss(embd_q, embd_k, q_local, k_local, q_global, k_global, queue, queue2, tau, tau, 0.5)[0].backward()

    def head(pos, neg):
        logits = torch.cat((pos, neg), dim=1) / tau
        return F.cross_entropy(logits, torch.zeros(len(pos), dtype=torch.long, device=dev))

    def eager():
        q_local.grad = q_global.grad = None
        lg = head((q_global * k_global).sum(1, keepdim=True), q_global @ queue.clone())
        idx = torch.einsum("ncx,ncy->nxy", embd_q, embd_k).max(dim=2)[1]
        pos = torch.gather(torch.einsum("ncx,ncy->nxy", q_local, k_local), 2, idx.unsqueeze(2)).reshape(-1, 1)
        rows = q_local.permute(0, 2, 1).reshape(-1, C)
        ll = head(pos, rows @ queue2.clone())
        (0.5 * lg + 0.5 * ll).backward()

    t_ours, t_eager = timeit(ours, iters), timeit(eager, 5)
    return [{"B": B, "P": P, "C": C, "K": K, "rows": B * P, "densecl_loss_fwd_bwd_us": round(t_ours, 1),
             "eager_torch_us": round(t_eager, 1), "speedup_vs_eager": round(t_eager / t_ours, 2),
             "note": "queue sums of the 6272-row local InfoNCE on the dense-loss tcgen05 kernels (11 blocks of "
                     "6272 queue entries), global InfoNCE on cp2_global_nce, backbone match = cuBLAS bmm in both arms"}]


# ------------------------------------------------------------------------------ own arm
def algorithmic_work(batch, n_params, world):
    """Per-CALL algorithmic bytes / flops of each C-ABI entry point on the step."""
    P, C, hw = (IMG // OS_) ** 2, DIM, IMG * IMG
    return {
        "cp2_composite": ("hbm", 3 * batch * 3 * hw * 4 + batch * P * 4),
        "cp2_prep_features": ("hbm", 6 * batch * C * P + 4 * batch * P),
        "cp2_dense_lse": ("tensor", 2.0 * batch * P * P * C),
        "cp2_dense_grad": ("tensor", 2.0 * batch * P * P * C),
        "cp2_global_nce": ("hbm", 4 * C * QUEUE + 8 * batch * C),
        "cp2_ema": ("hbm", 12 * n_params),
        "cp2_enqueue": ("hbm", 2 * world * batch * C * 4),
        "cp2_loss_combine": ("hbm", batch * (9 * C + 2 * P) * 4),   # latency-bound: [B,C] / [B,P] operands once
        "cp2_id_iou": ("hbm", 2 * batch * P * (8 + 4)),       # latency-bound: 2P ids + masks per image
        # f4 arg-max meters, fp32 CUDA cores: counted by the bytes they must read (features / queue once)
        "cp2_dense_top1": ("hbm", 2 * batch * C * P * 4 + 2 * batch * P * 4),
        "cp2_global_top1": ("hbm", 4 * C * QUEUE + 8 * batch * C),
    }


def own_arm(args):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the cp2_b200 path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    from cp2_b200 import _lib, ops
    from cp2_b200.builder import MODEL
    from cp2_b200.encoder import DeepLabV3Encoder
    _lib.load()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    peaks = load_peaks()

    if args.sweep_only:
        if rank == 0:
            print(json.dumps({"dense_microbench": dense_sweep(args, peaks)}), flush=True)
        return 0

    torch.manual_seed(0)
    torch.backends.cudnn.benchmark = True
    model = MODEL(None, rank=rank, dim=DIM, K=QUEUE, m=0.999, instance_logits_temp=TAU_G,
                  pretrain_from_scratch=True, lmbd_cp2_dense_loss=LMBD, dense_logits_temp=TAU_D, device=dev,
                  encoder_factory=lambda: DeepLabV3Encoder(dim=DIM), output_stride=OS_).to(dev)
    if not args.nchw:
        # cuDNN's tensor-core convolutions are NHWC kernels: NCHW weights/activations cost an
        # nchwToNhwc / nhwcToNchw transpose around every convolution (11 % of the r01b step)
        model = model.to(memory_format=torch.channels_last)
    n_params = sum(p.numel() for p in model.encoder_k.parameters())
    net = model
    if world > 1:
        # queues are identical on every rank by construction -> no per-forward buffer broadcast
        # conv_seg never receives a gradient in contrast mode (the reason the reference passes
        # find_unused_parameters=True, main.py:456-461): freezing it is behaviour-neutral and spares DDP the
        # per-step graph traversal; gradients are reduced in place in the buckets
        for p in model.encoder_q.decode_head.conv_seg.parameters():
            p.requires_grad_(False)
        net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local], find_unused_parameters=False,
                                                        broadcast_buffers=False, gradient_as_bucket_view=True)
    opt = torch.optim.SGD(model.encoder_q.parameters(), 1e-3, momentum=0.9, weight_decay=1e-4)

    B = args.batch
    host = [synth_batch(B, seed=1234 + rank + 100 * i, pinned=True) for i in range(2)]
    resident = [{k: v.to(dev) for k, v in hb.items()} for hb in host]
    h2d = sum(v.numel() * v.element_size() for v in host[0].values())

    def step(batch):
        loss = net(visualize=False, step=0, new_epoch=False, **batch)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    for i in range(max(args.warmup, 3)):
        step(resident[i % 2])
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = ops.LAUNCHES
    ncu_range = os.environ.get("CP2_BENCH_CUDA_PROFILER") == "1"    # ncu --profile-from-start off
    if ncu_range:
        torch.cuda.profiler.start()
    with ops.KernelTimer() as kt:
        ms_value = timed(lambda i: step(resident[i % 2]), args.steps)
    if ncu_range:
        torch.cuda.profiler.stop()
    launches = ops.LAUNCHES - launches0
    ksum = kt.summary()

    # e2e: every step's inputs come from pinned host memory (77 MB per step) and the loss is read back.
    # As in a pin_memory data loader, the copy of step i+1 is issued on a copy stream before step i's loss
    # is waited for, so it overlaps the compute; each timed step still contains exactly one full H2D copy.
    copy_stream = torch.cuda.Stream(device=dev)
    pending = {}

    def prefetch(i):
        with torch.cuda.stream(copy_stream):
            b = {k: v.to(dev, non_blocking=True) for k, v in host[i % 2].items()}
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        pending["next"] = (b, ev)

    def e2e_step(i):
        batch, ev = pending.pop("next")
        cur = torch.cuda.current_stream()
        cur.wait_event(ev)
        for t in batch.values():
            t.record_stream(cur)
        prefetch(i + 1)
        return step(batch).item()           # device -> host read of the step's loss

    prefetch(0)
    for i in range(2):
        e2e_step(i)
    ms_e2e = timed(e2e_step, args.steps)
    clocks = sampler.stop()

    # ---- per-kernel figures from the CUDA events recorded inside the timed region
    work = algorithmic_work(B, n_params, world)
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")     # dram bytes per launch from `ncu --set full`
    if os.path.isfile(tpath):
        with open(tpath) as f:
            traffic = json.load(f)
    kernels, per_step_us = {}, 0.0
    for name, (calls, us) in ksum.items():
        bound, amount = work[name]
        per_step = us * calls / max(args.steps, 1)
        per_step_us += per_step
        if bound == "hbm":
            ach, peak, unit = amount / (us * 1e-6) / 1e9, peaks["hbm_gbs"], "GB/s"
        else:
            ach, peak, unit = amount / (us * 1e-6) / 1e12, peaks["tf_sustained"], "TFLOP/s"
        kernels[name] = {"calls_per_step": calls / max(args.steps, 1), "us_per_call": round(us, 2),
                         "us_per_step": round(per_step, 2), "bound": bound, "achieved": round(ach, 2),
                         "peak": peak, "unit": unit, "frac": round(ach / peak, 4)}
    top = max(kernels, key=lambda n: kernels[n]["us_per_step"]) if kernels else None
    roofline = None
    if top:
        k = kernels[top]
        roofline = {"kernel": top, "bound": k["bound"], "achieved": k["achieved"], "peak": k["peak"],
                    "unit": k["unit"], "frac": k["frac"], "traffic": traffic.get(top),
                    "algorithmic_per_launch": work[top][1], "us_per_launch": k["us_per_call"],
                    "peak_source": peaks["source"] + (" (sustained)" if k["bound"] == "tensor" else "")}

    line = {
        "metric": METRIC, "value": B * world * args.steps / (ms_value * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_value / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32+bf16",
        "data": "synthetic", "config": workload_config(B, world, args.nchw),
        "e2e": {"value": B * world * args.steps / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "kernels": kernels,
        "cp2_kernels_us_per_step": round(per_step_us, 1),
        "backbone_and_optimizer_ms_per_step": round(ms_value / args.steps - per_step_us * 1e-3, 3),
    }
    if world == 1:
        if not args.no_sweep:
            # free the training state first: the 56x56 / C=256 point needs ~3.5 GB
            line["dense_microbench"] = dense_sweep(args, peaks)
            line["mirror_microbench"] = mirror_sweep(peaks)
            line["weighted_microbench"] = weighted_sweep(peaks)
            line["densecl_microbench"] = densecl_sweep(peaks)
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_leg()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    return own_arm(args)


if __name__ == "__main__":
    sys.exit(main())
