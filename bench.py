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
import math
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
    ap.add_argument("--no-fold", action="store_true",
                    help="run the ASPP's dilated 3x3 convolutions as cuDNN runs them instead of tap by tap on the "
                         "windows where a tap sees data (MODEL.fold_dilated_convs)")
    ap.add_argument("--fold-cover", type=float, default=0.15)
    ap.add_argument("--no-graphs", action="store_true",
                    help="launch the (out-of-scope) encoders kernel by kernel instead of from CUDA graphs")
    ap.add_argument("--no-sweep", action="store_true", help="skip the dense-loss microbench sweep")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sweep-only", action="store_true", help="run only the dense-loss sweep (profiling)")
    ap.add_argument("--sweep-points", default="196,784,3136", help="P values of the sweep")
    ap.add_argument("--sweep-dims", default="128,256")
    ap.add_argument("--sweep-batch", type=int, default=256)
    ap.add_argument("--sweep-iters", type=int, default=10)
    ap.add_argument("--cpu-budget", type=float, default=200.0, help="--impl reference: seconds of CPU work allowed")
    ap.add_argument("--no-eager-baseline", action="store_true", help="skip the same-GPU eager (reference) legs")
    ap.add_argument("--workload", default="cp2", choices=["cp2", "mirror"],
                    help="cp2: configs[1] (the metric); mirror: configs[4] MirrorModule.shared_step at 512x512")
    return ap.parse_args()


def workload_config(batch, n_gpus, nchw=False):
    return {
        "backbone_memory_format": "NCHW" if nchw else "channels_last (NHWC; fp32 values unchanged)",
        "workload": "BASELINE.json configs[1]: CP2 pretrain, ResNet-50 DeepLabV3 OS16 (torchvision stand-in "
                    "for the mmseg segmentor), 224x224 views (14x14 map), C=128, queue 65536",
        "batch_per_gpu": batch, "global_batch": batch * n_gpus, "image": IMG, "feature_map": IMG // OS_,
        "dim": DIM, "queue": QUEUE, "temp_global": TAU_G, "temp_dense": TAU_D, "lmbd_dense": LMBD,
        "optimizer": "SGD(lr=1e-3, momentum=0.9, wd=1e-4)", "precision": "fp32 parameters / activations; the "
        "(out-of-scope) cuDNN convolutions run as TF32 tensor-op kernels (PyTorch's default, what the reference "
        "gets too); bf16 operands / fp32 accumulate in the two contraction kernels of the loss",
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

    def __init__(self, index, enabled=True):
        # enabled: rank 0 only.  One `nvidia-smi -lms` poller per rank (8 on an 8-GPU box, every query taking the
        # driver's lock) cost 4.7 ms/step at N=8 (29.1 vs 24.4 ms/step with and without them, r02d); rank 0's GPU
        # stands for the box: all ranks run the same step behind the same barrier.
        self.index, self.proc, self.enabled = index, None, enabled

    def start(self):
        if not self.enabled:
            return
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if not self.enabled:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["sampled on rank 0 only"]}
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


def feature_encoder(batch, dim, hw, seed):
    """'Encoder' that returns a learnable stored feature map whatever the image: isolates the loss slice
    (composite, id maps, normalise, both InfoNCE losses, enqueue, backward to the features) of either
    implementation for the same-GPU comparison."""
    import torch
    import torch.nn as nn

    class _BB(nn.Module):
        def forward(self, x):
            y = x.new_zeros(x.shape[0], 1, hw, hw)
            return (y, y, y, y)

        def init_weights(self):
            pass

    class FeatureEncoder(nn.Module):
        def __init__(self):
            super().__init__()
            g = torch.Generator().manual_seed(seed)
            self.feat = nn.Parameter(torch.randn(batch, dim, hw, hw, generator=g))
            self.backbone = _BB()

        def forward(self, x):
            if x.shape[0] != self.feat.shape[0]:      # the constructor's 2-image stride probe (builder.py:392-402)
                return x.new_zeros(x.shape[0], self.feat.shape[1], x.shape[2] // 16, x.shape[3] // 16)
            return self.feat * 1.0

    return FeatureEncoder()


# ----------------------------------------------------------------- CPU legs (the reference's own code)
def _host_threads():
    """All host threads, set explicitly: torch.distributed.run exports OMP_NUM_THREADS=1, which would otherwise
    pin the reference arm to one core at N > 1."""
    import torch
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        pass
    torch.set_num_threads(n)
    return torch.get_num_threads()


def _reference_stepper(batch, device="cpu", channels_last=False, encoder="deeplab", seed=0):
    """-> (step_fn(batch_dict) -> loss float, kind).  kind "_ref": the UNMODIFIED reference builder.MODEL
    (source tree or oracle/_ref/*.pyc, through oracle/ref_harness.py) driven as main.py:590-644 drives it;
    kind "port": oracle/cp2_oracle.CP2StepPort, only when no reference module is reachable (CPU only)."""
    import torch
    from cp2_b200.encoder import DeepLabV3Encoder
    from oracle import ref_harness as rh
    if encoder == "deeplab":
        factory = lambda: DeepLabV3Encoder(dim=DIM)
    else:
        seeds = iter((11, 12))
        factory = lambda: feature_encoder(batch, DIM, IMG // OS_, next(seeds))
    if rh.reference_available():
        from oracle.ref_step import ReferenceStep
        ref = ReferenceStep(factory, device=device, dim=DIM, K=QUEUE, temp_global=TAU_G, temp_local=TAU_D,
                            lmbd_dense=LMBD, lr=1e-3, seed=seed, channels_last=channels_last)
        return ref.step, "_ref", ref
    if device != "cpu":
        return None, None, None
    from oracle import cp2_oracle as O
    port = O.CP2StepPort(factory, dim=DIM, K=QUEUE, temp_global=TAU_G, temp_local=TAU_D, lmbd_dense=LMBD,
                         output_stride=OS_, lr=1e-3, seed=seed)
    return (lambda b: port.step(**b)), "port", port


def reference_arm(args):
    """--impl reference: the reference's own CPU implementation of the step, all host threads, the FULL per-GPU
    batch of configs[1] every step (the step COUNT is what shrinks when the budget would be exceeded)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = _host_threads()
    step, kind, _ = _reference_stepper(args.batch)
    b = synth_batch(args.batch, seed=1234)
    budget = float(args.cpu_budget)
    t0 = time.perf_counter()
    step(b)                                             # warm-up 1 (also sizes the rest)
    first = time.perf_counter() - t0
    want = args.steps + args.warmup
    fit = max(2, int(budget / max(first, 1e-3)))       # steps that fit the budget, incl. the one just run
    warm = max(0, min(args.warmup, fit - 2) - 1)
    timed = max(1, min(args.steps, fit - 1 - warm))
    for _ in range(warm):
        step(b)
    t0 = time.perf_counter()
    for _ in range(timed):
        step(b)
    dt = (time.perf_counter() - t0) / timed
    value = args.batch / dt
    what = ("the unmodified reference builder.MODEL.forward_cp2 + backward + SGD (main.py:616-644) via "
            "oracle/ref_harness.py" if kind == "_ref" else "oracle/cp2_oracle.CP2StepPort (reference modules not reachable)")
    sample = (f"{timed} timed steps after {warm + 1} warm-up (asked {args.steps}+{args.warmup}; budget {budget:.0f} s) "
              f"of the full pre-training step at the full per-GPU batch {args.batch}: {what}; segmentor = the same "
              f"torchvision stand-in as the GPU arm; {dt * 1e3:.0f} ms/step; {os.cpu_count()} logical CPUs, "
              f"torch threads {threads} (set explicitly; OMP_NUM_THREADS in env = {os.environ.get('OMP_NUM_THREADS')})")
    cfg = workload_config(args.batch, args.gpus)
    cfg["note"] = ("CPU arm: one process on rank 0's host cores runs ONE GPU's batch; at N > 1 the driver's ratio "
                   "compares it with N GPUs' aggregate")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "steps_executed": timed, "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads,
                             "kind": "reference" if kind == "_ref" else "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def cpu_baseline_leg(batch, budget_s=25.0):
    """`cpu_baseline` of the own arm: the reference arm in a child process (its `.cuda()` -> identity patch and
    thread settings stay out of this process), bounded to ~budget_s of CPU work."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "5", "--warmup", "1",
           "--batch", str(batch), "--cpu-budget", str(budget_s)]
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "OMP_NUM_THREADS", "CUDA_VISIBLE_DEVICES"):
        env.pop(k, None)
    env["CUDA_VISIBLE_DEVICES"] = ""                    # host cores only
    try:
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600, env=env)
        for ln in reversed(r.stdout.strip().splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)["cpu_baseline"]
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "failed: " + r.stderr[-300:]}
    except Exception as e:  # noqa: BLE001
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"failed: {e}"}


# ------------------------------------------------------------------ dense-loss microbench
def _eager_dense_us(q, k, ma, mb, inv_tau, bf16, chunk, max_chunks=2):
    """The eager torch formulation of the dense loss (builder.py:1261,1279,1289-1292,1431-1437) + autograd to q on
    the same GPU, fp32 (`bf16=False`: what the reference runs, cuBLAS fp32/TF32 bmm + separate softmax kernels)
    or with the bmm in bf16 (`bf16=True`).  The [b,P,P] tensors force chunking at 56x56: `max_chunks` chunks of
    `chunk` images are timed and the figure is scaled to the whole batch."""
    import torch
    import torch.nn.functional as F
    B = q.shape[0]
    n = min(max_chunks, (B + chunk - 1) // chunk)
    ts = []
    for it in range(n + 1):
        i0 = (it % n) * chunk
        qs = q[i0:i0 + chunk].clone().requires_grad_()
        ks, a, b = k[i0:i0 + chunk], ma[i0:i0 + chunk], mb[i0:i0 + chunk]
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        qh, kh = F.normalize(qs, dim=1), F.normalize(ks, dim=1)
        if bf16:
            logits = torch.bmm(qh.transpose(1, 2).to(torch.bfloat16), kh.to(torch.bfloat16)).float()
        else:
            logits = torch.einsum("ncx,ncy->nxy", qh, kh)
        labels = torch.einsum("nx,ny->nxy", a, b)
        nl = -torch.log_softmax(logits * inv_tau, dim=1)
        loss = ((nl * labels).flatten(1).sum(1) / labels.flatten(1).sum(1)).mean()
        loss.backward()
        e1.record()
        torch.cuda.synchronize()
        if it > 0:
            ts.append(e0.elapsed_time(e1) * 1e3)
        del qs, logits, labels, nl, loss
    return statistics.mean(ts) * (B / chunk), n * chunk


def dense_sweep(args, peaks, eager=True):
    """configs[3]: dense pixel InfoNCE fwd+bwd at B=256 images/GPU, P in {14^2,28^2,56^2}, C in {128,256}.
    Algorithmic work 4*B*P^2*C FLOP (Q^T K forward + dQ backward; the backward's recomputation of Q^T K is
    executed but not counted).  Two mask variants: all-ones (nothing can be skipped: the honest
    roofline point) and Bernoulli(0.65) fg masks (CP2's copy-paste masks are 50-80 % foreground), where the
    key compaction skips background key columns.  `frac` is the EXECUTED-work fraction of the bf16 burst peak
    (counted FLOP x share of key columns actually visited); `frac_dense_algorithmic` counts the skipped columns
    as if they had been computed.  `us` = cp2_dense_lse + cp2_dense_grad; `us_with_prep` adds the two
    cp2_prep_features launches.  L2 is flushed between iterations.  `eager_*`: the eager torch formulation on
    the same GPU (fp32 and bf16-bmm), chunked, scaled to B."""
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
            for variant in ("ones", "bernoulli(0.65)"):
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
                    lse, dq = ops.dense_fwd_bwd(qh, kh, cnt, ma, q_inv, alpha, v1, None, inv_tau)
                    e[2].record()
                    torch.cuda.synchronize()
                    if it >= 3:
                        t_core.append(e[1].elapsed_time(e[2]) * 1e3)
                        t_all.append(e[0].elapsed_time(e[2]) * 1e3)
                us, us_all = statistics.mean(t_core), statistics.mean(t_all)
                flop = 4.0 * B * P * P * C
                tf = flop / (us * 1e-6) / 1e12
                executed = float(cnt.float().mean()) / P
                row = {"P": P, "C": C, "B": B, "masks": variant, "us": round(us, 2),
                       "us_with_prep": round(us_all, 2), "flop": flop, "executed_frac": round(executed, 4),
                       "tflops_executed": round(tf * executed, 2), "peak": peaks["tf_burst"],
                       "frac": round(tf * executed / peaks["tf_burst"], 4),
                       "frac_dense_algorithmic": round(tf / peaks["tf_burst"], 4),
                       "finite": bool(torch.isfinite(dq).all())}
                del qh, kh, dq, lse
                if eager and variant == "ones":
                    chunk = max(1, min(B, (1 << 31) // (P * P * 4 * 8)))      # ~8 live [b,P,P] f32 tensors <= 2 GiB
                    for tag, bf in (("eager_fp32_us", False), ("eager_bf16_bmm_us", True)):
                        t, n_img = _eager_dense_us(q, k, ma, mb, inv_tau, bf, chunk)
                        row[tag] = round(t, 1)
                        row[tag.replace("_us", "_timed_images")] = n_img
                    row["speedup_vs_eager_fp32"] = round(row["eager_fp32_us"] / us_all, 1)
                    row["speedup_vs_eager_bf16_bmm"] = round(row["eager_bf16_bmm_us"] / us_all, 1)
                rows.append(row)
            del q, k
    return rows


def headline_dense(rows):
    """The north-star point (56x56, C=128, all-ones masks) as its own object."""
    for r in rows:
        if r["P"] == 3136 and r["C"] == 128 and r["masks"] == "ones":
            return {"P": 3136, "C": 128, "B": r["B"], "masks": "ones", "fwd_bwd_us": r["us"],
                    "tflops": r["tflops_executed"], "peak_tflops": r["peak"], "frac": r["frac"],
                    "target_frac": 0.60, "eager_fp32_us": r.get("eager_fp32_us"),
                    "eager_bf16_bmm_us": r.get("eager_bf16_bmm_us")}
    return None


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

        small = ops.dense_small_preferred(B, C, P, dev)          # maps that fit TMEM: the tcgen05 kernel

        stream = (not small) and ops.dense_weighted_stream_ok(C, P, inv_tau)    # larger maps: streaming tcgen05 pair

        def ours():
            if small:
                _, _, _, cnt, _, acc = ops.dense_weighted_small(qh, kh, ma, mb, pw, sq, nsq, sk, nsk, inv_tau)
                return ops.dense_small_bwd(acc, qh, ma, q_inv, alpha, v1, None, key_count=cnt)
            if stream:
                lse, _, wstats, _, _, ids_pad, b_pad = ops.dense_weighted_stream(qh, kh, ma, mb, pw, sq, nsq, sk, nsk,
                                                                                 inv_tau)
                return ops.dense_weighted_stream_grad(qh, kh, ma, mb, pw, lse, wstats, ids_pad, b_pad, q_inv, alpha, v1,
                                                      None, inv_tau)
            lse, dense_n, wstats, _, _ = ops.dense_weighted_lse(qh, kh, ma, mb, pw, sq, nsq, sk, nsk, inv_tau)
            return ops.dense_weighted_grad(qh, kh, ma, mb, pw, lse, wstats, q_inv, alpha, v1, None, inv_tau)

        def simt():
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

        t_ours, t_eager, t_simt = timeit(ours, iters), timeit(eager, 5), timeit(simt, iters)
        rows.append({"B": B, "P": P, "C": C, "negative_type": "FIXED", "weights": list(wts),
                     "weighted_lse_grad_us": round(t_ours, 2), "launches": 3 if small else (6 if stream else 4),
                     "path": "tcgen05 (cp2_dense_weighted_small + cp2_dense_small_bwd)" if small
                             else "tcgen05 streaming (cp2_dense_weighted_stream + cp2_dense_weighted_stream_grad)" if stream
                             else "fp32 SIMT (cp2_dense_weighted_lse + cp2_dense_weighted_grad)",
                     "fp32_simt_pair_us": round(t_simt, 2),
                     "eager_torch_us": round(t_eager, 1), "speedup_vs_eager": round(t_eager / t_ours, 1),
                     "executed_tflops": round(6.0 * B * P * P * C / t_ours / 1e6, 2)})
    return rows


# ------------------------------------------------------------------ forward_densecl loss microbench
def densecl_sweep(peaks, iters=10):
    """Row f2 at the DenseCL scripts' shape (B=32, 14x14 backbone map, 2048-d backbone features, dim 128,
    K=65536, tau 0.2): ops.densecl_loss forward + backward (backbone match = cp2_densecl_match)
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

    sh1, sh2 = ops.QueueShadow(), ops.QueueShadow()      # as MODEL keeps them: the bf16 shadows persist across steps

    def ours():
        q_local.grad = q_global.grad = None
        ops.densecl_loss(embd_q, embd_k, q_local, k_local, q_global, k_global, queue, queue2, tau, tau, 0.5,
                         shadow=sh1, shadow2=sh2)[0].backward()

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
             "note": "queue sums of the 6272-row local InfoNCE in one launch of the dense loss's tcgen05 backward "
                     "kernel (cp2_queue_sums, rectangular mode), global InfoNCE on cp2_global_nce, backbone match = "
                     "cp2_densecl_match (tcgen05, bf16 hi+lo split of the fp32 features); eager arm: cuBLAS bmm"}]


# ------------------------------------------------------------------------------ own arm
def algorithmic_work(batch, n_params, world):
    """Per-CALL algorithmic bytes / flops of each C-ABI entry point on the step."""
    P, C, hw = (IMG // OS_) ** 2, DIM, IMG * IMG
    comp = 3 * batch * 3 * hw * 4 + batch * P * 4
    return {
        "cp2_composite": ("hbm", comp),
        "cp2_composite2": ("hbm", 2 * comp),
        "cp2_prep_features": ("hbm", 6 * batch * C * P + 4 * batch * P),
        "cp2_prep_features2": ("hbm", 2 * (6 * batch * C * P + 4 * batch * P)),
        "cp2_dense_lse": ("tensor", 2.0 * batch * P * P * C),
        "cp2_dense_grad": ("tensor", 2.0 * batch * P * P * C),
        "cp2_dense_small": ("tensor", 4.0 * batch * P * P * C),
        "cp2_dense_small_bwd": ("hbm", batch * P * C * (4 + 2 + 4)),
        "cp2_dense_rescore": ("hbm", 2 * batch * C * P * 4),
        "cp2_global_nce": ("hbm", 4 * C * QUEUE + 8 * batch * C),
        "cp2_ema": ("hbm", 12 * n_params),
        "cp2_enqueue": ("hbm", 2 * world * batch * C * 4),
        "cp2_loss_combine": ("hbm", batch * (9 * C + 2 * P) * 4),   # latency-bound: [B,C] / [B,P] operands once
        "cp2_id_iou": ("hbm", 2 * batch * P * (8 + 4)),       # latency-bound: 2P ids + masks per image
        # f4 arg-max meters, fp32 CUDA cores: counted by the bytes they must read (features / queue once)
        "cp2_dense_top1": ("hbm", 2 * batch * C * P * 4 + 2 * batch * P * 4),
        "cp2_global_top1": ("hbm", 4 * C * QUEUE + 8 * batch * C),
    }


def multi_rank_parity(model, dev, rank, world):
    """World > 1, before the timed region: the NCCL paths against the CPU oracle (the untimed checker leg may
    import oracle/).  3 enqueues through `_dequeue_and_enqueue` (all-gather + cp2_enqueue) and one
    `_batch_shuffle_ddp` / `_batch_unshuffle_ddp` round trip (perm broadcast + all_to_all_single):
      queue / queue_ptr equal on every rank and equal to oracle.dequeue_and_enqueue on the gathered keys,
      the shuffled rows equal to oracle.batch_shuffle for every rank, the round trip the identity.
    The model's queue and pointer are restored afterwards.  -> dict of booleans (identical on every rank)."""
    import torch
    import torch.distributed as dist
    from oracle import cp2_oracle as O

    def gather(t):
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t.contiguous())
        return out

    res = {}
    with torch.no_grad():
        queue0, ptr0 = model.queue.clone(), model.queue_ptr.clone()
        B, C, K = 32, model.queue.shape[0], model.queue.shape[1]
        model.queue_ptr.fill_(K - 2 * B * world - 5)            # the third enqueue wraps around
        ref_q, ref_ptr = model.queue.cpu().clone(), int(model.queue_ptr.item())
        for stp in range(3):
            g = torch.Generator().manual_seed(1000 + 17 * stp + rank)
            keys = torch.nn.functional.normalize(torch.randn(B, C, generator=g), dim=1).to(dev)
            model._dequeue_and_enqueue(keys)
            keys_all = torch.cat([t.cpu() for t in gather(keys)])
            ref_ptr = O.dequeue_and_enqueue(ref_q, ref_ptr, keys_all)
        qs, ps = gather(model.queue), gather(model.queue_ptr)
        res["queue_equal_across_ranks"] = all(torch.equal(qs[0], t) for t in qs)
        res["queue_ptr_equal_across_ranks"] = all(torch.equal(ps[0], t) for t in ps)
        res["queue_matches_oracle"] = bool(torch.equal(model.queue.cpu(), ref_q))
        res["queue_ptr_matches_oracle"] = int(model.queue_ptr.item()) == ref_ptr
        model.queue.copy_(queue0)
        model.queue_ptr.copy_(ptr0)
        # shuffle / unshuffle: rows tagged with their global index
        x = (torch.arange(B, device=dev, dtype=torch.float32) + rank * B).reshape(B, 1, 1, 1).expand(B, 3, 4, 4).contiguous()
        x = x + torch.arange(3 * 16, device=dev, dtype=torch.float32).reshape(1, 3, 4, 4) / 64.0
        y, idx_unshuffle = model._batch_shuffle_ddp(x)
        z = model._batch_unshuffle_ddp(y, idx_unshuffle)
        x_all = torch.cat([t.cpu() for t in gather(x)])
        ys = [t.cpu() for t in gather(y)]
        idx_shuffle = torch.argsort(idx_unshuffle).cpu()
        res["shuffle_matches_oracle"] = all(torch.equal(ys[r], O.batch_shuffle(x_all, idx_shuffle, r, world)[0])
                                            for r in range(world))
        perms = gather(idx_unshuffle)
        res["perm_equal_across_ranks"] = all(torch.equal(perms[0], t) for t in perms)
        res["unshuffle_roundtrip"] = bool(torch.equal(z, x))
    flags = torch.tensor([int(bool(v)) for v in res.values()], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)                # a failure on any rank fails everywhere
    return {k_: bool(f) for k_, f in zip(res.keys(), flags.tolist())}


def ddp_weights_equal(model, dev, world):
    import torch
    import torch.distributed as dist
    with torch.no_grad():
        sig = torch.stack([p.detach().double().sum() for p in model.encoder_q.parameters()] +
                          [p.detach().double().sum() for p in model.encoder_k.parameters()])
    out = [torch.empty_like(sig) for _ in range(world)]
    dist.all_gather(out, sig)
    return all(torch.equal(out[0], t) for t in out)


def gpu_busy_breakdown(step_fn, n, dev):
    """Where one step's wall time goes on this rank's GPU (torch.profiler / CUPTI over n steps, untimed leg):
    union of all kernel intervals (GPU busy), the NCCL kernels' share, and what is left (GPU idle: host-bound
    launch gaps and waits on peers).  -> dict, or {"unavailable": why}."""
    import torch
    try:
        from torch.profiler import ProfilerActivity, profile
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            e0.record()
            for i in range(n):
                step_fn(i)
            e1.record()
            torch.cuda.synchronize(dev)
        wall_ms = e0.elapsed_time(e1) / n
        iv, nccl_us, memcpy_us = [], 0.0, 0.0
        for ev in prof.events():
            if str(getattr(ev, "device_type", "")).endswith("CUDA") and ev.time_range is not None:
                t0, t1 = ev.time_range.start, ev.time_range.end
                if t1 <= t0:
                    continue
                iv.append((t0, t1))
                if "nccl" in ev.name.lower():
                    nccl_us += t1 - t0
                if "memcpy" in ev.name.lower() or "memset" in ev.name.lower():
                    memcpy_us += t1 - t0
        if not iv:
            return {"unavailable": "profiler returned no device events"}
        iv.sort()
        busy, cs, ce = 0.0, iv[0][0], iv[0][1]
        for a, b in iv[1:]:
            if a > ce:
                busy += ce - cs
                cs, ce = a, b
            else:
                ce = max(ce, b)
        busy += ce - cs
        return {"step_ms": round(wall_ms, 3), "gpu_busy_ms_per_step": round(busy / n / 1e3, 3),
                "gpu_idle_ms_per_step": round(wall_ms - busy / n / 1e3, 3),
                "nccl_kernels_ms_per_step": round(nccl_us / n / 1e3, 3),
                "memcpy_memset_ms_per_step": round(memcpy_us / n / 1e3, 3), "kernels_per_step": len(iv) // n,
                "how": "torch.profiler (CUPTI) over %d steps on rank 0's GPU: union of all kernel / copy intervals" % n}
    except Exception as e:                                    # profiler not usable on this box: say so, do not fail
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


def collectives_summary(times, steps):
    """model.collective_times -> {name: us per step}; device-side CUDA-event durations on the issuing stream
    (host seconds for the gloo permutation broadcast)."""
    import torch
    torch.cuda.synchronize()
    out = {}
    for name, recs in times.items():
        if name.endswith("_s"):
            out[name[:-2] + "_us_per_step"] = round(1e6 * sum(recs) / max(steps, 1), 1)
        else:
            out[name + "_us_per_step"] = round(1e3 * sum(a.elapsed_time(b) for a, b in recs) / max(steps, 1), 1)
    return out


def gpu_eager_baseline(args, dev, own_step_ms, own_slice_fn):
    """Same-GPU eager baselines: the UNMODIFIED reference builder.MODEL on this B200 through
    its stock eager PyTorch path (cuBLAS bmm + separate softmax kernels + per-step host syncs), (a) the full step
    with the same stand-in segmentor, as the reference runs it (NCHW) and with the backbone in channels_last
    like the own arm, (b) the loss slice alone (a stored-feature 'encoder': composite x2, id maps, normalise,
    both InfoNCE losses, enqueue, backward to the features) for both implementations, (c) the EMA over the full
    66 M-parameter model.  CUDA events, 3 warm-up + 10 timed iterations each."""
    import torch
    from oracle import ref_harness as rh
    if not rh.reference_available():
        return {"unavailable": "reference modules not reachable (oracle/_ref not built)"}
    B = args.batch
    resident = {k: v.to(dev) for k, v in synth_batch(B, seed=1234).items()}

    def timeit(fn, n=10, warm=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    out = {"what": "unmodified reference builder.MODEL (oracle/ref_harness.py), eager PyTorch on the same B200",
           "batch": B}
    for tag, cl in (("step_ms_nchw", False), ("step_ms_channels_last", True)):
        step, kind, ref = _reference_stepper(B, device=str(dev), channels_last=cl)
        out[tag] = round(timeit(lambda: step(resident)), 3)
        if not cl:
            m = ref.model
            out["ema_eager_us"] = round(1e3 * timeit(m._momentum_update_key_encoder), 1)
        del step, ref
        torch.cuda.empty_cache()
    out["own_step_ms"] = round(own_step_ms, 3)
    out["speedup_vs_eager_gpu_step"] = round(out["step_ms_channels_last"] / own_step_ms, 3)
    out["speedup_vs_eager_gpu_step_nchw_reference"] = round(out["step_ms_nchw"] / own_step_ms, 3)
    step, kind, ref = _reference_stepper(B, device=str(dev), encoder="features")
    out["loss_slice_eager_us"] = round(1e3 * timeit(lambda: step(resident)), 1)
    del step, ref
    torch.cuda.empty_cache()
    out["loss_slice_own_us"] = round(1e3 * timeit(own_slice_fn(resident)), 1)
    out["speedup_vs_eager_gpu_loss_slice"] = round(out["loss_slice_eager_us"] / out["loss_slice_own_us"], 2)
    return out


def mirror_arm(args, dev, rank, world, peaks):
    """--workload mirror: BASELINE.json configs[4], one `MirrorModule.shared_step` training iteration
    (mirror_pretrain.py:212-245, networks/mirror_network.py:40-63) per step: 512x512 source / target images,
    B=8 per GPU, MirrorVariant.OUTPUT, Adam (segment_network.py:294-310), the segmentor = the same ResNet-50
    DeepLabV3 stand-in with a 2-class head (OS16 -> 32x32 logits; the loss kernel upsamples on chip)."""
    import torch
    import torch.distributed as dist
    from cp2_b200 import ops
    from cp2_b200.encoder import DeepLabV3Encoder
    from cp2_b200.mirror_network import MirrorModule, MirrorVariant, Stage
    torch.manual_seed(0)
    torch.backends.cudnn.benchmark = True
    B, H, Cls = (8 if args.batch == 32 else args.batch), 512, 2
    mod = MirrorModule(None, "RANDOM", 1e-4, 1e-4, Cls, (3, H, H), 0.01, 2.0, MirrorVariant.OUTPUT,
                       segmentor_factory=lambda: DeepLabV3Encoder(dim=Cls)).to(dev)
    if not args.nchw:
        mod = mod.to(memory_format=torch.channels_last)
    for p in mod.model.decode_head.conv_seg.parameters():      # unused head of the stand-in: no gradient ever
        p.requires_grad_(False)
    opt = mod.configure_optimizers()["optimizer"]
    g = torch.Generator().manual_seed(4321 + rank)
    host = [(torch.rand(B, 3, H, H, generator=g).pin_memory(), torch.rand(B, 3, H, H, generator=g).pin_memory(),
             torch.randint(0, Cls, (B, H, H), generator=g).pin_memory()) for _ in range(2)]
    resident = [tuple(t.to(dev) for t in hb) for hb in host]
    h2d = sum(t.numel() * t.element_size() for t in host[0])

    class _Wrap(torch.nn.Module):          # DDP drives forward(); route it to shared_step
        def __init__(self, m):
            super().__init__()
            self.m = m

        def forward(self, batch):
            return self.m.shared_step(batch, Stage.TRAIN)

    if world > 1:
        net = torch.nn.parallel.DistributedDataParallel(_Wrap(mod), device_ids=[dev.index], gradient_as_bucket_view=True)
        call = lambda b: net(b)
    else:
        call = lambda b: mod.shared_step(b, Stage.TRAIN)

    def step(batch):
        loss = call(batch)
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
    sampler = ClockSampler(dev.index, enabled=(rank == 0))
    sampler.start()
    l0 = ops.LAUNCHES
    with ops.KernelTimer() as kt:
        ms_value = timed(lambda i: step(resident[i % 2]), args.steps)
    launches = ops.LAUNCHES - l0
    ksum = kt.summary()

    def e2e_step(i):
        batch = tuple(t.to(dev, non_blocking=True) for t in host[i % 2])
        return step(batch).item()
    e2e_step(0)
    ms_e2e = timed(e2e_step, args.steps)
    clocks = sampler.stop()
    name = "cp2_mirror_loss_upsampled"
    calls, us = ksum.get(name, (0, 0.0))
    label_bytes = 8 * B * H * H + 4 * 2 * B * Cls * 32 * 32 * 4
    roof = {"kernel": name, "bound": "hbm", "achieved": round(label_bytes / max(us, 1e-9) / 1e3, 2),
            "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": round(label_bytes / max(us, 1e-9) / 1e3 / peaks["hbm_gbs"], 4),
            "traffic": None, "algorithmic_per_launch": label_bytes, "us_per_launch": round(us, 2),
            "peak_source": peaks["source"]}
    line = {"metric": "mirror_pretrain_images_per_sec", "value": B * world * args.steps / (ms_value * 1e-3),
            "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_value / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE.json configs[4]: mirror_pretrain, MirrorModule.shared_step (OUTPUT "
                                   "variant), 512x512 inputs, ResNet-50 DeepLabV3 stand-in (OS16, 32x32 logits), Adam",
                       "batch_per_gpu": B, "global_batch": B * world, "classes": Cls, "lmbd_compare": 0.01,
                       "softmax_temp": 2.0, "parallelism": f"dp{world}",
                       "l2": "no flush: one step streams several GB of activations through L2"},
            "e2e": {"value": B * world * args.steps / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "clocks": clocks, "roofline": roof,
            "kernels": {name: {"calls_per_step": calls / max(args.steps, 1), "us_per_call": round(us, 2)}}}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


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
            rows = dense_sweep(args, peaks, eager=not args.no_eager_baseline)
            print(json.dumps({"dense_microbench": rows, "dense_loss": headline_dense(rows)}), flush=True)
        return 0
    if args.workload == "mirror":
        return mirror_arm(args, dev, rank, world, peaks)

    torch.manual_seed(0)
    torch.backends.cudnn.benchmark = True

    def build_model(factory):
        m = MODEL(None, rank=rank, dim=DIM, K=QUEUE, m=0.999, instance_logits_temp=TAU_G,
                  pretrain_from_scratch=True, lmbd_cp2_dense_loss=LMBD, dense_logits_temp=TAU_D, device=dev,
                  encoder_factory=factory, output_stride=OS_).to(dev)
        return m

    model = build_model(lambda: DeepLabV3Encoder(dim=DIM))
    if not args.nchw:
        # cuDNN's tensor-core convolutions are NHWC kernels: NCHW weights/activations cost an
        # nchwToNhwc / nhwcToNchw transpose around every convolution (11 % of the r01b step)
        model = model.to(memory_format=torch.channels_last)
    n_params = sum(p.numel() for p in model.encoder_k.parameters())
    # The ~1100 cuDNN / ATen launches of the two encoders are the host-bound part of the step: capture them in CUDA
    # graphs (MODEL.graph_encoders).  Under DDP the query encoder is captured stage by stage (one autograd node per
    # stage): a single graph would serialise the gradient all-reduce behind the whole backward.
    folded = 0
    if not args.no_fold:
        try:
            folded = model.fold_dilated_convs(args.fold_cover)
        except Exception as e:                                   # noqa: BLE001 - the stock convolutions are the fallback
            print(f"[bench] fold_dilated_convs failed, stock convolutions kept: {type(e).__name__}: {e}", file=sys.stderr)
    graphs = {"encoder_q": False, "encoder_k": False, "folded_dilated_convs": folded}
    if not args.no_graphs:
        try:
            model.graph_encoders((args.batch, 3, IMG, IMG), query=True, key=True,
                                 query_stages=model.encoder_q.graph_stages() if world > 1 else None)
            graphs.update(model.graphed)
        except Exception as e:                                   # noqa: BLE001 - eager launches are the fallback
            graphs["error"] = f"{type(e).__name__}: {e}"[:300]
            print(f"[bench] CUDA-graph capture of the encoders failed, running them eagerly: {graphs['error']}",
                  file=sys.stderr)
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

    def step(batch, module=None):
        loss = (module or net)(visualize=False, step=0, new_epoch=False, **batch)
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

    parity = None
    if world > 1:
        parity = multi_rank_parity(model, dev, rank, world)
        # ... and the replicas after real all-reduced steps: 4 full training steps (forward through both encoders,
        # loss, backward with DDP's gradient all-reduce, SGD), then every parameter of both encoders must be
        # bit-identical on all ranks.  (Untimed checker leg; it also lets NCCL, DDP's bucket rebuild and the cuDNN
        # autotuner settle on every rank before the W warm-up steps.)
        for i in range(4):
            step(resident[i % 2])
        parity["ddp_weights_equal_after_4_steps"] = ddp_weights_equal(model, dev, world)

    for i in range(max(args.warmup, 3)):
        step(resident[i % 2])
    sampler = ClockSampler(local, enabled=(rank == 0))
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

    # e2e: every step's inputs come from pinned host memory and the loss is read back.
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

    def e2e_step_sync(i):
        batch, ev = pending.pop("next")
        cur = torch.cuda.current_stream()
        cur.wait_event(ev)
        for t in batch.values():
            t.record_stream(cur)
        prefetch(i + 1)
        return step(batch).item()           # device -> host read of the step's loss, blocking (main.py:644-660)

    # The same step with the read-back a training loop that does not want to stall uses: every step's loss is copied
    # into pinned host memory by an asynchronous D2H copy issued in that step and consumed on the host one step later,
    # so the host keeps queueing work while the GPU finishes the step.  Every step still moves its own inputs H2D and
    # its own loss D2H inside the timed region (the last copy completes before the closing event).
    loss_host = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event(), torch.cuda.Event()]
    seen = []

    def e2e_step(i):
        batch, ev = pending.pop("next")
        cur = torch.cuda.current_stream()
        cur.wait_event(ev)
        for t in batch.values():
            t.record_stream(cur)
        prefetch(i + 1)
        loss = step(batch)
        loss_host[i % 2].copy_(loss.detach(), non_blocking=True)
        loss_ev[i % 2].record(cur)
        if i > 0:
            loss_ev[(i - 1) % 2].synchronize()
            seen.append(float(loss_host[(i - 1) % 2]))

    prefetch(0)
    for i in range(2):
        e2e_step_sync(i)
    ms_e2e_sync = timed(e2e_step_sync, args.steps)
    ms_e2e = timed(e2e_step, args.steps)
    seen.append(float(loss_host[(args.steps - 1) % 2]))
    assert len(seen) == args.steps and all(math.isfinite(v) for v in seen), "e2e: a step's loss did not come back"
    clocks = sampler.stop()

    # ---- N > 1: where the step time goes between ranks (untimed extra legs, device-side event timings)
    collectives = None
    if world > 1:
        # after warm-up + both timed regions of all-reduced steps, and BEFORE the no_sync() A/B below, which steps the
        # optimizer on rank-local gradients and lets the replicas drift on purpose
        parity["ddp_weights_equal_across_ranks"] = ddp_weights_equal(model, dev, world)
        n_c = min(args.steps, 10)
        model.collective_times = {}
        timed(lambda i: step(resident[i % 2]), n_c)
        collectives = collectives_summary(model.collective_times, n_c)
        model.collective_times = None

        def step_no_sync(i):                # same step without DDP's gradient all-reduce
            with net.no_sync():
                step(resident[i % 2])
        ms_sync = timed(lambda i: step(resident[i % 2]), n_c) / n_c
        ms_nosync = timed(step_no_sync, n_c) / n_c
        model.overlap_exchange = False
        ms_serial = timed(lambda i: step(resident[i % 2]), n_c) / n_c
        model.overlap_exchange = True
        collectives["gpu_busy"] = gpu_busy_breakdown(lambda i: step(resident[i % 2]), 3, dev)
        collectives.update({"ddp_all_reduce_overhead_ms_per_step": round(ms_sync - ms_nosync, 3),
                            "step_ms": round(ms_sync, 3), "step_ms_without_grad_all_reduce": round(ms_nosync, 3),
                            "step_ms_with_serial_image_exchange": round(ms_serial, 3),
                            "note": "all_to_all / all_gather: CUDA events on the issuing stream (the image exchange "
                                    "runs on a side stream under encoder_q); all-reduce overhead from a no_sync() A/B"})

    if world == 1 and os.environ.get("CP2_BENCH_BREAKDOWN"):          # same breakdown at N=1, for comparison
        collectives = {"gpu_busy": gpu_busy_breakdown(lambda i: step(resident[i % 2]), 3, dev)}

    # ---- per-kernel figures from the CUDA events recorded inside the timed region
    work = algorithmic_work(B, n_params, world)
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")     # dram bytes per launch from `ncu --set full`
    if os.path.isfile(tpath):
        with open(tpath) as f:
            traffic = json.load(f)
    kernels, per_step_us = {}, 0.0
    for name, (calls, us) in ksum.items():
        bound, amount = work.get(name, ("hbm", 0))
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
        "data": "synthetic", "config": dict(workload_config(B, world, args.nchw), encoder_cuda_graphs=graphs),
        "e2e": {"value": B * world * args.steps / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / args.steps,
                "readback": "async D2H into pinned memory, consumed one step later",
                "blocking_item_value": B * world * args.steps / (ms_e2e_sync * 1e-3)},
        "gpu_launches": launches, "launches_per_step": launches / max(args.steps, 1), "clocks": clocks,
        "roofline": roofline, "kernels": kernels,
        "cp2_kernels_us_per_step": round(per_step_us, 1),
        "backbone_and_optimizer_ms_per_step": round(ms_value / args.steps - per_step_us * 1e-3, 3),
    }
    if parity is not None:
        line["parity_check"] = parity
    if collectives is not None:
        line["collectives"] = collectives
    if world == 1:
        own_ms = ms_value / args.steps
        net = opt = None
        if not args.no_eager_baseline:
            def own_slice_fn(batch):
                seeds = iter((11, 12))
                m2 = build_model(lambda: feature_encoder(B, DIM, IMG // OS_, next(seeds)))
                o2 = torch.optim.SGD(m2.encoder_q.parameters(), 1e-3, momentum=0.9, weight_decay=1e-4)

                def run():
                    loss = m2(visualize=False, step=0, new_epoch=False, **batch)
                    o2.zero_grad(set_to_none=True)
                    loss.backward()
                    o2.step()
                    return loss.item()
                return run
            try:
                line["gpu_eager_baseline"] = gpu_eager_baseline(args, dev, own_ms, own_slice_fn)
            except Exception as e:  # noqa: BLE001 - a baseline leg must not take the bench line down
                line["gpu_eager_baseline"] = {"failed": f"{type(e).__name__}: {e}"[:400]}
        model = None
        torch.cuda.empty_cache()
        if not args.no_sweep:
            rows = dense_sweep(args, peaks, eager=not args.no_eager_baseline)
            line["dense_loss"] = headline_dense(rows)
            line["dense_microbench"] = rows
            line["mirror_microbench"] = mirror_sweep(peaks)
            line["weighted_microbench"] = weighted_sweep(peaks)
            line["densecl_microbench"] = densecl_sweep(peaks)
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_leg(B)
    ok = parity is None or all(parity.values())
    if rank == 0:
        print(json.dumps(line), flush=True)
        if not ok:
            sys.stderr.write(f"bench.py: multi-rank parity check FAILED: {parity}\n")
    if world > 1:
        dist.destroy_process_group()
    return 0 if ok else 3


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    return own_arm(args)


if __name__ == "__main__":
    sys.exit(main())
