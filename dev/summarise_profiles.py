#!/usr/bin/env python
"""Turns the scratch captures under gpurun_out/ into the small tracked summaries under profiles/.

    python dev/summarise_profiles.py r01a        # tag = round + letter of the capture

Reads (whatever exists): launches_step.csv (ncu --metrics gpu__time_duration.sum over the timed region of
bench.py), prof_step.ncu-rep / prof_dense.ncu-rep (ncu --set full), bench.json, bench_ref.json.
Writes profiles/<tag>_launches_step.txt, profiles/<tag>_ncu_<name>.csv, profiles/<tag>_bench.json and
refreshes profiles/traffic.json (dram bytes per launch of each cp2 kernel, read by bench.py)."""
import collections
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "gpurun_out")
DST = os.environ.get("CP2_PROFILES_DST") or os.path.join(ROOT, "profiles")   # on the GPU box: a directory under gpurun_out/

METRICS = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "dram__bytes_read.sum",
    "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct",
]
# ncu kernel name fragment -> C-ABI entry point (bench.py's kernel names)
ENTRY = {"composite_kernel": "cp2_composite", "ema_kernel": "cp2_ema", "prep_kernel": "cp2_prep_features",
         "global_nce_kernel": "cp2_global_nce", "dense_lse_kernel": "cp2_dense_lse",
         "dense_grad_kernel": "cp2_dense_grad", "enqueue_kernel": "cp2_enqueue",
         "loss_combine_kernel": "cp2_loss_combine", "id_iou_kernel": "cp2_id_iou",
         "dense_top1_kernel": "cp2_dense_top1", "global_top1_kernel": "cp2_global_top1",
         "dense_small_kernel": "cp2_dense_small", "dense_small_bwd_kernel": "cp2_dense_small_bwd",
         "dense_wsmall_kernel": "cp2_dense_weighted_small", "densecl_match_kernel": "cp2_densecl_match"}
UNIT = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE,
                         stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if len(rows) < 3:
        return [], [], []
    return rows[0], rows[1], rows[2:]


def summarise_rep(rep, dst, traffic):
    hdr, units, rows = raw_rows(rep)
    if not rows:
        return
    cols = [i for i, h in enumerate(hdr) if h in METRICS]
    kn = hdr.index("Kernel Name")
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + [f"{hdr[i]} [{units[i]}]" for i in cols])
        for r in rows:
            w.writerow([r[kn][:90]] + [r[i] for i in cols])
    ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    acc = collections.defaultdict(list)
    for r in rows:
        for frag, entry in ENTRY.items():
            if frag in r[kn]:
                b = float(r[ir]) * UNIT.get(units[ir], 1.0) + float(r[iw]) * UNIT.get(units[iw], 1.0)
                acc[entry].append(b)
    for entry, v in acc.items():
        traffic[entry] = round(sum(v) / len(v))


def summarise_launches(src, dst):
    rows = list(csv.reader(open(src)))
    hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    hdr = rows[hi]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot, cnt = collections.Counter(), collections.Counter()
    for r in rows[hi + 1:]:
        if len(r) <= mv:
            continue
        v = float(r[mv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[mu], 1.0)
        tot[r[kn][:80]] += v
        cnt[r[kn][:80]] += 1
    T = sum(tot.values())
    with open(dst, "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum over the timed region of `bench.py --steps 2` "
                f"(cold-cache, serialised: compare shares)\n# total {T:.1f} us, {sum(cnt.values())} launches\n")
        f.write("# --- cp2_b200 kernels\n")
        for n, v in tot.most_common():
            if any(k in n for k in list(ENTRY) + ["prep_reduce", "global_finalize", "advance_ptr", "loss_finalize", "top1_finalize",
                                                "dense_nl_kernel", "weighted_", "mirror_"]):
                f.write(f"{v:10.1f} us {100 * v / T:6.2f}% x{cnt[n]:4d}  {n}\n")
        f.write("# --- top 25 overall (backbone = cuDNN / ATen, out of scope)\n")
        for n, v in tot.most_common(25):
            f.write(f"{v:10.1f} us {100 * v / T:6.2f}% x{cnt[n]:4d}  {n}\n")


def main():
    tag = sys.argv[1]
    os.makedirs(DST, exist_ok=True)
    tpath = os.path.join(DST, "traffic.json")
    traffic = json.load(open(tpath)) if os.path.isfile(tpath) else {}
    for name in ("prof_step", "prof_dense", "prof_global", "prof_mirror", "prof_f2f3", "prof_small"):
        rep = os.path.join(SRC, name + ".ncu-rep")
        if os.path.isfile(rep):
            summarise_rep(rep, os.path.join(DST, f"{tag}_ncu_{name[5:]}.csv"), traffic if name == "prof_step" else {})
    if os.path.isfile(os.path.join(SRC, "launches_step.csv")):
        summarise_launches(os.path.join(SRC, "launches_step.csv"), os.path.join(DST, f"{tag}_launches_step.txt"))
    for name in ("bench.json", "bench_ref.json", "bench_nographs.json", "mirror_times.log"):
        if os.path.isfile(os.path.join(SRC, name)) and os.path.getsize(os.path.join(SRC, name)) > 0:
            shutil.copy(os.path.join(SRC, name), os.path.join(DST, f"{tag}_{name}"))
    with open(tpath, "w") as f:
        json.dump(traffic, f, indent=1, sort_keys=True)
    print("wrote", sorted(os.listdir(DST)))


if __name__ == "__main__":
    main()
