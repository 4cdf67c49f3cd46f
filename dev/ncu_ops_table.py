import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/ncu_ops.csv")))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
kn, mn, mv = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value")
acc = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= mv:
        continue
    acc.setdefault((r[kn][:60], r[mn]), []).append(float(r[mv].replace(",", "")))
for (k, m), v in acc.items():
    print(f"{k:60s} {m:28s} n={len(v):2d} min={min(v):10.1f} med={sorted(v)[len(v)//2]:10.1f}")
