"""profiles/r02_launches_by_kernel.csv from the raw ncu launch list of one timed epoch (scripts/run_launchlist.sh):
launches, summed and average duration per kernel, and each kernel's share of the summed kernel time."""
import collections
import csv
import sys

raw = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/r02_launches_raw.csv"
rows = [l for l in open(raw) if l.startswith('"')]
agg = collections.OrderedDict()
for r in csv.DictReader(rows):
    if r["Metric Name"] != "gpu__time_duration.sum":
        continue
    ns = float(r["Metric Value"]) * {"ns": 1.0, "us": 1e3, "ms": 1e6}[r["Metric Unit"]]
    a = agg.setdefault(r["Kernel Name"], [0, 0.0])
    a[0] += 1
    a[1] += ns
total = sum(a[1] for a in agg.values())
w = csv.writer(sys.stdout)
w.writerow(["kernel", "launches", "total_us", "avg_us", "share_of_epoch"])
for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    w.writerow([k, n, round(ns / 1e3, 1), round(ns / 1e3 / n, 2), round(ns / total, 4)])
print(f"# {sum(a[0] for a in agg.values())} launches, {total / 1e6:.2f} ms summed (cold-cache, serialised: shares, not absolutes)", file=sys.stderr)
