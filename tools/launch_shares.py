"""Per-kernel launches / mean ns / share of listed time from an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: python tools/launch_shares.py profiles/<capture>_launches.csv"""
import csv, sys, collections, re
rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
rd = csv.DictReader(rows)
acc = collections.OrderedDict()
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    v = float(r["Metric Value"].replace(",", ""))
    if r.get("Metric Unit", "ns") in ("us", "usecond"):
        v *= 1e3
    n, t = acc.get(name, (0, 0.0))
    acc[name] = (n + 1, t + v)
tot = sum(t for _, t in acc.values())
print("| kernel | launches | mean ns | share of listed time |\n|---|---|---|---|")
for k, (n, t) in sorted(acc.items(), key=lambda kv: -kv[1][1]):
    print("| %s | %d | %.0f | %.1f%% |" % (k, n, t / n, 100 * t / tot))
