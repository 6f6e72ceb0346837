"""Turn an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel totals and shares.
usage: python scripts/summarise_launches.py launches.csv "header comment" > profiles/<name>_summary.csv"""
import csv, io, sys
from collections import OrderedDict

txt = open(sys.argv[1]).read()
txt = txt[txt.index('"ID"'):]
tot, cnt = OrderedDict(), {}
for row in csv.DictReader(io.StringIO(txt)):
    if row["Metric Name"] != "gpu__time_duration.sum":
        continue
    v = float(row["Metric Value"].replace(",", ""))
    unit = row["Metric Unit"]
    ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
    k = row["Kernel Name"][:100]
    tot[k] = tot.get(k, 0.0) + ms
    cnt[k] = cnt.get(k, 0) + 1
total = sum(tot.values())
print("# " + (sys.argv[2] if len(sys.argv) > 2 else "ncu launch list"))
print("# gpu__time_duration.sum per kernel name; cold-cache, serialised: compare SHARES")
print("kernel,launches,total_ms,share")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f'"{k}",{cnt[k]},{v:.3f},{v / total:.4f}')
