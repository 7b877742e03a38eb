"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals of the last step + per-launch table."""
import collections
import csv
import re
import sys

path, per_step, out = sys.argv[1], int(sys.argv[2]), sys.argv[3] if len(sys.argv) > 3 else None
rows = list(csv.DictReader([l for l in open(path) if not l.startswith("==")]))[-per_step:]
agg, tot, lines = collections.OrderedDict(), 0.0, []
for i, r in enumerate(rows):
    n = re.sub(r"\(.*", "", r["Kernel Name"]).replace("dtb::", "").replace("void ", "")[:40]
    v = float(r["Metric Value"].replace(",", "")) / 1000.0
    a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += v; tot += v
    lines.append(f"{i:3d} {v:8.1f} us grid={r['Grid Size']:>14s} blk={r['Block Size']:>12s} {n}")
summ = [f"{t:9.1f} us {c:4d} launches {t / tot * 100:5.1f}%  {n}" for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])]
summ.append(f"total {tot:.1f} us over {len(rows)} launches (cold-cache, serialised: compare shares, not absolutes)")
print("\n".join(summ))
if out:
    open(out, "w").write("# " + " ".join(sys.argv) + "\n" + "\n".join(summ) + "\n\n" + "\n".join(lines) + "\n")
