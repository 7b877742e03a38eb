"""ncu CSV (`ncu -i x.ncu-rep --page raw --csv` or `ncu --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum ...`)
-> per-kernel-family DRAM traffic table, merged into profiles/r02_traffic.json (read by bench.py's `roofline.traffic`).

    python tools/ncu_traffic.py <workload> <csv> [--source "how the CSV was captured"]

Families follow bench.PROF_FAMILIES; a launch's family is decided from its kernel name (GEMM launches cannot be split into the tensor /
HBM-bound families from the name alone, so both GEMM families get the mean over all GEMM launches of the capture, stated in `source`).
"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RULES = [(r"gemm_tcgen05", ["gemm_tcgen05_kernel (+pair)", "gemm_tcgen05_kernel, HBM-bound shapes"]), (r"conv1_fwd", ["conv1_fwd_kernel"]),
         (r"conv1_wgrad", ["conv1_wgrad_kernel"]), (r"attn_fwd", ["attn_fwd_tc_kernel"]), (r"attn_bwd", ["attn_bwd_dq+dkv_tc_kernel"]),
         (r"ln_fwd", ["ln_fwd_kernel"]), (r"ln_bwd_kernel", ["ln_bwd_kernel"]), (r"embed_fwd", ["embed_fwd_kernel"]), (r"colsum", ["colsum_kernel"]),
         (r"im2col", ["im2col_nhwc_kernel"]), (r"col2im", ["col2im_nhwc_relu_kernel"]), (r"convn_kernel<\(bool\)0>|convn_kernel<false>|convn_kernel<0>", ["convn_kernel<fwd>"]),
         (r"convn_kernel<\(bool\)1>|convn_kernel<true>|convn_kernel<1>", ["convn_kernel<dgrad>"]), (r"convn_wgrad_kernel", ["convn_wgrad_kernel"]),
         (r"opt_adamw", ["opt_adamw"]), (r"heads_", ["heads"]), (r"get_batch", ["get_batch_kernel"])]


def parse(path):
    """Returns {kernel name: [(read bytes, write bytes, duration ns), ...]} from either ncu CSV flavour."""
    rows = list(csv.reader(l for l in open(path, newline="") if l.startswith('"')))
    hdr = rows[0]
    per = {}
    if "Metric Name" in hdr:  # long format: one row per (launch, metric)
        ik, iid, im, iv, iu = hdr.index("Kernel Name"), hdr.index("ID"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
        launches = {}
        for r in rows[1:]:
            if len(r) <= iv or not r[iid].isdigit():
                continue
            d = launches.setdefault(int(r[iid]), {"name": r[ik]})
            try:
                v = float(r[iv].replace(",", ""))
            except ValueError:
                continue
            scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "usecond": 1e3, "msecond": 1e6, "nsecond": 1.0, "second": 1e9}.get(r[iu], 1.0)
            d[r[im]] = v * scale
        for d in launches.values():
            per.setdefault(d["name"], []).append((d.get("dram__bytes_read.sum", 0.0), d.get("dram__bytes_write.sum", 0.0), d.get("gpu__time_duration.sum", 0.0)))
    else:  # --page raw: one row per launch, one column per metric, second row = units
        units = rows[1]
        ik = hdr.index("Kernel Name")
        cols = {m: hdr.index(m) for m in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum") if m in hdr}
        sc = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "usecond": 1e3, "msecond": 1e6, "nsecond": 1.0, "second": 1e9}
        for r in rows[2:]:
            vals = []
            for m in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"):
                i = cols.get(m)
                vals.append(float(r[i].replace(",", "")) * sc.get(units[i], 1.0) if i is not None and r[i] else 0.0)
            per.setdefault(r[ik], []).append(tuple(vals))
    return per


def main():
    workload, path = sys.argv[1], sys.argv[2]
    source = sys.argv[sys.argv.index("--source") + 1] if "--source" in sys.argv else os.path.relpath(path, ROOT)
    per = parse(path)
    fam = {}
    for name, ls in per.items():
        for pat, fams in RULES:
            if re.search(pat, name):
                for f in fams:
                    fam.setdefault(f, []).extend(ls)
                break
    out_path = os.path.join(ROOT, "profiles", "r02_traffic.json")
    table = json.load(open(out_path)) if os.path.exists(out_path) else {}
    w = table.setdefault(workload, {})
    for f, ls in sorted(fam.items()):
        n = len(ls)
        w[f] = {"dram_bytes_per_launch": sum(a + b for a, b, _ in ls) / n, "dram_read_per_launch": sum(a for a, _, _ in ls) / n,
                "dram_write_per_launch": sum(b for _, b, _ in ls) / n, "avg_duration_us_under_ncu": sum(c for _, _, c in ls) / n / 1e3,
                "launches_in_capture": n, "source": source}
        print(f"{workload:4s} {f:42s} {n:4d} launches  {w[f]['dram_bytes_per_launch'] / 1e6:9.3f} MB/launch  {w[f]['avg_duration_us_under_ncu']:8.1f} us")
    json.dump(table, open(out_path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
