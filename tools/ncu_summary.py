"""ncu full capture (.ncu-rep) -> a short per-kernel summary of the metrics the roofline argument uses (B200_PROFILING.md recipe):
    python tools/ncu_summary.py gpurun_out/x.ncu-rep [out.txt] [--all]
One block per distinct (kernel, grid) — the mean over its launches in the capture; --all prints every launch."""
import collections
import csv
import subprocess
import sys

KEYS = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
        ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
        ("sm__inst_executed_pipe_tensor.sum", "tensor instructions"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active % (occupancy)"),
        ("launch__registers_per_thread", "registers / thread"), ("launch__shared_mem_per_block_dynamic", "dynamic smem / block"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_scoreboard"),
        ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_scoreboard"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts")]


def main():
    rep = sys.argv[1]
    out_path = next((a for a in sys.argv[2:] if not a.startswith("--")), None)
    every = "--all" in sys.argv
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(l for l in txt.splitlines() if l.startswith('"')))
    hdr, units = rows[0], rows[1]
    ik, ig, ib = hdr.index("Kernel Name"), hdr.index("Grid Size"), hdr.index("Block Size")
    groups = collections.OrderedDict()
    for n, r in enumerate(rows[2:]):
        name = r[ik].replace("dtb::", "").split("(")[0].replace("void ", "")
        groups.setdefault((name, r[ig], r[ib]) + ((n,) if every else ()), []).append(r)
    lines = [f"# {rep}: {len(rows) - 2} profiled launches, ncu --set full --clock-control none (cold caches, serialised: read fractions, not absolutes)"]
    for key, rs in groups.items():
        lines.append(f"## {key[0]}  grid {key[1]} x block {key[2]}  ({len(rs)} launch{'es' if len(rs) > 1 else ''})")
        for m, label in KEYS:
            if m not in hdr:
                continue
            i = hdr.index(m)
            vals = []
            for r in rs:
                try:
                    vals.append(float(r[i].replace(",", "")))
                except ValueError:
                    pass
            if vals:
                lines.append(f"   {label:34s} {sum(vals) / len(vals):14.3f} {units[i]}")
    text = "\n".join(lines) + "\n"
    print(text)
    if out_path:
        open(out_path, "w").write(text)


if __name__ == "__main__":
    main()
