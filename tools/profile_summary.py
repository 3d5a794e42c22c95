#!/usr/bin/env python
"""Turns an ncu launch list (--metrics gpu__time_duration.sum --csv) and a raw-page CSV of an `ncu --set full`
capture into the markdown summary kept under profiles/.
    python tools/profile_summary.py launches.csv raw.csv > profiles/rNN_summary.md
"""
import collections
import csv
import sys


def launch_table(path):
    rows = list(csv.reader(open(path)))
    hdr = None
    agg = collections.OrderedDict()
    for r in rows:
        if len(r) > 5 and r[0] == "ID":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            if d.get("Metric Name") == "gpu__time_duration.sum":
                v = float(d["Metric Value"].replace(",", ""))
                u = d["Metric Unit"]
                v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
                key = d["Kernel Name"][:90]
                a = agg.setdefault(key, [0, 0.0])
                a[0] += 1
                a[1] += v
    tot = sum(v[1] for v in agg.values())
    out = ["| kernel | launches | total us | share | avg us |", "|---|---:|---:|---:|---:|"]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append("| `%s` | %d | %.1f | %.1f%% | %.1f |" % (k, v[0], v[1], 100 * v[1] / tot, v[1] / v[0]))
    return "\n".join(out)


KEYS = [
    ("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of peak"),
    ("sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active", "DMMA pipe active %"),
    ("sm__ops_path_tensor_src_fp64.avg.pct_of_peak_sustained_elapsed", "FP64 tensor ops % of peak"),
    ("launch__registers_per_thread", "registers/thread"), ("launch__block_size", "threads/CTA"), ("launch__grid_size", "CTAs"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/CTA"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math_pipe_throttle"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_scoreboard"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_scoreboard"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
]


def raw_tables(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    seen, out = set(), []
    for r in rows[2:]:
        name = r[idx["Kernel Name"]]
        if name in seen:
            continue
        seen.add(name)
        out.append("### `%s`\n\n| metric | value |\n|---|---:|" % name[:90])
        for k, label in KEYS:
            if k in idx:
                out.append("| %s | %s %s |" % (label, r[idx[k]], units[idx[k]]))
        out.append("")
    return "\n".join(out)


if __name__ == "__main__":
    print("## Launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`), whole process incl. data set-up\n")
    print(launch_table(sys.argv[1]))
    print("\n## `ncu --set full --clock-control none --import-source on`, one launch per kernel\n")
    for p in sys.argv[2:]:
        print(raw_tables(p))
