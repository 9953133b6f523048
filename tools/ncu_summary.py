#!/usr/bin/env python
"""Turns ncu outputs brought back in gpurun_out/ into the small text summaries committed under profiles/.

    python tools/ncu_summary.py launches gpurun_out/<launches>.csv          > profiles/<name>.md
    python tools/ncu_summary.py full     gpurun_out/<report>.ncu-rep [regex] > profiles/<name>.md
"""
import collections
import csv
import re
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
           "launch__block_size", "launch__shared_mem_per_block_dynamic",
           "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum"]


def to_ns(v, u):
    v = float(v.replace(",", ""))
    return v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9, "second": 1e9}.get(u, 1)


def launches(path):
    lines = [l for l in open(path) if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for row in r:
        a = agg.setdefault(row[ki], [0, 0.0])
        a[0] += 1
        a[1] += to_ns(row[vi], row[ui])
    tot = sum(a[1] for a in agg.values())
    # this library's kernels are the only CamelCase "...Kernel" names in the process (torch's end in "_kernel");
    # ncu prints anonymous-namespace kernels without the vktb:: prefix
    is_ours = lambda k: "vktb::" in k or re.search(r"[a-z0-9]Kernel\b", k) is not None  # noqa: E731
    ours = sum(a[1] for k, a in agg.items() if is_ours(k))
    print(f"# ncu launch list: {path}\n")
    print("`ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare SHARES, not absolutes)\n")
    print(f"total {tot / 1e6:.3f} ms over {sum(a[0] for a in agg.values())} launches; vktb:: kernels {ours / 1e6:.3f} ms "
          f"({100 * ours / tot:.1f} %); anything else is torch input synthesis outside the timed regions (excluded when the "
          f"capture was filtered with -k regex:Kernel)\n")
    print("| launches | total ms | avg ms | share | share of vktb | kernel |\n|---:|---:|---:|---:|---:|---|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        mine = is_ours(k)
        print(f"| {a[0]} | {a[1] / 1e6:.3f} | {a[1] / a[0] / 1e6:.4f} | {100 * a[1] / tot:.1f} % | "
              f"{(f'{100 * a[1] / ours:.1f} %') if mine else '-'} | `{k[:120]}` |")


def full(path, pattern=None):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    cols = [(m, hdr.index(m)) for m in METRICS if m in hdr]
    ki = hdr.index("Kernel Name")
    print(f"# ncu --set full: {path}\n")
    print("per launch (cold cache, clocks not controlled); traffic = dram__bytes_read.sum + dram__bytes_write.sum\n")
    print("| kernel | " + " | ".join(f"{m} [{units[i]}]" for m, i in cols) + " |")
    print("|---|" + "---:|" * len(cols))
    for row in rows[2:]:
        if pattern and not re.search(pattern, row[ki]):
            continue
        print(f"| `{row[ki][:70]}` | " + " | ".join(row[i] for _, i in cols) + " |")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        full(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
