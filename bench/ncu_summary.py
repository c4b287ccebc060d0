#!/usr/bin/env python
"""Summarise an `ncu --set full` report into markdown + keep its raw metric page.

    python bench/ncu_summary.py gpurun_out/lenet_full.ncu-rep profiles/ncu_lenet.md "title"

Writes the table to the .md and the full `--page raw --csv` dump next to it as .raw.csv.gz (the .ncu-rep
itself is tens of MB and stays in gpurun_out/).  Runs on the CPU box (ncu -i needs no GPU).
"""
from __future__ import annotations

import csv
import gzip
import io
import re
import subprocess
import sys

COLS = [
    ("time us", "gpu__time_duration.sum", 1.0),
    ("grid", "launch__grid_size", None),
    ("block", "launch__block_size", None),
    ("regs", "launch__registers_per_thread", None),
    ("dyn smem KB", "launch__shared_mem_per_block_dynamic", None),
    ("waves/SM", "launch__waves_per_multiprocessor", None),
    ("SM busy %", "sm__throughput.avg.pct_of_peak_sustained_elapsed", None),
    ("tensor pipe % (active)", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", None),
    ("warps active %", "sm__warps_active.avg.pct_of_peak_sustained_active", None),
    ("DRAM %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", None),
    ("DRAM rd MB", "dram__bytes_read.sum", None),
    ("DRAM wr MB", "dram__bytes_write.sum", None),
    ("L2 hit %", "lts__t_sector_hit_rate.pct", None),
    ("issue slots busy %", "sm__inst_issued.avg.pct_of_peak_sustained_active", None),
    ("inst (M)", "smsp__inst_executed.sum", None),
]


def to_unit(val: str, unit: str, want: str) -> str:
    try:
        v = float(val.replace(",", ""))
    except ValueError:
        return val
    scale = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}
    if want == "time us":
        return f"{v * scale.get(unit, 1.0):.2f}"
    if want.endswith("MB"):
        b = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
        return f"{v * b / 1e6:.2f}"
    if want.endswith("KB"):
        b = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6}.get(unit, 1)
        return f"{v * b / 1e3:.1f}"
    if want == "inst (M)":
        return f"{v / 1e6:.2f}"
    return f"{v:.2f}" if "." in val else val


def main(rep: str, out_md: str, title: str) -> int:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    with gzip.GzipFile(re.sub(r"\.md$", ".raw.csv.gz", out_md), "wb", mtime=0) as f:
        f.write(raw.encode())
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(out_md, "w") as f:
        f.write(f"# {title}\n\n`ncu --set full --clock-control none --import-source on`, one launch of each kernel of one "
                f"training step (after warm-up). Source report: `{rep}`; every metric of the report: "
                f"`{re.sub(r'.*/', '', re.sub(r'[.]md$', '.raw.csv.gz', out_md))}`. Durations under ncu are serialised and "
                "cold-cache: compare shares, not absolutes.\n\n")
        names = [c[0] for c in COLS if c[1] in idx]
        f.write("| # | kernel | " + " | ".join(names) + " |\n|---:|---|" + "---:|" * len(names) + "\n")
        for n, d in enumerate(data):
            k = re.sub(r"\(.*", "", d[idx["Kernel Name"]]).replace("void ", "")
            cells = [to_unit(d[idx[m]], units[idx[m]], lab) for lab, m, _ in COLS if m in idx]
            f.write(f"| {n} | `{k}` | " + " | ".join(cells) + " |\n")
    print(f"{len(data)} kernels -> {out_md}")
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "ncu summary"))
