"""One markdown table row per captured launch of an ``.ncu-rep`` file (``ncu -i ... --page raw --csv``).

    python tools/ncu_summary.py gpurun_out/prof_levels.ncu-rep [more.ncu-rep ...] > profiles/r2/ncu_summary.md
"""
import csv
import io
import subprocess
import sys

COLS = [
    ("dur", "gpu__time_duration.sum", 1.0, "{:.3f} ms"),
    ("dram_rd", "dram__bytes_read.sum", 1.0, "{:.3g} GB"),
    ("dram_wr", "dram__bytes_write.sum", 1.0, "{:.3g} GB"),
    ("dram%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1.0, "{:.1f}"),
    ("sm%", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1.0, "{:.1f}"),
    ("issue%", "sm__issue_active.avg.pct_of_peak_sustained_elapsed", 1.0, "{:.1f}"),
    ("alu%", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", 1.0, "{:.1f}"),
    ("lsu wavefronts%", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", 1.0, "{:.1f}"),
    ("fp64%", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", 1.0, "{:.1f}"),
    ("warps%", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0, "{:.1f}"),
    ("regs", "launch__registers_per_thread", 1.0, "{:.0f}"),
    ("grid", "launch__grid_size", 1.0, "{:.0f}"),
    ("block", "launch__block_size", 1.0, "{:.0f}"),
    ("cluster", "launch__cluster_size", 1.0, "{:.0f}"),
]
UNIT_SCALE = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0}


def rows_of(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(io.StringIO(out)))
    if len(rd) < 3:
        return
    head, units = rd[0], rd[1]
    idx = {h: i for i, h in enumerate(head)}
    for r in rd[2:]:
        name = r[idx["Kernel Name"]] if "Kernel Name" in idx else "?"
        cells = []
        for _, key, _, fmt in COLS:
            if key not in idx or not r[idx[key]]:
                cells.append("-")
                continue
            try:
                v = float(r[idx[key]].replace(",", "")) * UNIT_SCALE.get(units[idx[key]], 1.0)
                cells.append(fmt.format(v))
            except ValueError:
                cells.append(r[idx[key]])
        yield name.split("(")[0][:60], cells


def main(paths):
    for p in paths:
        print(f"\n## {p.split('/')[-1]}\n")
        print("| kernel | " + " | ".join(c[0] for c in COLS) + " |")
        print("|---|" + "---|" * len(COLS))
        for name, cells in rows_of(p):
            print(f"| `{name}` | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main(sys.argv[1:])
