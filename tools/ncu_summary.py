#!/usr/bin/env python
"""Summarise an `ncu --set full` report (read here, no GPU needed) into a markdown table for profiles/.

    python tools/ncu_summary.py gpurun_out/X.ncu-rep profiles/X_summary.md "title"
"""
import csv
import io
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1 %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid")]


def to_bytes(v, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1)
    return float(v.replace(",", "")) * scale


def to_us(v, unit):
    scale = {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(unit, 1)
    return float(v.replace(",", "")) * scale


def main():
    rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    lines = [f"# {title}", "", "`ncu --set full --clock-control none --import-source on`; one row per captured launch. "
             "`traffic` = dram read + write per launch.", "",
             "| kernel | " + " | ".join(n for _, n in COLS) + " | traffic MB | achieved DRAM GB/s |",
             "|---|" + "---:|" * (len(COLS) + 2)]
    for r in data:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("gmg::", "")
        cells = []
        t_us = to_us(r[idx["gpu__time_duration.sum"]], units[idx["gpu__time_duration.sum"]])
        rd = to_bytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]])
        wr = to_bytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]])
        for key, _ in COLS:
            v = r[idx[key]]
            if key == "gpu__time_duration.sum":
                cells.append(f"{t_us:.1f} us")
            elif key.startswith("dram__bytes"):
                cells.append(f"{to_bytes(v, units[idx[key]]) / 1e6:.1f} MB")
            else:
                try:
                    cells.append(f"{float(v.replace(',', '')):.1f}" if "." in v else v)
                except ValueError:
                    cells.append(v)
        lines.append(f"| `{name}` | " + " | ".join(cells) + f" | {(rd + wr) / 1e6:.1f} | {(rd + wr) / t_us / 1e3:.0f} |")
    with open(out, "w") as f:
        f.write("\n".join(lines) + "\n")
    if len(sys.argv) > 4:      # also a machine-readable copy: kernel -> dram bytes per launch (bench.py's roofline.traffic)
        import json
        js = {}
        for r in data:
            name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("gmg::", "")
            rd = to_bytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]])
            wr = to_bytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]])
            if name in js:      # several launches of one kernel: keep the first (level 0, the largest)
                continue
            js[name] = {"dram_bytes": int(rd + wr),
                        "time_us": to_us(r[idx["gpu__time_duration.sum"]], units[idx["gpu__time_duration.sum"]])}
        with open(sys.argv[4], "w") as f:
            json.dump({"capture": title, "kernels": js}, f, indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
