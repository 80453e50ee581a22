"""Summarise an ncu --set full report (read on the CPU box) into a small CSV/markdown table for profiles/.
    python tools/ncu_summary.py gpurun_out/r01_full.ncu-rep profiles/r01_ncu_full_summary.md
"""
import csv, io, subprocess, sys

WANT = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("launch__grid_size", "grid")]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    names = [(m, s) for m, s in WANT if m in col]
    lines = ["| kernel | " + " | ".join(s for _, s in names) + " |", "|---|" + "---|" * len(names)]
    for r in data:
        k = r[col["Kernel Name"]].split("(")[0].split("::")[-1]
        cells = []
        for m, s in names:
            v, u = r[col[m]], units[col[m]]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:.1f}" if "%" in s or s == "time" else (f"{f:.1f}" if "byte" in u.lower() else f"{f:g}")
            except ValueError:
                pass
            cells.append(f"{v} {u}" if s in ("time", "dram_rd", "dram_wr") else v)
        lines.append(f"| {k} | " + " | ".join(cells) + " |")
    open(out, "w").write(f"ncu --set full --clock-control none summary of {rep} (one step of bench.py config 3)\n\n"
                         + "\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])


# ---- profiles/traffic.json: measured DRAM bytes per launch of each bench stage -----------------
STAGE_OF = [("preprocess_backward", "preprocess_backward"), ("preprocess_", "preprocess"),
            ("depth_hist_kernel", "depth_sort"), ("radix_pass_kernel<unsigned int, 1", "depth_sort"),
            ("radix_pass_kernel<unsigned int, 0, 8", "depth_sort"), ("emit_scan", "emit_keys"),
            ("tile_offsets", "tile_ranges"), ("radix_pass_kernel<unsigned short", "radix_sort"),
            ("radix_pass_kernel<unsigned int", "radix_sort"),
            ("blend_forward", "blend_forward"), ("blend_backward", "blend_backward")]


def traffic(rep, out):
    """Sum dram__bytes_read.sum + dram__bytes_write.sum over the kernels of each stage of ONE step."""
    import json
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    acc = {}
    for r in data:
        name = r[col["Kernel Name"]]
        stage = next((s for pat, s in STAGE_OF if pat in name), None)
        if stage is None:
            continue
        b = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            b += float(r[col[m]].replace(",", "")) * scale[units[col[m]]]
        acc[stage] = acc.get(stage, 0.0) + b
    json.dump({k: round(v) for k, v in acc.items()}, open(out, "w"), indent=1)
    print(acc)


if __name__ == "__main__" and len(sys.argv) > 3 and sys.argv[3] == "traffic":
    traffic(sys.argv[1], sys.argv[4])
