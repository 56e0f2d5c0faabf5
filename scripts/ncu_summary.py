"""Summarise ncu reports (gpurun_out/prof_*.ncu-rep) into profiles/ncu_summary.md."""
import csv
import io
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("sm__cycles_elapsed.avg.per_second", "SM clock"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2->SM bytes"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
]


def read(rep):
    out = subprocess.run(["ncu", "-i", str(rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if len(rows) < 3:
        return None
    hdr, units, vals = rows[0], rows[1], rows[-1]
    d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    return d


def main():
    src = Path(sys.argv[1]) if len(sys.argv) > 1 else ROOT / "gpurun_out"
    dst = Path(sys.argv[2]) if len(sys.argv) > 2 else ROOT / "profiles" / "ncu_summary.md"
    reps = sorted(src.glob("prof_*.ncu-rep"))
    lines = ["# ncu summaries (ncu --set full --clock-control none, one instance per kernel)", "",
             f"Measured peaks used for fractions: HBM {peaks.get('hbm_gbs')} GB/s, bf16 "
             f"{peaks.get('bf16_tflops')} TFLOP/s burst / {peaks.get('bf16_tflops_sustained')} sustained.", ""]
    for rep in reps:
        d = read(rep)
        if d is None:
            continue
        name = d.get("Kernel Name", ("?", ""))[0]
        lines += [f"## {rep.stem}", "", f"`{name[:140]}`", "", "| metric | value |", "|---|---|"]
        for key, label in WANT:
            if key in d:
                v, u = d[key]
                lines.append(f"| {label} | {v} {u} |")
        lines.append("")
    dst.parent.mkdir(parents=True, exist_ok=True)
    dst.write_text("\n".join(lines))
    print("\n".join(lines))


if __name__ == "__main__":
    main()
