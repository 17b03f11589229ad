"""Summarise an `ncu --set full` report into profiles/:
python scripts/ncu_summary.py <rep> <tag> [--traffic <workload>]
Writes profiles/<tag>_summary.md and, with --traffic, the measured DRAM bytes per launch of the
decode kernel on bench workload <workload> (1 GPU) into profiles/decode_kernel_traffic.json, from
where bench.py reads `roofline.traffic`."""
import csv
import io
import json
import subprocess
import sys
from pathlib import Path

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
    "launch__shared_mem_per_block_dynamic", "sm__inst_executed_pipe_tensor.sum",
    "smsp__inst_executed.sum",
]
TO_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    rep, tag = sys.argv[1], sys.argv[2]
    root = Path(__file__).resolve().parent.parent
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_i = hdr.index("Kernel Name")
    lines = [f"# ncu --set full summary: {tag}", "",
             f"source report: `{Path(rep).name}` (gpurun_out/, scratch); command: see profiles/README.md", ""]
    traffic = []
    for r in data:
        lines.append(f"## {r[name_i]}")
        lines.append("")
        lines.append("| metric | value | unit |")
        lines.append("|---|---|---|")
        rd = wr = 0.0
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                lines.append(f"| {k} | {r[i]} | {units[i]} |")
                if k == "dram__bytes_read.sum":
                    rd = float(r[i]) * TO_BYTES.get(units[i], 1)
                if k == "dram__bytes_write.sum":
                    wr = float(r[i]) * TO_BYTES.get(units[i], 1)
        lines.append("")
        traffic.append(rd + wr)
    out = root / "profiles" / f"{tag}_summary.md"
    out.write_text("\n".join(lines) + "\n")
    print(f"wrote {out}")
    if "--traffic" in sys.argv and traffic:
        wl = sys.argv[sys.argv.index("--traffic") + 1]
        t = sum(traffic) / len(traffic)
        path = root / "profiles" / "decode_kernel_traffic.json"
        table = json.loads(path.read_text()) if path.exists() else {}
        table[wl] = {"dram_bytes_per_launch": t, "from": out.name, "launches_averaged": len(traffic)}
        path.write_text(json.dumps(table, indent=1, sort_keys=True) + "\n")
        print(f"{wl}: traffic per launch {t:.0f} B")


if __name__ == "__main__":
    main()
