"""Turns ncu reports into the tracked summaries: a per-kernel table (markdown) and profiles-style traffic.json.

    python scripts/ncu_summary.py OUT.md TRAFFIC.json  key=report.ncu-rep [key=report.ncu-rep ...]

`key` is "<workload>:<grad_mode>" as bench.py looks it up; the traffic figure is dram__bytes_read.sum +
dram__bytes_write.sum of the fused update kernel, per launch (mean over the captured launches).
Needs only `ncu -i` (no GPU)."""
import csv
import io
import json
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU %"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU %"),
    ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
]
TO_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
TO_US = {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}


def rows_of(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    res = []
    for r in rows[2:]:
        d = {"name": r[idx["Kernel Name"]]}
        for m, _ in WANT:
            if m in idx:
                v, u = r[idx[m]].replace(",", ""), units[idx[m]]
                try:
                    v = float(v)
                except ValueError:
                    continue
                if "bytes" in m:
                    v *= TO_BYTES.get(u, 1)
                if m.startswith("gpu__time"):
                    v *= TO_US.get(u, 1)
                d[m] = v
        res.append(d)
    return res


def main():
    out_md, out_json = sys.argv[1], sys.argv[2]
    traffic, md = {}, ["| capture | kernel | " + " | ".join(t for _, t in WANT) + " |", "|---|---|" + "---|" * len(WANT)]
    for spec in sys.argv[3:]:
        key, rep = spec.split("=", 1)
        rows = rows_of(rep)
        fused = [r for r in rows if "fused_update_kernel" in r["name"] or "reduce_update_kernel" in r["name"]]
        if fused:
            traffic[key] = sum(r["dram__bytes_read.sum"] + r["dram__bytes_write.sum"] for r in fused) / len(fused)
        seen = set()
        for r in rows:
            short = r["name"].split("(")[0].replace("void ", "").replace("dpk::", "")
            if short in seen:
                continue
            seen.add(short)
            cells = []
            for m, _ in WANT:
                v = r.get(m)
                if v is None:
                    cells.append("")
                elif "bytes" in m:
                    cells.append(f"{v / 1e6:.1f} MB")
                elif m.startswith("gpu__time"):
                    cells.append(f"{v:.1f} us")
                elif m.startswith("launch__"):
                    cells.append(f"{int(v)}")
                else:
                    cells.append(f"{v:.1f}")
            md.append(f"| {key} | `{short}` | " + " | ".join(cells) + " |")
    with open(out_md, "w") as f:
        f.write("\n".join(md) + "\n")
    with open(out_json, "w") as f:
        json.dump(traffic, f, indent=1)
    print("\n".join(md))
    print(json.dumps(traffic))


if __name__ == "__main__":
    main()
