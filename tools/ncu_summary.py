"""Markdown summary of ncu --set full captures: python tools/ncu_summary.py out.md rep1.ncu-rep [rep2.ncu-rep ...]
One block per captured launch: duration, occupancy, issue utilisation, DRAM bytes, cache hit rates, registers, and the
warp-stall breakdown (warps stalled per issue slot, the reasons above 0.3)."""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration", "", 1.0),
    ("launch__grid_size", "grid", "", 1.0),
    ("launch__block_size", "block", "", 1.0),
    ("launch__registers_per_thread", "registers/thread", "", 1.0),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "sm__warps_active", "% of peak (while active)", 1.0),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__issue_active", "% (while active)", 1.0),
    ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__issue_active", "% (of elapsed)", 1.0),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput", "% of peak", 1.0),
    ("smsp__inst_executed.sum", "warp instructions", "", 1.0),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "active threads / instruction", "", 1.0),
    ("dram__bytes_read.sum", "dram read", "", 1.0),
    ("dram__bytes_write.sum", "dram write", "", 1.0),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate", "%", 1.0),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit rate", "%", 1.0),
    ("sm__cycles_active.avg", "SM cycles active (avg)", "", 1.0),
    ("sm__cycles_elapsed.avg", "SM cycles elapsed", "", 1.0),
]


def rows_of(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    return rows[0], rows[1], rows[2:]


def main():
    out = ["# ncu --set full summaries (clock-control none; every launch is profiled with flushed caches and serialised, so the",
           "# absolute durations are cold-cache ones: compare shares, not microseconds, with the CUDA-event times of bench.py)", ""]
    for rep in sys.argv[2:]:
        H, U, rows = rows_of(rep)
        out.append(f"## {rep.split('/')[-1]}")
        for r in rows:
            name = r[H.index("Kernel Name")]
            out.append(f"### `{name[:110]}`")
            for key, label, unit, _ in KEYS:
                if key in H and r[H.index(key)] not in ("", "n/a"):
                    u = U[H.index(key)] if not unit else unit
                    out.append(f"* {label}: {r[H.index(key)]} {u}")
            st = []
            for i, h in enumerate(H):
                if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
                    try:
                        v = float(r[i])
                    except ValueError:
                        continue
                    if v >= 0.3:
                        st.append((v, h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
            out.append("* warps stalled per issue slot: " + ", ".join(f"{n} {v:.2f}" for v, n in sorted(st, reverse=True)))
            out.append("")
    open(sys.argv[1], "w").write("\n".join(out))
    print("wrote", sys.argv[1])


if __name__ == "__main__":
    main()
