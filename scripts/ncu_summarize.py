#!/usr/bin/env python
"""Turns `ncu -i X.ncu-rep --page raw --csv` into the short summaries kept under profiles/.

    python scripts/ncu_summarize.py RAW.csv OUT.txt [--traffic rows,dim,queries,k[,world]]

OUT.txt gets, per captured launch, the counters the design argues from (duration, SM clock, tensor-pipe
activity, DRAM / L2 / shared-memory traffic, registers, stall reasons).  With --traffic, the per-kernel
DRAM bytes of the LAST launch of each kernel are merged into profiles/ncu_traffic.json, which bench.py
reads for `roofline.traffic`."""
import csv
import json
import re
import sys
from pathlib import Path

WANT = [
    ("duration_us", r"gpu__time_duration\.sum$"),
    ("sm_clock_mhz", r"sm__cycles_elapsed\.avg\.per_second$"),
    ("dram_bytes_read", r"dram__bytes_read\.sum$"),
    ("dram_bytes_write", r"dram__bytes_write\.sum$"),
    ("dram_throughput_pct", r"dram__throughput\.avg\.pct_of_peak_sustained_elapsed$"),
    ("l2_to_sm_bytes", r"l1tex__m_xbar2l1tex_read_bytes\.sum$"),
    ("lts_throughput_pct", r"lts__throughput\.avg\.pct_of_peak_sustained_elapsed$"),
    ("tensor_pipe_pct", r"sm__pipe_tensor.*cycles_active.*pct_of_peak_sustained_(active|elapsed)$"),
    ("tensor_inst_pct", r"sm__inst_executed_pipe_tensor.*pct"),
    ("issue_active_pct", r"sm__inst_issued\.avg\.pct_of_peak_sustained_active$|smsp__issue_active\.avg\.pct"),
    ("registers_per_thread", r"launch__registers_per_thread$"),
    ("smem_per_block_dyn", r"launch__shared_mem_per_block_dynamic$"),
    ("achieved_occupancy_pct", r"sm__warps_active\.avg\.pct_of_peak_sustained_active$"),
    ("smem_wavefronts", r"l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum$"),
    ("stall_long_scoreboard", r"smsp__average_warp.*issue_stalled_long_scoreboard.*ratio$|smsp__average_warps_issue_stalled_long_scoreboard"),
    ("stall_barrier", r"smsp__average_warps_issue_stalled_barrier"),
    ("stall_membar", r"smsp__average_warps_issue_stalled_membar"),
    ("stall_sleeping", r"smsp__average_warps_issue_stalled_sleeping"),
]


def to_float(x: str):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def main():
    raw, out = sys.argv[1], sys.argv[2]
    traffic = None
    if "--traffic" in sys.argv:
        traffic = [int(v) for v in sys.argv[sys.argv.index("--traffic") + 1].split(",")]
    rows = list(csv.reader(open(raw, newline="")))
    # ncu writes banner lines before the header when the log is shared with the application
    h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    header, units, body = rows[h], rows[h + 1], rows[h + 2:]
    cols = {}
    for name, pat in WANT:
        for j, c in enumerate(header):
            if re.search(pat, c.split(".", 2)[-1] if c.count(".") >= 2 and c.split(".")[1].startswith("Triage") else c):
                cols.setdefault(name, j)
    kcol = header.index("Kernel Name")
    lines, last = [], {}
    for r in body:
        if len(r) <= kcol:
            continue
        kern = re.sub(r"\(.*", "", r[kcol]).split("::")[-1]
        vals = {n: to_float(r[j]) for n, j in cols.items() if j < len(r)}
        unit = {n: units[j] for n, j in cols.items()}
        lines.append(f"{kern}  grid={r[header.index('Grid Size')]} block={r[header.index('Block Size')]}")
        for n, v in vals.items():
            if v is not None:
                lines.append(f"    {n:28s} {v:,.3f} {unit[n]}")
        last[kern] = (vals, unit)
    Path(out).write_text("\n".join(lines) + "\n")
    print("\n".join(lines))
    if traffic:
        p = Path(__file__).resolve().parent.parent / "profiles" / "ncu_traffic.json"
        cur = json.loads(p.read_text()) if p.exists() else []
        for kern, (vals, unit) in last.items():
            def to_bytes(n):
                v, u = vals.get(n), unit.get(n, "byte").lower()
                if v is None:
                    return None
                return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
            rd, wr = to_bytes("dram_bytes_read"), to_bytes("dram_bytes_write")
            if rd is None:
                continue
            e = {"kernel": kern, "rows": traffic[0], "dim": traffic[1], "queries": traffic[2], "k": traffic[3],
                 "world": traffic[4] if len(traffic) > 4 else 1, "dram_bytes_read": rd, "dram_bytes_write": wr or 0.0,
                 "source": Path(out).name}
            cur = [c for c in cur if (c["kernel"], c["rows"], c["dim"], c["queries"], c["k"], c.get("world", 1)) !=
                   (e["kernel"], e["rows"], e["dim"], e["queries"], e["k"], e["world"])]
            cur.append(e)
        p.write_text(json.dumps(cur, indent=1) + "\n")


if __name__ == "__main__":
    main()
