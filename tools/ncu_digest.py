"""
profiles/ncu_<name>.json from an `ncu --set full` raw page (csv) and, if given, the opcode digest
tools/ncu_summary.py printed for the same report: the handful of hardware counters bench.py
quotes beside its algorithmic roofline (it reads the file; nothing is typed into bench.py).

  python tools/ncu_digest.py <name> <raw.csv> <voices> <frames> [summary.txt] [kernel substring]
"""
import csv
import json
import os
import re
import sys

KEYS = {
    "gpu__time_duration.sum": "gpu_time",
    "launch__registers_per_thread": "registers_per_thread",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_slots_busy_pct",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "xu_pipe_pct",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active": "fma_pipe_pct",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active": "fp64_pipe_pct",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "alu_pipe_pct",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active": "lsu_pipe_pct",
    "smsp__inst_executed.sum": "warp_instructions",
    "dram__bytes_read.sum": "dram_bytes_read",
    "dram__bytes_write.sum": "dram_bytes_write",
}
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def main():
    name, raw, voices, frames = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
    summary = sys.argv[5] if len(sys.argv) > 5 else None
    want = sys.argv[6] if len(sys.argv) > 6 else None
    rows = list(csv.reader(open(raw)))
    head, units = rows[0], rows[1]
    col = {c: i for i, c in enumerate(head)}
    vals = next(r for r in rows[2:] if want is None or want in r[col["Kernel Name"]])
    out = {"kernel": vals[col["Kernel Name"]], "voices": voices, "frames": frames,
           "source": "profiles/" + os.path.basename(raw)}
    for k, short in KEYS.items():
        if k in col:
            v, u = float(vals[col[k]]), units[col[k]]
            if short.startswith("dram_bytes"):
                v = int(v * SCALE.get(u, 1.0))
            elif short == "gpu_time":
                short, v = "gpu_time_ms", v * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(u, 1.0)
            out[short] = v
    if summary:
        runs = voices * frames / 512.0                 # warp-level (voice, 16-frame run) units in the launch
        for line in open(summary):
            m = re.match(r"\s+MUFU\.SIN\s+([0-9.]+)", line)
            if m:
                n = float(m.group(1))
                out["mufu_sin_per_voice_sample"] = (n if n < 1e4 else n / runs) / 16.0
        out["opcode_source"] = "profiles/" + os.path.basename(summary)
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_%s.json" % name)
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
