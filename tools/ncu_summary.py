"""
Summarises one kernel of an `ncu --set full --import-source on` report on the CPU box:
the headline counters from the raw page and the dynamic opcode mix from the source page
(executed warp instructions per opcode, optionally divided by a unit count such as the
number of warp-level voice-runs in the launch).

  python tools/ncu_summary.py gpurun_out/r1j_bank_fast.ncu-rep [units]
"""
import collections
import csv
import io
import subprocess
import sys

RAW = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
       "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
       "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
       "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
       "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
       "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active",
       "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
       "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
       "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
       "smsp__issue_active.avg.pct_of_peak_sustained_active",
       "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    rows = page(rep, "raw")
    head, units, vals = rows[0], rows[1], rows[-1]
    col = {c: i for i, c in enumerate(head)}
    for vals in rows[2:]:                    # every profiled launch; the stall / opcode digest below is the last one's
        print(vals[col["Kernel Name"]])
        for k in RAW:
            if k in col:
                print("  %-70s %s %s" % (k, vals[col[k]], units[col[k]]))
    stalls = []
    for c, i in col.items():
        if "issue_stalled" in c and c.endswith("per_issue_active.ratio") or ("issue_stalled" in c and "ratio" in c):
            if "not_issued" in c:
                continue
            try:
                stalls.append((float(vals[i]), c.split("issue_stalled_")[1].split("_per_")[0]))
            except ValueError:
                pass
    print("  stalls per issue:", ", ".join("%s %.2f" % (n, v) for v, n in sorted(set(stalls), reverse=True)[:8]))

    rows = page(rep, "source")
    # a report of several kernels lists them one after the other ("Kernel Name" row, header row,
    # instructions): the opcode mix below is of the LAST one (the kernel the capture was made for)
    marks = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    if len(marks) > 1:
        rows = rows[marks[-1]:]
        print("  (opcode mix of the report's last kernel: %s)" % rows[0][1])
    head = rows[1]
    col = {c: i for i, c in enumerate(head)}
    ins = []
    for r in rows[2:]:
        try:
            ins.append((r[col["Source"]], int(r[col["Instructions Executed"]]), int(r[col["# Samples"]])))
        except (ValueError, IndexError):
            pass
    total = sum(x[1] for x in ins)
    units = float(sys.argv[2]) if len(sys.argv) > 2 else None      # e.g. warp-level voice-runs in the launch
    print("  warp instructions executed: %d%s" % (total, " = %.1f per unit" % (total / units) if units else ""))
    ops, smp = collections.Counter(), collections.Counter()
    for src, ex, sm in ins:
        parts = src.split()
        if not parts:
            continue
        op = parts[1] if parts[0].startswith("@") and len(parts) > 1 else parts[0]
        op = op.rstrip(";")
        key = op.split(".")[0]
        if key in ("MUFU", "F2F", "I2F", "F2I"):
            key = ".".join(op.split(".")[:2])
        ops[key] += ex
        smp[key] += sm
    allsmp = sum(smp.values()) or 1
    print("  dynamic opcode mix (executed warp instructions%s, share of stall samples):" % (" per unit" if units else ""))
    for k, v in ops.most_common(28):
        print("    %-10s %10.2f   %5.1f %%" % (k, v / units if units else v, 100.0 * smp[k] / allsmp))
    print("  most-sampled instructions:")
    for src, ex, sm in sorted(ins, key=lambda x: -x[2])[:14]:
        print("    %6d  %s" % (sm, src[:110]))


if __name__ == "__main__":
    main()
