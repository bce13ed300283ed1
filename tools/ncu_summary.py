#!/usr/bin/env python3
"""Key figures of one `ncu --set full` capture (the numbers profiles/README.md quotes), as text.

usage: tools/ncu_summary.py <report.ncu-rep> [--json]
"""
import csv
import io
import json
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_static", "static smem/block"),
    ("launch__occupancy_limit_registers", "occupancy limit (registers), blocks/SM"),
    ("launch__occupancy_limit_shared_mem", "occupancy limit (smem), blocks/SM"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__warps_active.avg.per_cycle_active", "warps active / scheduler"),
    ("smsp__issue_active.avg.per_cycle_active", "issue slots busy / cycle / scheduler"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "FP64 pipe % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "threads / instruction"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
    ("smsp__inst_executed_op_local_ld.sum", "local loads (warp inst)"),
    ("smsp__inst_executed_op_local_st.sum", "local stores (warp inst)"),
    ("launch__stack_size", "stack bytes"),
]


def main():
    rep = sys.argv[1]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    h, u, v = rows[0], rows[1], rows[2]
    d = {k: (val, un) for k, un, val in zip(h, u, v)}
    out = {"kernel": d.get("Kernel Name", ("?",))[0]}
    for k, label in KEYS:
        if k in d:
            out[label] = "%s %s" % d[k]
    stalls = {k.replace("smsp__pcsamp_warps_issue_stalled_", ""): int(float(d[k][0].replace(",", "")))
              for k in d if k.startswith("smsp__pcsamp_warps_issue_stalled_") and "not_issued" not in k}
    tot = sum(stalls.values()) or 1
    out["stall samples"] = {k: "%d (%.0f%%)" % (n, 100.0 * n / tot) for k, n in sorted(stalls.items(), key=lambda kv: -kv[1]) if n * 50 >= tot}
    # executed FP64 work from the report's own SASS listing (thread instructions of DFMA/DADD/DMUL; an FMA counts 2 FLOP)
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    srows = list(csv.reader(io.StringIO(src)))
    hi = next((i for i, r in enumerate(srows) if r and r[0] == "Address"), None)
    if hi is not None:
        col = {n: i for i, n in enumerate(srows[hi])}
        ops = {"DFMA": 0, "DADD": 0, "DMUL": 0}
        n_inst = 0
        for r in srows[hi + 1:]:
            if len(r) != len(srows[hi]):
                continue
            n_inst += 1
            ins = r[col["Source"]].split()
            op = (ins[1] if ins and ins[0].startswith("@") and len(ins) > 1 else (ins[0] if ins else "")).split(".")[0]
            if op in ops:
                ops[op] += int(r[col["Thread Instructions Executed"]] or 0)
        out["SASS instructions in the kernel"] = n_inst
        out["FP64 thread instructions"] = ops
        out["executed FP64 FLOP per launch"] = 2 * ops["DFMA"] + ops["DADD"] + ops["DMUL"]
    if "--json" in sys.argv:
        print(json.dumps(out, indent=1))
    else:
        for k, val in out.items():
            print("%-44s %s" % (k, val))


if __name__ == "__main__":
    main()
