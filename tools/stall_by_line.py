#!/usr/bin/env python3
"""Attribute the warp-stall samples of an ncu capture to source lines of the fused tick.

usage: tools/stall_by_line.py <report.ncu-rep> <object.o> <kernel-mangled-substring> [top]

ncu's `--page source --csv` lists the SASS of the profiled kernel with per-instruction stall samples; `nvdisasm -g` of the
object file lists the same SASS with `//## File "...", line N` markers.  The two listings are zipped by instruction index.
"""
import collections
import csv
import io
import re
import subprocess
import sys


def sass_lines(obj, kernel):
    import os
    obj = os.path.abspath(obj)
    cubin = "/tmp/_stall_by_line.cubin"
    subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd="/tmp", check=True, capture_output=True)
    name = subprocess.run(["cuobjdump", "-lelf", obj], capture_output=True, text=True).stdout.split()[-1]
    subprocess.run(["mv", "/tmp/" + name, cubin], check=True)
    txt = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout
    out, inside, cur = [], False, (None, 0)
    for l in txt.splitlines():
        if l.lstrip().startswith(".section"):
            inside = (".text." in l) and (kernel in l)
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            out.append((cur, m.group(2).strip()))
    return out


def main():
    rep, obj, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    sass = sass_lines(obj, kernel)
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    h = rows[hi]
    col = {n: i for i, n in enumerate(h)}
    body = [r for r in rows[hi + 1:] if len(r) == len(h)]
    if len(body) != len(sass):
        print("warning: %d instructions in the report, %d in the object (stale build?)" % (len(body), len(sass)), file=sys.stderr)
    stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
    per_line = collections.defaultdict(lambda: collections.Counter())
    per_op = collections.defaultdict(lambda: collections.Counter())
    tot = collections.Counter()
    for (loc, ins), r in zip(sass, body):
        op = ins.split()[1] if ins.startswith("@") else ins.split()[0]
        n = int(r[col["# Samples"]] or 0)
        per_line[loc]["samples"] += n
        per_line[loc]["inst"] += int(r[col["Instructions Executed"]] or 0)
        per_op[op.split(".")[0]]["samples"] += n
        per_op[op.split(".")[0]]["inst"] += int(r[col["Instructions Executed"]] or 0)
        for s in stalls:
            v = int(r[col[s]] or 0)
            per_line[loc][s] += v
            per_op[op.split(".")[0]][s] += v
            tot[s] += v
    total = sum(c["samples"] for c in per_line.values())
    print("total samples", total, dict(tot.most_common(8)))
    print("\n-- by source line (file:line samples% inst | top stalls)")
    for loc, c in sorted(per_line.items(), key=lambda kv: -kv[1]["samples"])[:top]:
        st = ", ".join("%s %d" % (k[6:], v) for k, v in collections.Counter({k: c[k] for k in stalls}).most_common(3) if v)
        print("%-28s %5.1f%% %8d | %s" % ("%s:%d" % loc, 100.0 * c["samples"] / max(total, 1), c["inst"], st))
    print("\n-- by opcode")
    for op, c in sorted(per_op.items(), key=lambda kv: -kv[1]["samples"])[:16]:
        st = ", ".join("%s %d" % (k[6:], v) for k, v in collections.Counter({k: c[k] for k in stalls}).most_common(3) if v)
        print("%-10s %5.1f%% %9d | %s" % (op, 100.0 * c["samples"] / max(total, 1), c["inst"], st))
    # phase table: cumulative samples in source-line order of this kernel's own file
    print("\n-- cumulative by line range of osc_fused2.cuh (20-line buckets)")
    bucket = collections.Counter()
    for (f, ln), c in per_line.items():
        bucket[(f, ln // 20 * 20)] += c["samples"]
    for (f, ln), v in sorted(bucket.items()):
        if v * 100 >= total:
            print("%-24s %4d-%4d %5.1f%%" % (f, ln, ln + 19, 100.0 * v / total))


if __name__ == "__main__":
    main()
