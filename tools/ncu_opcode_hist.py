#!/usr/bin/env python
"""Dynamic opcode histogram of one kernel from an ncu report captured with --import-source on: executed thread instructions
per SASS opcode (and per key with --keys), grouped by the pipe they issue on. `ncu -i REP --page source --csv` is the input."""
import argparse
import collections
import csv
import io
import re
import subprocess

ALU = ("IADD3", "LOP3", "SHF", "PRMT", "SEL", "LEA", "ISETP", "PLOP3", "VIADD", "IABS", "FLO", "POPC", "BREV", "ISCADD", "MOV", "VIMNMX", "IMNMX", "SGXT", "BMSK", "I2I", "CS2R", "P2R", "R2P", "VABSDIFF")
FMA = ("IMAD", "IMUL", "FFMA", "FMUL", "FADD")


def pipe_of(op):
    base = op.split(".")[0]
    if base in ("IMAD",):
        return "fma"
    if base in ALU:
        return "alu"
    if base in ("LDG", "STG", "LDS", "STS", "LDL", "STL", "LD", "ST", "ATOMG", "ATOM", "RED", "LDC", "LDCU", "ATOMS", "MATCH", "VOTE", "SHFL", "BAR", "S2R", "S2UR", "REDUX"):
        return "lsu/other"
    if base in ("BRA", "BSSY", "BSYNC", "EXIT", "CALL", "RET", "WARPSYNC", "NOP", "BREAK", "JMP"):
        return "branch"
    if base.startswith("U"):
        return "uniform"
    return "other"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--keys", type=int, default=0)
    ap.add_argument("--top", type=int, default=25)
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    ci, si = hdr.index("Thread Instructions Executed"), hdr.index("Source")
    ops, pipes = collections.Counter(), collections.Counter()
    for r in rows[hdr_i + 1:]:
        if len(r) <= ci:
            continue
        m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[si])
        if not m:
            continue
        try:
            n = int(r[ci])
        except ValueError:
            continue
        op = m.group(1)
        ops[op] += n
        pipes[pipe_of(op)] += n
    tot = sum(ops.values())
    div = a.keys or 1
    print("%s: %d thread instructions%s" % (rows[0][1] if rows[0] else "", tot, " (%.1f per key)" % (tot / div) if a.keys else ""))
    for p, n in pipes.most_common():
        print("  pipe %-10s %8.1f %5.1f%%" % (p, n / div, 100.0 * n / tot))
    for op, n in ops.most_common(a.top):
        print("  %-22s %8.1f %5.1f%%  [%s]" % (op, n / div, 100.0 * n / tot, pipe_of(op)))


if __name__ == "__main__":
    main()
