#!/usr/bin/env python3
"""Summarise `ncu -i X.ncu-rep --page source --csv` of one scatter-pass launch.

usage: ncu_source_summary.py src.csv [keys_per_launch] [top_n]
Prints (1) shared/L1 wavefronts and L1 tag requests per warp-row (32 keys) by opcode,
(2) the warp-stall sampling breakdown, (3) the hottest SASS lines with their main stall reasons.
"""
import csv
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    nkeys = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 28
    topn = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    rows = list(csv.reader(open(path)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    col = {n: i for i, n in enumerate(hdr)}
    body = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]
    nrows = nkeys / 32.0

    def num(r, name):
        try:
            return float(r[col[name]])
        except (ValueError, KeyError):
            return 0.0

    by_op = defaultdict(lambda: [0.0, 0.0, 0.0, 0.0])  # instrs, smem wf, smem ideal, tag req
    total_instr = 0.0
    for r in body:
        sass = r[col["Source"]].strip()
        toks = sass.split()
        if not toks:
            continue
        op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
        op = op.split(".")[0]
        ex = num(r, "Instructions Executed")
        total_instr += ex
        if op in ("LDS", "STS", "ATOMS", "LDG", "STG", "ATOMG", "RED", "ST", "LD", "LDSM", "UBLKCP", "REDG", "LDL", "STL"):
            b = by_op[op]
            b[0] += ex
            b[1] += num(r, "L1 Wavefronts Shared")
            b[2] += num(r, "L1 Wavefronts Shared Ideal")
            b[3] += num(r, "L1 Tag Requests Global")
    tot = 0.0
    for op, b in sorted(by_op.items()):
        wf = b[1] + b[3]
        tot += wf
        print("%-7s instr/row %6.2f  smem wf/row %6.2f (ideal %5.2f)  wf/instr %5.2f  global tag req/row %5.2f" % (
            op, b[0] / nrows, b[1] / nrows, b[2] / nrows, (b[1] / b[0]) if b[0] else 0, b[3] / nrows))
    print("total LSU wavefronts+tags per row %.2f ; warp instrs %.0f = %.1f per row" % (tot / nrows, total_instr, total_instr / nrows))

    stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
    tot_s = {n: sum(num(r, n) for r in body) for n in stall_cols}
    all_s = sum(tot_s.values()) or 1.0
    print("stall samples %d:" % all_s, {k: round(100 * v / all_s, 1) for k, v in sorted(tot_s.items(), key=lambda kv: -kv[1]) if v / all_s > 0.004})
    hot = sorted(body, key=lambda r: -num(r, "# Samples"))[:topn]
    hot.sort(key=lambda r: body.index(r))
    for r in hot:
        reasons = {n[6:]: int(num(r, n)) for n in stall_cols if num(r, n) > 0.15 * max(1.0, num(r, "# Samples"))}
        print("%5d %-52s %6d %9d %s" % (body.index(r), r[col["Source"]].strip()[:52], num(r, "# Samples"), num(r, "Instructions Executed"), reasons))


if __name__ == "__main__":
    main()
