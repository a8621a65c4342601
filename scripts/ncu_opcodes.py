#!/usr/bin/env python
"""Dynamic opcode histogram + stall summary from an ncu report's source page.
   ncu_opcodes.py <report.ncu-rep> [units]   (units = cell-tracer count to normalise by)"""
import csv, subprocess, sys, collections, re
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
units = float(sys.argv[2]) if len(sys.argv) > 2 else None
rows = list(csv.reader(out.splitlines()))
hdr = None; ops = collections.Counter(); stalls = collections.Counter(); samples = 0
for r in rows:
    if len(r) > 5 and r[0] == "Address": hdr = r; continue
    if hdr is None or len(r) < len(hdr) - 2: continue
    src = r[hdr.index("Source")].strip()
    src = re.sub(r"^@!?U?P\d+\s+", "", src)
    op = src.split()[0] if src else "?"
    op = ".".join(op.split(".")[:2]) if op.startswith(("IMAD", "LDG", "STG", "SHFL")) else op.split(".")[0]
    n = int(r[hdr.index("Thread Instructions Executed")] or 0)
    ops[op] += n
    for i, h in enumerate(hdr):
        if h.startswith("stall_") and "Not Issued" not in h:
            try: stalls[h] += int(r[i])
            except ValueError: pass
tot = sum(ops.values())
print("total thread-instructions %.4g%s" % (tot, ("  = %.1f per unit" % (tot / units)) if units else ""))
for op, n in ops.most_common(28):
    print("  %-14s %6.2f %%%s" % (op, 100.0 * n / tot, ("  %7.2f/unit" % (n / units)) if units else ""))
ts = sum(stalls.values())
print("stall samples:", ", ".join("%s %.0f%%" % (k[6:], 100.0 * v / ts) for k, v in stalls.most_common(8)))
