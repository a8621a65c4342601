"""Offline instruction-mix of the hot loop of a kernel, from cuobjdump -sass (no GPU needed).

    python scripts/sass_loop.py <lib.so> <mangled-name-substring> [--all]

Finds the largest backward branch (loop) of the function and prints the opcode histogram of
its body (static count per iteration); --all prints the histogram of the whole function.
The static count per loop iteration / cells per iteration is the issue-slot budget per
cell*tracer that DESIGN.md section 4 argues about (FP64 instructions hold their pipe 2 cycles
per warp on B200, everything else 1 issue slot).
"""
import collections
import re
import subprocess
import sys


def sass(lib, name):
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    cur, funcs = None, {}
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1); funcs[cur] = []; continue
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if m and cur:
            funcs[cur].append((int(m.group(1), 16), m.group(2).strip()))
    hits = [f for f in funcs if name in f]
    if not hits:
        raise SystemExit("no function matching %r" % name)
    return hits[0], funcs[hits[0]]


def find_loop(ins):
    """Largest backward branch that is not one of the BRA.DIV fallback stubs at the end of the function."""
    stub = min([a for a, t in ins if "WARPSYNC.COLLECTIVE" in t] or [1 << 30])
    best = None
    for addr, txt in ins:
        m = re.search(r"\bBRA\b.*?0x([0-9a-f]+)", txt)
        if m and addr < stub:
            tgt = int(m.group(1), 16)
            if tgt < addr and (best is None or addr - tgt > best[1] - best[0]):
                best = (tgt, addr)
    return best


def opcode(txt):
    t = txt.split()
    if t[0].startswith("@"):
        t = t[1:]
    op = t[0]
    parts = op.split(".")
    base = parts[0]
    if base in ("IMAD",) and len(parts) > 1 and parts[1] in ("MOV", "X", "WIDE", "SHL", "IADD"):
        base += "." + parts[1]
    if base in ("SHFL", "LDG", "STG", "LDS", "STS") and len(parts) > 1:
        base += "." + parts[1]
    return base


FP64 = ("DADD", "DMUL", "DFMA", "DSETP", "DMNMX")


def main():
    lib, name = sys.argv[1], sys.argv[2]
    fn, ins = sass(lib, name)
    lo, hi = ins[0][0], ins[-1][0]
    if "--all" not in sys.argv:
        best = find_loop(ins)
        if best:
            lo, hi = best
    body = [(a, t) for a, t in ins if lo <= a <= hi]
    h = collections.Counter(opcode(t) for _, t in body)
    n = sum(h.values())
    f64 = sum(v for k, v in h.items() if k in FP64)
    print("%s\n  range 0x%x..0x%x  instructions %d  fp64 %d  other %d" % (fn, lo, hi, n, f64, n - f64))
    for k, v in h.most_common():
        print("  %-12s %4d" % (k, v))




def hot_trace(lib, name, keep=()):
    """Walk the loop once taking every forward conditional branch (the rare regions are written as
    `if (rare) {...}`, compiled to `@!P BRA skip`), except those whose address is listed in keep."""
    fn, ins = sass(lib, name)
    lo, hi = find_loop(ins)
    if "--range" in sys.argv:
        k = sys.argv.index("--range")
        lo, hi = int(sys.argv[k + 1], 16), int(sys.argv[k + 2], 16)
    idx = {a: n for n, (a, _) in enumerate(ins)}
    h = collections.Counter()
    pc = idx[lo]
    taken = []
    while ins[pc][0] <= hi:
        a, t = ins[pc]
        h[opcode(t)] += 1
        m = re.search(r"^(@!?U?P\d+\s+)?BRA(\.U)?\s.*?0x([0-9a-f]+)", t)
        if m and "BRA.DIV" not in t:
            tgt = int(m.group(3), 16)
            if tgt <= a and (not m.group(1) or a == hi):
                break                      # the loop's back edge
            if tgt > a and tgt <= hi + 0x10 and a not in keep:
                taken.append((a, tgt, bool(m.group(1))))
                if tgt > hi:
                    break
                pc = idx[tgt]
                continue
        pc += 1
    n = sum(h.values())
    f64 = sum(v for k, v in h.items() if k in FP64)
    print("HOT TRACE %s\n  instructions %d  fp64 %d  other %d" % (fn, n, f64, n - f64))
    print("  skipped regions:", ", ".join("%x->%x(+%d)" % (a, t, (t - a) // 16 - 1) for a, t, _ in taken))
    for k, v in h.most_common():
        print("  %-12s %4d" % (k, v))


if "--hot" in sys.argv:
    args = sys.argv[sys.argv.index("--hot") + 1:]
    if "--range" in args:
        args = args[:args.index("--range")]
    keep = tuple(int(x, 16) for x in args if re.fullmatch(r"[0-9a-f]+", x))
    hot_trace(sys.argv[1], sys.argv[2], keep)

if __name__ == "__main__" and "--hot" not in sys.argv:
    main()
