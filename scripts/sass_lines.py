"""Instruction counts of a kernel's hot loop per SOURCE line (nvdisasm line info; build with -lineinfo).

    python scripts/sass_lines.py <lib.so> <cubin-name-substring> <mangled-function-substring> [--loop]
"""
import collections, os, re, subprocess, sys, tempfile
lib, cub, fn = sys.argv[1:4]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if cub in f and not f.startswith("kernels-")][0]
out = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
# split into functions
cur, body = None, collections.defaultdict(list)
line = None
for l in out.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m:
        cur = m.group(1); continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m and cur:
        inl = re.findall(r'inlined at "([^"]+)", line (\d+)', m.group(3))
        line = (os.path.basename(m.group(1)), int(m.group(2)), tuple((os.path.basename(a), int(b)) for a, b in inl))
        continue
    m = re.match(r"\s+(?:/\*[0-9a-f]+\*/\s+)?(@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)\s", l)
    if m and cur and line:
        body[cur].append((m.group(2), line, l.strip()))
f = [k for k in body if fn in k][0]
ins = body[f]
agg = collections.Counter(); ops = collections.defaultdict(collections.Counter)
for op, ln, txt in ins:
    key = ln[2][-1] if ln[2] else (ln[0], ln[1])      # outermost (kernel-level) line
    key = "%s:%d" % key
    agg[key] += 1; ops[key][op.split(".")[0]] += 1
print(f, len(ins), "instructions")
for k, n in sorted(agg.items(), key=lambda kv: (kv[0].split(":")[0], int(kv[0].split(":")[1]))):
    print("%-22s %4d  %s" % (k, n, dict(ops[k].most_common(8))))
