"""Count the SASS instructions of the timed loop of every k_issue instantiation (no GPU needed)."""
import re, subprocess, sys
from collections import Counter
txt = subprocess.run(['cuobjdump', '-sass', sys.argv[1]], capture_output=True, text=True).stdout
for f in re.split(r'\n\s*Function : ', txt)[1:]:
    name = f.split('\n')[0]
    dem = subprocess.run(['c++filt', name], capture_output=True, text=True).stdout.strip()
    if 'k_issue' not in dem:
        continue
    ins = []
    for l in f.split('\n'):
        m = re.search(r'/\*([0-9a-f]{4})\*/\s+(.*?);', l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2)))
    best = None
    for a, t in ins:
        m = re.search(r'BRA(?:\.U)?\s+(?:!?U?P\d+,\s*)?0x([0-9a-f]+)', t)
        if m and int(m.group(1), 16) < a:
            body = [x for x in ins if int(m.group(1), 16) <= x[0] <= a]
            if best is None or len(body) > len(best):
                best = body
    if best:
        c = Counter(re.sub(r'@!?U?P\d+\s+', '', t).split()[0].split('.')[0] for _, t in best)
        print(dem.split('(')[0], dict(c))
