"""Build differently-flagged variants of the CUDA library for A/B timing on one box.

    python scripts/variants.py name1="-DYTW=1 -DYW=13" name2="..."     -> mom6_b200/variants/libmom6adv_b200_<name>.so
Run one with MOM6ADV_LIB=<path> python bench.py ...
"""
import os, subprocess, sys
from concurrent.futures import ThreadPoolExecutor
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mom6_b200 import build as b

def one(arg):
    name, flags = arg.split("=", 1)
    out = os.path.join(ROOT, "mom6_b200", "variants", "libmom6adv_b200_%s.so" % name)
    os.makedirs(os.path.dirname(out), exist_ok=True)
    cmd = [b._nvcc()] + b.NVCC_FLAGS + flags.split() + ["-o", out] + b.sources() + b._nccl_flags() + ["-ccbin", "/usr/bin/g++"]
    r = subprocess.run(cmd, cwd=b.CSRC, capture_output=True, text=True)
    return name, r.returncode, (r.stdout + r.stderr)[-2000:]

if __name__ == "__main__":
    with ThreadPoolExecutor(4) as ex:
        for name, rc, log in ex.map(one, sys.argv[1:]):
            print(name, "OK" if rc == 0 else "FAILED\n" + log)
