"""Opcode histogram per kernel of the built library (evidence that the hot kernels are tcgen05 / TMEM / TMA code):
    python tools/sass_hist.py [lib.so] > profiles/sass_r2.txt
Runs `cuobjdump -sass` and counts, per kernel, the tensor-core / tensor-memory / TMA / mbarrier opcodes plus the legacy
HMMA (mma.sync), and the total instruction count."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "embeddings.cpp_b200", "lib", "libbert.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCIMMA", "UTCBAR", "UTCCP", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "SYNCS", "HMMA", "IMMA", "MUFU",
        "LDG", "STG", "LDS", "STS", "ATOM", "RED", "BAR"]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s+Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(anonymous namespace\)::", "", kern)
        kern = re.sub(r"\(.*$", "", kern)                      # drop the argument list
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)", line)
    if m and kern:
        op, mods = m.group(1), m.group(2)
        hist[kern]["_total"] += 1
        hist[kern][op] += 1
        if op in ("UTCHMMA", "UTMALDG", "UTCBAR") and ".2CTA" in mods:
            hist[kern][op + ".2CTA"] += 1
print(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)}: opcode counts per kernel (static instructions)")
tot = collections.Counter()
for k, h in hist.items():
    cols = [f"{key}={h[key]}" for key in KEYS + ["UTCHMMA.2CTA", "UTMALDG.2CTA", "UTCBAR.2CTA"] if h[key]]
    print(f"{k}\n    total={h['_total']}  " + "  ".join(cols))
    tot.update(h)
print("# whole library")
print("    " + "  ".join(f"{key}={tot[key]}" for key in KEYS + ["UTCHMMA.2CTA", "UTMALDG.2CTA", "UTCBAR.2CTA"] if tot[key]))
