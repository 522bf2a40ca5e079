"""Interleaved A/B timing of two builds of libbert.so on the same box (clocks drift between boxes and
over a run, so only back-to-back alternation is comparable).
    python tools/ab.py libA.so libB.so rounds -- <command ...>      (command prints a line with a number)
The command is run with BERT_B200_LIB pointing at each library in turn; the first float on its last
output line is collected."""
import os
import re
import statistics
import subprocess
import sys

a, b, rounds = sys.argv[1], sys.argv[2], int(sys.argv[3])
cmd = sys.argv[sys.argv.index("--") + 1:]
res = {a: [], b: []}
for r in range(rounds):
    for lib in (a, b):
        env = dict(os.environ, BERT_B200_LIB=os.path.abspath(lib))
        out = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=300)
        lines = [l for l in out.stdout.strip().splitlines() if l.strip()]
        m = re.search(r"[-+]?\d+\.\d+(?:[eE][-+]?\d+)?", lines[-1]) if lines else None
        if out.returncode != 0 or not m:
            print("FAILED", lib, out.stdout[-300:], out.stderr[-300:])
            continue
        res[lib].append(float(m.group(0)))
for lib in (a, b):
    v = res[lib]
    if v:
        print(f"{os.path.basename(lib):24s} n={len(v)} median={statistics.median(v):.5f} min={min(v):.5f} max={max(v):.5f}  {['%.4f' % x for x in v]}")
