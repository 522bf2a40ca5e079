"""SM clock, power draw and throttle reasons while ONE kernel runs back to back (is it power-capped?):
    python tools/clock_probe.py attn|qkv|ffn_in [seconds]"""
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
what = sys.argv[1] if len(sys.argv) > 1 else "attn"
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 4.0
rows = []
proc = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,power.limit,clocks_event_reasons.sw_power_cap,"
                         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown,temperature.gpu",
                         "--format=csv,noheader,nounits", "-lms", "100", "-i", "0"], stdout=subprocess.PIPE, text=True)
thr = threading.Thread(target=lambda: rows.extend(l.strip() for l in proc.stdout), daemon=True)
thr.start()
t0 = time.time()
n = 0
while time.time() - t0 < secs:
    if what == "attn":
        ms, tf = pkg.bench_attention(512, 512, 12, 64, reps=200)
    else:
        N, K, epi = {"qkv": (2304, 768, 1), "ffn_in": (3072, 768, 11)}[what]
        ms, tf = pkg.bench_gemm(131072, N, K, 2, epi, reps=200)
    n += 1
proc.terminate()
thr.join(timeout=2)
print(f"{what}: {ms:.4f} ms per launch, {tf:.1f} TFLOP/s (last of {n} batches of 200)")
for r in rows[len(rows) // 3::max(1, len(rows) // 12)]:
    print("   sm MHz, max, W, limit W, power cap, hw slowdown, sw thermal, temp:", r)
