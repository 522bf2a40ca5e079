#!/bin/bash
# round 2, GPU call 11: no-alloc quantiser entry point (alone in a fresh process, then at the end of the whole suite),
# latency path on one stream, device time of small-model forwards against the micro-batch size
mkdir -p gpurun_out
echo "== no-alloc alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4 | cut -c1-200
echo "== whole suite"; timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | cut -c1-200
echo "== latency"; timeout 300 python tools/latency.py > gpurun_out/r2c11_latency.log 2>&1; grep -v "^bert_load" gpurun_out/r2c11_latency.log | head -12
echo "== forward time against micro-batch size"
timeout 300 python - <<'PY' 2>&1 | grep -v "^bert_load"
import sys, numpy as np
sys.path.insert(0, '.')
import __graft_entry__ as ge
from tools import synth
import bench
pkg = ge.load_package()
for wl in ("minilm-q4_0-b1738-var", "bge-small-f16-b256-s128", "bge-small-zh-q4_1-b1024-var"):
    shape, ftype, n, n_tok, kind, _ = bench.WORKLOADS[wl]
    texts = bench.make_texts(shape, kind, n, n_tok)
    m = pkg.BertModel(synth.model_path(shape, ftype))
    ids, lens = m.tokenize_batch(texts)
    cu = np.concatenate([[0], np.cumsum(lens)])
    for frac in (8, 4, 2, 1):
        k = n // frac
        m.stage(ids[:cu[k]], lens[:k]); m.run_staged(reps=5); ms, _ = m.run_staged(reps=20)
        print(f"{wl}: {k} sentences {int(cu[k])} tokens: {ms/20*1e3:.0f} us per forward, {cu[k]/(ms/20)/1e3:.1f} M tokens/s", flush=True)
    m.close()
PY
