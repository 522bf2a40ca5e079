"""Accuracy guard beyond cosine-to-oracle (reference harness benchmarks/run_mteb.py:34-106): the GPU library
must rank sentence pairs like the compiled reference.  See tests/pair_similarity.py."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import reflib  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not reflib.available(), reason="compiled reference (oracle/_ref) not built")
def test_pair_similarities_rank_like_the_reference():
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import pair_similarity
    rows = pair_similarity.run("all-MiniLM-L6-v2", n_pairs=300, modes=("f32", "q4_0", "q4_1"))
    for r in rows:
        assert r["similarity_range"][1] - r["similarity_range"][0] > 0.3, r       # the pairs do spread
        assert r["pearson_gpu_vs_reference"] >= 0.9995, r
        assert r["spearman_gpu_vs_reference"] >= 0.999, r
        assert r["max_abs_similarity_diff"] <= 5e-3, r
        s = r["sts_style_spearman_vs_f32_reference"]
        assert abs(s["gpu"] - s["reference"]) <= 2e-3, r
