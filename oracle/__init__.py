"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the bert.h hot path.

This package is a numpy / pure-Python restatement of the reference's algorithm
(snowyu/embeddings.cpp: bert.cpp + the ggml CPU kernels it reaches) plus a ctypes
driver for the compiled, unmodified reference under ``oracle/_ref`` (built by
``oracle/Makefile`` from /root/reference, git-ignored).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu-baseline /
``--impl reference`` legs may import it.  The product (``embeddings.cpp_b200/``,
``libbert.so``) never imports, links or executes anything in here and has no CPU
fallback.

Parity status: PINNED.  The restatement is checked (tests/test_oracle_*.py)
against (a) outputs of the compiled reference run in the build container and
committed as fixtures under tests/golden/ (generator: tests/golden/make_golden.py),
and (b) the live ``oracle/_ref/libbert_ref.so`` wherever it is present.
"""
