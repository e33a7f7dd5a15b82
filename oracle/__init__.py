"""oracle -- TEST INFRASTRUCTURE, not product code.

CPU restatement of the reference's algorithm for the hot path (plain C for the
arithmetic in ``restated/kernels.c``, numpy for the integer plans in
``plan.py``), plus loaders for ``oracle/_ref`` (the reference's own sources
compiled where they lie under /root/reference).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs
(``cpu_baseline`` / ``--impl reference``) may import this package.  The product
(``spmv_b200/``) never does and fails loudly when its CUDA library is missing.
"""
