"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it, and there only as the checker / the timed CPU baseline.

Parity pin: the reference ships no tests ("parity unpinned by tests").
The oracle is instead pinned against the reference's own scripted+frozen SevenNet-0
graph (``deployed_serial.pt`` / ``deployed_parallel_*.pt``) run in this container;
see ``oracle/make_golden.py`` and ``tests/golden/``.
"""
