"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU restatement (numpy/scipy) of the SimuBox density-field analysis hot path,
plus a loader for the live reference (``/root/reference``, only in the build
container).  Nothing in ``simubox_b200/`` may import this package: it is the
checker for ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs, never the product path.

Parity pin: the restatement in ``oracle/scatter_np.py`` / ``oracle/voronoi_np.py``
is checked against the LIVE reference (``oracle/ref_live.py``) by
``oracle/make_golden.py``, which also writes the committed fixtures under
``tests/golden/``.  The reference ships no tests of its own,
so the live run is the only pin.
"""
