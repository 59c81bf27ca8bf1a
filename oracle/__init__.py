"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference hot path.

Nothing under ``muscle_b200/`` may import this package.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs use it, and only as the checker / the timed CPU baseline.
"""
