"""CPU oracle for the CSS hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``css_b200/`` may import this package.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs use it, and only as the checker / reported CPU baseline.
"""
