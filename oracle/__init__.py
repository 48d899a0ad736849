"""CPU oracle of the paged decode-attention path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``ltp-sglang_b200/`` imports this package.  Allowed importers:
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py`` (``cpu_baseline`` leg and
``--impl reference``).  See ``decode_oracle.py`` for the parity status.
"""
