"""CPU oracle for the NCNet matching head.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / reference arm may
import it, and there only as the checker / the timed CPU baseline.
"""
