"""CPU parity checkers for the hg-c-benchmarks hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product
(``hg_c_benchmarks_b200``) never does.

* ``oracle.port``  -- our C restatement (``hg_oracle.c`` -> ``libhgoracle.so``)
* ``oracle.ref``   -- the unmodified reference compiled into ``oracle/_ref/``
* ``oracle.npref`` -- a numpy restatement (third opinion, small sizes)
"""
from . import port, ref, npref  # noqa: F401
