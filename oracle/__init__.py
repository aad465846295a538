"""TEST INFRASTRUCTURE ONLY — CPU oracle for the AcinoSet TRI + SBA hot path.

Nothing under ``oracle/`` is part of the product. Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import it, and only as the checker (or the timed CPU baseline), never as
the thing shipped. The product (``acinoset_b200``) never imports this package.

Parity status: the reference ships no golden vectors / tests, so by
the reference's own tests parity is unpinned; it is pinned here by executing the
reference itself (``/root/reference/src/lib``) in the build container and committing
its outputs as fixtures under ``tests/golden/`` (generator: ``oracle/make_golden.py``).
"""
