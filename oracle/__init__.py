"""CPU oracle for the shifted-window attention hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl
reference`` legs of ``bench.py`` may import it, and only as the checker (or as
the CPU arm that is timed *beside* the product), never as the thing shipped.

Parity status: the reference repository holds no golden vectors or tests for
this path, so the oracle is pinned against outputs of the
reference modules themselves, generated in the build container by
``oracle/make_golden.py`` (which imports ``/root/reference``) and committed as
``tests/golden/*.pt``.
"""
