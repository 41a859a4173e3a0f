"""oracle/ -- CPU checkers for the EKAT tridiag + LinInterp hot path.

TEST INFRASTRUCTURE ONLY. Importable from ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs; the product
package ``ekat_b200`` never imports it (tests/test_boundary.py enforces that).

Two checkers with one Python face (:class:`oracle.api.CpuOracle`):

* ``port()``  -- ``libekat_oracle.so``: plain-C restatement (oracle/ekat_oracle.c),
  buildable anywhere gcc exists.
* ``ref()``   -- ``_ref/libekat_ref.so``: the UNCHANGED reference headers compiled
  where they lie under /root/reference against oracle/standin (Kokkos is an
  absent submodule). Prebuilt here; travels to the GPU box; ``None`` if missing.
"""
from .api import CpuOracle, build, port, ref, SOLVERS  # noqa: F401
