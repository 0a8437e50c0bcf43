"""CPU oracle for the point-cloud codec hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU, the semantics of the two un-vendored third-party
packages in which the reference's hot path executes (MinkowskiEngine master / v0.5.4 and
compressai 1.2.4), plus the reference's own host helpers (`utils.sort_tensor`,
`metrics/metric.py`).  Nothing under ``oracle/`` is product code: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` leg may
import it, and only as the checker / CPU baseline.

PARITY UNPINNED: the reference ships no tests, golden vectors or weights, and neither
MinkowskiEngine nor compressai is installable here (no source, no network), so no
reference-owned vector pins this oracle.  The independent pins we *do* hold are in
``tests/test_oracle_*.py``: dense ``torch.nn.functional.conv3d / conv_transpose3d`` on
densified grids (geometry + weight-index convention), ``scipy.stats.norm`` for the
Gaussian likelihood, ``scipy.special.expit`` chains for the factorised prior, and
``scipy.spatial.cKDTree`` for the PSNR metric; and ``tests/golden/*.npz`` hold outputs of
the reference's *own* ``model/*.py`` executed (unmodified, from /root/reference) on top of
this oracle via ``oracle/shim.py`` (generator: ``oracle/make_golden.py``).
"""
