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

ME conventions that stay UNVERIFIABLE here, and why none of them is observable through the reference's call sites:
  * which duplicate the default RANDOM_SUBSAMPLE quantisation keeps (we keep the first row): the reference de-duplicates with
    `sparse_quantize` before every SparseTensor it builds from data (model/model.py:152-161, train.py:204-211), and every other
    constructor call passes coordinates that are unique by construction;
  * the row order of strided / generated coordinate sets (ours: first occurrence; ME: hash-map iteration order): every consumer
    is row-order invariant -- convolutions, pruning by coordinate key (transforms.py:257-282), `features_at_coordinates`, and the
    bitstream order is re-derived by `sort_tensor` / `sort_points` (utils.py:142-187) on both sides of the channel;
  * the trilinear corner rule of `features_at_coordinates` off the grid: all call sites query on-grid points
    (entropy_models.py:294,381,446; loss.py:105,254-271), where interpolation is an exact row lookup;
  * the kernel-offset enumeration order (ix fastest) and the even-kernel offset range {0..k-1}: they fix which slice of `kernel`
    multiplies which neighbour, i.e. the meaning of a checkpoint.  Pinned against dense conv3d / conv_transpose3d under the
    stated weight-index mapping (tests/test_oracle_dense.py); a real-ME checkpoint would be the only further evidence, and the
    reference ships none (README.md:122).
"""
