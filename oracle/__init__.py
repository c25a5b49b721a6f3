"""CPU oracle for the reducer hot path of openeo-processes-dask.

TEST INFRASTRUCTURE ONLY.  Nothing under ``openeo_processes_dask_b200/`` may
import this package; only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s CPU-baseline / ``--impl reference`` legs use it, as the checker
or as the timed CPU reference -- never as the product path.

What it is: a numpy/pandas restatement of the reference's array-level reducers
(``process_implementations/math.py:118-209,341-398,412-414``,
``arrays.py:587-606``, ``comparison.py:24-33``) and of the cube-level glue of
``cubes/reduce.py:13-62``, ``cubes/aggregate.py:29-101`` and
``cubes/indices.py:15-59``.  The arithmetic of the reference lives in numpy
(pinned as ``numpy>=1.26.3,<3`` in the reference's ``pyproject.toml:31``; the
numpy installed here is 2.3.5), so the oracle calls the same numpy functions
the reference calls.  The orchestration the reference delegates to xarray
(``DataArray.reduce``, ``groupby_bins``) is restated on plain
``(array, dims, coords)`` triples because xarray/dask/pg-parser are not
installable in this environment.

Parity pinning (see DESIGN.md "Oracle"):
  * array-level reducers: pinned against (i) every ``examples`` entry of the
    openEO process specs the reference ships and (ii) outputs of the
    reference's *own* ``math.py`` / ``arrays.py`` functions, executed in the
    build container with stub ``dask``/``xarray`` modules
    (``tests/golden/make_golden.py`` -> ``tests/golden/*.npz|json``).
  * cube-level glue (dimension -> axis, right-closed bins, labels): pinned
    against the reference tests' shape/label assertions and ``pandas.cut``;
    *values* of ``aggregate_temporal`` are not pinned by any reference test
    ("parity unpinned" for that boundary, documented in DESIGN.md).
"""

from . import reducers, cubes, synth  # noqa: F401
