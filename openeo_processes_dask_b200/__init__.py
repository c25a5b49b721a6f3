"""B200-native reducers for openEO datacubes.

Drop-in for ONE path of Open-EO/openeo-processes-dask: the nan-aware reducers that
``reduce_dimension`` / ``aggregate_temporal`` run over ``t`` / ``bands`` (mean, sum, min, max, count,
sd, variance, median, quantiles, product) plus ``ndvi`` / ``normalized_difference`` feeding such a
reduce.  The arithmetic runs in hand-written CUDA kernels for sm_100a behind a C ABI
(``include/eoc.h`` -> ``_native/libeoc_sm100.so``); this package is the thin Python host layer that
keeps the reference's process signatures.  See DESIGN.md and INTEGRATION.md.
"""

__version__ = "0.1.0"

from . import process_implementations  # noqa: F401,E402
