"""Golden vectors for `order` / `rearrange` from the REFERENCE's own functions (arrays.py:498-555).

    python tests/golden/make_golden_order.py        # build container only (needs /root/reference)

Same loader as make_golden.py (unmodified reference sources under inert stubs).  Output: reference_order.npz
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402


def cases():
    rng = np.random.default_rng(99)
    out = {"ref_test": np.array([6, -1, 2, np.nan, 7, 4, np.nan, 8, 3, 9, 9])}   # tests/test_arrays.py:414-440
    a = rng.integers(0, 6, size=100).astype(np.float32)      # many ties: stability matters
    a[rng.random(100) < 0.2] = np.nan
    out["ties_f32"] = a
    z = np.array([0.0, -0.0, 1.0, -0.0, 0.0, np.nan, -np.inf, np.inf, -1.0, np.nan, -0.0], dtype=np.float32)
    out["zeros_inf_f32"] = z
    out["f64"] = rng.normal(size=37)
    out["i16"] = rng.integers(-5, 5, size=50).astype(np.int16)
    c = rng.integers(0, 9, size=(20, 3, 7)).astype(np.float32)
    c[rng.random(c.shape) < 0.25] = np.nan
    out["cube_f32"] = c
    return out


def main():
    math, arrays, logic = mg.load_reference_modules()
    warnings.simplefilter("ignore")
    store = {}
    for name, x in cases().items():
        store[f"in/{name}"] = x
        if x.ndim == 1:
            for asc in (True, False):
                for nd in (None, True, False):
                    store[f"order/{name}/asc{int(asc)}/nodata{nd}"] = np.asarray(arrays.order(x.copy(), asc=asc, nodata=nd))
            idx = np.array([3, 0, 0, len(x) - 1, -1, 2])
            store[f"rearrange/{name}/idx"] = idx
            store[f"rearrange/{name}/out"] = np.asarray(arrays.rearrange(x.copy(), idx))
        else:
            for axis in range(x.ndim):
                for asc in (True, False):
                    store[f"order/{name}/axis{axis}/asc{int(asc)}/nodataTrue"] = np.asarray(
                        arrays.order(x.copy(), asc=asc, nodata=True, axis=axis))
                idx = np.array([1, 0, 2, 2, -1])
                store[f"rearrange/{name}/axis{axis}/idx"] = idx
                store[f"rearrange/{name}/axis{axis}/out"] = np.asarray(arrays.rearrange(x.copy(), idx, axis=axis))
    np.savez_compressed(os.path.join(HERE, "reference_order.npz"), **store)
    print("wrote", len(store), "arrays")


if __name__ == "__main__":
    main()
