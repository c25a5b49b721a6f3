"""`order` / `rearrange` (reference arrays.py:498-555) against outputs of the reference's own
functions (tests/golden/reference_order.npz, written by make_golden_order.py), incl. the reference-test vectors of
tests/test_arrays.py:414-470: stability on ties, -0.0 == +0.0, NaNs last / first / dropped, np.flip quirk of
asc=False, negative indices in rearrange."""

import os

import numpy as np
import pytest

from oracle import reducers as R

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "reference_order.npz"), allow_pickle=False)


def _cases(gold):
    for key in gold.files:
        parts = key.split("/")
        if parts[0] == "order":
            x = gold["in/" + parts[1]]
            if x.ndim == 1:
                asc, nd = parts[2] == "asc1", {"nodataNone": None, "nodataTrue": True, "nodataFalse": False}[parts[3]]
                yield key, "order", x, dict(asc=asc, nodata=nd), gold[key]
            else:
                yield key, "order", x, dict(asc=parts[3] == "asc1", nodata=True, axis=int(parts[2][4:])), gold[key]
        elif parts[0] == "rearrange" and parts[-1] == "out":
            x = gold["in/" + parts[1]]
            idx = gold[key[:-3] + "idx"]
            kw = {} if x.ndim == 1 else {"axis": int(parts[2][4:])}
            yield key, "rearrange", x, dict(order=idx, **kw), gold[key]


def test_oracle_matches_reference(gold):
    n = 0
    for key, fn, x, kw, want in _cases(gold):
        got = getattr(R, fn)(x.copy(), **kw)
        np.testing.assert_array_equal(got, want, err_msg=key)
        n += 1
    assert n >= 40


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["numpy", "cuda"])
def test_gpu_order_rearrange_match_reference(gold, kind):
    import torch

    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200 import engine

    for key, fn, x, kw, want in _cases(gold):
        data = torch.from_numpy(x).cuda() if kind == "cuda" else x.copy()
        engine.launch_count(reset=True)
        got = getattr(P, fn)(data, **kw)
        assert engine.launch_count() >= 1, key
        got = got.cpu().numpy() if hasattr(got, "cpu") else np.asarray(got)
        np.testing.assert_array_equal(got, want, err_msg=f"{key}/{kind}")
        assert got.dtype == want.dtype, (key, got.dtype, want.dtype)
    # the reference's own parametrised lists (tests/test_arrays.py:414-470) given as python lists
    data = [6, -1, 2, np.nan, 7, 4, np.nan, 8, 3, 9, 9]
    np.testing.assert_array_equal(P.order(data=data, asc=True, nodata=None), [1, 2, 8, 5, 0, 4, 7, 9, 10])
    np.testing.assert_array_equal(P.order(data=data, asc=False, nodata=False), [6, 3, 10, 9, 7, 4, 0, 5, 8, 2, 1])
    np.testing.assert_array_equal(P.rearrange(data=[5, 4, 3], order=[2, 1, 0]), [3, 4, 5])
    np.testing.assert_array_equal(P.rearrange(data=[[5, 4, 3, 2], [5, 4, 3, 2]], order=[1, 3], axis=1), [[4, 2], [4, 2]])
    with pytest.raises(ValueError):
        P.rearrange(data=[[5, 4, 3, 2], [5, 4, 3, 2]], order=[[]], axis=1)
    with pytest.raises(ValueError):
        P.order(np.zeros((3, 4), np.float32), nodata=None, axis=0)


@pytest.mark.gpu
def test_gpu_order_full_columns_vs_oracle():
    """A (t, bands, y, x) cube: order along t == stable argsort per pixel; rearrange(order) along t sorts the cube;
    'value of band X at the date of the maximum of band Y' composes from the two."""
    import torch

    import openeo_processes_dask_b200.process_implementations as P

    rng = np.random.default_rng(5)
    for T_, dt in ((100, np.float32), (37, np.float64), (200, np.float32), (64, np.uint16)):
        x = rng.integers(0, 50, size=(T_, 2, 5, 33)).astype(dt)
        if np.dtype(dt).kind == "f":
            x[rng.random(x.shape) < 0.3] = np.nan
        d = torch.from_numpy(x).cuda()
        perm = P.order(d, nodata=True, axis=0)
        want = np.argsort(x, kind="mergesort", axis=0)
        np.testing.assert_array_equal(perm.cpu().numpy(), want)
        y = np.ascontiguousarray(np.moveaxis(x, 0, 2))       # t in the middle, inner = 33
        np.testing.assert_array_equal(P.order(torch.from_numpy(y).cuda(), nodata=True, axis=2).cpu().numpy(),
                                      np.argsort(y, kind="mergesort", axis=2))
