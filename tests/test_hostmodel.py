"""The kernels' per-column arithmetic, compiled for the host, against the oracle (CPU only).

``tests/hostmodel`` builds ``csrc/column_math.cuh`` -- the header every CUDA kernel includes -- with
g++.  For a reduction over a non-contiguous axis (the (t,b,y,x) cube reduced over t) the results must
equal eager numpy BIT FOR BIT: sequential float32 sums, two-pass variance, sorting-network median,
float64 quantile interpolation.  This pins the arithmetic before any GPU time is spent; the GPU
tests then check that the device build of the same header agrees.
"""

import warnings

import numpy as np
import pytest

from oracle import reducers as R
from tests import hostmodel

warnings.simplefilter("ignore", RuntimeWarning)


@pytest.fixture(scope="module")
def hm():
    return hostmodel.load()


def _cube(rng, R_, inner, dtype, nan_prob, scale=1000.0):
    x = rng.normal(3000, scale, size=(R_, inner)).astype(dtype)
    x[rng.random(x.shape) < nan_prob] = np.nan
    if inner >= 8:
        x[:, 0] = np.nan            # all-NaN column
        x[1:, 1] = np.nan           # exactly one valid sample
        x[2:, 2] = np.nan           # two valid samples
        x[0, 3] = np.inf
        x[min(1, R_ - 1), 4] = -np.inf
        x[:, 5] = np.inf            # constant inf
        x[:, 6] = 7.25              # constant column
        x[:, 7] = 0.0
        x[::2, 7] = -0.0
    return x


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("R_,nan_prob", [(1, 0.0), (2, 0.5), (7, 0.3), (32, 0.0), (100, 0.3), (365, 0.4)])
@pytest.mark.parametrize("ignore", [True, False])
def test_streaming_reducers_bit_exact(hm, dtype, R_, nan_prob, ignore):
    rng = np.random.default_rng(R_ * 7 + int(ignore))
    x = _cube(rng, R_, 257, dtype, nan_prob)
    got = hostmodel.reduce_columns(hm, x, ignore=ignore, ddof=1)
    for name in ("sum", "mean", "min", "max", "product", "variance", "sd"):
        want = R.BY_NAME[name](x, ignore_nodata=ignore, axis=0)
        assert want.dtype == got[name].dtype
        np.testing.assert_array_equal(got[name], want, err_msg=name)
    np.testing.assert_array_equal(got["count"], R.count(x, axis=0))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("ignore", [True, False])
def test_first_last_any_all(hm, dtype, ignore):
    """G_ORDER accumulators vs the oracle (arrays.py:461-495, logic.py:58-87).  The oracle's inf clamp quirk
    of first / last lives in the host layer, so the comparison here is on columns without an infinity."""
    rng = np.random.default_rng(3 + int(ignore))
    x = rng.choice(np.array([0.0, -0.0, 1.0, -2.5, np.nan], dtype=dtype), size=(9, 300), p=[0.3, 0.05, 0.3, 0.1, 0.25])
    x[:, 0] = np.nan
    x[:, 1] = 0.0
    x[:, 2] = 1.0
    x[1:, 3] = np.nan
    x[:-1, 4] = np.nan
    got = hostmodel.order_columns(hm, x, ignore=ignore)
    for name, fn in (("first", R.first), ("last", R.last), ("any", R._any), ("all", R._all)):
        want = np.asarray(fn(x, ignore_nodata=ignore, axis=0)).astype(dtype)
        np.testing.assert_array_equal(got[name], want, err_msg=name)
    y = x.copy()
    y[0, 10] = np.inf   # +-inf: truthy for both, and a valid first / last sample
    y[-1, 11] = -np.inf
    got = hostmodel.order_columns(hm, y, ignore=ignore)
    np.testing.assert_array_equal(got["any"], np.asarray(R._any(y, ignore_nodata=ignore, axis=0)).astype(dtype))
    np.testing.assert_array_equal(got["all"], np.asarray(R._all(y, ignore_nodata=ignore, axis=0)).astype(dtype))
    assert got["first"][10] == np.inf and got["last"][11] == -np.inf


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_signed_zero_totals(hm, dtype):
    """The kernels skip NaN slots where numpy adds +0: the sign of a zero total must still match
    (sum_fixup in column_math.cuh)."""
    R_ = 6
    x = np.full((R_, 8), -0.0, dtype=dtype)
    x[0, 1] = np.nan            # NaN first, then -0.0: numpy 0 + -0 = +0
    x[3, 2] = np.nan            # NaN in the middle
    x[:, 3] = np.nan            # all NaN: +0
    x[:, 4] = 0.0
    x[1, 5] = 0.0               # mixed zeros, no NaN
    x[:, 6] = [1.5, -1.5, np.nan, -0.0, -0.0, -0.0]
    got = hostmodel.reduce_columns(hm, x, ignore=True, ddof=1)
    for name in ("sum", "mean", "variance", "sd"):
        want = R.BY_NAME[name](x, ignore_nodata=True, axis=0)
        np.testing.assert_array_equal(got[name], want, err_msg=name)
        np.testing.assert_array_equal(np.signbit(got[name]), np.signbit(want), err_msg=name + " sign")


@pytest.mark.parametrize("dtype,N,R_", [
    (np.float32, 8, 5), (np.float32, 8, 8), (np.float32, 16, 9), (np.float32, 32, 31), (np.float32, 64, 33),
    (np.float32, 100, 65), (np.float32, 100, 99), (np.float32, 100, 100), (np.float32, 128, 101),
    (np.float32, 128, 128), (np.float64, 8, 1), (np.float64, 16, 12), (np.float64, 32, 20), (np.float64, 64, 64),
])
@pytest.mark.parametrize("nan_prob", [0.0, 0.3, 0.9])
@pytest.mark.parametrize("ignore", [True, False])
def test_median_network_bit_exact(hm, dtype, N, R_, nan_prob, ignore):
    rng = np.random.default_rng(N * 1000 + R_)
    x = _cube(rng, R_, 300, dtype, nan_prob)
    # integer-valued data: many ties, like Sentinel-2 digital numbers
    x[:, 100:200] = np.where(np.isnan(x[:, 100:200]), np.nan, np.round(x[:, 100:200] / 500))
    got = hostmodel.median_columns(hm, x, N, ignore=ignore)
    want = R.median(x, ignore_nodata=ignore, axis=0)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("dtype,N,R_", [
    (np.float32, 8, 3), (np.float32, 32, 20), (np.float32, 100, 100), (np.float32, 128, 111),
    (np.float64, 16, 16), (np.float64, 64, 50),
])
@pytest.mark.parametrize("ignore", [True, False])
def test_quantiles_bit_exact(hm, dtype, N, R_, ignore):
    rng = np.random.default_rng(N + R_)
    x = _cube(rng, R_, 200, dtype, 0.3)
    x[:, 0] = 1.0  # numpy sizes its result buffer from the first slice; all-NaN there would make it float32
    probs = [0.1, 0.5, 0.9, 0.0, 1.0, 1 / 3, 0.005, 0.999]
    got, med = hostmodel.quantile_columns(hm, x, N, probs, ignore=ignore)
    want = R.quantiles(x, probabilities=probs, ignore_nodata=ignore, axis=0)
    assert want.dtype == np.float64
    np.testing.assert_array_equal(got, want)
    np.testing.assert_array_equal(med, R.median(x, ignore_nodata=ignore, axis=0))


def test_golden_fixtures_axis0(hm, golden):
    """The committed reference outputs (tests/golden) for reductions over the leading axis."""
    for cname in ("int_valued_f32", "nan30_f32", "nan40_f64", "even_odd_f32"):
        x = golden[f"in/{cname}"]
        flat = np.ascontiguousarray(x.reshape(x.shape[0], -1))
        N = next(n for n in (8, 16, 32, 64, 100, 128) if n >= x.shape[0])
        for ign in (1, 0):
            got = hostmodel.reduce_columns(hm, flat, ignore=bool(ign))
            for name in ("sum", "mean", "min", "max", "product", "variance", "sd"):
                want = golden[f"out/{cname}/{name}/axis0/ign{ign}"].reshape(-1)
                np.testing.assert_array_equal(got[name], want, err_msg=f"{cname}/{name}/ign{ign}")
            med = hostmodel.median_columns(hm, flat, N, ignore=bool(ign))
            np.testing.assert_array_equal(med, golden[f"out/{cname}/median/axis0/ign{ign}"].reshape(-1))
            q, _ = hostmodel.quantile_columns(hm, flat, N, [0.1, 0.5, 0.9], ignore=bool(ign))
            want = golden[f"out/{cname}/quantiles3/axis0/ign{ign}"].reshape(3, -1)
            np.testing.assert_allclose(q, want, rtol=1e-6, equal_nan=True)  # float32 buffer quirk, see oracle
        np.testing.assert_array_equal(hostmodel.reduce_columns(hm, flat)["count"],
                                      golden[f"out/{cname}/count/axis0/ign1"].reshape(-1))


@pytest.mark.parametrize("NH,R_", [(16, 17), (16, 32), (25, 26), (25, 50), (32, 33), (32, 64), (50, 65), (50, 99),
                                   (50, 100), (64, 101), (64, 128), (50, 20), (64, 64)])
@pytest.mark.parametrize("nan_prob", [0.0, 0.3, 0.9])
@pytest.mark.parametrize("ignore", [True, False])
def test_split_selection_bit_exact(hm, NH, R_, nan_prob, ignore):
    """Round-2 selection kernels: the column as TWO sorted runs of NH plus a partition search for the ranks.
    Median and quantiles must equal numpy bit for bit: ties, infinities, all-NaN, one / two valid samples, runs
    that are empty (R <= NH, or every sample of a half is NaN)."""
    rng = np.random.default_rng(NH * 1000 + R_)
    x = _cube(rng, R_, 400, np.float32, nan_prob)
    x[:, 100:200] = np.where(np.isnan(x[:, 100:200]), np.nan, np.round(x[:, 100:200] / 500))   # many ties
    x[: R_ // 2, 200:220] = np.nan          # first half all NaN
    x[NH:, 220:240] = np.nan                # second run empty
    x[:, 8] = 1.0                           # (numpy sizes the quantile buffer's dtype from the first slices)
    probs = [0.1, 0.5, 0.9, 0.0, 1.0, 1 / 3, 0.005, 0.999]
    q, med = hostmodel.split_columns(hm, x, NH, probs, ignore=ignore)
    np.testing.assert_array_equal(med, R.median(x, ignore_nodata=ignore, axis=0))
    x[:, 0] = 1.0
    q, _ = hostmodel.split_columns(hm, x, NH, probs, ignore=ignore)
    want = R.quantiles(x, probabilities=probs, ignore_nodata=ignore, axis=0)
    np.testing.assert_array_equal(q, want)


@pytest.mark.parametrize("NH,R_", [(6, 17), (6, 24), (8, 25), (8, 32), (12, 33), (12, 48), (13, 49), (13, 52), (14, 53), (14, 56), (16, 57), (16, 64), (20, 65), (20, 77), (20, 80), (25, 81), (25, 99), (25, 100),
                                   (32, 101), (32, 127), (32, 128), (25, 30)])
@pytest.mark.parametrize("nan_prob", [0.0, 0.3, 0.9])
@pytest.mark.parametrize("ignore", [True, False])
def test_quad_selection_bit_exact(hm, NH, R_, nan_prob, ignore):
    """k_select_quad: the column as FOUR runs sorted by the raw network (non-canonical NaN payloads in, no
    canonicalisation pass), NaN sentinels behind the runs, rank elimination without bounds checks == numpy's median."""
    rng = np.random.default_rng(NH * 1000 + R_)
    inner = 400
    x = _cube(rng, R_, inner, np.float32, nan_prob)
    x[:, 100:200] = np.where(np.isnan(x[:, 100:200]), np.nan, np.round(x[:, 100:200] / 500))   # many ties
    x[: R_ // 2, 200:220] = np.nan          # leading runs all NaN
    x[NH:, 220:240] = np.nan                # only the first run holds samples
    x[:, 240:250] = np.nan                  # all NaN
    x[::3, 250:260] = np.inf
    x[1::3, 250:260] = -np.inf
    x[1:, 260:270] = np.nan                 # one valid sample
    # (column 7 of _cube alternates -0.0 / +0.0: which zero numpy's introselect leaves at the middle depends on the
    # input order; this selection works in the total order -0.0 < +0.0 like the device's FMNMX -- not comparable)
    x[:, 7] = 1.0
    # NaNs with payloads and sign bits: the first-touch comparators must canonicalise them (no signalling NaN here:
    # glibc's fminf answers NaN for those, the device's FMNMX does not -- the GPU test covers them)
    bits = x.view(np.uint32)
    nanmask = np.isnan(x)
    bits[nanmask] = rng.choice(np.array([0x7FC00000, 0xFFC00000, 0x7FC12345, 0xFFFFFFFF], dtype=np.uint32),
                               size=int(nanmask.sum()))
    med = hostmodel.quad_columns(hm, x, NH, ignore=ignore)
    np.testing.assert_array_equal(med, R.median(x, ignore_nodata=ignore, axis=0))


@pytest.mark.parametrize("NH,R_", [(20, 129), (20, 160), (25, 161), (25, 200), (32, 201), (32, 256), (40, 257), (40, 320), (50, 321), (50, 365),
                                   (50, 400), (64, 401), (64, 512), (128, 513), (128, 730), (128, 1024)])
@pytest.mark.parametrize("nan_prob", [0.0, 0.3, 0.9])
@pytest.mark.parametrize("ignore", [True, False])
def test_multi_selection_bit_exact(hm, NH, R_, nan_prob, ignore):
    """k_select_multi: long columns as EIGHT raw-sorted runs with NaN sentinels + rank elimination == numpy's median."""
    rng = np.random.default_rng(NH * 1000 + R_)
    inner = 200
    x = _cube(rng, R_, inner, np.float32, nan_prob)
    x[:, 100:150] = np.where(np.isnan(x[:, 100:150]), np.nan, np.round(x[:, 100:150] / 500))   # many ties
    x[: R_ // 2, 150:160] = np.nan          # leading runs all NaN
    x[NH:, 160:170] = np.nan                # only the first run holds samples
    x[:, 170:175] = np.nan                  # all NaN
    x[::3, 175:180] = np.inf
    x[1:, 180:185] = np.nan                 # one valid sample
    x[:, 7] = 1.0                           # (see test_quad_selection_bit_exact)
    med = hostmodel.multi_columns(hm, x, NH, ignore=ignore)
    np.testing.assert_array_equal(med, R.median(x, ignore_nodata=ignore, axis=0))


@pytest.mark.parametrize("NH,K,R_", [(16, 3, 33), (16, 3, 48), (16, 4, 49), (16, 4, 64), (25, 3, 75), (100, 2, 200),
                                     (100, 3, 201), (100, 3, 300), (128, 3, 365), (100, 4, 399), (128, 4, 512),
                                     (128, 4, 400), (128, 2, 256), (128, 3, 130)])
@pytest.mark.parametrize("nan_prob", [0.0, 0.3, 0.9])
@pytest.mark.parametrize("ignore", [True, False])
def test_k_run_selection_bit_exact(hm, NH, K, R_, nan_prob, ignore):
    """Long columns (a year of daily scenes): K sorted runs + rank elimination == numpy, bit for bit."""
    rng = np.random.default_rng(NH * 1000 + K * 100 + R_)
    inner = 160
    x = _cube(rng, R_, inner, np.float32, nan_prob)
    x[:, 40:80] = np.where(np.isnan(x[:, 40:80]), np.nan, np.round(x[:, 40:80] / 500))   # many ties
    x[: R_ // 2, 80:90] = np.nan            # leading runs empty
    x[NH:, 90:100] = np.nan                 # only the first run holds samples
    x[:, 8] = 1.0
    probs = [0.1, 0.5, 0.9, 0.0, 1.0, 1 / 3, 0.005, 0.999]
    q, med = hostmodel.kruns_columns(hm, x, NH, K, probs, ignore=ignore)
    np.testing.assert_array_equal(med, R.median(x, ignore_nodata=ignore, axis=0))
    x[:, 0] = 1.0
    q, _ = hostmodel.kruns_columns(hm, x, NH, K, probs, ignore=ignore)
    np.testing.assert_array_equal(q, R.quantiles(x, probabilities=probs, ignore_nodata=ignore, axis=0))
