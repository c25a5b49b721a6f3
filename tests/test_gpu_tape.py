"""Elementwise math processes and fused ``apply`` chains on the GPU against numpy (what the reference computes:
math.py:212-319 are numpy one-liners).

Bars: bit-exact for + - * / clip abs sign sqrt floor ceil trunc round and linear_scale_range (IEEE operations, one
rounding each, in the array's dtype); rtol 1e-6 (float32 results) / 1e-12 (float64) for the libm functions.
"""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

UNARY = ["exp", "ln", "cos", "arccos", "cosh", "arcosh", "sin", "arcsin", "sinh", "arsinh", "tan", "arctan", "tanh",
         "artanh"]
NUMPY_OF = {"ln": np.log, "arcosh": np.arccosh, "arsinh": np.arcsinh, "artanh": np.arctanh}


def _np(name):
    return NUMPY_OF.get(name) or getattr(np, name)


def _close(got, want, name, single=False, half=False):
    got = np.asarray(got)
    assert got.dtype == want.dtype, (name, got.dtype, want.dtype)
    assert got.shape == want.shape, (name, got.shape, want.shape)
    if want.dtype == np.float16 or half:
        rtol, atol = 2e-3, 1e-3
    else:
        single = single or want.dtype == np.float32
        rtol = 1e-6 if single else 1e-12
        atol = (1e-6 if single else 1e-13) * max(1.0, float(np.nanmax(np.abs(want[np.isfinite(want)]), initial=1.0)))
    np.testing.assert_allclose(got, want, rtol=rtol, atol=atol, equal_nan=True, err_msg=name)
    np.testing.assert_array_equal(np.isnan(got), np.isnan(want), err_msg=name)
    np.testing.assert_array_equal(np.isinf(got), np.isinf(want), err_msg=name)


@pytest.mark.parametrize("dtype", [np.float32, np.float64, np.int16, np.uint8, np.int32])
def test_unary_math_processes_match_numpy(dtype):
    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.device_array import DeviceArray

    rng = np.random.default_rng(11)
    if np.dtype(dtype).kind == "f":
        x = rng.uniform(-3, 3, (5, 37, 41)).astype(dtype)
        x[0, 0, :8] = [np.nan, 0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, 0.5]
    else:
        lo = 0 if np.dtype(dtype).kind == "u" else -4
        x = rng.integers(lo, 5, (5, 37, 41)).astype(dtype)
    for name in UNARY:
        with np.errstate(all="ignore"):
            want = _np(name)(x)
        engine.launch_count(reset=True)
        _close(getattr(P, name)(x), want, f"{name}/numpy")
        assert engine.launch_count() >= 1
        out = getattr(P, name)(DeviceArray.from_numpy(x))
        assert isinstance(out, DeviceArray)
        _close(out, want, f"{name}/device")
    with np.errstate(all="ignore"):
        # np.log(x) is float32 (float16 for uint8) for narrow input, the division by np.float64 widens it afterwards
        narrow = np.log(x).dtype != np.float64
        half = np.log(x).dtype == np.float16
        _close(P.log(x, 10), np.log(x) / np.log(10), "log10", single=narrow, half=half)
        _close(P.log(DeviceArray.from_numpy(x), 2), np.log(x) / np.log(2), "log2/device", single=narrow, half=half)
    for p in (0, 2, -1):
        want = np.around(x * (1 if np.dtype(dtype).kind != "f" else 37.123), decimals=p)
        src = x * (1 if np.dtype(dtype).kind != "f" else dtype(37.123))
        got = np.asarray(P._round(src, p))
        assert got.dtype == want.dtype
        np.testing.assert_array_equal(got, want, err_msg=f"round p={p}")
        np.testing.assert_array_equal(np.asarray(P._round(DeviceArray.from_numpy(src), p)), want)


def test_binary_math_processes_match_numpy():
    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200.device_array import DeviceArray

    rng = np.random.default_rng(12)
    for dtype in (np.float32, np.float64):
        a = rng.uniform(-4, 4, (6, 33, 20)).astype(dtype)
        b = rng.uniform(-4, 4, (6, 33, 20)).astype(dtype)
        a[0, 0, :4] = [np.nan, 0.0, 2.0, -2.0]
        b[0, 0, :4] = [1.0, 0.0, np.nan, 0.0]
        row = b[0:1, :, 0:1]                                  # broadcasting (1, 33, 1)
        with np.errstate(all="ignore"):
            cases = [("arctan2", np.arctan2, a, b), ("arctan2", np.arctan2, a, 0.5), ("arctan2", np.arctan2, 2.0, b),
                     ("arctan2", np.arctan2, a, row), ("mod", np.mod, a, b), ("mod", np.mod, a, 3), ("mod", np.mod, a, -1.5),
                     ("mod", np.mod, 7.5, b), ("power", np.power, np.abs(a), b), ("power", np.power, a, 2),
                     ("power", np.power, a, 0.5), ("power", np.power, 2.0, b), ("power", np.power, a, row)]
            for name, fn, p, q in cases:
                want = fn(p, q)
                _close(getattr(P, name)(p, q), want, f"{name}/numpy/{dtype.__name__}")
                wp = DeviceArray.from_numpy(p) if isinstance(p, np.ndarray) else p
                wq = DeviceArray.from_numpy(q) if isinstance(q, np.ndarray) else q
                out = getattr(P, name)(wp, wq)
                assert isinstance(out, DeviceArray)
                _close(out, want, f"{name}/device/{dtype.__name__}")
    # mixed dtypes promote like numpy; integers give numpy's integer results
    f32, f64 = rng.uniform(1, 3, 50).astype(np.float32), rng.uniform(1, 3, 50)
    _close(P.power(f32, f64), np.power(f32, f64), "power f32^f64")
    i = rng.integers(-20, 20, 200).astype(np.int16)
    j = rng.integers(1, 7, 200).astype(np.int16)
    for name, fn, p, q in [("mod", np.mod, i, j), ("mod", np.mod, i, -3), ("power", np.power, i.astype(np.int32), 2),
                           ("arctan2", np.arctan2, i, j)]:
        want = fn(p, q)
        got = np.asarray(getattr(P, name)(p, q))
        assert got.dtype == want.dtype, (name, got.dtype, want.dtype)
        if want.dtype.kind in "iu":
            np.testing.assert_array_equal(got, want, err_msg=name)
        else:
            _close(got, want, name)
    assert float(np.asarray(P.mod(27, 5))) == 2 and float(np.asarray(P.power(2, 10))) == 1024
    assert float(np.asarray(P.mod(-1, 3))) == 2 and float(np.asarray(P.mod(3.2, -1.5))) == np.mod(3.2, -1.5)


def _cube(x):
    from openeo_processes_dask_b200.device_array import DeviceArray
    from openeo_processes_dask_b200.process_implementations.cubes.cube import DataCube

    dims = ("t", "y", "x")
    return DataCube(DeviceArray.from_numpy(x) if isinstance(x, np.ndarray) else x, dims,
                    {d: np.arange(n) for d, n in zip(dims, x.shape)})


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_apply_fuses_the_child_chain_into_one_launch(dtype):
    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.device_array import DeviceArray
    from openeo_processes_dask_b200.lazy_elementwise import LazyElementwise
    from openeo_processes_dask_b200.process_implementations.cubes.apply import apply

    rng = np.random.default_rng(13)
    x = rng.uniform(-0.5, 1.5, (7, 129, 67)).astype(dtype)          # odd sizes: vector body + scalar tail
    x[rng.random(x.shape) < 0.05] = np.nan
    cube = _cube(x)
    seen = {}

    def chain(data, **kw):                                           # a child process graph, as the parser calls it
        seen["type"] = type(data)
        nd = P.divide(P.subtract(data, 0.25), P.add(data, 0.75))
        return P.linear_scale_range(P.clip(P.multiply(nd, 1.5), -1, 1), -1, 1, 0, 255)

    with np.errstate(all="ignore"):
        nd = (x - 0.25) / (x + 0.75)
        c = np.clip(nd * 1.5, -1, 1)
        want = ((np.clip(c, -1, 1) - (-1)) / (1 - (-1))) * (255 - 0) + 0
    engine.launch_count(reset=True)
    out = apply(cube, chain)
    assert engine.launch_count() == 1, "the chain must run as one kernel"
    assert seen["type"] is LazyElementwise and type(out.data) is DeviceArray
    got = np.asarray(out.data)
    assert got.dtype == want.dtype and out.dims == cube.dims
    np.testing.assert_array_equal(got, want)
    # node by node on the plain DeviceArray (no apply): same bits, more launches
    engine.launch_count(reset=True)
    eager = np.asarray(chain(DeviceArray.from_numpy(x)))
    assert engine.launch_count() >= 5
    np.testing.assert_array_equal(eager, want)

    def poly(data, **kw):   # exact ops and libm ops mixed; both sides of the root are expressions
        return P.subtract(P.sqrt(P.absolute(P.multiply(data, data))), P.exp(P.multiply(P.sin(data), -0.5)))

    with np.errstate(all="ignore"):
        want = np.sqrt(np.abs(x * x)) - np.exp(np.sin(x) * dtype(-0.5))
    engine.launch_count(reset=True)
    got = np.asarray(apply(cube, poly).data)
    assert engine.launch_count() == 1
    _close(got, want, "poly")

    def rounding(data, **kw):
        return P.add(P._round(P.multiply(data, 100), 1), P.subtract(P.floor(data), P.multiply(P.ceil(data), P._int(data))))

    with np.errstate(all="ignore"):
        want = np.around(x * 100, 1) + (np.floor(x) - np.ceil(x) * np.trunc(x))
    np.testing.assert_array_equal(np.asarray(apply(cube, rounding).data), want)

    def scalar_left(data, **kw):
        return P.divide(1, P.subtract(2, P.power(data, 2)))

    with np.errstate(all="ignore"):
        want = 1 / (2 - x ** 2)
    _close(np.asarray(apply(cube, scalar_left).data), want, "scalar_left")


def test_apply_on_a_lazy_cast_cube_and_on_views():
    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.process_implementations.cubes.apply import apply

    rng = np.random.default_rng(14)
    dn = rng.integers(0, 10000, (5, 64, 96)).astype(np.uint16)
    cube = _cube(dn).astype("float32")                                # LazyCast: stays uint16 in memory

    def scale(data, **kw):
        return P.clip(P.subtract(P.multiply(data, 0.0001), 0.1), 0, 1)

    want = np.clip(dn.astype(np.float32) * 0.0001 - 0.1, 0, 1)
    engine.launch_count(reset=True)
    out = apply(cube, scale)
    assert engine.launch_count() == 1
    got = np.asarray(out.data)
    assert got.dtype == np.float32
    np.testing.assert_array_equal(got, want)
    # a transposed view of a float cube: evaluated in the view's own memory order, result indexed like the view
    x = rng.uniform(0, 1, (6, 40, 50)).astype(np.float32)
    from openeo_processes_dask_b200.device_array import DeviceArray
    from openeo_processes_dask_b200.process_implementations.cubes.cube import DataCube

    view = DeviceArray.from_numpy(x).transpose(2, 0, 1)
    vc = DataCube(view, ("x", "t", "y"), {"x": np.arange(50), "t": np.arange(6), "y": np.arange(40)})
    got = np.asarray(apply(vc, scale).data)
    np.testing.assert_array_equal(got, np.clip(np.transpose(x, (2, 0, 1)) * np.float32(0.0001) - np.float32(0.1), 0, 1))
    sl = DataCube(DeviceArray.from_numpy(x)[:, 3:17, ::2], ("t", "y", "x"), {"t": np.arange(6), "y": np.arange(14), "x": np.arange(25)})
    np.testing.assert_array_equal(np.asarray(apply(sl, scale).data), np.clip(x[:, 3:17, ::2] * np.float32(0.0001) - np.float32(0.1), 0, 1))


def test_expressions_the_tape_cannot_hold_are_split():
    import openeo_processes_dask_b200.process_implementations as P
    from openeo_processes_dask_b200 import engine
    from openeo_processes_dask_b200.lazy_elementwise import LazyElementwise
    from openeo_processes_dask_b200.device_array import DeviceArray

    rng = np.random.default_rng(15)
    x = rng.uniform(0.5, 2.0, (3, 50, 50)).astype(np.float32)
    y = rng.uniform(0.5, 2.0, (3, 50, 50)).astype(np.float32)
    z = rng.uniform(0.5, 2.0, (3, 50, 50)).astype(np.float32)
    lx, ly = LazyElementwise(DeviceArray.from_numpy(x)), LazyElementwise(DeviceArray.from_numpy(y))
    deep = lx
    want = x.copy()
    for k in range(40):                                               # 40 steps > TAPE_MAX_OPS: split, not refused
        deep = P.add(P.multiply(deep, 1.01), 0.125 * k)
        want = want * np.float32(1.01) + np.float32(0.125 * k)
    np.testing.assert_array_equal(np.asarray(deep), want)
    leafy = lambda a, c: P.multiply(P.add(a, 1.0), P.add(c, 2.0))
    three = P.subtract(leafy(lx, ly), P.subtract(leafy(ly, lx), leafy(lx, lx)))      # needs a third temporary
    assert isinstance(three, LazyElementwise)
    np.testing.assert_array_equal(np.asarray(three), (x + 1) * (y + 2) - ((y + 1) * (x + 2) - (x + 1) * (x + 2)))
    # a third distinct input cannot join the expression: evaluated eagerly, same values
    engine.launch_count(reset=True)
    out = P.add(P.multiply(lx, ly), DeviceArray.from_numpy(z))
    np.testing.assert_array_equal(np.asarray(out), x * y + z)
    # comparison processes take a lazy like any DeviceArray
    np.testing.assert_array_equal(np.asarray(P.gt(P.add(lx, 1.0), 2.0)), np.where(x + 1 > 2.0, 1.0, 0.0))


def test_apply_tape_rejects_bad_programs():
    import ctypes as C

    import torch

    from openeo_processes_dask_b200 import _lib as L

    lib = L.lib()
    x = torch.zeros(16, dtype=torch.float64, device="cuda")
    out = torch.zeros(16, dtype=torch.float64, device="cuda")
    ops = (L.TapeOp * 2)()
    ops[0].code, ops[0].operand = L.TAPE_ADD, 2
    rc = lib.eoc_apply_tape(C.c_void_p(x.data_ptr()), None, L.F64, 16, ops, 1, C.c_void_p(out.data_ptr()), L.F64, None)
    assert rc == -1 and b"y is NULL" in lib.eoc_last_error()
    ops[0].code, ops[0].operand = 99, 0
    assert lib.eoc_apply_tape(C.c_void_p(x.data_ptr()), None, L.F64, 16, ops, 1, C.c_void_p(out.data_ptr()), L.F64, None) == -1
    rc = lib.eoc_apply_tape(C.c_void_p(x.data_ptr()), None, L.F64, 16, ops, 0, C.c_void_p(out.data_ptr()), L.F32, None)
    assert rc == -2 and b"round the samples" in lib.eoc_last_error()
    assert lib.eoc_apply_tape(C.c_void_p(x.data_ptr()), None, L.F64, 16, ops, L.TAPE_MAX_OPS + 1, C.c_void_p(out.data_ptr()), L.F64, None) == -1
    torch.cuda.synchronize()
