"""The expression compiler of lazy_elementwise.py against numpy, with a python interpreter of the accumulator
machine of eoc_apply_tape (include/eoc.h) standing in for the kernel: checks the host logic on CPU."""

import numpy as np
import pytest

from openeo_processes_dask_b200 import _lib as L
from openeo_processes_dask_b200.lazy_elementwise import TapeOverflow, compile_expression


def interpret(steps, x, y=None):
    """numpy model of k_tape (csrc/tape.cu): same step semantics, arithmetic in x's dtype"""
    A = x.dtype.type
    acc, t0, t1 = x.copy(), np.zeros_like(x), np.zeros_like(x)
    binary = {
        L.TAPE_ADD: lambda a, o: a + o, L.TAPE_SUB: lambda a, o: a - o, L.TAPE_RSUB: lambda a, o: o - a,
        L.TAPE_MUL: lambda a, o: a * o, L.TAPE_DIV: lambda a, o: a / o, L.TAPE_RDIV: lambda a, o: o / a,
        L.TAPE_MAXNUM: np.maximum, L.TAPE_MINNUM: np.minimum, L.TAPE_POW: np.power, L.TAPE_RPOW: lambda a, o: np.power(o, a),
        L.TAPE_MOD: np.mod, L.TAPE_RMOD: lambda a, o: np.mod(o, a), L.TAPE_ATAN2: np.arctan2,
        L.TAPE_RATAN2: lambda a, o: np.arctan2(o, a)}
    unary = {L.TAPE_ABS: np.abs, L.TAPE_SGN: np.sign, L.TAPE_SQRT: np.sqrt, L.TAPE_FLOOR: np.floor, L.TAPE_CEIL: np.ceil,
             L.TAPE_TRUNC: np.trunc, L.TAPE_EXP: np.exp, L.TAPE_LN: np.log, L.TAPE_SIN: np.sin, L.TAPE_COS: np.cos}
    with np.errstate(all="ignore"):
        for code, which, imm in steps:
            o = [np.full_like(x, A(imm)), x, y, t0, t1][which]
            if code == L.TAPE_LOAD:
                acc = o.copy()
            elif code == L.TAPE_STORE_T0:
                t0 = acc.copy()
            elif code == L.TAPE_STORE_T1:
                t1 = acc.copy()
            elif code in binary:
                acc = binary[code](acc, o).astype(x.dtype)
            elif code == L.TAPE_ROUND:
                acc = (np.rint(acc * A(imm)) / A(imm)).astype(x.dtype)
            else:
                acc = unary[code](acc).astype(x.dtype)
    return acc


X, Y, I = ("in", 0), ("in", 1), (lambda v: ("imm", v))


def b(code, l, r):
    return ("b", code, l, r)


def u(code, c, imm=0.0):
    return ("u", code, c, imm)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_compiled_expressions_match_numpy(dtype):
    rng = np.random.default_rng(5)
    x = rng.normal(0.3, 1.0, 1000).astype(dtype)
    y = rng.normal(0.1, 1.0, 1000).astype(dtype)
    x[::17] = np.nan
    with np.errstate(all="ignore"):
        cases = [
            (b(L.TAPE_DIV, b(L.TAPE_SUB, X, Y), b(L.TAPE_ADD, X, Y)), (x - y) / (x + y)),
            (b(L.TAPE_SUB, I(1.0), X), 1.0 - x),
            (b(L.TAPE_DIV, I(2.0), b(L.TAPE_ADD, X, I(0.5))), 2.0 / (x + 0.5)),
            (b(L.TAPE_SUB, b(L.TAPE_MUL, X, X), b(L.TAPE_MUL, b(L.TAPE_ADD, Y, I(1.0)), b(L.TAPE_SUB, Y, X))),
             x * x - (y + 1.0) * (y - x)),
            (u(L.TAPE_SQRT, u(L.TAPE_ABS, b(L.TAPE_MUL, Y, I(3.0)))), np.sqrt(np.abs(y * 3.0))),
            (b(L.TAPE_MINNUM, b(L.TAPE_MAXNUM, X, I(-0.5)), I(0.75)), np.clip(x, -0.5, 0.75)),
            (u(L.TAPE_ROUND, X, 100.0), np.around(x, 2)),
            (Y, y),
            # ((x+1)*(y+2)) - ((x-3)/(y*y+1)) : both sides of the root and of the right child need a temporary
            (b(L.TAPE_SUB, b(L.TAPE_MUL, b(L.TAPE_ADD, X, I(1.0)), b(L.TAPE_ADD, Y, I(2.0))),
               b(L.TAPE_DIV, b(L.TAPE_SUB, X, I(3.0)), b(L.TAPE_ADD, b(L.TAPE_MUL, Y, Y), I(1.0)))),
             (x + 1.0) * (y + 2.0) - (x - 3.0) / (y * y + 1.0)),
        ]
    for expr, want in cases:
        got = interpret(compile_expression(expr), x, y)
        assert got.dtype == want.dtype
        np.testing.assert_array_equal(got, want)


def test_tape_overflow_is_reported():
    deep = X
    for k in range(L.TAPE_MAX_OPS + 1):
        deep = b(L.TAPE_ADD, deep, I(float(k)))
    with pytest.raises(TapeOverflow):
        compile_expression(deep)
    # three nested "both sides are expressions" need a third temporary
    leafy = lambda a, c: b(L.TAPE_MUL, b(L.TAPE_ADD, a, I(1.0)), b(L.TAPE_ADD, c, I(2.0)))
    three = b(L.TAPE_SUB, leafy(X, Y), b(L.TAPE_SUB, leafy(Y, X), leafy(X, X)))
    with pytest.raises(TapeOverflow):
        compile_expression(three)


def test_tape_codes_match_the_header():
    import re
    from pathlib import Path

    text = (Path(__file__).resolve().parents[1] / "include" / "eoc.h").read_text()
    codes = dict(re.findall(r"(EOC_TAPE_[A-Z0-9_]+) = (\d+)", text))
    assert len(codes) == 39
    for name, value in codes.items():
        assert getattr(L, name[4:]) == int(value), name
    assert int(re.search(r"#define EOC_TAPE_MAX_OPS (\d+)", text).group(1)) == L.TAPE_MAX_OPS
