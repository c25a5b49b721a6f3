"""LazyElementwise: a chain of elementwise processes over one or two device arrays, recorded and run as ONE kernel.

``apply`` hands its child process graph the cube's array; the reference then evaluates the graph node by node, each
node a numpy expression over the whole cube (``cubes/apply.py:24-39`` with ``math.py:98-115, 212-338``).  Here
``apply`` wraps the array in a ``LazyElementwise`` and the elementwise processes of ``process_implementations.math``
extend its expression instead of launching; the first consumer that needs values (the end of ``apply``, a reducer,
``np.asarray``) compiles the expression to the accumulator program of ``eoc_apply_tape`` (include/eoc.h) -- one
read of each input and one write, whatever the length of the chain.

Arithmetic: every node rounds to the array's compute dtype, python scalars are weak (NEP 50), so the fused result
has the bits of the node-by-node evaluation for + - * / clip abs sign sqrt floor ceil trunc round, and agrees
within 1e-6 relative for the libm functions (exp, ln, trigonometry, power, arctan2).

An expression the accumulator machine cannot hold (more than two distinct inputs, more than two live
temporaries, more than ``TAPE_MAX_OPS`` steps) is split: sub-expressions are evaluated to device arrays first.
"""

from __future__ import annotations

import numpy as np

from . import _lib as L
from . import engine
from .device_array import DeviceArray, LazyCast

# expression nodes (tuples):  ("in", k) | ("imm", value) | ("u", code, child, imm) | ("b", code, left, right)
_REVERSED = {L.TAPE_ADD: L.TAPE_ADD, L.TAPE_MUL: L.TAPE_MUL, L.TAPE_MAXNUM: L.TAPE_MAXNUM, L.TAPE_MINNUM: L.TAPE_MINNUM,
             L.TAPE_SUB: L.TAPE_RSUB, L.TAPE_DIV: L.TAPE_RDIV, L.TAPE_POW: L.TAPE_RPOW, L.TAPE_MOD: L.TAPE_RMOD,
             L.TAPE_ATAN2: L.TAPE_RATAN2}


class TapeOverflow(Exception):
    """the expression needs more temporaries or steps than eoc_apply_tape has"""


def compile_expression(node, max_ops=L.TAPE_MAX_OPS):
    """expression tree -> [(code, operand, imm)] for ``eoc_apply_tape`` (operand: 0 imm, 1 x, 2 y, 3 t0, 4 t1)."""
    steps = []

    def leaf_operand(n):
        return (0, float(n[1])) if n[0] == "imm" else (1 + n[1], 0.0)

    def gen(n, temps):
        kind = n[0]
        if kind in ("in", "imm"):
            which, imm = leaf_operand(n)
            steps.append((L.TAPE_LOAD, which, imm))
        elif kind == "u":
            gen(n[2], temps)
            steps.append((n[1], 0, float(n[3])))
        else:
            code, left, right = n[1], n[2], n[3]
            if right[0] in ("in", "imm"):
                gen(left, temps)
                which, imm = leaf_operand(right)
                steps.append((code, which, imm))
            elif left[0] in ("in", "imm"):
                gen(right, temps)
                which, imm = leaf_operand(left)
                steps.append((_REVERSED[code], which, imm))
            else:
                if not temps:
                    raise TapeOverflow("more than two live temporaries")
                gen(left, temps)
                t = temps[0]
                steps.append((L.TAPE_STORE_T0 if t == 3 else L.TAPE_STORE_T1, 0, 0.0))
                gen(right, temps[1:])
                steps.append((_REVERSED[code], t, 0.0))  # acc holds the right side: left OP acc

    gen(node, [3, 4])
    if steps and steps[0] == (L.TAPE_LOAD, 1, 0.0):
        steps = steps[1:]  # the accumulator starts as x
    if len(steps) > max_ops:
        raise TapeOverflow(f"{len(steps)} steps")
    return steps


def _remap(node, mapping):
    if node[0] == "in":
        return ("in", mapping[node[1]])
    if node[0] == "imm":
        return node
    if node[0] == "u":
        return ("u", node[1], _remap(node[2], mapping), node[3])
    return ("b", node[1], _remap(node[2], mapping), _remap(node[3], mapping))


def _inputs_of(node, acc):
    if node[0] == "in":
        acc.add(node[1])
    elif node[0] == "u":
        _inputs_of(node[2], acc)
    elif node[0] == "b":
        _inputs_of(node[2], acc)
        _inputs_of(node[3], acc)
    return acc


class LazyElementwise(DeviceArray):
    """An elementwise expression over one or two same-shape device arrays (see the module docstring).  A
    ``DeviceArray``: ``.tensor`` evaluates it (once) with ``eoc_apply_tape``."""

    __slots__ = ("inputs", "expr", "compute", "_dense")

    def __init__(self, source):
        """wrap a dense device array as the expression ``x``.  ``LazyCast`` sources stay integer in memory and are
        converted in the kernel's load stage (float32 arithmetic, as ``cube.astype('float32')`` asked)."""
        self._dense = None
        if isinstance(source, LazyCast):
            self.inputs, self.compute = [source.source], np.dtype(np.float32)
        else:
            t = source.tensor if isinstance(source, DeviceArray) else source
            dt = engine.np_dtype_of(t)
            if dt.kind != "f" or dt.itemsize < 4:
                raise TypeError("LazyElementwise wraps float32 / float64 arrays (integers go through LazyCast)")
            self.inputs, self.compute = [t], dt
        self.expr = ("in", 0)

    @staticmethod
    def wraps(source) -> bool:
        if isinstance(source, LazyElementwise):
            return False
        if isinstance(source, LazyCast):
            return True
        if type(source) is not DeviceArray:
            return False
        return source.dtype in (np.dtype(np.float32), np.dtype(np.float64))

    def _new(self, inputs, expr):
        out = object.__new__(LazyElementwise)
        out.inputs, out.expr, out.compute, out._dense = inputs, expr, self.compute, None
        return out

    # ---- duck array ---------------------------------------------------------------------------------------
    @property
    def shape(self):
        return tuple(self.inputs[0].shape)

    @property
    def ndim(self):
        return self.inputs[0].ndim

    def __len__(self):
        return self.shape[0] if self.ndim else 0

    @property
    def dtype(self):
        return self.compute

    @property
    def size(self):
        return int(self.inputs[0].numel())

    def __getitem__(self, key):
        return self._new([t[key] for t in self.inputs], self.expr)

    def permute(self, *axes):
        if len(axes) == 1 and isinstance(axes[0], (tuple, list)):
            axes = tuple(axes[0])
        if not axes:
            axes = tuple(reversed(range(self.ndim)))
        return self._new([t.permute(*axes) for t in self.inputs], self.expr)

    transpose = permute

    def astype(self, dtype, copy=True):
        return DeviceArray(engine._astype(self.tensor, np.dtype(dtype)))

    def __array__(self, dtype=None, copy=None):
        host = self.tensor.cpu().numpy()
        return host if dtype is None else host.astype(dtype)

    def __repr__(self):
        return f"LazyElementwise(shape={self.shape}, dtype={self.dtype}, inputs={len(self.inputs)})"

    # ---- building -----------------------------------------------------------------------------------------
    def unary(self, code, imm=0.0):
        return self._new(self.inputs, ("u", code, self.expr, float(imm)))

    def _weak_scalar(self, v):
        """python scalars are weak; numpy scalars only when they do not widen the compute dtype"""
        if isinstance(v, (bool, np.bool_)):
            return float(int(v))
        if isinstance(v, (int, float)):
            return float(v)
        if isinstance(v, np.generic) and v.dtype.kind in "fiu" and np.result_type(self.compute, v.dtype) == self.compute:
            return float(v)
        return None

    def binary(self, code, other, reverse=False):
        """self OP other (``reverse``: other OP self) -> LazyElementwise, or None when ``other`` cannot join the
        expression (the caller then evaluates eagerly)."""
        if engine._is_host_scalar(other) and not isinstance(other, DeviceArray):
            v = self._weak_scalar(other if not isinstance(other, np.ndarray) else other[()])
            if v is None:
                return None
            mine, theirs, inputs = self.expr, ("imm", v), self.inputs
        else:
            if LazyElementwise.wraps(other):
                other = LazyElementwise(other)
            if not isinstance(other, LazyElementwise) or other.compute != self.compute or other.shape != self.shape:
                return None
            inputs = list(self.inputs)
            mapping = {}
            for k, t in enumerate(other.inputs):
                for j, mine_t in enumerate(inputs):
                    if mine_t is t or (mine_t.data_ptr() == t.data_ptr() and mine_t.dtype == t.dtype
                                       and mine_t.stride() == t.stride() and mine_t.shape == t.shape):
                        mapping[k] = j
                        break
                else:
                    inputs.append(t)
                    mapping[k] = len(inputs) - 1
            if len(inputs) > 2 or any(t.dtype != inputs[0].dtype for t in inputs):
                return None
            mine, theirs = self.expr, _remap(other.expr, mapping)
        return self._new(inputs, ("b", code, theirs, mine) if reverse else ("b", code, mine, theirs))

    # ---- evaluation ---------------------------------------------------------------------------------------
    def _evaluate(self, node):
        """node -> dense tensor in the compute dtype; splits what the tape cannot hold"""
        try:
            steps = compile_expression(node)
        except TapeOverflow:
            if node[0] == "u":
                child = self._new(self.inputs, node[2])._materialized_input()
                return child._evaluate(("u", node[1], ("in", 0), node[3]))
            left = self._new(self.inputs, node[2])._materialized_input()
            right = self._new(self.inputs, node[3])._materialized_input()
            both = left._new([left.inputs[0], right.inputs[0]], ("b", node[1], ("in", 0), ("in", 1)))
            return both._evaluate(both.expr)
        used = _inputs_of(node, set())
        x = self.inputs[0]
        y = self.inputs[1] if (1 in used and len(self.inputs) > 1) else None
        return engine.apply_tape(x, y, steps, self.compute)

    def _materialized_input(self):
        dense = self._evaluate(self.expr)
        out = object.__new__(LazyElementwise)
        out.inputs, out.expr, out.compute, out._dense = [dense], ("in", 0), self.compute, dense
        return out

    @property
    def tensor(self):
        if self._dense is None:
            self._dense = self._evaluate(self.expr)
        return self._dense

    def materialize(self):
        return DeviceArray(self.tensor)
