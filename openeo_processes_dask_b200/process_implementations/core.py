"""``@process``: the calling convention openeo-pg-parser-networkx uses for every process.

Behaviour mirrors the reference's decorator (``process_implementations/core.py:16-92``), which the
parser's ``ProcessRegistry(wrap_funcs=[process])`` applies to each registered function:

* ``positional_parameters={"data": 0}`` names positional arguments (``core.py:48-49``);
* every ``ParameterReference`` among args/kwargs is replaced by the entry of ``named_parameters`` it
  points to, ``ProcessParameterMissing`` if there is none (``core.py:51-69``); positional arguments
  that are not references are dropped, as in the reference;
* ``axis, keepdims, source_transposed_axis, context, dim_labels`` are removed from the kwargs when
  the target does not declare them (``core.py:71-81``);
* ``Running process <name>`` is logged at INFO, the resolved parameters at DEBUG (``core.py:83-88``).
"""

from __future__ import annotations

import functools
import inspect
import logging

from .exceptions import ProcessParameterMissing

logger = logging.getLogger(__name__)

try:  # the real class when the parser is installed, so isinstance() checks agree with its objects
    from openeo_pg_parser_networkx.pg_schema import ParameterReference  # type: ignore
except Exception:

    class ParameterReference:  # minimal stand-in: a named pointer into `named_parameters`
        def __init__(self, from_parameter: str):
            self.from_parameter = from_parameter

        def __repr__(self):
            return f"ParameterReference(from_parameter={self.from_parameter!r})"


_OPTIONAL_KWARGS = ("axis", "keepdims", "source_transposed_axis", "context", "dim_labels")


def _lookup(ref, named, fname):
    try:
        return named[ref.from_parameter]
    except KeyError:
        raise ProcessParameterMissing(
            f"Error: Process Parameter {ref.from_parameter} was missing for process {fname}") from None


def process(f):
    accepted = set(inspect.signature(f).parameters)
    fname = getattr(f, "__name__", repr(f))

    @functools.wraps(f)
    def wrapper(*args, positional_parameters=None, named_parameters=None, **kwargs):
        named = dict(named_parameters or {})
        for pname, index in (positional_parameters or {}).items():
            named[pname] = args[index]
        call_args = [_lookup(a, named, fname) for a in args if isinstance(a, ParameterReference)]
        call_kwargs = {
            k: (_lookup(v, named, fname) if isinstance(v, ParameterReference) else v) for k, v in kwargs.items()
        }
        for k in _OPTIONAL_KWARGS:
            if k not in accepted:
                call_kwargs.pop(k, None)
        if hasattr(f, "__name__"):
            logger.info("Running process %s", fname)
            if logger.isEnabledFor(logging.DEBUG):
                logger.debug("Running process %s with resolved parameters: %s", fname,
                             {k: repr(v)[:80] for k, v in call_kwargs.items()})
        return f(*call_args, **call_kwargs)

    return wrapper
