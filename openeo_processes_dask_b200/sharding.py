"""Multi-GPU layer: one process per GPU, the cube cut into contiguous y-strips.

* Reductions over ``t`` / ``bands`` (reduce_dimension, aggregate_temporal, fused ndvi): every output pixel
  depends on its own column only, so each rank reduces its own strip and NOTHING is exchanged
  (``reduce_strip``).  Throughput scales with the number of GPUs.
* Reductions over a spatial dimension (reduce_spatial, reduce_dimension over x / y): each rank reduces its
  strip to per-(t, band) partials with the same kernels, then ONE small collective combines them
  (``spatial_reduce``): all_reduce(SUM) on (sum, n), all_reduce(MIN / MAX) on extrema, all_gather of
  (n, mean, M2) followed by Chan's pairwise merge for variance / sd (NCCL has no custom reduction op).
  Payloads are T*B values -- latency-bound on NVLink.  median / quantiles over a spatial dimension do not
  decompose into mergeable partials: the strips are gathered to every rank and reduced there.

The collectives go through ``torch.distributed`` (NCCL on GPUs; the CPU test-suite drives the same code
over gloo with world_size 2).  No collective is issued when the process group has a single rank or
``torch.distributed`` is not initialised.
"""

from __future__ import annotations

from typing import Dict, Optional, Sequence, Tuple

import numpy as np

from . import _lib as L
from .views import strip_bounds

try:
    import torch
    import torch.distributed as dist
except Exception:  # pragma: no cover
    torch = None
    dist = None

__all__ = ["strip_bounds", "my_strip", "reduce_strip", "local_spatial_partials", "combine_spatial_partials",
           "spatial_reduce", "MERGEABLE"]

MERGEABLE = ("sum", "mean", "min", "max", "count", "variance", "sd")


def bind_rank_to_gpu_numa(device_index: int) -> dict:
    """Pin the calling process to the CPUs that are local to GPU ``device_index`` (sysfs ``local_cpulist`` of its
    PCI function), so that host buffers this rank allocates afterwards are first-touched -- and pinned -- on the
    memory node the GPU's PCIe root hangs off (with all host strips on node 0, the GPUs of another socket pull
    their H2D copies across the inter-socket link).  Best effort: returns what it did, never raises.  On the
    round-1 GPU box this is a no-op -- a VM with ONE virtual memory node and 32 vCPUs (`numa_node` -1, `nvidia-smi
    topo -m`: every GPU local to CPUs 0-31) -- and the end-to-end rate there is bounded by the host side of the VM:
    54 GB/s with one rank, 108 with two, 114 with four, 180 with eight (aggregate)."""
    import os

    info = {"bound": False}
    try:
        import torch

        pr = torch.cuda.get_device_properties(device_index)
        bdf = f"{getattr(pr, 'pci_domain_id', 0):04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bdf}"
        with open(f"{base}/local_cpulist") as f:
            cpulist = f.read().strip()
        try:
            with open(f"{base}/numa_node") as f:
                info["numa_node"] = int(f.read().strip())
        except Exception:
            pass
        cpus = set()
        for part in cpulist.split(","):
            if not part:
                continue
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        info.update({"pci": bdf, "local_cpus": len(cpus), "allowed_cpus": len(allowed)})
        if cpus and cpus != allowed:
            os.sched_setaffinity(0, cpus)
            info["bound"] = True
    except Exception as e:  # sysfs not exposed, no affinity syscall, ...
        info["error"] = repr(e)[:120]
    return info


def _world(group=None) -> Tuple[int, int]:
    if dist is None or not dist.is_available() or not dist.is_initialized():
        return 1, 0
    return dist.get_world_size(group), dist.get_rank(group)


def my_strip(n_rows: int, group=None) -> Tuple[int, int]:
    """[y0, y1) of the calling rank."""
    world, rank = _world(group)
    return strip_bounds(n_rows, world, rank)


def reduce_strip(process, strip, axis, **kwargs):
    """A t / bands reduction of the rank's own y-strip: no communication at all.  ``process`` is one of the
    process functions (``mean``, ``median``, ``quantiles`` ...); the result is the same strip of the result."""
    return process(strip, axis=axis, **kwargs)


# ---- spatial reductions: local partials + one combine step --------------------------------------------------
def local_spatial_partials(strip, axes: Sequence[int], ignore_nodata: bool = True) -> Dict[str, "torch.Tensor"]:
    """Per-rank partials of a reduction over spatial ``axes`` of the rank's strip, from the CUDA kernels:
    sum, n (not-NaN), count (finite), min, max, mean and M2 = sum((x - mean)^2), all float64 / int64."""
    from . import engine

    ops = [L.OP_SUM, L.OP_NOT_NAN, L.OP_COUNT, L.OP_MIN, L.OP_MAX, L.OP_MEAN, L.OP_VARIANCE]
    out = engine.reduce(strip, ops, tuple(axes), ignore_nodata=ignore_nodata, ddof=0)
    n = out[L.OP_NOT_NAN]
    mean = out[L.OP_MEAN].double()
    var0 = out[L.OP_VARIANCE].double()
    total = int(np.prod([strip.shape[a] for a in axes]))
    return {
        "sum": out[L.OP_SUM].double(), "n": n, "count": out[L.OP_COUNT], "min": out[L.OP_MIN].double(),
        "max": out[L.OP_MAX].double(), "mean": mean, "m2": var0 * n.double(),
        "len": torch.full_like(n, total),
    }


def _all_reduce(t, op, group):
    if _world(group)[0] > 1:
        dist.all_reduce(t, op=op, group=group)
    return t


def combine_spatial_partials(op: str, part: Dict[str, "torch.Tensor"], ignore_nodata: bool = True, ddof: int = 1,
                             group=None):
    """Combine the per-rank partials of ``local_spatial_partials`` into the reduction over the whole cube.

    Every rank receives the result.  Works on CPU tensors over gloo and CUDA tensors over NCCL alike.
    """
    world, _ = _world(group)
    nan = float("nan")
    n = _all_reduce(part["n"].clone(), dist.ReduceOp.SUM if dist else None, group)
    length = _all_reduce(part["len"].clone(), dist.ReduceOp.SUM if dist else None, group)
    poisoned = (n < length) if not ignore_nodata else torch.zeros_like(n, dtype=torch.bool)

    def finish(value, empty_value=nan):
        value = torch.where(n == 0, torch.full_like(value, empty_value), value)
        return torch.where(poisoned, torch.full_like(value, nan), value)

    if op == "count":
        return _all_reduce(part["count"].clone(), dist.ReduceOp.SUM if dist else None, group)
    if op in ("sum", "mean"):
        # an all-NaN strip already contributes 0 (np.nansum semantics of the kernel); a NaN partial is a real
        # inf - inf and must survive the exchange
        s = _all_reduce(part["sum"].clone(), dist.ReduceOp.SUM if dist else None, group)
        if op == "sum":
            return torch.where(poisoned, torch.full_like(s, nan), s)
        return finish(s / n.double())
    if op in ("min", "max"):
        # a strip without valid samples holds NaN: make it neutral for the exchange
        neutral = float("inf") if op == "min" else float("-inf")
        v = torch.where(torch.isnan(part[op]), torch.full_like(part[op], neutral), part[op])
        v = _all_reduce(v.clone(), (dist.ReduceOp.MIN if op == "min" else dist.ReduceOp.MAX) if dist else None, group)
        return finish(v)
    if op in ("variance", "sd"):
        # gather (n, mean, M2) of every rank, merge pairwise (Chan et al.): exact in real arithmetic, float64 here
        # (a strip without valid samples has NaN mean / M2: neutral element instead; a NaN that comes from
        # inf - inf inside a strip must survive, numpy yields NaN there too)
        empty = part["n"] == 0
        zero = torch.zeros_like(part["mean"])
        trip = torch.stack([part["n"].double(), torch.where(empty, zero, part["mean"]),
                            torch.where(empty, zero, part["m2"])])
        if world > 1:
            bucket = [torch.empty_like(trip) for _ in range(world)]
            dist.all_gather(bucket, trip, group=group)
        else:
            bucket = [trip]
        na, ma, m2a = bucket[0][0], bucket[0][1], bucket[0][2]
        for other in bucket[1:]:
            nb, mb, m2b = other[0], other[1], other[2]
            tot = na + nb
            safe = torch.where(tot > 0, tot, torch.ones_like(tot))
            delta = mb - ma
            ma = torch.where(tot > 0, ma + delta * nb / safe, ma)
            m2a = m2a + m2b + delta * delta * na * nb / safe
            na = tot
        dof = na - ddof
        var = torch.where(dof > 0, m2a / torch.where(dof > 0, dof, torch.ones_like(dof)), torch.full_like(m2a, nan))
        var = torch.where(poisoned, torch.full_like(var, nan), var)
        return var if op == "variance" else torch.sqrt(var)
    raise ValueError(f"{op!r} has no mergeable partials; gather the strips instead")


def spatial_reduce(op: str, strip, axes: Sequence[int], ignore_nodata: bool = True, ddof: int = 1, group=None,
                   shard_axis: Optional[int] = None, **kwargs):
    """reduce_spatial / reduce_dimension(x|y) of a y-strip-sharded cube; every rank gets the full result.

    Mergeable reducers exchange O(T*B) partials; ``median`` / ``quantiles`` gather the strips along
    ``shard_axis`` (default: the first of ``axes``) to every rank and reduce there ("replicas only").
    """
    if op in MERGEABLE:
        part = local_spatial_partials(strip, axes, ignore_nodata)
        return combine_spatial_partials(op, part, ignore_nodata, ddof, group)
    from .process_implementations import math as M

    world, _ = _world(group)
    full = strip
    if world > 1:
        ax = axes[0] if shard_axis is None else shard_axis
        sizes = [torch.zeros(1, dtype=torch.int64, device=strip.device) for _ in range(world)]
        dist.all_gather(sizes, torch.tensor([strip.shape[ax]], dtype=torch.int64, device=strip.device), group=group)
        parts = []
        for r, sz in enumerate(sizes):
            shape = list(strip.shape)
            shape[ax] = int(sz.item())
            buf = strip.contiguous() if r == dist.get_rank(group) else torch.empty(shape, dtype=strip.dtype,
                                                                                    device=strip.device)
            dist.broadcast(buf, src=dist.get_global_rank(group, r) if group is not None else r, group=group)
            parts.append(buf)
        full = torch.cat(parts, dim=ax)
    fn = {"median": M.median, "quantiles": M.quantiles}[op]
    return fn(full, ignore_nodata=ignore_nodata, axis=tuple(axes), **kwargs)


# ---- zonal statistics of a y-strip-sharded cube (aggregate_spatial) -----------------------------------------------
def local_zonal_partials(strip, ys_strip, xs, rings, ignore_nodata: bool = True) -> Dict[str, "torch.Tensor"]:
    """The partials of ``local_spatial_partials`` for the samples of the rank's strip (``[..., y_strip, x]``) whose
    pixel centre lies inside the polygon ``rings``: the polygon is rasterised over the strip's own rows only
    (``eoc_rasterize_polygon``) and reduced by ``eoc_zonal_reduce``.  ``len`` counts the strip's whole raster, the
    length the reference reduces over (pixels outside the polygon read NaN there)."""
    from . import engine

    ys_strip, xs = np.asarray(ys_strip, dtype=np.float64), np.asarray(xs, dtype=np.float64)
    allv = np.concatenate([np.asarray(r, dtype=np.float64).reshape(-1, 2) for r in rings], axis=0)

    def window(c, lo, hi):
        hit = np.nonzero((c >= lo) & (c <= hi))[0]
        return (0, 0) if len(hit) == 0 else (int(hit[0]), int(hit[-1]) + 1)

    x0, x1 = window(xs, allv[:, 0].min(), allv[:, 0].max())
    y0, y1 = window(ys_strip, allv[:, 1].min(), allv[:, 1].max())
    mask = engine.rasterize_polygon(xs[x0:x1], ys_strip[y0:y1], rings, device=strip.device)
    ops = [L.OP_SUM, L.OP_NOT_NAN, L.OP_COUNT, L.OP_MIN, L.OP_MAX, L.OP_MEAN, L.OP_VARIANCE]
    out = engine.zonal_reduce(strip, mask, (y0, y1, x0, x1), ops, ignore_nodata=True, ddof=0)
    n = out[L.OP_NOT_NAN]
    total = int(strip.shape[-2]) * int(strip.shape[-1])
    return {
        "sum": out[L.OP_SUM].double(), "n": n, "count": out[L.OP_COUNT], "min": out[L.OP_MIN].double(),
        "max": out[L.OP_MAX].double(), "mean": out[L.OP_MEAN].double(), "m2": out[L.OP_VARIANCE].double() * n.double(),
        "len": torch.full_like(n, total),
    }


def zonal_reduce(op: str, strip, ys_strip, xs, rings, ignore_nodata: bool = True, ddof: int = 1, group=None):
    """aggregate_spatial of ONE geometry on a y-strip-sharded cube: every rank reduces the part of the polygon that
    falls on its rows, then one small exchange (all_reduce of sums / counts / extrema, all_gather + Chan merge of
    (n, mean, M2)) gives every rank the statistic of the whole polygon.  Payload: one value per (t, band) plane."""
    if op not in MERGEABLE:
        raise ValueError(f"{op!r} has no mergeable partials")
    part = local_zonal_partials(strip, ys_strip, xs, rings, ignore_nodata)
    return combine_spatial_partials(op, part, ignore_nodata, ddof, group)
