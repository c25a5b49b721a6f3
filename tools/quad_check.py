#!/usr/bin/env python3
"""k_select_quad (EOC_SELECT_QUAD=1) against the round-1 median cone (EOC_SELECT_QUAD=0) on the GPU: bit-equal medians
for T = 65 .. 128, ragged widths, both nodata modes, +-0.0 / inf / all-NaN columns; then timings at T = 100."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import openeo_processes_dask_b200.process_implementations as P
from openeo_processes_dask_b200 import engine


def med(x, quad, ignore=True):
    os.environ["EOC_SELECT_QUAD"] = "1" if quad else "0"
    os.environ["EOC_SELECT_TMA"] = "1"
    return P.median(x, axis=0, ignore_nodata=ignore)


def others(x, mask, quad, ign):
    """quantiles, fused quantiles + sd / median + mean + max, mask -> median through the process functions"""
    from openeo_processes_dask_b200.process_implementations.math import reduce_many

    os.environ["EOC_SELECT_QUAD"] = "1" if quad else "0"
    os.environ["EOC_SELECT_TMA"] = "1"
    q3 = [0.1, 0.5, 0.9]
    out = [P.quantiles(x, probabilities=q3, axis=0, ignore_nodata=ign),
           P.quantiles(x, probabilities=[0.9, 0.25, 0.5, 1.0], axis=0, ignore_nodata=ign)]
    qs, sd = reduce_many(x, [("quantiles", {"probabilities": q3}), "sd"], axis=0, ignore_nodata=ign)
    md, mean, mx = reduce_many(x, ["median", "mean", "max"], axis=0, ignore_nodata=ign)
    out += [qs, sd, md, mean, mx]
    T = x.shape[0]
    cube = P.DataCube(x, ("t", "bands", "y", "x"), {"t": list(range(T))})
    mcube = P.DataCube(mask, ("t", "y", "x"), {"t": list(range(T))})
    masked = P.mask(data=cube, mask=mcube).data
    out.append(P.median(masked, axis=0, ignore_nodata=ign))
    out.append(P.quantiles(masked, probabilities=q3, axis=0, ignore_nodata=ign))
    return [o if isinstance(o, torch.Tensor) else torch.as_tensor(o) for o in out]


def same(a, b):
    """bit-equal, except that the sign of a zero result may differ: the two-run kernels break -0.0 / +0.0 ties by
    run, k_select_quad and the median cone by the total order of FMNMX (numpy's own choice depends on its
    partition's input order)"""
    return bool(((a == b) | (torch.isnan(a) & torch.isnan(b))).all())


def same_bits(a, b):
    return bool(((a.view(torch.int32) == b.view(torch.int32)) | (torch.isnan(a) & torch.isnan(b))).all())


def main():
    bad = 0
    g = torch.Generator(device="cuda").manual_seed(7)
    for T in (65, 72, 77, 80, 81, 96, 99, 100, 101, 112, 127, 128):
        for X in (1024, 1000, 36):
            x = torch.randn((T, 3, 7, X), device="cuda", generator=g)
            x[torch.rand(x.shape, device="cuda", generator=g) < 0.3] = float("nan")
            x[:, 0, 0, :8] = float("nan")            # all-NaN columns
            x[: T // 2, 0, 1, :8] = 0.0               # signed zeros around the middle
            x[T // 2:, 0, 1, :8] = -0.0
            x[::3, 0, 2, :8] = float("inf")
            x[1::3, 0, 2, :8] = float("-inf")
            x[:, 1, 0, :4] = 1.5                      # constant column without NaN
            # NaN payloads, negative and signalling NaNs: the first-touch comparators canonicalise them
            bits = x.view(torch.int32)
            nanmask = torch.isnan(x)
            pay = torch.tensor([0x7FC00000, -0x400000, 0x7FC12345, -1, 0x7F800001], dtype=torch.int32, device="cuda")
            bits[nanmask] = pay[torch.randint(0, 5, (int(nanmask.sum()),), device="cuda", generator=g)]
            mask = torch.rand((T, 7, X), device="cuda", generator=g) < 0.25
            for ign in (True, False):
                if X % 4 == 0:
                    for j, (a, b) in enumerate(zip(others(x, mask, True, ign), others(x, mask, False, ign))):
                        if not same(a, b):
                            bad += 1
                            d = ~((a == b) | (torch.isnan(a) & torch.isnan(b)))
                            idx = torch.nonzero(d)[:3].tolist()
                            print("MISMATCH other", j, T, X, ign, "value-level diffs:", int(d.sum()), idx,
                                  a[d][:3].tolist(), b[d][:3].tolist())
                a, b = med(x, True, ign), med(x, False, ign)
                ok = same_bits(a, b)
                if not ok:
                    bad += 1
                    d = (a.view(torch.int32) != b.view(torch.int32)) & ~(torch.isnan(a) & torch.isnan(b))
                    print("MISMATCH", T, X, ign, int(d.sum()), a[d][:4].tolist(), b[d][:4].tolist())
    print("quad_check mismatching cases:", bad)
    T, B, rows, X = 100, 4, int(os.environ.get("ROWS", "1024")), 10980
    strip = torch.empty((T, B, rows, X), dtype=torch.float32, device="cuda")
    engine.synth_fill(strip.view(T * B, rows * X), T * B, 10980 * X, 0, rows * X, 43, 0.30)
    for quad in (False, True):
        for _ in range(2):
            med(strip, quad)
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(5)]
        for a, b in ev:
            a.record()
            med(strip, quad)
            b.record()
        torch.cuda.synchronize()
        ms = sum(a.elapsed_time(b) for a, b in ev) / len(ev)
        print(f"quad={int(quad)} W={os.environ.get('EOC_QUAD_W', '128')} median t=100: {strip.numel() * 4 / ms / 1e6:.0f} GB/s")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
