#!/usr/bin/env python3
"""Small invocations of every kernel that hand-rolls mbarrier phases, cp.async groups or shared-memory reuse
between the TMA (async proxy) and generic writes -- the command compute-sanitizer wraps:

    compute-sanitizer --tool memcheck  python tools/sanitize_small.py
    compute-sanitizer --tool racecheck python tools/sanitize_small.py

EOC_SELECT_TMA=1 forces the TMA-staged kernels on these small cubes.  Results are checked against numpy so a run
that "passes" the sanitizer with wrong answers still fails.
"""
import os
import sys
import warnings

os.environ.setdefault("EOC_SELECT_TMA", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
warnings.simplefilter("ignore", RuntimeWarning)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import openeo_processes_dask_b200.process_implementations as P  # noqa: E402
from openeo_processes_dask_b200.process_implementations.math import reduce_many  # noqa: E402


def main():
    rng = np.random.default_rng(1)
    done = []
    for T in (100, 77, 200, 365):
        x = rng.normal(3000, 1000, size=(T, 2, 3, 300)).astype(np.float32)
        x[rng.random(x.shape) < 0.3] = np.nan
        d = torch.from_numpy(x).cuda()
        np.testing.assert_array_equal(P.median(d, axis=0).cpu().numpy(), np.nanmedian(x, axis=0))
        done.append(f"median t={T}")
        q = P.quantiles(d, probabilities=[0.1, 0.5, 0.9], axis=0).cpu().numpy()
        np.testing.assert_allclose(q, np.nanquantile(x, [0.1, 0.5, 0.9], axis=0), rtol=1e-6, equal_nan=True)
        done.append(f"quantiles t={T}")
        np.testing.assert_array_equal(P.sd(d, axis=0).cpu().numpy(), np.nanstd(x, axis=0, ddof=1))
        done.append(f"sd (k_tile) t={T}")
        if T <= 128:
            qs, sd = reduce_many(d, [("quantiles", {"probabilities": [0.1, 0.5, 0.9]}), "sd"], axis=0)
            np.testing.assert_array_equal(sd.cpu().numpy(), np.nanstd(x, axis=0, ddof=1))
            done.append(f"fused quantiles+sd t={T}")
            m = rng.random((T, 3, 300)) < 0.3
            cube = P.DataCube(d, ("t", "bands", "y", "x"), {"t": np.arange(T)})
            mc = P.DataCube(torch.from_numpy(m).cuda(), ("t", "y", "x"), {"t": np.arange(T)})
            got = P.median(P.mask(data=cube, mask=mc).data, axis=0).cpu().numpy()
            np.testing.assert_array_equal(got, np.nanmedian(np.where(m[:, None], np.nan, x), axis=0))
            done.append(f"masked median t={T}")
            os.environ["EOC_SELECT_SPLIT"] = "0"
            np.testing.assert_array_equal(P.median(d, axis=0).cpu().numpy(), np.nanmedian(x, axis=0))
            P.quantiles(d, probabilities=[0.1, 0.5, 0.9], axis=0)
            del os.environ["EOC_SELECT_SPLIT"]
            done.append(f"k_select_tile median+quantiles t={T}")
    torch.cuda.synchronize()
    print("sanitize_small ok:", "; ".join(done))


if __name__ == "__main__":
    main()
