import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import openeo_processes_dask_b200.process_implementations as P
from openeo_processes_dask_b200 import engine
for T in (730, 1024):
    x = torch.empty((T, 2, 64, 10980), dtype=torch.float32, device="cuda")
    engine.synth_fill(x.view(T * 2, 64 * 10980), T * 2, 10980 * 10980, 0, 64 * 10980, 43, 0.30)
    for flag in ("1", "0"):
        os.environ["EOC_SELECT_MULTI"] = flag
        for _ in range(2): out = P.median(x, axis=0)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); out = P.median(x, axis=0); b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        print(f"t={T} EOC_SELECT_MULTI={flag}: {ms:.2f} ms = {x.numel()*4/ms/1e6:.0f} GB/s")
