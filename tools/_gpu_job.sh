set -u
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -x -q 2>&1 | tail -30 > $OUT/r2i_pytest.log
tail -12 $OUT/r2i_pytest.log
python - <<'PY'
import time, torch, numpy as np
import openeo_processes_dask_b200.process_implementations as P
for R in (730, 2000):
    x = torch.randn((R, 4, 64, 1024), device="cuda"); x[torch.rand_like(x) < 0.3] = float("nan")
    for _ in range(2): P.median(x, axis=0)
    torch.cuda.synchronize(); t0 = time.perf_counter(); P.median(x, axis=0); torch.cuda.synchronize()
    print(f"general median R={R}: {x.numel()*4/(time.perf_counter()-t0)/1e9:.1f} GB/s")
PY
