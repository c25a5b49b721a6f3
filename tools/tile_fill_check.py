import os, sys, torch
sys.path.insert(0, os.getcwd())
import openeo_processes_dask_b200.process_implementations as P
for shape in ((4000,100,8,12),(4000,100,8,16),(2000,100,10,20),(1000,100,16,24),(8000,100,48)):
    x = torch.randn(shape, device="cuda"); x[torch.rand(shape, device="cuda") < 0.3] = float("nan")
    res=[]
    for env in ("1","0"):
        os.environ["EOC_SELECT_TMA"]=env
        for _ in range(2): P.median(x, axis=1)
        torch.cuda.synchronize()
        a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        a.record(); P.median(x, axis=1); b.record(); torch.cuda.synchronize()
        res.append(x.numel()*4/a.elapsed_time(b)/1e6)
    inner = 1
    for s in shape[2:]: inner*=s
    print(shape, "inner", inner, "fill", inner/(-(-inner//128)*128), "TMA %.0f regs %.0f GB/s" % tuple(res))
