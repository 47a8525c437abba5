import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vcm_ts_b200 import _lib
L = _lib.lib()
out = torch.zeros(2, dtype=torch.int64, device="cuda")
res = {}
for N in (16, 32, 64, 96, 128, 192, 256):
    for iters in (64, 1024):
        L.vcm_umma_rate(N, iters, out.data_ptr(), None)
        torch.cuda.synchronize()
        c = out[0].item()
        res[f"N{N}_it{iters}"] = c / iters
        print(f"N={N:4d} iters={iters:5d}: {c/iters:8.1f} cycles/MMA  -> {128*N*16*2/(c/iters):8.0f} FLOP/cycle/SM")
json.dump(res, open("gpurun_out/umma_rate.json", "w"), indent=1)
