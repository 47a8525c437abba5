import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vcm_ts_b200 import _probe
L = _probe.lib()
out = torch.zeros(2, dtype=torch.int64, device="cuda")
res = {}
for N in (16, 32, 64, 128, 256):
    for start, sbo in ((0, 1024), (1, 1024), (0, 1280), (1, 1280), (11, 1280), (0, 2048), (3, 2304)):
        L.vcm_umma_rate2(N, 1024, start, sbo, out.data_ptr(), None)
        torch.cuda.synchronize()
        c = out[0].item() / 1024
        res[f"N{N}_start{start}_sbo{sbo}"] = c
        print(f"N={N:4d} start_row={start:2d} sbo={sbo:5d}: {c:7.1f} cycles/MMA -> {128*N*16*2/c:7.0f} FLOP/cycle/SM")
for N in (16, 32, 64, 128, 256):
    for nacc in (1, 2, 4, 8):
        if nacc * N > 512:
            continue
        L.vcm_umma_rate3(N, 1024, nacc, out.data_ptr(), None)
        torch.cuda.synchronize()
        c = out[0].item() / 1024
        res[f"N{N}_nacc{nacc}"] = c
        print(f"N={N:4d} accumulators={nacc}: {c:7.1f} cycles/MMA -> {128*N*16*2/c:7.0f} FLOP/cycle/SM")
json.dump(res, open("gpurun_out/umma_rate.json", "w"), indent=1)
