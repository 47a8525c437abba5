import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vcm_ts_b200 import _probe
L = _probe.lib()
H, W = 1088, 1920
img = torch.randn((H, W, 64), device="cuda").to(torch.bfloat16)
out = torch.zeros(2, dtype=torch.int64, device="cuda")
res = []
for bw, bh in ((8, 16), (10, 18), (16, 8), (18, 10), (32, 4), (34, 6), (64, 2), (66, 4), (128, 1), (16, 16), (18, 18), (32, 8), (34, 10)):
    for stages, nsplit in ((2, 1), (4, 1), (6, 1), (4, 2) if bh % 2 == 0 else (4, 1)):
        sb = (bw * bh * 128 + 1023) // 1024 * 1024
        if stages * sb > 220 * 1024:
            continue
        for _ in range(2):
            rc = L.vcm_tma_probe(img.data_ptr(), H, W, bw, bh, stages, nsplit, out.data_ptr(), None)
            assert rc == 0, rc
            torch.cuda.synchronize()
        cyc, n = out[0].item(), out[1].item()
        r = dict(bw=bw, bh=bh, stages=stages, nsplit=nsplit, cycles_per_box=cyc / n, bytes_per_cycle_sm=bw * bh * 128 * n / cyc,
                 cycles_per_row=cyc / n / (bw * bh), cycles_per_line=cyc / n / bh)
        res.append(r)
        print(r)
json.dump(res, open("gpurun_out/tma_probe.json", "w"), indent=1)
