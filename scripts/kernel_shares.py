"""Device-time share of every kernel in one eager 1080p P-frame encode+decode (torch profiler, CUPTI):
    python scripts/kernel_shares.py [bf16|parity|fp32] -> gpurun_out/kernel_shares_<mode>.json"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import bench as Bn  # noqa: E402
from vcm_ts_b200.models import DMC, IntraNoAR  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "bf16"
dev = torch.device("cuda:0")
sd_p, sd_i = Bn.make_state_dicts(0)
dmc, intra = DMC(precision=mode).eval(), IntraNoAR(precision=mode).eval()
dmc.load_state_dict(sd_p)
intra.load_state_dict(sd_i)
dmc, intra = dmc.to(dev), intra.to(dev)
dmc.update(force=True)
intra.update(force=True)
seq = torch.nn.functional.pad(Bn.synth_sequence(4, 0).to(dev), (0, Bn.WP - Bn.W, 0, Bn.HP - Bn.H))
r, _ = intra.encode_decode_bytes(seq[0:1], 1.0, Bn.H, Bn.W)
dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
for t in (1, 2):
    r, _ = dmc.encode_decode_bytes(seq[t:t + 1], dpb, Bn.H, Bn.W, 1.0, 1.0)
    dpb = r["dpb"]
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    dmc.encode_decode_bytes(seq[3:4], dpb, Bn.H, Bn.W, 1.0, 1.0)
    torch.cuda.synchronize()
rows = [e for e in prof.key_averages() if e.device_time_total > 0]
tot = sum(e.device_time_total for e in rows)
out = {"mode": mode, "frame_device_ms": tot / 1e3,
       "kernels": [{"name": e.key[:90], "ms": e.device_time_total / 1e3, "share": e.device_time_total / tot, "calls": e.count}
                   for e in sorted(rows, key=lambda e: -e.device_time_total)[:40]]}
for k in out["kernels"][:25]:
    print(f"{k['ms']:8.3f} ms {100 * k['share']:5.1f}% n={k['calls']:4d} {k['name']}")
print("total device ms", out["frame_device_ms"])
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open(f"gpurun_out/kernel_shares_{mode}.json", "w"), indent=1)
