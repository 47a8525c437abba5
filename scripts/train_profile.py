"""Where a training step's time goes (BASELINE config 5 shape: 4 anchors x 256x256, bf16):
    python scripts/train_profile.py -> gpurun_out/train_profile.json"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench as Bn  # noqa: E402
from vcm_ts_b200.models import DMC  # noqa: E402
from vcm_ts_b200.trainer import Trainer  # noqa: E402

dev = torch.device("cuda:0")
sd_p, _ = Bn.make_state_dicts(0)
dmc = DMC(anchor_num=4, precision=sys.argv[1] if len(sys.argv) > 1 else "bf16")
dmc.load_state_dict(sd_p)
dmc = dmc.to(dev)
tr = Trainer(dmc, Bn.TRAIN_LAMBDAS, lr=1e-5)
batch = torch.rand((4, 7, 3, 256, 256), device=dev)
for i in range(3):
    tr.single_step(batch, batch, t_i=i, p_frames=1)
torch.cuda.synchronize()
B = dmc._rt.B
acc = {"wgrad": [0.0, 0], "conv": [0.0, 0], "pack": [0.0, 0]}


def wrap(name, key):
    orig = getattr(B, name)

    def f(*a, **k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = orig(*a, **k)
        e1.record()
        acc[key].append((e0, e1))
        return r
    setattr(B, name, f)


for n, k in (("wgrad", "wgrad"), ("conv", "conv"), ("pack_conv", "pack")):
    wrap(n, k)
x, dpb = batch[:, 1], {"ref_frame": batch[:, 0], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
tr.optimizer.zero_grad()
torch.cuda.synchronize()
t0 = time.perf_counter()
res = tr.model(x, x, dpb)
torch.cuda.synchronize()
t1 = time.perf_counter()
res["loss_to_opt"].backward()
torch.cuda.synchronize()
t2 = time.perf_counter()
tr.optimizer.step()
torch.cuda.synchronize()
t3 = time.perf_counter()
out = {"forward_ms": (t1 - t0) * 1e3, "backward_ms": (t2 - t1) * 1e3, "optimizer_ms": (t3 - t2) * 1e3}
for k, v in acc.items():
    ev = v[2:]
    out[k + "_device_ms"] = sum(a.elapsed_time(b) for a, b in ev)
    out[k + "_calls"] = len(ev)
from torch.profiler import profile, ProfilerActivity  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    tr.single_step(batch, batch, t_i=2, p_frames=1)
    torch.cuda.synchronize()
rows = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:25]
out["top_device_kernels"] = [{"name": e.key[:80], "device_ms": e.device_time_total / 1e3, "calls": e.count} for e in rows]
print(json.dumps(out, indent=1))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/train_profile.json", "w"), indent=1)
