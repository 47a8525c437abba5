"""HBM roofline check of the enhancement-layer kernels: N x 1080p uint8 frames (default 32, so a launch outlasts the host-side call overhead), boxes uploaded once."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from vcm_ts_b200 import enhancement as E

N, H, W = int(sys.argv[1]) if len(sys.argv) > 1 else 32, 1080, 1920
a = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, device="cuda")
b = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, device="cuda")
lp = [[[100, 100, 400, 300], [900, 500, 1300, 900]]] * N
fc = [[[600, 200, 800, 500]]] * N
out = torch.empty_like(a)
peak = 6530.6
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
res = {}
for name, fn, plan in (("vcm_enh_residual", E.compute_residual, E.prepare_boxes(lp, fc, N, H, W, "cuda")),
                       ("vcm_enh_fuse", E.fuse, E.prepare_boxes(lp, fc, N, H, W, "cuda", 5, 9))):
    for _ in range(5):
        fn(a, b, out=out, plan=plan)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        fn(a, b, out=out, plan=plan)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 50
    gbs = 3 * a.numel() / ms / 1e6
    res[name] = {"ms_per_launch": ms, "frames": N, "algorithmic_bytes": 3 * a.numel(), "achieved_gbs": gbs,
                 "peak_gbs": peak, "frac": gbs / peak}
    print(name, res[name])
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "enh_bench.json"), "w"), indent=1)
