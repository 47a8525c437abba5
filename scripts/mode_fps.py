"""Per-mode device time of one 1080p P-frame encode+decode (eager launches, one sequence):
    python scripts/mode_fps.py [fp32 parity bf16]  -> gpurun_out/mode_fps.json"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench as Bn  # noqa: E402
from vcm_ts_b200.models import DMC, IntraNoAR  # noqa: E402

modes = sys.argv[1:] or ["fp32", "parity", "bf16"]
dev = torch.device("cuda:0")
sd_p, sd_i = Bn.make_state_dicts(0)
seq = Bn.synth_sequence(4, 0)
pad = lambda x: torch.nn.functional.pad(x, (0, Bn.WP - Bn.W, 0, Bn.HP - Bn.H))  # noqa: E731
fr = pad(seq).to(dev)
out = {}
for mode in modes:
    dmc, intra = DMC(precision=mode).eval(), IntraNoAR(precision=mode).eval()
    dmc.load_state_dict(sd_p)
    intra.load_state_dict(sd_i)
    dmc, intra = dmc.to(dev), intra.to(dev)
    dmc.update(force=True)
    intra.update(force=True)
    torch.cuda.synchronize()
    t0 = time.time()
    r, blob = intra.encode_decode_bytes(fr[0:1], 1.0, Bn.H, Bn.W)
    torch.cuda.synchronize()
    t_i = time.time() - t0
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    times = []
    for t in range(1, 4):
        torch.cuda.synchronize()
        t0 = time.time()
        res, blob = dmc.encode_decode_bytes(fr[t:t + 1], dpb, Bn.H, Bn.W, 1.0, 1.0)
        torch.cuda.synchronize()
        times.append(time.time() - t0)
        dpb = res["dpb"]
    # device-only time of the encoder side (no host coder in between): events around compress_symbols
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    dmc.compress_symbols(fr[3:4], dpb, 1.0, 1.0)
    e1.record()
    torch.cuda.synchronize()
    out[mode] = {"i_frame_wall_s_first_call": t_i, "p_frame_wall_s": times, "p_frame_fps_serial": 1.0 / min(times),
                 "encoder_side_device_ms": e0.elapsed_time(e1), "p_bytes": len(blob)}
    print(mode, out[mode], flush=True)
    del dmc, intra
    torch.cuda.empty_cache()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/mode_fps.json", "w"), indent=1)
