"""bpp / host-coder time of the synthetic workload as a function of the scale applied to the
analysis transforms' last convs (random-init weights give absurd 5-8 bpp streams; the reference's
checkpoints operate at 0.03-0.2 bpp)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench as Bn
from vcm_ts_b200.models import DMC, IntraNoAR

dev = torch.device("cuda:0")
F = torch.nn.functional
seq = F.pad(Bn.synth_sequence(4, 0).to(dev), (0, Bn.WP - Bn.W, 0, Bn.HP - Bn.H))
res = []
for f in [float(v) for v in sys.argv[1:]] or [1.0, 0.2, 0.05]:
    sd_p, sd_i = Bn.make_state_dicts(0)
    for k in ("mv_encoder.6.weight", "mv_encoder.6.bias", "contextual_encoder.conv4.weight", "contextual_encoder.conv4.bias"):
        sd_p[k] = sd_p[k] * f
    for k in ("enc.6.weight", "enc.6.bias"):
        sd_i[k] = sd_i[k] * f
    dmc, intra = DMC(precision="bf16").eval(), IntraNoAR(precision="bf16").eval()
    dmc.load_state_dict(sd_p); intra.load_state_dict(sd_i)
    dmc, intra = dmc.to(dev), intra.to(dev)
    dmc.update(force=True); intra.update(force=True)
    r, blob = intra.encode_decode_bytes(seq[0:1], 1.0, Bn.H, Bn.W)
    ibpp = len(blob) * 8 / (Bn.H * Bn.W)
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    for t in (1, 2, 3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        host, _ = dmc.compress_symbols(seq[t:t + 1], dpb, 1.0, 1.0, want_dpb=False)
        t1 = time.perf_counter()
        bs = dmc.encode_symbols(host)
        t2 = time.perf_counter()
        dpb = dmc.decompress(dpb, bs, Bn.H, Bn.W, 1.0, 1.0)["dpb"]
        torch.cuda.synchronize(); t3 = time.perf_counter()
    zero = sum(int((h == 0).sum()) for h in (host[1], host[2], host[6], host[7])) / sum(h.numel() for h in (host[1], host[2], host[6], host[7]))
    res.append({"factor": f, "i_bpp": ibpp, "p_bpp": len(bs) * 8 / (Bn.H * Bn.W), "zero_frac": zero,
                "rans_encode_ms": (t2 - t1) * 1e3, "decompress_ms": (t3 - t2) * 1e3,
                "max_abs_sym": max(int(h.abs().max()) for h in (host[1], host[2], host[6], host[7]))})
    print(res[-1], flush=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "bpp_calib.json"), "w"), indent=1)
