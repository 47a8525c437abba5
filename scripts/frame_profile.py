"""Host-side breakdown of one 1080p P-frame encode+decode: wall time with and without a device
sync per stage (cProfile of the Python/ctypes side + CUDA-event totals)."""
import cProfile, pstats, io, os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench as Bn
from vcm_ts_b200.models import DMC, IntraNoAR
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
sd_p, sd_i = Bn.make_state_dicts(0)
dev = torch.device("cuda:0")
dmc, intra = DMC(precision=prec).eval(), IntraNoAR(precision=prec).eval()
dmc.load_state_dict(sd_p); intra.load_state_dict(sd_i)
dmc, intra = dmc.to(dev), intra.to(dev)
dmc.update(force=True); intra.update(force=True)
seq = Bn.synth_sequence(4, 0).to(dev)
F = torch.nn.functional
def pad(x): return F.pad(x, (0, Bn.WP - Bn.W, 0, Bn.HP - Bn.H))
r, _ = intra.encode_decode_bytes(pad(seq[0:1]), 1.0, Bn.H, Bn.W)
dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
for t in (1, 2):
    r, _ = dmc.encode_decode_bytes(pad(seq[t:t+1]), dpb, Bn.H, Bn.W, 1.0, 1.0); dpb = r["dpb"]
torch.cuda.synchronize()
x = pad(seq[3:4])
out = {}
# stage timing with syncs
t0 = time.time(); enc = dmc.compress(x, dpb, 1.0, 1.0); torch.cuda.synchronize(); t1 = time.time()
dec = dmc.decompress(dpb, enc["bit_stream"], Bn.H, Bn.W, 1.0, 1.0); torch.cuda.synchronize(); t2 = time.time()
out["compress_ms"], out["decompress_ms"], out["bytes"] = (t1 - t0) * 1e3, (t2 - t1) * 1e3, len(enc["bit_stream"])
# rANS alone
rt = dmc._rt; c = rt.coder; s = dmc._last_symbols
t0 = time.time(); c.reset_encoder()
c.encode_channels(s["mv_z"], rt.tables["bit_estimator_z_mv"]); c.encode_planes(s["mv_y_q_w_0"], s["mv_idx_0"], rt.tables["y"])
c.encode_planes(s["mv_y_q_w_1"], s["mv_idx_1"], rt.tables["y"]); c.encode_channels(s["z"], rt.tables["bit_estimator_z"])
c.encode_planes(s["y_q_w_0"], s["y_idx_0"], rt.tables["y"]); c.encode_planes(s["y_q_w_1"], s["y_idx_1"], rt.tables["y"])
bs = c.flush_encoder(); out["rans_encode_ms"] = (time.time() - t0) * 1e3
# GPU-only time of compress (events) without host coder in between is not separable; use launch-free estimate:
pr = cProfile.Profile(); pr.enable()
r, _ = dmc.encode_decode_bytes(x, dpb, Bn.H, Bn.W, 1.0, 1.0); torch.cuda.synchronize()
pr.disable()
sio = io.StringIO(); pstats.Stats(pr, stream=sio).sort_stats("tottime").print_stats(22)
print(sio.getvalue()[:5000])
print(json.dumps(out))
