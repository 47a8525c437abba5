"""Where one 1080p P-frame's wall time goes in graph mode: per captured segment the device time
(replay + sync) and the host-coder time that follows it.  Serial single-sequence view."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench as Bn
from vcm_ts_b200 import graphs
from vcm_ts_b200.models import DMC, IntraNoAR

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
sd_p, sd_i = Bn.make_state_dicts(0)
dev = torch.device("cuda:0")
dmc, intra = DMC(precision=prec).eval(), IntraNoAR(precision=prec).eval()
dmc.load_state_dict(sd_p); intra.load_state_dict(sd_i)
dmc, intra = dmc.to(dev), intra.to(dev)
dmc.update(force=True); intra.update(force=True)
dmc.set_cuda_graphs(True)
F = torch.nn.functional
seq = F.pad(Bn.synth_sequence(6, 0).to(dev), (0, Bn.WP - Bn.W, 0, Bn.HP - Bn.H))
log = []


def timed_replay(self, service):
    B = self.B
    for g, req, up, n in self.segs:
        t0 = time.perf_counter()
        g.replay()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        host = None
        if req is not None:
            host = service(req)
            t2 = time.perf_counter()
            up.copy_(host)
            torch.cuda.synchronize()
        t3 = time.perf_counter()
        log.append({"kernels": n, "device_ms": (t1 - t0) * 1e3, "request": req[0] if req else None,
                    "host_ms": ((t2 - t1) * 1e3) if req else 0.0, "upload_ms": ((t3 - t2) * 1e3) if req else 0.0})
        B.launches += n
    return self.result


s = torch.cuda.Stream(dev)
with torch.cuda.stream(s):
    r, _ = intra.encode_decode_bytes(seq[0:1], 1.0, Bn.H, Bn.W)
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    for t in range(1, 5):
        r, _ = dmc.encode_decode_bytes(seq[t:t + 1], dpb, Bn.H, Bn.W, 1.0, 1.0)
        dpb = r["dpb"]
    torch.cuda.synchronize()
    graphs.Program.replay = timed_replay
    x = seq[5:6]
    t0 = time.perf_counter()
    host, _ = dmc.compress_symbols(x, dpb, 1.0, 1.0, want_dpb=False)
    t1 = time.perf_counter()
    bs = dmc.encode_symbols(host)
    t2 = time.perf_counter()
    n_enc = len(log)
    dmc.decompress(dpb, bs, Bn.H, Bn.W, 1.0, 1.0)
    torch.cuda.synchronize()
    t3 = time.perf_counter()
out = {"compress_symbols_ms": (t1 - t0) * 1e3, "rans_encode_ms": (t2 - t1) * 1e3, "decompress_ms": (t3 - t2) * 1e3,
       "bytes": len(bs), "enc_segments": log[:n_enc], "dec_segments": log[n_enc:]}
print(json.dumps(out, indent=1))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "frame_timeline.json"), "w"), indent=1)
