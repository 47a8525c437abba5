#!/usr/bin/env python
"""bench.py -- 1080p DCVC-HEM encode+decode throughput (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--precision bf16|parity|fp32]
                    [--scaling weak|strong] [--impl reference]

One "step" = one synthetic GOP of `--gop` frames (1 I-frame through IntraNoAR + P-frames through DMC),
1920x1080 padded to 1920x1088, of every sequence in flight; each frame goes through the public
`encode_decode`-equivalent call (compress -> real rANS bitstream -> decompress).  The path shards by
independent coding units (a GOP of a sequence: every I-frame resets the decoded-picture buffer), with no
data-path collective:

  --scaling weak   (default; the contract's line): `--seqs-per-gpu S` sequences PER GPU, N*S in total;
  --scaling strong (BASELINE config 4 as written): `--total-seqs 8` sequences in total, their GOPs dealt
                   round-robin to the ranks.

Per GPU, S worker threads (own model instance + CUDA stream each) pull units from a queue, so the
sequential host rANS coder of one unit overlaps the convolutions of the others.  Two timed passes:

  value : frames already resident in HBM, reconstruction left in HBM;
  e2e   : frames start as uint8 HWC in pinned host memory (what a PNG decoder yields; H2D, /255 and
          padding inside the timed region) and every reconstructed frame is converted back to uint8 HWC
          and copied to pinned host memory (D2H inside the timed region).

Both include the D2H of symbols/indexes, the CPU rANS coder and the H2D of decoded symbols: that hand-off
is part of the path (the bitstream is produced and consumed on the host).  Beside the headline (plain-bf16
tensor-core mode, self-decodable streams) the line carries bounded samples of the two symbol-exact modes
(`parity_mode`: tcgen05 on split-bf16 operands; `fp32_mode`: fp32 FMA on CUDA cores), of the same workload
with bpp-calibrated weights, and of one sequence alone.  Prints ONE JSON line (rank 0).
`--impl reference` times the CPU oracle port instead (rank 0 only).
"""
import argparse
import json
import os
import queue
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

H, W = 1080, 1920
HP, WP = 1088, 1920
METRIC = "1080p GOP encode+decode frames/sec"
# algorithmic conv work per padded 1080p frame, FLOP
P_FRAME_FLOP = 12.101e12
I_FRAME_FLOP = 4.275e12


def synth_sequence(n_frames, seed, device="cpu"):
    """Smooth moving texture in [0,1] + 1 % noise, [n,3,1080,1920] fp32."""
    g = torch.Generator(device="cpu").manual_seed(1234 + seed)
    margin = 4 * n_frames + 16
    ch, cw = (H + margin) // 16 + 2, (W + margin) // 16 + 2
    coarse = torch.rand((1, 3, ch, cw), generator=g)
    mid = torch.rand((1, 3, ch * 4, cw * 4), generator=g)
    F = torch.nn.functional
    base = 0.75 * F.interpolate(coarse, scale_factor=16, mode="bicubic", align_corners=False) + \
        0.25 * F.interpolate(mid, scale_factor=4, mode="bicubic", align_corners=False)
    base = base.clamp(0, 1)
    out = torch.empty((n_frames, 3, H, W))
    ox = oy = 0
    for t in range(n_frames):
        crop = base[0, :, oy:oy + H, ox:ox + W]
        out[t] = (crop + 0.01 * torch.randn(crop.shape, generator=g)).clamp(0, 1)
        ox += 1 + (t % 3)
        oy += 1 + ((t + 1) % 2)
    return out.to(device)


def make_state_dicts(seed=0, calibrate=None):
    """Random-init weights of the reference architecture: its own init (xavier_normal gain sqrt 2,
    bias 0.01) with conv weights x0.7 (un-scaled random init is numerically chaotic).
    calibrate=f scales the last conv of the three analysis transforms by f: random-init latents are
    large (7-13 bpp, most symbols on the escape path of the coder); f = 0.05 gives the ~0.2-0.3 bpp
    the reference's trained checkpoints operate at (scripts/bpp_calib.py)."""
    from vcm_ts_b200.models import DMC, IntraNoAR
    torch.manual_seed(seed)
    sds = []
    for cls in (DMC, IntraNoAR):
        m = cls()
        sd = m.state_dict()
        for k, v in sd.items():
            if k.endswith(".weight") and v.dim() == 4:
                v.mul_(0.7)
        sds.append(sd)
    if calibrate is not None:
        for k in ("mv_encoder.6.weight", "mv_encoder.6.bias", "contextual_encoder.conv4.weight",
                  "contextual_encoder.conv4.bias"):
            sds[0][k] = sds[0][k] * calibrate
        for k in ("enc.6.weight", "enc.6.bias"):
            sds[1][k] = sds[1][k] * calibrate
    return sds


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


class HostLoad:
    """Host CPU utilisation over a region (all cores, from /proc/stat) -- the limiter of the sharded path
    is the host (Python launch threads + the sequential rANS coder), not NVLink."""

    @staticmethod
    def _read():
        with open("/proc/stat") as f:
            v = [int(x) for x in f.readline().split()[1:9]]
        idle = v[3] + v[4]
        return sum(v), idle

    def start(self):
        self.t0 = self._read()
        return self

    def stop(self):
        t1 = self._read()
        tot, idle = t1[0] - self.t0[0], t1[1] - self.t0[1]
        return None if tot <= 0 else round(1.0 - idle / tot, 4)


def code_gop(dmc, intra, frames_src, gop, q, host_out=None, pinned_in=False, dev=None):
    """Encode+decode one GOP.  frames_src[t]: padded float [3,1088,1920] on the device, or (pinned_in)
    a pinned-host uint8 [1080,1920,3] frame -- what a PNG decoder hands over -- that is copied to
    the device, normalised and padded there (vcm_ts_b200.frame_io).  host_out: pinned uint8
    [gop,1080,1920,3] that receives every reconstruction.  Returns (bits, last reconstruction)."""
    from vcm_ts_b200 import frame_io
    bits = 0
    dpb = None
    for t in range(gop):
        if pinned_in:
            x = frame_io.frames_from_u8(frames_src[t], device=dev)
        else:
            x = frames_src[t][None]
        if t == 0:
            r, _ = intra.encode_decode_bytes(x, q, H, W)
            rec = r["x_hat"]
            dpb = {"ref_frame": rec, "ref_feature": None, "ref_y": None, "ref_mv_y": None}
        else:
            r, _ = dmc.encode_decode_bytes(x, dpb, H, W, q, q)
            dpb = r["dpb"]
            rec = dpb["ref_frame"]
        bits += r["bit"]
        if host_out is not None:
            frame_io.frames_to_u8(rec, H, W, out=host_out[t:t + 1])
    return bits, rec


PEAK_TC, PEAK_HBM = 1400.0, 6500.0      # TFLOP/s, GB/s; replaced by MEASURED_PEAKS.json in main()


def conv_roofline(dmc, x, dpb, q, passes=3):
    """Instrumented passes: CUDA events around every conv launch of one P-frame encode+decode, on the
    launching stream.  Launches are issued eagerly (graphs off); ONE untimed eager pass first (the eager
    allocator's cudaMalloc / lazy module loads must not be timed), then `passes` timed passes of which the
    per-launch MINIMUM is kept (a host stall between the start event and the launch inflates single
    samples).  Returns None-able fields plus the wall time of the frame for the sanity check."""
    B = dmc._rt.B
    orig = B.conv
    graphs = dmc._graphs
    dmc.set_cuda_graphs(False)      # per-launch events need the launches issued one by one
    all_recs = []
    walls = []
    try:
        dmc.encode_decode_bytes(x, dpb, H, W, q, q)       # untimed eager pass
        torch.cuda.synchronize()
        for _ in range(passes):
            recs = []

            def timed_conv(xv, pc, *a, _recs=recs, **kw):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                # the host needs ~15 us of Python per launch; when the device has caught up with it the gap
                # between the event and the kernel would be timed: keep the device behind the host
                if not _recs or _recs[-1][1].query():
                    torch.cuda._sleep(10_000_000)
                e0.record(torch.cuda.current_stream(B.device))
                out = orig(xv, pc, *a, **kw)
                e1.record(torch.cuda.current_stream(B.device))
                stride = kw.get("stride", 1)
                ho, wo = (xv.H + stride - 1) // stride, (xv.W + stride - 1) // stride
                flop = 2.0 * xv.N * ho * wo * pc.K * pc.K * pc.Cin * pc.Cout
                on_tc = pc.w_umma is not None and (B.precision == "parity" or
                                                   (xv.dtype == torch.bfloat16 and kw.get("pre") is None))
                _recs.append((e0, e1, flop, on_tc, (pc.K, stride, pc.Cin, pc.Cout, xv.H, xv.W)))
                return out

            B.conv = timed_conv
            w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            dmc.encode_decode_bytes(x, dpb, H, W, q, q)
            torch.cuda.synchronize()
            walls.append((time.perf_counter() - t0) * 1e3)
            B.conv = orig
            all_recs.append([(e0.elapsed_time(e1), f, tc, shp) for e0, e1, f, tc, shp in recs])
    finally:
        B.conv = orig
        dmc.set_cuda_graphs(graphs)
    n = len(all_recs[0])
    if any(len(r) != n for r in all_recs):
        return {"error": "launch sequences of the instrumented passes differ"}
    recs = [(min(p[i][0] for p in all_recs),) + all_recs[0][i][1:] for i in range(n)]
    tc_t = sum(ms for ms, _, tc, _ in recs if tc) * 1e-3
    tc_f = sum(f for _, f, tc, _ in recs if tc)
    all_t = sum(ms for ms, _, _, _ in recs) * 1e-3
    all_f = sum(f for _, f, _, _ in recs)
    by_shape = {}
    for ms, f, tc, shp in recs:
        k = ("tc" if tc else "simt") + ":k%d s%d %d->%d @%dx%d" % shp
        a = by_shape.setdefault(k, [0.0, 0.0, 0])
        a[0] += ms
        a[1] += f
        a[2] += 1
    # SURVEY 8(d) "which roofline": per layer max(flops / tensor peak, bytes / HBM peak), activations read once
    # and written once in bf16 (weights and residual operands not counted)
    roof_s = 0.0
    for _, f, _, (k, stride, cin, cout, hh, ww) in recs:
        ho, wo = (hh + stride - 1) // stride, (ww + stride - 1) // stride
        nbytes = 2.0 * (hh * ww * cin + ho * wo * cout)
        roof_s += max(f / (PEAK_TC * 1e12), nbytes / (PEAK_HBM * 1e9))
    return dict(tc_s=tc_t, tc_flop=tc_f, all_s=all_t, all_flop=all_f, n=n, per_layer_roof_s=roof_s,
                n_tc=sum(1 for r in recs if r[2]), by_shape=by_shape, frame_wall_ms=min(walls), passes=passes)


# ------------------------------------------------------------------------------------ CPU oracle arm
ORACLE_CROPS = ((576, 960), (384, 640), (256, 448))


def oracle_sample(sd_p, sd_i, threads, hh=576, ww=960, with_i=True):
    """CPU oracle port (reference algorithm, PyTorch CPU ops + C rANS) on an hh x ww crop of the synthetic
    sequence: seconds per I-frame and per P-frame encode+decode, converted to 1080p-equivalent seconds by
    the pixel ratio (the nets are fully convolutional, cost is linear in pixels)."""
    from oracle import dcvc_oracle as O
    torch.set_num_threads(threads)
    odmc, ointra = O.OracleDMC(sd_p), O.OracleIntra(sd_i)
    odmc.update()
    ointra.update()
    seq = synth_sequence(2, 0)[:, :, :hh, :ww].contiguous()
    t0 = time.time()
    r, _ = ointra.encode_decode_bytes(seq[0:1], 1.0, hh, ww)
    t_i = time.time() - t0
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    t0 = time.time()
    odmc.encode_decode_bytes(seq[1:2], dpb, hh, ww, 1.0, 1.0)
    t_p = time.time() - t0
    ratio = (HP * WP) / float(hh * ww)
    return dict(p_frame_s_1080p=t_p * ratio, i_frame_s_1080p=t_i * ratio, cpu_s=t_i + t_p,
                sample=f"1 I-frame + 1 P-frame encode+decode (real rANS streams) at {ww}x{hh}, f32, "
                       f"{threads} thread(s), scaled x{ratio:.2f} to 1920x1088; GOP value = gop / (I + (gop-1) P)")


def reference_arm(a, gop):
    """`--impl reference`: the reference's algorithm on the host cores (kind "port": the reference has no
    installable package -- no setup.py / pyproject -- so the timed code is oracle/dcvc_oracle.py, the
    torch-CPU restatement that is pinned bit-exactly to the reference)."""
    sd_p, sd_i = make_state_dicts(0)
    threads = os.cpu_count() or 1
    # calibration = the warm-up: choose the crop so that W + K samples end within ~3 minutes
    crop = ORACLE_CROPS[0]
    t0 = time.time()
    first = oracle_sample(sd_p, sd_i, threads, *crop)
    t_first = time.time() - t0
    n_runs = max(1, a.steps) + max(0, a.warmup - 1)
    for c in ORACLE_CROPS[1:]:
        if t_first * n_runs <= 170.0:
            break
        t_first *= (c[0] * c[1]) / float(crop[0] * crop[1])
        crop = c
    for _ in range(max(0, a.warmup - 1)):
        oracle_sample(sd_p, sd_i, threads, *crop)
    vals, t_all, last = [], [], first
    for _ in range(max(1, a.steps)):
        t0 = time.time()
        last = oracle_sample(sd_p, sd_i, threads, *crop)
        t_all.append(time.time() - t0)
        vals.append(gop / (last["i_frame_s_1080p"] + (gop - 1) * last["p_frame_s_1080p"]))
    v = sum(vals) / len(vals)
    cfg = {"workload": f"DCVC-HEM GOP-{gop} 1080p encode+decode, CPU oracle port of the reference path: per step "
                       f"1 I-frame + 1 P-frame at a {crop[1]}x{crop[0]} crop, extrapolated by the pixel ratio",
           "gop": gop, "crop": f"{crop[1]}x{crop[0]}", "precision": "f32", "threads": threads,
           "cuda_graphs": False, "parallelism": "host cores only (torch intra-op threads)",
           "weights": "random init (reference init, conv x0.7)", "same_workload_as": "bench.py default arm"}
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * sum(t_all) / len(t_all),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": "port",
                         "sample": last["sample"], "cores_note": "os.cpu_count() of this box"},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------------------------ GPU arm helpers
def build_workers(n, precision, sd_p, sd_i, dev, graphs, gop):
    from vcm_ts_b200.models import DMC, IntraNoAR
    workers = []
    for _ in range(n):
        dmc, intra = DMC(precision=precision).eval(), IntraNoAR(precision=precision).eval()
        dmc.load_state_dict(sd_p)
        intra.load_state_dict(sd_i)
        dmc, intra = dmc.to(dev), intra.to(dev)
        dmc.update(force=True)
        intra.update(force=True)
        dmc.set_cuda_graphs(bool(graphs))
        # the I-frame (1 frame in 32) stays eager here: its captured programs (192-channel activations) would add
        # ~4 GB of private graph pool per worker, and 8 workers x (DMC + IntraNoAR) pools overflow the 180 GB
        intra.set_cuda_graphs(bool(graphs) and os.environ.get("VCM_BENCH_INTRA_GRAPHS", "0") == "1")
        workers.append(dict(dmc=dmc, intra=intra, stream=torch.cuda.Stream(dev), bits=0,
                            host_out=torch.empty((gop, H, W, 3), dtype=torch.uint8).pin_memory()))
    return workers


def run_pool(workers, units, fn, dev):
    """Every worker thread (own CUDA stream) pulls units from one queue until it is empty."""
    torch.cuda.synchronize()
    qu = queue.SimpleQueue()
    for u in units:
        qu.put(u)
    errs = []

    def body(w):
        try:
            torch.cuda.set_device(dev)
            with torch.cuda.stream(w["stream"]):
                while True:
                    try:
                        u = qu.get_nowait()
                    except queue.Empty:
                        break
                    fn(w, u)
                w["stream"].synchronize()
        except BaseException as e:   # noqa: BLE001
            errs.append(e)
    n = min(len(workers), len(units))
    if n <= 1:
        body(workers[0])
    else:
        ts = [threading.Thread(target=body, args=(w,)) for w in workers[:n]]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
    if errs:
        raise errs[0]
    torch.cuda.synchronize()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--gop", type=int, default=32)
    ap.add_argument("--precision", default=os.environ.get("VCM_B200_PRECISION", "bf16"))
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the bounded samples of the other numerics modes / calibrated weights / one sequence")
    ap.add_argument("--cuda-graphs", type=int, default=1,
                    help="replay each frame's launch sequence from captured CUDA graphs (vcm_ts_b200/graphs.py)")
    ap.add_argument("--seqs-per-gpu", type=int, default=8,
                    help="worker threads per GPU (own model instance + CUDA stream each); with --scaling weak "
                         "also the number of independent sequences per GPU")
    ap.add_argument("--scaling", default="weak", choices=("weak", "strong"))
    ap.add_argument("--total-seqs", type=int, default=8, help="--scaling strong: sequences in total (config 4)")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    gop = a.gop

    if a.impl == "reference":
        if rank == 0:
            reference_arm(a, gop)
        return

    S = max(1, a.seqs_per_gpu)
    strong = a.scaling == "strong"
    n_seqs = a.total_seqs if strong else world * S
    cfg = {"workload": f"DCVC-HEM GOP-{gop} 1080p (1920x1088 padded) sequence encode+decode: "
                       f"1 I-frame (IntraNoAR) + {gop - 1} P-frames (DMC), real rANS bitstreams; "
                       + (f"{n_seqs} sequences in total, GOPs dealt round-robin to the ranks (config 4)" if strong
                          else f"{S} sequences per GPU"),
           "gop": gop, "sequences_total": n_seqs, "workers_per_gpu": S, "precision": a.precision,
           "cuda_graphs": bool(a.cuda_graphs), "parallelism": f"gop-shard x{world}",
           "weights": "random init (reference init, conv x0.7)",
           "l2": "working set per frame >> 126 MB L2 (1080p activations), no explicit flush"}
    sd_p, sd_i = make_state_dicts(0)

    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", init_method="env://", device_id=torch.device(f"cuda:{local}"))
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    from vcm_ts_b200 import frame_io, sharding
    q = 1.0

    # ---- units: (sequence id, step).  weak: this rank owns S sequences; strong: global unit u -> rank u % world
    if strong:
        all_units = [(s, g) for g in range(a.steps) for s in range(n_seqs)]
        my_units = [all_units[u] for u in sharding.assign_units(len(all_units), world, rank)]
    else:
        my_seqs = sharding.assign_units(n_seqs, world, rank)
        my_units = [(s, g) for g in range(a.steps) for s in my_seqs]
    my_seq_ids = sorted({s for s, _ in my_units}) or [rank % max(1, n_seqs)]
    workers = build_workers(S, a.precision, sd_p, sd_i, dev, a.cuda_graphs, gop)
    # 8-bit frames, as video is: uint8 HWC in pinned host memory for the e2e arm, and the same
    # frames normalised + padded on the device for the resident arm
    src = {}
    for sid in my_seq_ids:
        sh = (synth_sequence(gop, sid) * 255.0).round().to(torch.uint8).permute(0, 2, 3, 1).contiguous().pin_memory()
        sd = torch.cat([frame_io.frames_from_u8(sh[t0:t0 + 4], device=dev) for t0 in range(0, gop, 4)])
        src[sid] = (sh, sd)
    dmc, intra = workers[0]["dmc"], workers[0]["intra"]
    seq_dev = src[my_seq_ids[0]][1]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item()

    def one_resident(w, u, n=None):
        w["bits"], _ = code_gop(w["dmc"], w["intra"], src[u[0]][1], n or gop, q)

    def one_e2e(w, u):
        w["bits"], _ = code_gop(w["dmc"], w["intra"], src[u[0]][0], gop, q, host_out=w["host_out"],
                                pinned_in=True, dev=dev)

    def warm(ws, n_gops):
        # at least 3 short GOPs (I + 2 P) per worker: every per-frame program has its eager pass and its
        # capture pass (vcm_ts_b200/graphs.py) before anything is timed
        for _ in range(n_gops):
            run_pool(ws, [(my_seq_ids[i % len(my_seq_ids)], -1) for i in range(len(ws))],
                     lambda w, u: one_resident(w, u, min(gop, 3)), dev)

    warm(workers, max(a.warmup, 3))
    backends = [w[m]._rt.B for w in workers for m in ("dmc", "intra")]
    clocks = ClockSampler(local)
    clocks.start()
    l0 = sum(b.launches for b in backends)
    load = HostLoad().start()
    ms_res = timed(lambda: run_pool(workers, my_units, one_resident, dev))
    host_util = load.stop()
    launches = sum(b.launches for b in backends) - l0
    ms_e2e = timed(lambda: run_pool(workers, my_units, one_e2e, dev))
    clk = clocks.stop()
    frames_total = gop * a.steps * n_seqs
    value = frames_total / (ms_res * 1e-3)
    e2e = frames_total / (ms_e2e * 1e-3)
    bpp = workers[0]["bits"] / float(gop * H * W)
    extras = {}

    train = None
    if not a.no_extras:
        train = train_sample(dev, rank, world, local, sd_p)      # every rank takes part (DDP all-reduce)
    if rank == 0 and not a.no_extras:
        extras = run_extras(a, workers, src, my_seq_ids, dev, gop, q, sd_p, sd_i)
        extras["train_step"] = train

    roof = None
    cpu = None
    if rank == 0:
        x1, x0 = seq_dev[1:2], seq_dev[0:1]
        r0, _ = intra.encode_decode_bytes(x0, q, H, W)
        dpb = {"ref_frame": r0["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        global PEAK_TC, PEAK_HBM
        PEAK_TC = peaks.get("bf16_tflops_sustained", PEAK_TC)
        PEAK_HBM = peaks.get("hbm_gbs", PEAK_HBM)
        roof = roofline_entry(conv_roofline(dmc, x1, dpb, q), peaks, a.precision)
        if not a.no_cpu_baseline:
            threads = os.cpu_count() or 1
            ob = oracle_sample(sd_p, sd_i, threads)
            per_gop = ob["i_frame_s_1080p"] + (gop - 1) * ob["p_frame_s_1080p"]
            cpu = {"value": gop / per_gop, "unit": "frames/s", "cores": threads, "kind": "port",
                   "sample": ob["sample"], "cores_note": "os.cpu_count() of this box"}
            # the reference's own inference setting (video_coder.py:177 torch.set_num_threads(1)), on a smaller crop
            o1 = oracle_sample(sd_p, sd_i, 1, *ORACLE_CROPS[2])
            cpu["threads_1"] = {"value": gop / (o1["i_frame_s_1080p"] + (gop - 1) * o1["p_frame_s_1080p"]),
                                "unit": "frames/s", "cores": 1, "sample": o1["sample"]}
            torch.set_num_threads(threads)
    if rank == 0:
        seq_bytes = n_seqs * gop * 3 * H * W      # whole job, per step: uint8 frames in, uint8 reconstructions out
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms_res / a.steps, "higher_is_better": True,
            "scaling": a.scaling, "vs_baseline": None,
            "dtype": "bf16" if a.precision == "bf16" else ("bf16x3" if a.precision == "parity" else "f32"),
            "data": "synthetic", "config": cfg, "clocks": clk, "bpp": bpp,
            "parity": ("benched mode is plain bf16: streams are self-decodable only, symbols differ from the fp32 "
                       "reference (see tests/test_gpu_e2e.py); parity_mode / fp32_mode below are the symbol-exact modes")
            if a.precision == "bf16" else "symbol-exact mode (tests/test_gpu_e2e.py exact_modes)",
            "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": seq_bytes,
                    "d2h_bytes_per_step": seq_bytes, "ms_per_step": ms_e2e / a.steps},
            "gpu_launches": launches, "host": {"cpu_count": os.cpu_count(), "threads_per_rank": S,
                                               "cpu_utilisation_timed_region": host_util},
            "roofline": roof, "cpu_baseline": cpu}
        line.update(extras)
        print(json.dumps(line))
    barrier()
    if world > 1:
        dist.destroy_process_group()


def roofline_entry(rf, peaks, precision):
    peak = peaks.get("bf16_tflops_sustained", 1400.0)
    if rf.get("error"):
        return {"bound": "tensor", "achieved": None, "peak": peak, "unit": "TFLOP/s", "frac": None, "traffic": None,
                "reason": rf["error"]}
    use_tc = rf["tc_s"] > 0
    ach = (rf["tc_flop"] / rf["tc_s"] if use_tc else rf["all_flop"] / max(rf["all_s"], 1e-9)) / 1e12
    roof = {"bound": "tensor",
            "kernel": "conv_halo_kernel / conv_rows_kernel / conv_tc_kernel (tcgen05 implicit GEMM, all launches of one P-frame)"
            if use_tc else "conv_simt_kernel",
            "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
            "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback 1400 (sustained)",
            "flop_convention": "algorithmic conv FLOPs (2*MAC of the fp32 convolution); the parity mode issues 6 "
                               "bf16 MMAs per product, so its tensor-pipe work is 6x this" if precision == "parity"
            else "algorithmic conv FLOPs (2*MAC)",
            "traffic": None, "launches_timed": rf["n_tc"] if use_tc else rf["n"],
            "timing": f"per-launch minimum over {rf['passes']} eager passes after one untimed eager pass",
            "per_layer_bound": {"what": "sum over the conv launches of one P-frame of max(flops/tensor peak, "
                                        "activation bytes/HBM peak) / measured conv time (SURVEY 8(d))",
                                "roofline_ms": rf["per_layer_roof_s"] * 1e3, "measured_ms": rf["all_s"] * 1e3,
                                "frac": rf["per_layer_roof_s"] / max(rf["all_s"], 1e-9), "hbm_peak_gbs": PEAK_HBM},
            "conv_share_of_p_frame": {"tc_ms": rf["tc_s"] * 1e3, "all_conv_ms": rf["all_s"] * 1e3,
                                      "frame_wall_ms": rf["frame_wall_ms"],
                                      "all_conv_tflops": rf["all_flop"] / max(rf["all_s"], 1e-9) / 1e12}}
    if rf["all_s"] * 1e3 > rf["frame_wall_ms"]:
        # the launches of a frame cannot take longer than the frame: the samples are instrument artefacts
        roof.update({"achieved": None, "frac": None,
                     "reason": f"sum of timed conv launches {rf['all_s'] * 1e3:.1f} ms exceeds the frame's own wall "
                               f"time {rf['frame_wall_ms']:.1f} ms"})
    top = sorted(rf["by_shape"].items(), key=lambda kv: -kv[1][0])
    if use_tc and top:
        # the launch shape that takes most of the frame (3x3 64->64 at 1080p), live-timed here; its DRAM
        # traffic per launch is NOT measured in this run: it comes from the committed `ncu --set full` capture
        k0, v0 = top[0]
        dom = {"shape": k0, "launches": v0[2], "ms_per_launch": v0[0] / v0[2],
               "achieved": v0[1] / max(v0[0], 1e-9) / 1e9, "unit": "TFLOP/s",
               "frac": v0[1] / max(v0[0], 1e-9) / 1e9 / peak, "traffic": None}
        for name in ("ncu_conv64_rows_r02.json", "ncu_conv64_halo_r02.json", "ncu_conv64_halo_r01.json"):
            try:
                prof = json.load(open(os.path.join(ROOT, "profiles", name)))[0]
                mb = lambda key: float(next(v for k, v in prof.items() if k.startswith(key)))   # noqa: E731
                dom["traffic"] = (mb("dram__bytes_read.sum") + mb("dram__bytes_write.sum")) * 1e6
                dom["traffic_source"] = f"from file profiles/{name} (ncu --set full of the 3x3 64->64 + residual " \
                                        "launch at 1088x1920), not measured in this run"
                dom["algorithmic_bytes"] = 3 * 64 * 2 * HP * WP
                break
            except Exception:
                pass
        roof["dominant_launch"] = dom
        roof["traffic"] = dom["traffic"]
        roof["traffic_source"] = dom.get("traffic_source")
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"conv_by_shape_{precision}.json"), "w") as f:
        json.dump([{"shape": k, "ms": v[0], "tflops": v[1] / max(v[0], 1e-9) / 1e9, "calls": v[2]}
                   for k, v in top], f, indent=1)
    return roof


TRAIN_LAMBDAS = (85.0, 170.0, 380.0, 840.0)


def train_sample(dev, rank, world, local, sd_p, steps=6, warm=3):
    """BASELINE config 5: the `trainer_multi` step on the drop-in DMC -- a [4,7,3,256,256] random batch per rank
    (4 anchors x 7 frames of 256x256 crops), stage ['1','all','single_multi','rec','all',lr,.,'false'] = one
    P-frame forward / backward / AdamW step per window, bf16 training precision, DistributedDataParallel over
    NCCL with bf16 gradient buckets (the path's only collective).  The perceptual detector loss is off
    (torchvision / ultralytics code outside the hot path; no pretrained weights on the box).  Times `steps`
    optimizer steps with CUDA events (max over ranks), and the same without the gradient all-reduce
    (DDP no_sync) to expose the collective's cost."""
    import torch.distributed as dist
    from vcm_ts_b200.models import DMC
    from vcm_ts_b200.trainer import Trainer
    torch.manual_seed(1000 + rank)
    dmc = DMC(anchor_num=4, precision="bf16")
    dmc.load_state_dict(sd_p)
    dmc = dmc.to(dev)
    tr = Trainer(dmc, TRAIN_LAMBDAS, lr=1e-5, world_size=world, device_ids=[local] if world > 1 else None)
    batch = torch.rand((4, 7, 3, 256, 256), device=dev)

    def run(n, sync):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0.record()
        for i in range(n):
            tr.single_step(batch, batch, t_i=i % 6, p_frames=1, sync=sync)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item() / n

    run(warm, True)
    l0 = dmc._rt.B.launches
    ms_sync = run(steps, True)
    launches = (dmc._rt.B.launches - l0) // steps
    ms_nosync = run(steps, False) if world > 1 else ms_sync
    nparam = sum(p.numel() for p in dmc.parameters())
    res = {"value": world * 4 * 1e3 / ms_sync, "unit": "training frames/s (4 anchors x 256x256 per step per GPU)",
           "ms_per_step": ms_sync, "ms_per_step_without_allreduce": ms_nosync,
           "allreduce_ms": max(0.0, ms_sync - ms_nosync), "n_gpus": world, "precision": "bf16",
           "gradient_bytes_bf16": 2 * nparam, "own_kernel_launches_per_step": launches,
           "sample": f"{steps} optimizer steps (forward + backward + DDP all-reduce + AdamW) after {warm} warm-up "
                     "steps on a random [4,7,3,256,256] batch per rank; perceptual loss off"}
    del tr, dmc
    torch.cuda.empty_cache()
    return res


def run_extras(a, workers, src, my_seq_ids, dev, gop, q, sd_p, sd_i):
    """Bounded samples next to the headline (rank 0, after the timed passes): one sequence alone (config 2),
    the two symbol-exact numerics modes, and bpp-calibrated weights."""
    out = {}
    S = len(workers)
    dmc, intra = workers[0]["dmc"], workers[0]["intra"]
    seq_dev = src[my_seq_ids[0]][1]

    def wall_fps(fn, frames):
        torch.cuda.synchronize()
        t0 = time.time()
        fn()
        torch.cuda.synchronize()
        return frames / (time.time() - t0)

    # ---- ONE sequence alone on the GPU (BASELINE config 2).  serial: frame-by-frame encode_decode, the GPU idles
    # while the frame's symbols are in the host coder; pipelined: vcm_ts_b200.pipeline.GopPipeline overlaps
    # encoder / host coder / decoder across the frames of the GOP (same bytes, same frames)
    single = {"serial": None, "pipelined": None}
    n1 = min(gop, 16)
    single["serial"] = wall_fps(lambda: code_gop(dmc, intra, seq_dev, n1, q), n1)
    if S > 1:
        from vcm_ts_b200.pipeline import GopPipeline
        w0, w1 = workers[0], workers[1]
        pipe = GopPipeline((w1["dmc"], w1["intra"]), (w0["dmc"], w0["intra"]), dev)
        frames = [seq_dev[t:t + 1] for t in range(n1)]
        for _ in range(2):      # eager pass, then the pass that captures the encoder-chain graphs
            pipe.code_gop(frames[:4], H, W, q, q, q)
        single["pipelined"] = wall_fps(lambda: pipe.code_gop(frames, H, W, q, q, q), n1)
        # the same ONE sequence when it is long: its GOPs are independent units (every I-frame resets the decoded
        # picture buffer, video_coder.py:122-130), so S of them are coded concurrently by the S workers
        sid = my_seq_ids[0]

        def one(w, u):
            w["bits"], _ = code_gop(w["dmc"], w["intra"], src[sid][1], n1, q)
        single["gop_interleaved"] = wall_fps(lambda: run_pool(workers, [(sid, g_) for g_ in range(S)], one, dev),
                                             S * n1)
        single["note"] = (f"serial / pipelined: one GOP of {n1} frames alone on the GPU; gop_interleaved: {S} GOPs of "
                          f"{n1} frames of the same sequence in flight")
    out["single_sequence_frames_per_s"] = single

    def sample_mode(precision, n_workers, n_frames, sds, label):
        ws = build_workers(n_workers, precision, sds[0], sds[1], dev, a.cuda_graphs, n_frames)
        units = [(my_seq_ids[i % len(my_seq_ids)], 0) for i in range(n_workers)]

        def one(w, u, n=n_frames):
            w["bits"], _ = code_gop(w["dmc"], w["intra"], src[u[0]][1], n, q)
        for _ in range(3):
            run_pool(ws, units, lambda w, u: one(w, u, 3), dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        run_pool(ws, units, one, dev)
        e1.record()
        torch.cuda.synchronize()
        fps = n_workers * n_frames / (e0.elapsed_time(e1) * 1e-3)
        res = {"value": fps, "unit": "frames/s", "precision": precision, "bpp": ws[0]["bits"] / float(n_frames * H * W),
               "sample": f"{n_workers} sequences x GOP-{n_frames} (1 I + {n_frames - 1} P), frames resident in HBM, "
                         f"CUDA graphs {'on' if a.cuda_graphs else 'off'}; {label}"}
        del ws
        torch.cuda.empty_cache()
        return res

    if a.precision != "parity":
        out["parity_mode"] = sample_mode("parity", min(S, 4), 8, (sd_p, sd_i),
                                         "tcgen05 convs on split-bf16 operands (6 MMAs per product), symbol-exact")
    if a.precision != "fp32":
        out["fp32_mode"] = sample_mode("fp32", min(S, 2), 4, (sd_p, sd_i),
                                       "fp32 FMA convs on CUDA cores (conv_simt_kernel), symbol-exact")
    cal = make_state_dicts(0, calibrate=0.05)
    out["calibrated_bpp"] = sample_mode(a.precision, S, 8, cal,
                                        "last analysis convs x0.05: the bpp range trained checkpoints code at, "
                                        "so the host rANS coder sees mostly in-table symbols")
    return out


if __name__ == "__main__":
    main()
