"""HBM-bound kernels of the path at BASELINE config 3 shapes (1920x1088, batch 8 for the SpyNet glue):
achieved algorithmic GB/s against the measured copy bandwidth (MEASURED_PEAKS.json hbm_gbs).

    python scripts/motion_bench.py      -> gpurun_out/motion_bench.json
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from vcm_ts_b200.backend import CudaBackend, View  # noqa: E402

H, W = 1088, 1920
bf = torch.bfloat16


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    B = CudaBackend("cuda:0", "bf16")
    peak = 6500.0
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    g = torch.Generator().manual_seed(0)
    res = []

    def rec(name, ms, nbytes):
        r = dict(kernel=name, ms=ms, algorithmic_gb=nbytes / 1e9, gbs=nbytes / ms / 1e6, frac_of_measured_hbm=nbytes / ms / 1e6 / peak)
        res.append(r)
        print("%-44s %8.4f ms %7.3f GB %8.0f GB/s  %.2f of measured HBM" % (name, ms, r["algorithmic_gb"], r["gbs"], r["frac_of_measured_hbm"]))

    # working sets > 126 MB L2 everywhere (or rotated buffers)
    flow = View((torch.randn((1, H, W, 2), generator=g) * 3).cuda())
    feats = [View(torch.randn((1, H, W, 64), device="cuda").to(bf)) for _ in range(3)]
    outs = [View(torch.empty((1, H, W, 64), device="cuda", dtype=bf)) for _ in range(3)]
    i = [0]

    def warp64():
        i[0] += 1
        B.warp(feats[i[0] % 3], flow, out=outs[i[0] % 3])
    rec("flow_warp 64ch bf16 1088x1920 (context1)", timeit(warp64), (2 * 64 * 2 + 8) * H * W)

    flow8 = View((torch.randn((8, H, W, 2), generator=g) * 3).cuda())
    im8 = View(torch.rand((8, H, W, 3), generator=g).cuda())
    buf8 = View(torch.empty((8, H, W, 8), device="cuda"))
    rec("flow_warp 3ch fp32 B=8 (SpyNet level 3)", timeit(lambda: B.warp(im8, flow8, out=buf8.sl(3, 6))),
        (2 * 3 * 4 + 8) * 8 * H * W)
    rec("avgpool2 3ch fp32 B=8 (pyramid)", timeit(lambda: B.avgpool(im8)), int(1.25 * 3 * 4 * 8 * H * W))
    half = View((torch.randn((8, H // 2, W // 2, 2), generator=g)).cuda())
    rec("flow_up2 x2 B=8 (544x960 -> 1088x1920)", timeit(lambda: B.flow_up2(half, out=buf8.sl(6, 8))),
        int(1.25 * 2 * 4 * 8 * H * W))
    rec("flow_down2 /2 B=8", timeit(lambda: B.flow_down2(flow8)), int(1.25 * 2 * 4 * 8 * H * W))
    x32 = [View(torch.randn((1, H, W, 32), device="cuda").to(bf)) for _ in range(4)]

    def mp():
        i[0] += 1
        B.maxpool(x32[i[0] % 4])
    rec("maxpool2 32ch bf16 1088x1920", timeit(mp), int(1.25 * 32 * 2 * H * W))
    w1 = torch.randn((2, 32), device="cuda")
    w2 = torch.randn((32, 2), device="cuda")

    def se():
        i[0] += 1
        B.se_scale(x32[i[0] % 4], w1, w2)
    rec("SE mean + fc 32ch bf16 1088x1920", timeit(se), 32 * 2 * H * W)
    nchw = torch.randn((1, 64, H, W), device="cuda")
    rec("nchw fp32 -> nhwc bf16 64ch", timeit(lambda: B.from_nchw(nchw)), 64 * 6 * H * W)
    rec("nhwc bf16 -> nchw fp32 64ch", timeit(lambda: B.to_nchw(feats[0])), 64 * 6 * H * W)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump({"peak_gbs": peak, "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy, read+write)", "rows": res},
              open(os.path.join(ROOT, "gpurun_out", "motion_bench.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
