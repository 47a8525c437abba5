"""Micro-benchmark of the tcgen05 conv engine on the dominant 1080p shapes (for ncu and tuning).

    python scripts/conv_bench.py [--iters 20] [--shapes all|top]

Prints one line per shape: ms / launch (CUDA events on the launching stream, L2 flushed by cycling
through distinct input/output buffers > 126 MB) and achieved algorithmic TFLOP/s.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from vcm_ts_b200.backend import CudaBackend, View  # noqa: E402

SHAPES = {
    # name: (K, stride, Cin, Cout, H, W, res, ps)
    "c64_64_full": (3, 1, 64, 64, 1088, 1920, True, False),
    "c64_64_full_nores": (3, 1, 64, 64, 1088, 1920, False, False),
    "k1_64_64_full_nores": (1, 1, 64, 64, 1088, 1920, False, False),
    "c64_64_q": (3, 1, 64, 64, 272, 480, False, False),
    "c64_64_half": (3, 1, 64, 64, 544, 960, True, False),
    "c128_128_q": (3, 1, 128, 128, 272, 480, True, False),
    "c128_64_full": (3, 1, 128, 64, 1088, 1920, False, False),
    "k7_32_64_full": (7, 1, 32, 64, 1088, 1920, False, False),
    "k7_64_32_full": (7, 1, 64, 32, 1088, 1920, False, False),
    "k7_16_2_full": (7, 1, 16, 2, 1088, 1920, False, False),
    "k1_64_64_full": (1, 1, 64, 64, 1088, 1920, True, False),
    "c64_256ps_q": (3, 1, 64, 256, 272, 480, False, True),
    "c192_192_16": (3, 1, 192, 192, 68, 120, False, False),
    "c64_64_8": (3, 1, 64, 64, 136, 240, True, False),
    "c128_128_q2": (3, 1, 128, 128, 272, 480, False, False),
    "c384_288_16": (3, 1, 384, 288, 68, 120, False, False),
    "c64_32_full": (3, 1, 64, 32, 1088, 1920, False, False),
    "c32_32_full": (3, 1, 32, 32, 1088, 1920, False, False),
    "c16_64_full": (3, 1, 16, 64, 1088, 1920, False, False),
    "c64_64_half_nores": (3, 1, 64, 64, 544, 960, False, False),
    "c32_64_half": (3, 1, 32, 64, 544, 960, False, False),
    "k7_16_32_full": (7, 1, 16, 32, 1088, 1920, False, False),
    "k7_32_16_full": (7, 1, 32, 16, 1088, 1920, False, False),
    "c64_3_full": (3, 1, 64, 3, 1088, 1920, False, False),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--shapes", default="all")
    ap.add_argument("--hot", action="store_true", help="single buffer set (L2-resident where it fits)")
    a = ap.parse_args()
    B = CudaBackend("cuda:0", "bf16")
    bf = torch.bfloat16
    names = list(SHAPES) if a.shapes == "all" else a.shapes.split(",")
    out = []
    for name in names:
        K, stride, Cin, Cout, H, W, use_res, ps = SHAPES[name]
        g = torch.Generator().manual_seed(0)
        w = torch.randn((Cout, Cin, K, K), generator=g) / (Cin * K * K) ** 0.5
        pc = B.pack_conv(w.cuda(), torch.zeros(Cout).cuda(), ps=ps)
        nbuf = max(2, int(300e6 // (H * W * Cin * 2)) + 1)
        nbuf = min(nbuf, 8)
        if a.hot:
            nbuf = 1
        xs = [View(torch.randn((1, H, W, Cin), device="cuda").to(bf)) for _ in range(nbuf)]
        oc = Cout // 4 if ps else Cout
        oh, ow = (2 * H, 2 * W) if ps else (H // stride, W // stride)
        fp32_out = Cout < 16
        odt = torch.float32 if fp32_out else bf
        ys = [View(torch.empty((1, oh, ow, oc), device="cuda", dtype=odt)) for _ in range(nbuf)]
        rs = [View(torch.randn((1, oh, ow, oc), device="cuda").to(odt)) for _ in range(nbuf)] if use_res else None

        def run(i):
            B.conv(xs[i % nbuf], pc, out=ys[i % nbuf], stride=stride, post=0.01, ps=ps,
                   res=rs[i % nbuf] if use_res else None)

        for i in range(3):
            run(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(a.iters):
            run(i)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.iters
        flop = 2.0 * (H // stride) * (W // stride) * K * K * Cin * Cout
        gb = (H * W * Cin * 2 + oh * ow * oc * (4 if fp32_out else 2) * (2 if use_res else 1)) / 1e9
        rec = dict(shape=name, ms=ms, tflops=flop / ms / 1e9, hbm_gbs=gb / ms * 1e3)
        out.append(rec)
        print("%-16s %8.4f ms  %8.1f TFLOP/s  %7.0f GB/s(min traffic)" % (name, ms, rec["tflops"], rec["hbm_gbs"]))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "conv_bench.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
