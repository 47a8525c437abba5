"""`-m gpu` kernel-level parity tests: every C-ABI kernel of libvcmts_b200.so against the op of the
reference it replaces, evaluated with stock torch ops / the oracle on the same seeded inputs.

Integer outputs (symbols, scale indexes) are compared bit-exactly; fp32 kernels within a few ulp
(stated per test); the tcgen05 bf16 engine against an fp32 convolution of the same bf16-rounded
operands (only accumulation order and the final bf16 rounding differ).
"""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def B32():
    from vcm_ts_b200.backend import CudaBackend
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return CudaBackend("cuda:0", "fp32")


@pytest.fixture(scope="module")
def B16():
    from vcm_ts_b200.backend import CudaBackend
    return CudaBackend("cuda:0", "bf16")


@pytest.fixture(scope="module")
def BPAR():
    from vcm_ts_b200.backend import CudaBackend
    return CudaBackend("cuda:0", "parity")


def view_of(t_nchw, dtype=torch.float32, cs=None, c0=0):
    """NCHW tensor -> View (optionally a channel slice of a wider NaN-free buffer)."""
    from vcm_ts_b200.backend import View
    N, C, H, W = t_nchw.shape
    cs = cs or C
    buf = torch.zeros((N, H, W, cs), dtype=dtype, device=t_nchw.device)
    buf[..., c0:c0 + C] = t_nchw.permute(0, 2, 3, 1).to(dtype)
    return View(buf, c0, C)


def nchw(v):
    return v.dense().permute(0, 3, 1, 2).float().contiguous()


def ref_conv(x, w, b, stride, pre, post, ps, res, res_scale, res2):
    t = x if pre is None else F.leaky_relu(x, pre)
    y = F.conv2d(t, w, b, stride, w.shape[-1] // 2)
    if post is not None:
        y = F.leaky_relu(y, post)
    if ps:
        y = F.pixel_shuffle(y, 2)
    if res is not None and res_scale is None:
        y = res + y
    if res is not None and res_scale is not None:
        y = y + res * res_scale[:, :, None, None]
    if res2 is not None:
        y = res2 + y
    return y


CONV_CASES = [
    # K, stride, Cin, Cout, H, W, N, pre, post, ps, res, res_scale, res2, slope2
    (3, 1, 64, 64, 40, 56, 1, None, None, False, False, False, False, None),
    (3, 1, 64, 64, 33, 47, 2, 0.01, 0.01, False, True, False, True, 0.1),
    (3, 2, 67, 64, 32, 48, 1, None, 0.1, False, False, False, False, None),
    (7, 1, 8, 32, 24, 40, 1, None, 0.0, False, False, False, False, None),
    (7, 1, 16, 2, 24, 40, 1, None, None, False, True, False, False, None),
    (1, 1, 64, 256, 17, 30, 1, None, 0.01, True, False, False, False, None),
    (1, 2, 2, 64, 32, 48, 1, None, None, False, False, False, False, None),
    (1, 1, 32, 64, 16, 24, 2, None, None, False, True, True, False, 0.0),
    (3, 1, 128, 256, 16, 24, 1, None, None, True, True, False, False, 0.01),
    (3, 2, 3, 192, 64, 64, 1, None, None, False, False, False, False, None),
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_fp32_engine(B32, case):
    K, stride, Cin, Cout, H, W, N, pre, post, ps, use_res, use_scale, use_res2, slope2 = case
    g = torch.Generator(device="cpu").manual_seed(CONV_CASES.index(case))
    dev = "cuda:0"
    x = torch.randn((N, Cin, H, W), generator=g).to(dev)
    w = (torch.randn((Cout, Cin, K, K), generator=g) / (Cin * K * K) ** 0.5).to(dev)
    b = torch.randn((Cout,), generator=g).to(dev)
    Ho, Wo = (H + 2 * (K // 2) - K) // stride + 1, (W + 2 * (K // 2) - K) // stride + 1
    oshape = (N, Cout // 4, 2 * Ho, 2 * Wo) if ps else (N, Cout, Ho, Wo)
    res = torch.randn(oshape, generator=g).to(dev) if use_res else None
    res2 = torch.randn(oshape, generator=g).to(dev) if use_res2 else None
    scale = torch.rand((N, oshape[1]), generator=g).to(dev).contiguous() if use_scale else None
    pc = B32.pack_conv(w, b)
    xv = view_of(x, cs=Cin + 5, c0=3) if Cin % 4 else view_of(x, cs=Cin + 8, c0=4)
    out = B32.conv(xv, pc, stride=stride, pre=pre, post=post, ps=ps,
                   res=view_of(res, cs=oshape[1] + 4, c0=4) if use_res else None, res_scale=scale,
                   res2=view_of(res2) if use_res2 else None, slope2=slope2)
    out2 = None
    if slope2 is not None:
        out, out2 = out
    want = ref_conv(x, w, b, stride, pre, post, ps, res, scale, res2)
    got = nchw(out)
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-4)
    if out2 is not None:
        torch.testing.assert_close(nchw(out2), F.leaky_relu(want, slope2), rtol=1e-4, atol=1e-4)


PARITY_EXTRA = [
    (3, 1, 128, 128, 34, 50, 1, None, 0.0, False, True, False, False, None),     # streamed, MH = 2
    (3, 1, 480, 384, 17, 30, 1, None, 0.2, False, False, False, False, None),    # N tiles
    (3, 1, 192, 192, 35, 41, 2, 0.01, 0.2, False, True, False, True, None),
    (7, 1, 64, 32, 37, 21, 1, None, 0.0, False, False, False, False, None),
    (3, 2, 128, 64, 48, 32, 2, None, None, False, False, False, False, None),
    (3, 1, 32, 32, 40, 56, 1, None, 0.01, False, False, False, False, None),     # resident split weights
]


@pytest.mark.parametrize("case", CONV_CASES + PARITY_EXTRA)
def test_conv_parity_engine(B32, BPAR, case):
    """Split-bf16 tensor-core engine (VCM_CONV_SPLIT3) against an fp64 convolution: its error must be
    fp32-class, i.e. of the size of the fp32 FMA engine's own rounding error -- not bf16-class."""
    import ctypes
    from vcm_ts_b200 import _lib
    K, stride, Cin, Cout, H, W, N, pre, post, ps, use_res, use_scale, use_res2, slope2 = case
    allc = CONV_CASES + PARITY_EXTRA
    g = torch.Generator(device="cpu").manual_seed(500 + allc.index(case))
    dev = "cuda:0"
    x = torch.randn((N, Cin, H, W), generator=g).to(dev)
    w = (torch.randn((Cout, Cin, K, K), generator=g) / (Cin * K * K) ** 0.5).to(dev)
    b = torch.randn((Cout,), generator=g).to(dev)
    Ho, Wo = (H + 2 * (K // 2) - K) // stride + 1, (W + 2 * (K // 2) - K) // stride + 1
    oshape = (N, Cout // 4, 2 * Ho, 2 * Wo) if ps else (N, Cout, Ho, Wo)
    res = torch.randn(oshape, generator=g).to(dev) if use_res else None
    res2 = torch.randn(oshape, generator=g).to(dev) if use_res2 else None
    scale = torch.rand((N, oshape[1]), generator=g).to(dev).contiguous() if use_scale else None
    xv = view_of(x, cs=Cin + 5, c0=3) if Cin % 4 else view_of(x, cs=Cin + 8, c0=4)

    def run(B):
        pc = B.pack_conv(w, b, ps=ps and B is BPAR)
        out = B.conv(xv, pc, stride=stride, pre=pre, post=post, ps=ps,
                     res=view_of(res, cs=oshape[1] + 4, c0=4) if use_res else None, res_scale=scale,
                     res2=view_of(res2) if use_res2 else None, slope2=slope2)
        return (out[0], out[1]) if slope2 is not None else (out, None)

    pcp = BPAR.pack_conv(w, b, ps=ps)
    assert pcp.w_umma is not None
    d3 = _lib.ConvDesc(N, H, W, pcp.cin_tc, Cout, K, stride, 3 * pcp.cin_tc, oshape[1], 0, 0, 0,
                       _lib.CONV_SPLIT3, 0.0, 0.0, 0.0, _lib.BF16, _lib.F32, pcp.cin_tc)
    on_tc = BPAR.L.vcm_conv2d_tc_supported(ctypes.byref(d3)) == 1
    assert on_tc or (H % 2 or W % 2), "case must exercise the tcgen05 engine"
    got, got2 = run(BPAR)
    simt, _ = run(B32)
    torch.cuda.synchronize()
    d64 = lambda t: None if t is None else t.double()   # noqa: E731
    want = ref_conv(x.double(), w.double(), b.double(), stride, pre, post, ps, d64(res),
                    d64(scale), d64(res2))
    e_par = (nchw(got).double() - want).abs().max().item()
    e_simt = (nchw(simt).double() - want).abs().max().item()
    mag = want.abs().max().item()
    print(f"parity conv K{K}s{stride} {Cin}->{Cout}: max|err| parity {e_par:.3e}  fp32-FMA {e_simt:.3e}  (|y|max {mag:.2f})")
    rms_par = (nchw(got).double() - want).pow(2).mean().sqrt().item()
    rms_simt = (nchw(simt).double() - want).pow(2).mean().sqrt().item()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "parity_conv_errors.jsonl"), "a") as f:
        f.write(json.dumps({"K": K, "stride": stride, "Cin": Cin, "Cout": Cout, "H": H, "W": W, "on_tc": on_tc,
                            "max_err_parity": e_par, "max_err_fp32_fma": e_simt, "rms_err_parity": rms_par,
                            "rms_err_fp32_fma": rms_simt, "max_abs_y": mag}) + "\n")
    assert e_par <= 4 * e_simt + 2e-7 * mag, (e_par, e_simt)
    if got2 is not None:
        assert torch.equal(nchw(got2), F.leaky_relu(nchw(got), slope2))


def test_split_bf16x3_is_exact(BPAR):
    """x0 + x1 + x2 == x bit for bit (normal fp32 range), pad channels are zero, pre-activation applied."""
    dev = "cuda:0"
    g = torch.Generator().manual_seed(9)
    x = (torch.randn((2, 19, 23, 67), generator=g) * torch.exp(torch.randn((2, 19, 23, 67), generator=g) * 4)).to(dev)
    x[0, 0, 0, :4] = torch.tensor([0.0, -0.0, 1e-30, 1e38])
    from vcm_ts_b200.backend import View
    v = View(x.contiguous(), 0, 67)
    for pre in (None, 0.1):
        cp = 80
        xs = torch.full((2, 19, 23, 3 * cp), float("nan"), dtype=torch.bfloat16, device=dev)
        rc = BPAR.L.vcm_split_bf16x3(v.ptr, xs.data_ptr(), 2 * 19 * 23, 67, cp, 67, 3 * cp, cp,
                                     1 if pre is not None else 0, float(pre or 0), BPAR.stream)
        assert rc == 0
        p = xs.float().reshape(2, 19, 23, 3, cp)
        want = x if pre is None else F.leaky_relu(x, pre)
        assert torch.equal((p[..., 0, :67] + p[..., 1, :67]) + p[..., 2, :67], want)
        assert float(p[..., 67:].abs().max()) == 0.0


def test_conv_mixed_dtypes(B32):
    """fp32 in -> bf16 out and bf16 in -> fp32 out (the flow / frame boundaries of bf16 mode)."""
    dev = "cuda:0"
    g = torch.Generator().manual_seed(3)
    x = torch.randn((1, 8, 20, 28), generator=g).to(dev)
    w = (torch.randn((32, 8, 7, 7), generator=g) / 20).to(dev)
    b = torch.randn((32,), generator=g).to(dev)
    pc = B32.pack_conv(w, b)
    y = B32.conv(view_of(x), pc, post=0.0, out_dtype=torch.bfloat16)
    want = F.relu(F.conv2d(x, w, b, 1, 3))
    torch.testing.assert_close(nchw(y), want.bfloat16().float(), rtol=1e-2, atol=1e-2)
    xb = view_of(want, dtype=torch.bfloat16)
    w2 = (torch.randn((2, 32, 7, 7), generator=g) / 40).to(dev)
    b2 = torch.randn((2,), generator=g).to(dev)
    r = torch.randn((1, 2, 20, 28), generator=g).to(dev)
    y2 = B32.conv(xb, B32.pack_conv(w2, b2), res=view_of(r), out_dtype=torch.float32)
    want2 = r + F.conv2d(nchw(xb), w2, b2, 1, 3)
    torch.testing.assert_close(nchw(y2), want2, rtol=1e-4, atol=1e-4)


def test_umma_descriptor_probe(B16):
    """Records how the tensor core addresses a 128B-swizzled A tile that starts a few rows into the
    buffer / whose 8-row groups are 1280 B apart (what a halo-tile implicit GEMM needs).  Only the
    canonical case is asserted; the table goes to gpurun_out/umma_probe.json."""
    from vcm_ts_b200 import _probe
    L = _probe.lib()
    out = torch.zeros((128, 64), dtype=torch.float32, device="cuda:0")
    table = {}
    for start_row in (0, 1, 3, 8, 11):
        for sbo in (1024, 1280, 2048):
            for use_bo in (0, 1):
                out.zero_()
                rc = L.vcm_umma_probe(start_row, sbo, use_bo, out.data_ptr(), B16.stream)
                assert rc == 0
                torch.cuda.synchronize()
                m = torch.arange(128)
                rows = start_row + (m // 8) * (sbo // 128) + m % 8
                want = ((rows[:, None] * 64 + torch.arange(64)[None, :]) % 251).float()
                ok = bool(torch.equal(out.cpu(), want))
                table[f"start{start_row}_sbo{sbo}_bo{use_bo}"] = ok
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "umma_probe.json"), "w") as f:
        json.dump(table, f, indent=1)
    print(table)
    assert table["start0_sbo1024_bo0"], "canonical SW128 K-major descriptor does not work"


TC_CASES = [
    # K, stride, Cin, Cout, H, W, N, post, ps, res, res_scale, res2, slope2
    (3, 1, 64, 64, 40, 56, 1, None, False, False, False, False, None),
    (3, 1, 64, 64, 64, 96, 2, 0.01, False, True, False, True, 0.1),
    (3, 1, 128, 64, 24, 48, 1, 0.2, False, False, False, False, None),
    (3, 1, 96, 64, 17, 30, 1, None, False, False, False, False, None),
    (3, 1, 144, 192, 17, 30, 1, 0.01, False, False, False, False, None),
    (7, 1, 32, 64, 32, 48, 1, 0.0, False, False, False, False, None),
    (7, 1, 64, 32, 32, 48, 1, 0.0, False, False, False, False, None),
    (7, 1, 32, 16, 16, 32, 1, 0.0, False, False, False, False, None),
    (1, 1, 64, 256, 17, 30, 1, 0.01, True, False, False, False, None),
    (3, 1, 64, 256, 16, 32, 1, None, True, True, False, False, 0.01),
    (1, 1, 64, 32, 32, 32, 1, None, False, True, True, False, 0.0),
    (3, 1, 480, 384, 17, 30, 1, 0.2, False, False, False, False, None),
    (3, 2, 64, 64, 32, 64, 1, 0.01, False, False, False, False, 0.01),
    (3, 2, 128, 64, 48, 32, 2, None, False, False, False, False, None),
    (1, 2, 64, 64, 32, 32, 1, None, False, False, False, False, None),
    (3, 1, 64, 128, 19, 37, 1, 0.01, True, True, False, False, None),     # PS, 32 output channels
    (1, 1, 64, 64, 21, 13, 1, None, True, False, False, False, 0.1),      # PS, 16 output channels (BW 16)
    (3, 1, 96, 384, 9, 15, 1, 0.01, True, False, False, False, None),     # PS across two N tiles
    (3, 1, 192, 192, 35, 41, 2, 0.2, False, True, False, True, None),     # streamed weights, M = 256 items
    (7, 1, 64, 32, 37, 21, 1, 0.0, False, False, False, False, None),     # ragged 7x7
    (3, 1, 32, 48, 18, 26, 1, None, False, True, False, False, 0.0),      # BN = 48 (BW 16)
    # TMA-store epilogue (64- / 32-channel tiles): residual by TMA, SE-scaled residual, twin, ragged tile tails
    (3, 1, 64, 32, 37, 53, 2, 0.01, False, True, False, False, None),
    (1, 1, 64, 32, 21, 45, 1, None, False, True, True, False, None),
    (3, 1, 32, 32, 40, 56, 1, 0.01, False, False, False, False, 0.1),
    (3, 1, 64, 64, 37, 53, 2, 0.01, False, True, True, False, None),
    (3, 1, 64, 64, 35, 29, 1, None, False, False, False, False, 0.01),
    (3, 1, 128, 64, 70, 90, 1, 0.1, False, True, False, False, None),     # streamed weights, 256-pixel items
    (1, 1, 64, 128, 33, 47, 1, None, False, True, False, False, None),    # Cout 128 as two 64-channel N tiles on a small image
    # streamed weights, 256-pixel items, 3.6 items per CTA at the BASELINE 1/4-scale shape
    (3, 1, 128, 128, 272, 480, 1, 0.01, False, True, False, False, None),
    (3, 1, 128, 128, 272, 488, 1, None, False, False, False, False, None),
    (3, 1, 64, 128, 272, 488, 1, 0.1, False, False, False, False, None),
    # row-stacked engine (conv_rows.cu): Cout 64 -> 2 output rows per MMA, Cout 32 -> 4; ragged last strip
    (3, 1, 64, 64, 304, 512, 1, 0.01, False, True, False, False, None),
    (3, 1, 64, 64, 300, 480, 1, None, False, False, False, False, 0.01),
    (3, 1, 64, 32, 304, 512, 2, 0.01, False, True, True, False, None),
    (3, 1, 32, 32, 300, 460, 2, None, False, False, False, False, 0.1),
    (3, 1, 16, 64, 304, 500, 1, 0.1, False, True, False, False, None),
    (3, 1, 32, 64, 298, 640, 1, None, False, False, False, False, None),
    # ... and its 7x7 forms (SpyNet conv1 / conv4): 4 resp. 8 output rows per MMA, 16-channel TMA tiles
    (7, 1, 16, 32, 300, 460, 2, 0.0, False, False, False, False, None),
    (7, 1, 32, 16, 304, 500, 4, 0.0, False, False, False, False, None),
]


@pytest.fixture()
def rows_engine_small_images():
    """conv_rows.cu takes images at least 1024 wide by default; the test shapes are smaller."""
    os.environ["VCM_TC_ROWS_MIN_W"] = "448"
    yield
    os.environ.pop("VCM_TC_ROWS_MIN_W", None)


@pytest.mark.parametrize("case", TC_CASES)
def test_conv_tcgen05_engine(B16, case, rows_engine_small_images):
    K, stride, Cin, Cout, H, W, N, post, ps, use_res, use_scale, use_res2, slope2 = case
    import ctypes
    from vcm_ts_b200 import _lib
    g = torch.Generator(device="cpu").manual_seed(100 + TC_CASES.index(case))
    dev = "cuda:0"
    bf = torch.bfloat16
    x = torch.randn((N, Cin, H, W), generator=g).to(dev).to(bf)
    w = (torch.randn((Cout, Cin, K, K), generator=g) / (Cin * K * K) ** 0.5).to(dev)
    b = torch.randn((Cout,), generator=g).to(dev)
    Ho, Wo = (H + 2 * (K // 2) - K) // stride + 1, (W + 2 * (K // 2) - K) // stride + 1
    oshape = (N, Cout // 4, 2 * Ho, 2 * Wo) if ps else (N, Cout, Ho, Wo)
    res = torch.randn(oshape, generator=g).to(dev).to(bf) if use_res else None
    res2 = torch.randn(oshape, generator=g).to(dev).to(bf) if use_res2 else None
    scale = torch.rand((N, oshape[1]), generator=g).to(dev).contiguous() if use_scale else None
    pc = B16.pack_conv(w, b, ps=ps)
    assert pc.w_umma is not None
    xv = view_of(x.float(), dtype=bf, cs=Cin + 16, c0=8)
    rv = view_of(res.float(), dtype=bf, cs=oshape[1] + 8, c0=8) if use_res else None
    d = _lib.ConvDesc(N, H, W, Cin, Cout, K, stride, xv.cs, oshape[1], rv.cs if rv else 0,
                      oshape[1] if use_res2 else 0, oshape[1], 0, 0.0, 0.0, 0.0, _lib.BF16, _lib.BF16)
    assert B16.L.vcm_conv2d_tc_supported(ctypes.byref(d)) == 1, "case must exercise the tcgen05 engine"
    out = B16.conv(xv, pc, stride=stride, post=post, ps=ps, res=rv, res_scale=scale,
                   res2=view_of(res2.float(), dtype=bf) if use_res2 else None, slope2=slope2)
    out2 = None
    if slope2 is not None:
        out, out2 = out
    torch.cuda.synchronize()
    want = ref_conv(x.float(), w.to(bf).float(), b, stride, None, post, ps,
                    res.float() if use_res else None, scale, res2.float() if use_res2 else None)
    got = nchw(out)
    err = (got - want).abs()
    tol = 0.01 * want.abs() + 0.02
    assert bool((err <= tol).all()), f"max err {err.max().item():.4f} (rms of want {want.pow(2).mean().sqrt().item():.3f})"
    if out2 is not None:
        w2 = F.leaky_relu(want, slope2)
        assert bool(((nchw(out2) - w2).abs() <= 0.01 * w2.abs() + 0.02).all())


@pytest.mark.parametrize("shape", [(3, 64, 64, 1, 304, 640), (7, 32, 16, 4, 304, 512), (7, 16, 32, 2, 304, 512)])
def test_conv_rows_engine_is_taken_and_equals_the_halo_engine(B16, rows_engine_small_images, shape):
    """3x3 64->64 (+ residual) on a large image runs on conv_rows_kernel (torch profiler kernel names); with
    VCM_TC_NO_ROWS=1 the halo engine takes it.  Same products, fp32 accumulation in another order: the bf16
    outputs may differ by one rounding step on a few values, never more.  Same for the 7x7 forms."""
    from torch.profiler import ProfilerActivity, profile
    g = torch.Generator(device="cpu").manual_seed(5)
    dev, bf = "cuda:0", torch.bfloat16
    K, Ci, C, N, H, W = shape
    x = torch.randn((N, Ci, H, W), generator=g).to(dev).to(bf)
    w = (torch.randn((C, Ci, K, K), generator=g) / (Ci * K * K) ** 0.5).to(dev)
    b = torch.randn((C,), generator=g).to(dev)
    res = torch.randn((N, C, H, W), generator=g).to(dev).to(bf)
    pc = B16.pack_conv(w, b)
    xv, rv = view_of(x.float(), dtype=bf), view_of(res.float(), dtype=bf)

    def run():
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            out = B16.conv(xv, pc, post=0.01, res=rv)
            torch.cuda.synchronize()
        return nchw(out), [e.key for e in prof.key_averages()]

    os.environ.pop("VCM_TC_NO_ROWS", None)
    got, names = run()
    assert any("conv_rows_kernel" in k for k in names), names
    os.environ["VCM_TC_NO_ROWS"] = "1"
    try:
        halo, names2 = run()
    finally:
        os.environ.pop("VCM_TC_NO_ROWS", None)
    assert any("conv_halo_kernel" in k for k in names2) and not any("conv_rows_kernel" in k for k in names2), names2
    diff = (got - halo).abs()
    assert bool((diff <= 0.0079 * halo.abs() + 1e-6).all()), diff.max().item()      # one bf16 ulp
    assert float((diff > 0).float().mean()) < 0.02
    want = ref_conv(x.float(), w.to(bf).float(), b, 1, None, 0.01, False, res.float(), None, None)
    assert bool(((got - want).abs() <= 0.01 * want.abs() + 0.02).all())


@pytest.mark.parametrize("shape", [(64, 64, 160, 128), (32, 32, 304, 1024), (64, 32, 304, 1024), (64, 64, 37, 53),
                                   (128, 128, 160, 128)])
def test_conv_fused_se_channel_sums(B16, rows_engine_small_images, shape):
    """VCM_CONV_SE_SUM: the SELayer mean out of the producing conv's epilogue (halo and row engines) gives the
    scales of the stand-alone kernel (other summation order: 1e-5); ragged tiles / wide N tiles report
    'not available' and the caller keeps the stand-alone kernel."""
    Ci, C, H, W = shape
    g = torch.Generator(device="cpu").manual_seed(31)
    dev, bf = "cuda:0", torch.bfloat16
    x = torch.randn((1, Ci, H, W), generator=g).to(dev).to(bf)
    w = (torch.randn((C, Ci, 3, 3), generator=g) / (Ci * 9) ** 0.5).to(dev)
    b = torch.randn((C,), generator=g).to(dev)
    w1 = (torch.randn((C // 16, C), generator=g) / 8).to(dev)
    w2 = torch.randn((C, C // 16), generator=g).to(dev)
    pc = B16.pack_conv(w, b)
    x1, sums = B16.conv(view_of(x.float(), dtype=bf), pc, se_sums=True)
    want_sums = (H % 16 == 0 and W % 8 == 0 and C <= 64)
    assert (sums is not None) == want_sums, (shape, None if sums is None else tuple(sums.shape))
    ref_out = ref_conv(x.float(), w.to(bf).float(), b, 1, None, None, False, None, None, None)
    assert bool(((nchw(x1) - ref_out).abs() <= 0.01 * ref_out.abs() + 0.02).all())
    if sums is None:
        return
    got = B16.se_scale_from_sums(sums, x1, w1, w2)
    want = B16.se_scale(x1, w1, w2)
    torch.cuda.synchronize()
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-6)
    mean = nchw(x1).float().mean(dim=(2, 3))[0]
    torch.testing.assert_close(sums.sum(0) / (H * W), mean, rtol=1e-4, atol=1e-5)


def test_conv_tcgen05_unaligned_views_take_scalar_epilogue(B16):
    """Output / residual views whose channel stride or offset is not a multiple of 8 cannot use the
    16-byte vector epilogue: the tensor-core engine must still take the conv (scalar bf16 epilogue)."""
    import ctypes
    from vcm_ts_b200 import _lib
    from vcm_ts_b200.backend import View
    g = torch.Generator(device="cpu").manual_seed(77)
    dev, bf = "cuda:0", torch.bfloat16
    N, Cin, Cout, H, W = 1, 64, 64, 21, 35
    x = torch.randn((N, Cin, H, W), generator=g).to(dev).to(bf)
    w = (torch.randn((Cout, Cin, 3, 3), generator=g) / 24).to(dev)
    b = torch.randn((Cout,), generator=g).to(dev)
    res = torch.randn((N, Cout, H, W), generator=g).to(dev).to(bf)
    pc = B16.pack_conv(w, b)
    out_buf = View(torch.zeros((N, H, W, Cout + 6), dtype=bf, device=dev), 3, Cout)   # cs = 70, c0 = 3
    rv = view_of(res.float(), dtype=bf, cs=Cout + 4, c0=4)
    d = _lib.ConvDesc(N, H, W, Cin, Cout, 3, 1, Cin, out_buf.cs, rv.cs, 0, 0, _lib.CONV_RES | _lib.CONV_POST_ACT,
                      0.0, 0.1, 0.0, _lib.BF16, _lib.BF16)
    assert B16.L.vcm_conv2d_tc_supported(ctypes.byref(d)) == 1
    out = B16.conv(view_of(x.float(), dtype=bf), pc, out=out_buf, post=0.1, res=rv)
    want = ref_conv(x.float(), w.to(bf).float(), b, 1, None, 0.1, False, res.float(), None, None)
    got = nchw(out)
    assert bool(((got - want).abs() <= 0.01 * want.abs() + 0.02).all()), (got - want).abs().max().item()
    assert float(out_buf.t[..., :3].abs().max()) == 0.0 and float(out_buf.t[..., 3 + Cout:].abs().max()) == 0.0


def test_layout_and_pool_kernels(B32):
    dev = "cuda:0"
    g = torch.Generator().manual_seed(0)
    x = torch.randn((2, 37, 18, 26), generator=g).to(dev)
    v = B32.from_nchw(x)
    assert torch.equal(nchw(v), x)
    assert torch.equal(B32.to_nchw(v), x)
    assert torch.equal(B32.to_nchw(v, clamp=(0.0, 1.0)), x.clamp(0, 1))
    vb = B32.from_nchw(x, dtype=torch.bfloat16)
    assert torch.equal(nchw(vb), x.bfloat16().float())
    assert torch.equal(nchw(B32.maxpool(v)), F.max_pool2d(x, 2, 2))
    x3 = torch.rand((2, 3, 32, 48), generator=g).to(dev)
    torch.testing.assert_close(nchw(B32.avgpool(view_of(x3, cs=8))), F.avg_pool2d(x3, 2, 2), rtol=0, atol=1e-7)
    vec = (torch.rand((2, 37), generator=g) + 0.5).to(dev).contiguous()
    assert torch.equal(nchw(B32.scale(v, vec, "div")), x / vec[:, :, None, None])
    assert torch.equal(nchw(B32.scale(v, vec, "mul")), x * vec[:, :, None, None])
    dst = view_of(torch.zeros((2, 37, 18, 26), device=dev), cs=64, c0=16)
    B32.copy(v, dst)
    assert torch.equal(nchw(dst), x)


def test_se_layer(B32):
    dev = "cuda:0"
    g = torch.Generator().manual_seed(1)
    for C, H, W in ((32, 64, 96), (128, 70, 90), (64, 16, 16)):
        x = torch.randn((2, C, H, W), generator=g).to(dev)
        w1 = (torch.randn((C // 16, C), generator=g) / C ** 0.5).to(dev).contiguous()
        w2 = (torch.randn((C, C // 16), generator=g)).to(dev).contiguous()
        want = torch.sigmoid(F.linear(F.relu(F.linear(x.mean(dim=(-1, -2)), w1)), w2))
        got = B32.se_scale(view_of(x), w1, w2)
        torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-6)
        got2 = B32.se_scale(view_of(x), w1, w2)
        assert torch.equal(got, got2), "SE reduction must be deterministic"


def test_se_layer_bf16_vector_path(B16):
    """bf16 activations take the vectorised partial-sum + striped reduction kernels."""
    dev = "cuda:0"
    g = torch.Generator().manual_seed(11)
    for C, H, W in ((32, 128, 160), (64, 70, 90), (128, 33, 47)):
        x = torch.randn((2, C, H, W), generator=g).to(dev).bfloat16()
        w1 = (torch.randn((C // 16, C), generator=g) / C ** 0.5).to(dev).contiguous()
        w2 = (torch.randn((C, C // 16), generator=g)).to(dev).contiguous()
        want = torch.sigmoid(F.linear(F.relu(F.linear(x.float().mean(dim=(-1, -2)), w1)), w2))
        xv = view_of(x.float(), dtype=torch.bfloat16, cs=C + 8, c0=8)
        got = B16.se_scale(xv, w1, w2)
        torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
        assert torch.equal(got, B16.se_scale(xv, w1, w2)), "SE reduction must be deterministic"


def test_bf16_warp_and_maxpool_vector_paths(B16):
    from oracle import dcvc_oracle as O
    dev = "cuda:0"
    g = torch.Generator().manual_seed(12)
    feat = torch.randn((2, 64, 40, 56), generator=g).to(dev).bfloat16()
    flow = (torch.randn((2, 2, 40, 56), generator=g) * 4).to(dev)
    fv = view_of(feat.float(), dtype=torch.bfloat16, cs=80, c0=16)
    out = B16.warp(fv, view_of(flow))
    want = O.torch_warp(feat.float().cpu(), flow.cpu()).to(dev)
    torch.testing.assert_close(nchw(out), want.bfloat16().float(), rtol=2e-2, atol=2e-2)
    mp = B16.maxpool(fv)
    assert torch.equal(nchw(mp), F.max_pool2d(feat.float(), 2, 2))


def test_motion_kernels(B32):
    from oracle import dcvc_oracle as O
    dev = "cuda:0"
    g = torch.Generator().manual_seed(2)
    flow = (torch.randn((2, 2, 17, 30), generator=g) * 3).to(dev)
    up = B32.flow_up2(view_of(flow), gain=2.0)
    torch.testing.assert_close(nchw(up), O.bilinear_up2(flow) * 2.0, rtol=1e-6, atol=1e-6)
    big = (torch.randn((1, 2, 64, 96), generator=g) * 4).to(dev)
    dn = B32.flow_down2(view_of(big), gain=0.5)
    torch.testing.assert_close(nchw(dn), O.bilinear_down2(big) / 2, rtol=1e-6, atol=1e-6)
    for C, dt in ((3, torch.float32), (64, torch.float32), (64, torch.bfloat16)):
        feat = torch.randn((1, C, 64, 96), generator=g).to(dev)
        fl = big.clone()
        fl[:, :, :4] *= 20          # drive samples out of the image: border clamp
        want = O.torch_warp(feat.cpu(), fl.cpu()).to(dev) if dt == torch.float32 else \
            O.torch_warp(feat.bfloat16().float().cpu(), fl.cpu()).to(dev)
        got = B32.warp(view_of(feat, dtype=dt, cs=C + (8 if C > 3 else 5)),
                       view_of(fl))
        if dt == torch.float32:
            torch.testing.assert_close(nchw(got), want, rtol=0, atol=2e-4)
        else:
            torch.testing.assert_close(nchw(got), want, rtol=1e-2, atol=2e-2)


def _dp_inputs(C, h, w, seed):
    g = torch.Generator().manual_seed(seed)
    y = torch.randn((1, C, h, w), generator=g) * 3
    q_step = torch.rand((1, C, h, w), generator=g) * 1.5 + 0.1
    scales = torch.randn((1, C, h, w), generator=g).abs() * 2 - 0.2      # some <= 0
    means = torch.randn((1, C, h, w), generator=g)
    prior = torch.randn((1, 2 * C, h, w), generator=g)
    prior[:, :C // 2] = prior[:, :C // 2].abs() * 3
    prior[:, C:C + C // 2] = prior[:, C:C + C // 2].abs() * 0.05
    q_outer = torch.rand((1, C), generator=g) + 0.5
    return y, q_step, scales, means, prior, q_outer


@pytest.mark.parametrize("dist,C,h,w", [("laplace", 64, 18, 30), ("laplace", 96, 12, 20), ("gaussian", 192, 8, 14),
                                        # BASELINE shapes: 1080p latents are 68 x 120
                                        ("laplace", 64, 68, 120), ("laplace", 96, 68, 120), ("gaussian", 192, 68, 120)])
def test_dual_prior_kernels_bit_exact(B32, dist, C, h, w):
    """Teacher-forced: identical float inputs -> symbols, scale indexes and y_hat must equal
    common_model.py:104-217 exactly (oracle restatement), encode and decode sides."""
    from oracle import dcvc_oracle as O
    from vcm_ts_b200.entropy import ScaleIndexTable
    dev = "cuda:0"
    y, q_step, scales, means, prior_out, q_outer = _dp_inputs(C, h, w, 5)
    table = ScaleIndexTable(dist)
    otab = O.ScaleTable(dist)
    # oracle: forward_dual_prior with the spatial prior replaced by a fixed tensor
    m0, m1 = O.checkerboard(h, w)
    qs = O.lower_bound(q_step, 0.5)
    yq = y / qs
    y0, y1 = yq.chunk(2, 1)
    s0, s1 = scales.chunk(2, 1)
    u0, u1 = means.chunk(2, 1)
    _, q00, h00, sh00 = O._masked(y0, s0, u0, m0)
    _, q11, h11, sh11 = O._masked(y1, s1, u1, m1)
    ps0, pu0, ps1, pu1 = prior_out.chunk(4, 1)
    _, q01, h01, sh01 = O._masked(y0, ps0, pu0, m1)
    _, q10, h10, sh10 = O._masked(y1, ps1, pu1, m0)
    want_yhat = torch.cat((h00 + h01, h11 + h10), 1) * qs * q_outer[:, :, None, None]
    want = dict(sym0=(q00 + q11).to(torch.int16), sym1=(q01 + q10).to(torch.int16),
                idx0=otab.build_indexes(sh00 + sh11).to(torch.uint8),
                idx1=otab.build_indexes(sh01 + sh10).to(torch.uint8))
    fusion = view_of(torch.cat((q_step, scales, means), 1).to(dev))
    prior = view_of(prior_out.to(dev))
    yv = view_of(y.to(dev))
    qo = q_outer.to(dev).contiguous()
    params = B32.new(1, h, w, 4 * C)
    yqb = B32.empty((1, h, w, C), torch.float32)
    plane = (1, C // 2, h, w)
    sym0, sym1 = B32.empty(plane, torch.int16), B32.empty(plane, torch.int16)
    idx0, idx1 = B32.empty(plane, torch.uint8), B32.empty(plane, torch.uint8)
    yqf, scf = B32.empty((1, h, w, C), torch.float32), B32.empty((1, h, w, C), torch.float32)
    B32.dp_pass0(0, table, fusion, params, C, y=yv, yq=yqb, sym=sym0, idx=idx0, y_q_full=yqf, scales_full=scf)
    want_params = torch.cat((h00, h11, means, scales, qs), 1)
    assert torch.equal(nchw(params).cpu(), want_params)
    y_hat = B32.new(1, h, w, C)
    B32.dp_pass1(0, table, fusion, prior, params, C, q_outer=qo, y_hat=y_hat, yq=yqb, sym=sym1, idx=idx1,
                 y_q_full=yqf, scales_full=scf)
    assert torch.equal(sym0.cpu(), want["sym0"]) and torch.equal(sym1.cpu(), want["sym1"])
    assert torch.equal(idx0.cpu(), want["idx0"]) and torch.equal(idx1.cpu(), want["idx1"])
    assert torch.equal(nchw(y_hat).cpu(), want_yhat)
    assert torch.equal(yqf.permute(0, 3, 1, 2).cpu(), torch.cat((q00 + q01, q11 + q10), 1))
    assert torch.equal(scf.permute(0, 3, 1, 2).cpu(), torch.cat((sh00 + sh01, sh11 + sh10), 1))
    # decoder side: index-only passes give the same indexes, reconstruct passes the same y_hat
    params2 = B32.new(1, h, w, 4 * C)
    i0 = B32.empty(plane, torch.uint8)
    B32.dp_pass0(1, table, fusion, params2, C, idx=i0)
    assert torch.equal(i0, idx0)
    B32.dp_pass0(2, table, fusion, params2, C, sym=sym0)
    assert torch.equal(params2.dense(), params.dense())
    i1 = B32.empty(plane, torch.uint8)
    B32.dp_pass1(1, table, fusion, prior, params2, C, idx=i1)
    assert torch.equal(i1, idx1)
    y_hat2 = B32.new(1, h, w, C)
    B32.dp_pass1(2, table, fusion, prior, params2, C, q_outer=qo, y_hat=y_hat2, sym=sym1)
    assert torch.equal(y_hat2.dense(), y_hat.dense())


def test_round_and_bits_kernels(B32):
    from oracle import dcvc_oracle as O
    from oracle import fixture
    dev = "cuda:0"
    g = torch.Generator().manual_seed(9)
    z = torch.randn((2, 64, 9, 14), generator=g) * 4
    z[0, 0, 0, :4] = torch.tensor([0.5, 1.5, -0.5, 2.5])          # round-half-to-even
    sym = B32.empty((2, 64, 9, 14), torch.int16)
    zh = B32.round_latent(view_of(z.to(dev)), sym=sym)
    assert torch.equal(nchw(zh).cpu(), torch.round(z)) and torch.equal(sym.cpu(), torch.round(z).to(torch.int16))
    back = B32.symbols_to_latent(sym, 2, 9, 14, 64)
    assert torch.equal(nchw(back).cpu(), torch.round(z))
    # likelihood kernels vs the oracle's restatement of common_model.py:51-73
    yq = torch.round(torch.randn((2, 9, 14, 96), generator=g) * 3)
    sc = torch.randn((2, 9, 14, 96), generator=g).abs() * 2
    for dist, fn in (("laplace", O.laplace_bits), ("gaussian", O.gaussian_bits)):
        want = fn(yq, sc).reshape(2, -1).sum(1)
        got = B32.latent_bits(yq.to(dev).contiguous(), sc.to(dev).contiguous(), 2, dist)
        torch.testing.assert_close(got.cpu(), want, rtol=2e-5, atol=1e-3)
    P = fixture.fixture_state_dict({k: s for k, s in O.state_shapes("dmc").items() if k.startswith("bit_estimator_z.")})
    rows = [P[f"bit_estimator_z.{f}.{p}"].reshape(-1) for f in ("f1", "f2", "f3", "f4") for p in ("h", "b", "a")
            if not (f == "f4" and p == "a")]
    fparams = torch.stack(rows).to(dev).contiguous()
    zr = torch.round(z)
    want = O.z_bits(P, "bit_estimator_z", zr).reshape(2, -1).sum(1)
    got = B32.factorized_bits(view_of(zr.to(dev)), fparams)
    torch.testing.assert_close(got.cpu(), want, rtol=2e-5, atol=1e-3)
    a, b = torch.rand((2, 3, 40, 50), generator=g), torch.rand((2, 3, 40, 50), generator=g)
    torch.testing.assert_close(B32.sq_err_sum(a.to(dev), b.to(dev)).cpu(), ((a - b) ** 2).reshape(2, -1).sum(1),
                               rtol=1e-5, atol=1e-5)


SCALAR_TC_CASES = [
    # K, Cin(real), Cin_p, Cout, ps, fp32 residual, H, W   -- fp32 outputs / tiny Cout on tcgen05
    (3, 64, 64, 3, False, False, 40, 56),      # recon_conv
    (7, 16, 16, 2, False, True, 24, 48),       # SpyNet conv5 + flow residual
    (1, 64, 64, 8, True, False, 17, 30),       # mv_decoder subpel -> 2-channel flow
    (7, 8, 16, 32, False, False, 24, 40),      # SpyNet conv1, input padded 8 -> 16 (bf16 out)
    (3, 3, 16, 64, False, False, 32, 48),      # feature_adaptor_I, frame padded 3 -> 16 (bf16 out)
    # row-stacked engine with the scalar epilogue (conv_rows.cu): 4 / 8 output rows per MMA
    (3, 64, 64, 3, False, False, 304, 1000),   # recon_conv at a size the engine takes
    (7, 16, 16, 2, False, True, 608, 1000),    # SpyNet conv5 + flow residual
]


@pytest.mark.parametrize("case", SCALAR_TC_CASES)
def test_conv_tcgen05_padded_and_fp32_out(B16, case, rows_engine_small_images):
    K, Cin, Cin_p, Cout, ps, use_res, H, W = case
    g = torch.Generator(device="cpu").manual_seed(200 + SCALAR_TC_CASES.index(case))
    dev, bf = "cuda:0", torch.bfloat16
    x = torch.randn((1, Cin, H, W), generator=g).to(dev).to(bf)
    w = (torch.randn((Cout, Cin, K, K), generator=g) / (Cin * K * K) ** 0.5).to(dev)
    b = torch.randn((Cout,), generator=g).to(dev)
    fp32_out = Cout < 16
    oshape = (1, Cout // 4, 2 * H, 2 * W) if ps else (1, Cout, H, W)
    res = torch.randn(oshape, generator=g).to(dev) if use_res else None
    in_map = list(range(Cin)) + [-1] * (Cin_p - Cin) if Cin_p != Cin else None
    pc = B16.pack_conv(w, b, in_map, ps=ps)
    assert pc.w_umma is not None
    from vcm_ts_b200.backend import View
    xv = View(view_of(x.float(), dtype=bf, cs=Cin_p).t)      # all Cin_p channels; the pad is zero
    launches = B16.launches
    out = B16.conv(xv, pc, ps=ps, res=view_of(res) if use_res else None,
                   out_dtype=torch.float32 if fp32_out else bf)
    torch.cuda.synchronize()
    assert B16.launches == launches + 1
    want = ref_conv(x.float(), w.to(bf).float(), b, 1, None, None, ps, res, None, None)
    got = nchw(out)
    assert bool(((got - want).abs() <= 0.01 * want.abs() + 0.02).all()), (got - want).abs().max().item()


def test_frame_io_bit_exact_vs_reference_formulas():
    """uint8 HWC <-> float NCHW on the device against the reference's numpy formulas
    (png_reader.py:37-39 `asarray(...).astype('float32').transpose(2,0,1) / 255.`, zero pad;
    stream_helper.py:151-154 `clip(rint(img * 255), 0, 255).astype(uint8)`): bit-exact, incl. ragged
    sizes, half-way values and out-of-range reconstructions."""
    from vcm_ts_b200 import frame_io
    rng = np.random.default_rng(5)
    for N, H, W in ((1, 240, 416), (2, 37, 53), (1, 1080, 1920)):
        u8 = rng.integers(0, 256, size=(N, H, W, 3), dtype=np.uint8)
        got = frame_io.frames_from_u8(torch.from_numpy(u8).pin_memory(), device="cuda:0")
        want = u8.astype("float32").transpose(0, 3, 1, 2) / 255.
        Hp, Wp = (H + 63) // 64 * 64, (W + 63) // 64 * 64
        assert tuple(got.shape) == (N, 3, Hp, Wp)
        assert np.array_equal(got[:, :, :H, :W].cpu().numpy(), want)
        assert float(got[:, :, H:, :].abs().sum()) == 0.0 and float(got[:, :, :, W:].abs().sum()) == 0.0
        # way back: exact round trip for in-range frames
        back = frame_io.frames_to_u8(got, H, W)
        assert np.array_equal(back.cpu().numpy(), u8)
        # arbitrary floats incl. x.5 / 255 ties, negatives and > 1
        f = rng.normal(0.5, 0.6, size=(N, 3, Hp, Wp)).astype("float32")
        f[0, 0, 0, :8] = (np.arange(8, dtype="float32") + 0.5) / 255.0
        want_u8 = np.clip(np.rint(f.transpose(0, 2, 3, 1)[:, :H, :W] * 255), 0, 255).astype(np.uint8)
        host = torch.empty((N, H, W, 3), dtype=torch.uint8).pin_memory()
        frame_io.frames_to_u8(torch.from_numpy(f).cuda(), H, W, out=host)
        torch.cuda.synchronize()
        assert np.array_equal(host.numpy(), want_u8)


# ------------------------------------------------------------------------------------------------------------
# BASELINE shapes (padded 1080p = 1088 x 1920): tile-edge / 64-bit-index / N-tiling bugs live here, not at 64 x 96
# ------------------------------------------------------------------------------------------------------------
FULL_CASES = [
    # K, stride, Cin, Cout, post, res   (the shapes that carry the FLOPs)
    (3, 1, 64, 64, 0.01, True),
    (3, 1, 128, 64, None, False),
    (7, 1, 32, 64, 0.0, False),
    (3, 2, 67, 64, 0.1, False),
    (3, 1, 64, 32, 0.01, True),      # row-stacked engine, 4 output rows per MMA (UNet conv1)
    (3, 1, 32, 32, None, False),
]


@pytest.mark.parametrize("case", FULL_CASES)
def test_conv_tcgen05_engine_at_1080p(B16, case):
    """bf16 tcgen05 engine at 1088 x 1920 against cuDNN's fp32 convolution (TF32 off) of the same
    bf16-rounded operands: only accumulation order and the final bf16 rounding differ."""
    K, stride, Cin, Cout, post, use_res = case
    torch.backends.cudnn.allow_tf32 = False
    dev, bf = "cuda:0", torch.bfloat16
    H, W = 1088, 1920
    g = torch.Generator(device="cpu").manual_seed(900 + FULL_CASES.index(case))
    x = torch.randn((1, Cin, H, W), generator=g).to(dev).to(bf)
    w = (torch.randn((Cout, Cin, K, K), generator=g) / (Cin * K * K) ** 0.5).to(dev)
    b = torch.randn((Cout,), generator=g).to(dev)
    Ho, Wo = H // stride, W // stride
    res = torch.randn((1, Cout, Ho, Wo), generator=g).to(dev).to(bf) if use_res else None
    cin_p = (Cin + 15) // 16 * 16
    in_map = list(range(Cin)) + [-1] * (cin_p - Cin) if cin_p != Cin else None
    pc = B16.pack_conv(w, b, in_map=in_map)
    assert pc.w_umma is not None
    xv = view_of(x.float(), dtype=bf, cs=cin_p)
    xv = type(xv)(xv.t, 0, cin_p)
    out = B16.conv(xv, pc, stride=stride, post=post, res=view_of(res.float(), dtype=bf) if use_res else None)
    torch.cuda.synchronize()
    want = ref_conv(x.float(), w.to(bf).float(), b, stride, None, post, False, res.float() if use_res else None,
                    None, None)
    got = nchw(out)
    err = (got - want).abs()
    tol = 0.01 * want.abs() + 0.02
    bad = int((err > tol).sum())
    assert bad == 0, f"{bad} of {err.numel()} outputs off; max err {err.max().item():.4f}"


@pytest.mark.parametrize("case", FULL_CASES[:1] + FULL_CASES[3:])
def test_conv_parity_engine_at_1080p(BPAR, case):
    """Split-bf16 tensor-core engine at 1088 x 1920 against cuDNN fp32 (TF32 off): fp32-class agreement."""
    K, stride, Cin, Cout, post, use_res = case
    torch.backends.cudnn.allow_tf32 = False
    dev = "cuda:0"
    H, W = 1088, 1920
    g = torch.Generator(device="cpu").manual_seed(950 + FULL_CASES.index(case))
    x = torch.randn((1, Cin, H, W), generator=g).to(dev)
    w = (torch.randn((Cout, Cin, K, K), generator=g) / (Cin * K * K) ** 0.5).to(dev)
    b = torch.randn((Cout,), generator=g).to(dev)
    res = torch.randn((1, Cout, H // stride, W // stride), generator=g).to(dev) if use_res else None
    pc = BPAR.pack_conv(w, b)
    out = BPAR.conv(view_of(x), pc, stride=stride, post=post, res=view_of(res) if use_res else None)
    torch.cuda.synchronize()
    want = ref_conv(x, w, b, stride, None, post, False, res, None, None)
    err = (nchw(out) - want).abs().max().item()
    assert err <= 2e-5, err            # |y| ~ 1..6: a few fp32 ulps of either summation order


def test_flow_warp_config3_shapes(B32, B16):
    """BASELINE config 3: flow_warp at B = 8, 1920 x 1088 (64-channel features in fp32 and bf16, 3-channel
    frames) against the oracle's torch_warp (video_net.py:32-50) evaluated by ATen on the same device
    tensors -- including flows that leave the image (border clamp)."""
    from oracle import dcvc_oracle as O
    dev = "cuda:0"
    N, H, W = 8, 1088, 1920
    g = torch.Generator().manual_seed(31)
    flow = (torch.randn((N, 2, H // 8, W // 8), generator=g) * 6).to(dev)
    flow = F.interpolate(flow, size=(H, W), mode="bilinear", align_corners=False).contiguous()
    flow[:, :, :16] *= 30
    flow[:, :, :, -16:] *= 30
    fv = view_of(flow)
    # fp32 tolerance: sample coordinates reach ~2000 px here (ulp 1.2e-4 px) and ATen works in normalised
    # coordinates, so a few ulps of coordinate noise times the feature gradient (<= ~2 per pixel for these
    # 4x-upsampled random features) is the reference's own rounding noise; the 64 x 96 test uses 2e-4.
    for C, dt, B, atol in ((3, torch.float32, B32, 1e-3), (64, torch.float32, B32, 1e-3),
                           (64, torch.bfloat16, B16, 2e-2)):
        feat = F.interpolate(torch.randn((N, C, H // 4, W // 4), generator=g).to(dev), size=(H, W), mode="bilinear",
                             align_corners=False).contiguous()
        if dt == torch.bfloat16:
            feat = feat.bfloat16().float()
        want = O.torch_warp(feat, flow)
        got = B.warp(view_of(feat, dtype=dt, cs=C + (8 if C > 3 else 5)), fv)
        diff = (nchw(got) - (want if dt == torch.float32 else want.bfloat16().float())).abs().max().item()
        assert diff <= atol, (C, dt, diff)
        del feat, want, got
        torch.cuda.empty_cache()


def test_kernels_do_not_write_outside_their_outputs(B16, BPAR, B32):
    """compute-sanitizer is closed on this GPU pool, so out-of-bounds WRITES are caught with guard bands: every
    output lives inside a larger sentinel-filled allocation (whole guard tensors before / after it, and guard
    channels left and right of the view inside each pixel); after the kernel all sentinels must be intact.
    Ragged sizes on purpose (tile tails of the tcgen05 engines, odd channel counts)."""
    from vcm_ts_b200.backend import View
    dev = "cuda:0"
    g = torch.Generator().manual_seed(5)
    SENT = 12345.0

    def guarded(N, H, W, C, dtype, pad_c=8):
        big = torch.full((N + 2, H, W, C + 2 * pad_c), SENT, dtype=dtype, device=dev)
        return big, View(big[1:N + 1], pad_c, C)

    def intact(big, N, C, pad_c=8):
        s = torch.tensor(SENT, dtype=big.dtype, device=dev)
        return bool((big[0] == s).all() and (big[N + 1] == s).all() and (big[1:N + 1, ..., :pad_c] == s).all()
                    and (big[1:N + 1, ..., pad_c + C:] == s).all())

    cases = [(B16, torch.bfloat16, 3, 1, 64, 64, 37, 53), (B16, torch.bfloat16, 3, 2, 64, 96, 38, 54),
             (B16, torch.bfloat16, 7, 1, 32, 16, 21, 19), (B16, torch.bfloat16, 3, 1, 192, 192, 19, 23),
             (BPAR, torch.float32, 3, 1, 64, 64, 37, 53), (BPAR, torch.float32, 3, 2, 67, 64, 38, 54),
             (B32, torch.float32, 3, 1, 67, 3, 21, 19)]
    for Bk, dt, K, stride, Cin, Cout, H, W in cases:
        w = (torch.randn((Cout, Cin, K, K), generator=g) / (Cin * K * K) ** 0.5).to(dev)
        pc = Bk.pack_conv(w, torch.zeros(Cout, device=dev))
        x = view_of(torch.randn((2, Cin, H, W), generator=g).to(dev), dtype=dt,
                    cs=(Cin + 15) // 16 * 16 if dt == torch.bfloat16 else Cin)
        Ho, Wo = (H + 2 * (K // 2) - K) // stride + 1, (W + 2 * (K // 2) - K) // stride + 1
        big, out = guarded(2, Ho, Wo, Cout, dt)
        Bk.conv(x, pc, out=out, stride=stride, post=0.1)
        torch.cuda.synchronize()
        assert intact(big, 2, Cout), ("conv", Bk.precision, K, stride, Cin, Cout)
        assert torch.isfinite(out.dense().float()).all() and float(out.dense().float().abs().max()) < 1e3
    # warp / maxpool / copy into guarded views
    feat = view_of(torch.randn((2, 64, 37, 53), generator=g).to(dev), dtype=torch.bfloat16)
    flow = view_of((torch.randn((2, 2, 37, 53), generator=g) * 30).to(dev))
    big, out = guarded(2, 37, 53, 64, torch.bfloat16)
    B16.warp(feat, flow, out=out)
    torch.cuda.synchronize()
    assert intact(big, 2, 64), "warp"
    big, out = guarded(2, 18, 26, 64, torch.bfloat16)
    B16.maxpool(feat, out=out)
    torch.cuda.synchronize()
    assert intact(big, 2, 64), "maxpool"


def test_fused_spynet_glue_equals_the_separate_kernels(B32):
    """vcm_spynet_pyramid / vcm_spynet_level_input against chains of nchw->nhwc, avgpool2, flow_up2, flow_warp
    (each checked against ATen elsewhere): bit-identical, including the bf16 16-channel copy."""
    from vcm_ts_b200.backend import View
    dev = "cuda:0"
    g = torch.Generator().manual_seed(21)
    N, H, W = 2, 64, 96
    x = torch.rand((N, 3, H, W), generator=g).to(dev)
    ref = torch.rand((N, 3, H, W), generator=g).to(dev)
    lvl = [B32.new(N, H >> k, W >> k, 8, dtype=torch.float32, zero=True) for k in range(4)]
    im2 = [B32.new(N, H >> k, W >> k, 3, dtype=torch.float32) for k in range(4)]
    B32.spynet_pyramid(x, ref, lvl, im2)
    w1 = [B32.from_nchw(x)]
    w2 = [B32.from_nchw(ref)]
    for k in range(1, 4):
        w1.append(B32.avgpool(w1[k - 1]))
        w2.append(B32.avgpool(w2[k - 1]))
    for k in range(4):
        assert torch.equal(lvl[k].dense()[..., :3], w1[k].dense()), k
        assert torch.equal(im2[k].dense(), w2[k].dense()), k
        assert torch.equal(nchw(w1[k]), F.avg_pool2d(x, 2 ** k, 2 ** k)) or k > 0
    k = 1
    flow_prev = View((torch.randn((N, H >> 2, W >> 2, 2), generator=g) * 3).to(dev).contiguous())
    buf16 = B32.new(N, H >> k, W >> k, 16, dtype=torch.bfloat16)
    B32.spynet_level_input(flow_prev, im2[k], lvl[k], buf16)
    up = B32.flow_up2(flow_prev, gain=2.0)
    wr = B32.warp(w2[k], up)
    got = lvl[k].dense()
    assert torch.equal(got[..., 6:8], up.dense())
    assert torch.equal(got[..., 3:6], wr.dense())
    assert torch.equal(got[..., 0:3], w1[k].dense())
    assert torch.equal(buf16.dense()[..., :8].float(), got.bfloat16().float())
    assert float(buf16.dense()[..., 8:].float().abs().max()) == 0.0
