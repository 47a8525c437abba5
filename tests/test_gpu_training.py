"""`-m gpu` tests of the training path: the backward kernels behind the
C-ABI against ATen's autograd of the ops the reference trains through, and train-mode
`DMC.forward_one_frame` on the B200 against the golden written by the unmodified reference's own autograd
(tests/golden/train_64x128.npz).
"""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CL = torch.channels_last


def _backend(precision):
    from vcm_ts_b200.backend import CudaBackend
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return CudaBackend("cuda:0", precision)


def _view(t):
    from vcm_ts_b200.backend import View
    t = t.contiguous(memory_format=CL)
    return View(t.permute(0, 2, 3, 1))


WG_CASES = [
    # K, stride, Cin, Cout, H, W, N
    (3, 1, 64, 64, 24, 40, 2), (3, 2, 67, 64, 32, 48, 1), (7, 1, 8, 32, 20, 28, 1), (1, 1, 64, 256, 17, 30, 1),
    (1, 2, 2, 64, 32, 48, 2), (3, 1, 480, 384, 9, 15, 1), (7, 1, 16, 2, 24, 40, 1), (3, 1, 128, 64, 33, 47, 1),
    # tensor-core path (bf16, Cin % 16 == 0): strides, Cout tails, small and large images
    (3, 2, 64, 96, 32, 48, 2), (1, 2, 64, 64, 32, 32, 1), (3, 1, 64, 64, 256, 256, 4), (7, 1, 32, 64, 64, 64, 2),
    (3, 1, 96, 144, 16, 16, 4), (3, 1, 192, 192, 16, 16, 1), (3, 2, 144, 192, 32, 32, 1), (1, 1, 96, 384, 4, 4, 4),
]


@pytest.mark.parametrize("case", WG_CASES)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_wgrad_kernel(case, dtype):
    """vcm_conv2d_wgrad against autograd of F.conv2d (fp64 on the same, dtype-rounded operands)."""
    K, stride, Cin, Cout, H, W, N = case
    B = _backend("fp32")
    dev = "cuda:0"
    g = torch.Generator().manual_seed(WG_CASES.index(case))
    x = torch.randn((N, Cin, H, W), generator=g).to(dev).to(dtype)
    Ho, Wo = (H + 2 * (K // 2) - K) // stride + 1, (W + 2 * (K // 2) - K) // stride + 1
    dy = torch.randn((N, Cout, Ho, Wo), generator=g).to(dev).to(dtype)
    got = B.wgrad(_view(x), _view(dy), K, stride, Cin, Cout)
    w = torch.zeros((Cout, Cin, K, K), dtype=torch.float64, device=dev, requires_grad=True)
    y = F.conv2d(x.double(), w, None, stride, K // 2)
    (want,) = torch.autograd.grad(y, w, dy.double())
    err = (got.double() - want).abs().max().item()
    assert err <= 2e-5 * max(1.0, want.abs().max().item()) * (N * Ho * Wo) ** 0.5, err
    assert torch.equal(got, B.wgrad(_view(x), _view(dy), K, stride, Cin, Cout)), "wgrad must be deterministic"


@pytest.mark.parametrize("C,dtype", [(3, torch.float32), (64, torch.float32), (64, torch.bfloat16)])
def test_flow_warp_backward_kernel(C, dtype):
    from oracle import dcvc_oracle as O
    B = _backend("fp32")
    dev = "cuda:0"
    g = torch.Generator().manual_seed(7)
    N, H, W = 2, 40, 56
    feat = torch.randn((N, C, H, W), generator=g).to(dev).to(dtype)
    flow = (torch.randn((N, 2, H, W), generator=g) * 3).to(dev)
    flow[:, :, :3] *= 20          # border clamp: zero flow gradient there
    gout = torch.randn((N, C, H, W), generator=g).to(dev).to(dtype)
    gfeat, gflow = B.warp_bwd(_view(feat), _view(flow), _view(gout))
    f64 = feat.double().requires_grad_(True)
    fl64 = flow.double().requires_grad_(True)
    y = O.torch_warp(f64, fl64)
    wf, wfl = torch.autograd.grad(y, (f64, fl64), gout.double())
    tol = 1e-4 if dtype == torch.float32 else 1e-3
    assert (gfeat.permute(0, 3, 1, 2).double() - wf).abs().max().item() <= tol * max(1.0, wf.abs().max().item())
    # the reference computes the sample position in fp32 normalised coordinates; at |x| ~ 50 px one ulp of that
    # moves a bilinear weight by ~4e-6, times sum_c |g_c * dfeat| ~ C
    assert (gflow.permute(0, 3, 1, 2).double() - wfl).abs().max().item() <= 2e-3 * max(1.0, wfl.abs().max().item())


CONVFN_CASES = [(3, 1, 64, 64, 0.01), (3, 2, 64, 96, None), (1, 2, 64, 64, None), (7, 1, 32, 16, 0.0), (1, 1, 64, 256, 0.1)]


@pytest.mark.parametrize("precision", ["fp32", "parity", "bf16"])
@pytest.mark.parametrize("case", CONVFN_CASES)
def test_conv_function_gradients(precision, case):
    """training.ConvFn (forward engine, dgrad = forward engine on flipped/transposed weights, wgrad kernel)
    against ATen autograd of conv2d + leaky_relu."""
    from vcm_ts_b200.training import TrainGraph
    K, stride, Cin, Cout, post = case
    B = _backend(precision)
    dev = "cuda:0"

    class RT:
        pass
    rt = RT()
    rt.B, rt.packed = B, {}
    m = torch.nn.Module()
    m.c = torch.nn.Conv2d(Cin, Cout, K, padding=K // 2).to(dev)
    tg = TrainGraph(m, rt)
    g = torch.Generator().manual_seed(11)
    x = torch.randn((2, Cin, 24, 40), generator=g).to(dev).requires_grad_(True)
    # bf16 precision: the reference sees the same bf16-rounded operands, otherwise activation-mask flips of
    # near-zero outputs (a factor 1/slope in the gradient) dominate the comparison.  The reference runs in
    # fp64: cuDNN's fp32 algorithm choice (Winograd: ~1e-5 absolute error) is not a usable reference for the
    # sign of outputs near zero, and the output gradient is zeroed where |y| < 1e-4 for the same reason.
    rnd = (lambda t: t.bfloat16().float()) if precision == "bf16" else (lambda t: t)
    x2 = rnd(x.detach()).double().requires_grad_(True)
    w2 = rnd(m.c.weight.detach()).double().requires_grad_(True)
    b2 = m.c.bias.detach().double().requires_grad_(True)
    y2 = F.conv2d(x2, w2, b2, stride, K // 2)
    go = torch.randn(y2.shape, generator=g).to(dev)
    if post is not None:
        go = go * (y2.detach().abs() > 1e-4)
        y2 = F.leaky_relu(y2, post)
    gx2, gw2, gb2 = torch.autograd.grad(y2, (x2, w2, b2), go.double())
    y = tg.conv("c", x, stride=stride, post=post, out_dtype=torch.float32)
    gx, gw, gb = torch.autograd.grad(y, (x, m.c.weight, m.c.bias), go)
    y2, gx2, gw2, gb2 = (t.detach().float() for t in (y2, gx2, gw2, gb2))
    rel = lambda a, b: ((a.float() - b).norm() / b.norm()).item()   # noqa: E731
    tol = 3e-2 if precision == "bf16" else 2e-5
    errs = dict(y=rel(y, y2), gx=rel(gx, gx2), gw=rel(gw, gw2), gb=rel(gb, gb2))
    assert all(v <= tol for v in errs.values()), errs


@pytest.mark.parametrize("precision", ["fp32", "parity"])
def test_train_mode_matches_reference_autograd_on_gpu(precision, golden_dir):
    """Two chained train-mode P-frames (batch = 4 anchors, learned q-scale Parameters, fixed noise) on the
    B200 against the golden from the unmodified reference's autograd: outputs, loss, the gradient norm of
    every parameter, a few gradients in full."""
    from oracle import dcvc_oracle as O
    from oracle import fixture
    from oracle import gen_train_golden as T
    from tests.test_training_cpu import check_against_train_golden
    from vcm_ts_b200.models import DMC
    torch.backends.cudnn.allow_tf32 = False
    G = np.load(os.path.join(golden_dir, "train_64x128.npz"))
    dmc = DMC(anchor_num=T.N, precision=precision)
    dmc.load_state_dict(T.train_state_dict(), strict=True)
    dmc = dmc.to("cuda:0").train()
    rep = check_against_train_golden(dmc, G, device="cuda:0", rtol_out=2e-5)
    rep["precision"] = precision
    rep["launches"] = dmc._rt.B.launches
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"train_parity_{precision}.json"), "w") as f:
        json.dump(rep, f, indent=1)
    print(rep)


def test_train_mode_bf16_close_and_optimizer_step(golden_dir):
    """bf16 training precision (config 5: bf16 autocast-equivalent): outputs within 2 % of the fp32 reference,
    gradient direction preserved (cosine > 0.9 on the full-gradient samples), and an AdamW step on every
    parameter changes the loss (the whole loop is differentiable and the packed-weight caches follow the
    parameter versions)."""
    from oracle import dcvc_oracle as O
    from oracle import fixture
    from oracle import gen_train_golden as T
    from tests.test_training_cpu import _noise_feeder
    from vcm_ts_b200.models import DMC
    dev = "cuda:0"
    G = np.load(os.path.join(golden_dir, "train_64x128.npz"))
    dmc = DMC(anchor_num=T.N, precision="bf16")
    dmc.load_state_dict(T.train_state_dict(), strict=True)
    dmc = dmc.to(dev).train()
    opt = torch.optim.AdamW(dmc.parameters(), lr=1e-4, betas=(0.9, 0.99))     # core/solver/optimizer.py:13
    xs, ref = T.train_inputs()
    lam = torch.tensor(T.LAMBDAS, device=dev)
    dpb = {"ref_frame": ref.to(dev), "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    losses = []
    for it in range(2):
        opt.zero_grad()
        dmc._noise_fn = _noise_feeder(T.train_noise(0, T.noise_shapes()))
        out = dmc.forward_one_frame(xs[0].to(dev), dpb, dmc.mv_y_q_scale, dmc.y_q_scale)
        loss = torch.mean(out["bpp"] + lam * out["mse"])
        loss.backward()
        if it == 0:
            np.testing.assert_allclose(loss.item(), G["s0_loss"][0], rtol=2e-2)
            for k in ("bpp", "mse"):
                np.testing.assert_allclose(out[k].detach().cpu().numpy(), G["s0_" + k], rtol=2e-2, err_msg=k)
            params = dict(dmc.named_parameters())
            for key in G.files:
                if key.startswith("s0_grad:"):
                    g = params[key.split(":", 1)[1]].grad.detach().float().cpu().numpy().ravel()
                    w = G[key].ravel()
                    cos = float(np.dot(g, w) / (np.linalg.norm(g) * np.linalg.norm(w) + 1e-30))
                    assert cos > 0.9, (key, cos)
            assert all(p.grad is not None and torch.isfinite(p.grad).all() for k, p in params.items()
                       if not G["s0_grad_none"][list(G["param_names"]).index(k)])
        opt.step()
        losses.append(loss.item())
    assert losses[1] != losses[0] and np.isfinite(losses).all(), losses
