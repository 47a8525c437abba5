"""CPU tests of the TRAINING path's host logic (vcm_ts_b200/training.py): the differentiable graph, its
autograd Functions and the DDP step, with the device kernels replaced by tests/emul_backend.py.

  * train-mode `DMC.forward_one_frame` against the golden written by the UNMODIFIED reference's own
    autograd (tests/golden/train_64x128.npz, oracle/gen_train_golden.py): outputs, loss, the gradient
    norm of every parameter, a few gradients in full -- two chained P-frames, batch = 4 anchors with the
    `[4,1,1,1]` q-scale Parameters, exactly the call of core/model/dcvc_hem.py:283-291;
  * world_size-2 gloo DistributedDataParallel step (core/engine/train_multi.py:174-191 wraps the model the
    same way): gradients after backward are the mean over ranks and identical on both ranks.
The CUDA kernels behind the same Functions are checked in tests/test_gpu_training.py (-m gpu).
"""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import dcvc_oracle as O
from oracle import fixture
from oracle import gen_train_golden as T
from tests.emul_backend import EmulBackend
from vcm_ts_b200.models import DMC

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _train_model():
    dmc = DMC(anchor_num=T.N, precision="fp32")
    dmc.load_state_dict(T.train_state_dict(), strict=True)
    dmc._backend_factory = lambda dev, prec: EmulBackend(dev, prec)
    return dmc.train()


def _noise_feeder(tensors):
    it = iter(tensors)

    def fn(x):
        n = next(it)
        assert n.shape == x.shape, (n.shape, x.shape)
        return n.to(x.device)
    return fn


def check_against_train_golden(dmc, G, device="cpu", rtol_out=1e-5, rtol_grad=2e-3):
    """Shared with the GPU test: runs the two golden steps and compares (weights: gen_train_golden
    train_state_dict, which keeps the predicted entropy scales positive so the objective is smooth)."""
    xs, ref = T.train_inputs()
    names = [str(n) for n in G["param_names"]]
    params = dict(dmc.named_parameters())
    assert sorted(names) == sorted(params), "parameter tree differs from the reference's"
    dpb = {"ref_frame": ref.to(device), "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    lam = torch.tensor(T.LAMBDAS, device=device)
    report = {}
    for step, x in enumerate(xs):
        p = f"s{step}_"
        dmc.zero_grad(set_to_none=True)
        dmc._noise_fn = _noise_feeder(T.train_noise(step, T.noise_shapes()))
        out = dmc.forward_one_frame(x.to(device), dpb, dmc.mv_y_q_scale, dmc.y_q_scale)
        loss = torch.mean(out["bpp"] + lam * out["mse"])
        loss.backward()
        np.testing.assert_allclose(loss.item(), G[p + "loss"][0], rtol=rtol_out)
        for k in ("bpp_mv_y", "bpp_mv_z", "bpp_y", "bpp_z", "bpp", "mse", "me_mse"):
            np.testing.assert_allclose(out[k].detach().cpu().numpy(), G[p + k], rtol=rtol_out, err_msg=k)
        none = np.array([params[k].grad is None for k in names])
        assert np.array_equal(none, G[p + "grad_none"]), \
            [n for n, a, b in zip(names, none, G[p + "grad_none"]) if a != b]
        norms = np.array([0.0 if params[k].grad is None else params[k].grad.double().norm().item() for k in names])
        want = G[p + "grad_norms"]
        rel = np.abs(norms - want) / np.maximum(want, 1e-7 * want.max())
        report[p + "max_rel_grad_norm_err"] = float(rel.max())
        report[p + "worst_param"] = names[int(rel.argmax())]
        np.testing.assert_allclose(norms, want, rtol=rtol_grad, atol=1e-7 * want.max())
        for key in G.files:
            if key.startswith(p + "grad:"):
                g = params[key.split(":", 1)[1]].grad.detach().cpu().numpy()
                err = np.linalg.norm((g - G[key]).ravel()) / max(np.linalg.norm(G[key].ravel()), 1e-30)
                report[p + "relerr:" + key.split(":", 1)[1]] = float(err)
                assert err <= rtol_grad, (key, err)
        dpb = {k: v.detach() for k, v in out["dpb"].items()}
    return report


def test_train_mode_matches_reference_autograd(golden_dir):
    torch.set_num_threads(4)
    G = np.load(os.path.join(golden_dir, "train_64x128.npz"))
    rep = check_against_train_golden(_train_model(), G)
    print(rep)


def test_eval_mode_is_unchanged_by_the_training_hook():
    """eval() keeps running the fused inference graph and returns tensors without gradients."""
    dmc = _train_model().eval()
    xs, ref = T.train_inputs()
    dpb = {"ref_frame": ref[:1], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    out = dmc.forward_one_frame(xs[0][:1], dpb, 1.0, 1.0)
    assert not out["bpp"].requires_grad and not out["dpb"]["ref_frame"].requires_grad


def _ddp_worker(rank, world, port, q):
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    torch.manual_seed(0)

    class Wrapper(torch.nn.Module):          # core/model/dcvc_hem.py: the trainer wraps `dmc` in a module
        def __init__(self):
            super().__init__()
            self.dmc = _train_model()

        def forward(self, x, dpb):
            out = self.dmc.forward_one_frame(x, dpb, self.dmc.mv_y_q_scale, self.dmc.y_q_scale)
            lam = torch.tensor(T.LAMBDAS)
            return torch.mean(out["bpp"] + lam * out["mse"])

    model = Wrapper()
    ddp = DDP(model, find_unused_parameters=True)        # core/engine/train_multi.py:177
    xs, ref = T.train_inputs()
    x = xs[rank % 2]                                     # every rank its own data (DistributedSampler)
    model.dmc._noise_fn = _noise_feeder(T.train_noise(rank % 2, T.noise_shapes()))
    dpb = {"ref_frame": ref, "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    loss = ddp(x, dpb)
    loss.backward()
    grads = {k: (None if p.grad is None else p.grad.clone()) for k, p in model.dmc.named_parameters()}
    # the reference's single_step through vcm_ts_b200.trainer (DDP wrap + AdamW + detached dpb hand-over)
    from vcm_ts_b200.trainer import Trainer
    tr = Trainer(_train_model(), T.LAMBDAS, lr=1e-3, world_size=world)
    seq = torch.stack([ref, xs[0], xs[1]], dim=1)          # [N, T=3, C, H, W]; every rank a different window
    tr.module.dmc._noise_fn = lambda t: torch.zeros_like(t)
    before = tr.module.dmc.get_parameter("mv_encoder.6.bias").detach().clone()
    res = tr.single_step(seq, seq, t_i=rank, p_frames=1)
    after = tr.module.dmc.get_parameter("mv_encoder.6.bias").detach().clone()
    q.put((rank, loss.item(), {k: (None if g is None else g.numpy()) for k, g in grads.items()
                                if k in T.FULL_GRADS or k.endswith("conv1.bias")},
           dict(before=before.numpy(), after=after.numpy(), loss=res[0]["loss_to_opt"].item(),
                dpb_detached=not res[0]["dpb"]["ref_frame"].requires_grad)))
    dist.barrier()
    dist.destroy_process_group()


def test_ddp_gloo_two_ranks_average_gradients():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_ddp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=600) for _ in range(2)), key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    (_, l0, g0, t0), (_, l1, g1, t1) = res
    assert l0 != l1                                     # different data per rank
    # Trainer.single_step: both ranks applied the same (all-reduced) update, and it changed the weights
    assert t0["loss"] != t1["loss"] and t0["dpb_detached"] and t1["dpb_detached"]
    np.testing.assert_array_equal(t0["after"], t1["after"])
    assert not np.array_equal(t0["before"], t0["after"])
    for k in g0:
        if g0[k] is None:
            assert g1[k] is None
            continue
        np.testing.assert_array_equal(g0[k], g1[k], err_msg=k)     # all-reduced: identical on both ranks
    # and equal to the mean of the two single-process gradients
    singles = []
    for step in range(2):
        dmc = _train_model()
        xs, ref = T.train_inputs()
        dmc._noise_fn = _noise_feeder(T.train_noise(step, T.noise_shapes()))
        dpb = {"ref_frame": ref, "ref_feature": None, "ref_y": None, "ref_mv_y": None}
        out = dmc.forward_one_frame(xs[step], dpb, dmc.mv_y_q_scale, dmc.y_q_scale)
        torch.mean(out["bpp"] + torch.tensor(T.LAMBDAS) * out["mse"]).backward()
        singles.append({k: p.grad for k, p in dmc.named_parameters()})
    for k in g0:
        if g0[k] is None:
            continue
        want = 0.5 * (singles[0][k] + singles[1][k]).numpy()
        np.testing.assert_allclose(g0[k], want, rtol=1e-4, atol=1e-6 * np.abs(want).max(), err_msg=k)


def test_intra_train_mode_gradients_and_eval_consistency():
    """IntraNoAR.forward in train() mode (image_model.py:54-105): gradients reach every parameter, and with
    the noise switched off its rate / distortion stay close to the eval-mode (hard rounding) forward of the
    fused inference graph on the same weights (same graph, same quantised latents; only the rate term sees
    y_res instead of round(y_res))."""
    from vcm_ts_b200.models import IntraNoAR
    torch.set_num_threads(4)
    intra = IntraNoAR(precision="fp32")
    intra.load_state_dict(fixture.fixture_state_dict(O.state_shapes("intra")), strict=True)
    intra._backend_factory = lambda dev, prec: EmulBackend(dev, prec)
    x = fixture.fixture_frames(64, 128, 2, seed=3)
    intra.train()
    intra._noise_fn = lambda t: torch.zeros_like(t)
    out = intra(x, intra.q_scale[:2])
    (out["bpp"].sum() + 100.0 * out["mse"].sum()).backward()
    missing = [k for k, p in intra.named_parameters() if p.grad is None]
    assert not missing, missing
    assert all(torch.isfinite(p.grad).all() for p in intra.parameters())
    intra.eval()
    ref = intra(x, intra.q_scale[:2].detach())
    torch.testing.assert_close(out["x_hat"].detach(), ref["x_hat"], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(out["mse"].detach(), ref["mse"], rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(out["bpp_z"].detach(), ref["bpp_z"], rtol=0.2, atol=1e-3)
