"""oracle/gen_train_golden.py -- TEST INFRASTRUCTURE: TRAIN-mode golden vectors from the REAL reference.

Run in the build container only (needs /root/reference):

    python oracle/gen_train_golden.py        -> tests/golden/train_64x128.npz

The unmodified reference `DMC` (oracle/ref_import.py) in `.train()` mode, fixture weights, batch of
4 anchors with the learned `[4,1,1,1]` q-scale Parameters exactly as the trainer calls it
(core/model/dcvc_hem.py:283-289): two P-frames (the second from the detached dpb of the first,
:290-291), loss = mean(bpp + lambda * mse) (:305-307 without the perceptual term), `loss.backward()`.
The file holds the outputs, the loss, the L2 norm of EVERY parameter's gradient and a few small
gradients in full.  The reference draws its additive noise from the global RNG
(common_model.py:46-49); `train_noise()` below reproduces that stream so the oracle restatement and
the CUDA product can be fed the very same tensors.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import fixture  # noqa: E402

PH, PW, N = 64, 128, 4
LAMBDAS = (85.0, 170.0, 380.0, 840.0)
NOISE_SEED = (4242, 4243)
FULL_GRADS = ("mv_y_q_scale", "y_q_scale", "mv_y_q_basic", "y_q_basic", "bit_estimator_z.f1.h",
              "bit_estimator_z_mv.f4.b", "recon_generation_net.recon_conv.bias",
              "optic_flow.moduleBasic.3.conv5.bias", "mv_encoder.6.bias", "y_spatial_prior.4.bias",
              "feature_adaptor_I.weight")


def train_state_dict():
    """Fixture weights (oracle/fixture.py) with the biases of the four layers that predict the entropy SCALES
    raised to 1.5.  With plain random weights about half of the predicted scales are negative and the
    reference clamps them to sigma = 1e-5 (common_model.py:66), where d bits / d y is ~1e5 times steeper than
    anywhere a trained model operates: single gradient elements then depend on the last ulp of a convolution
    and no two correct implementations agree on them.  Positive scales keep every term of the training
    objective smooth, so gradients can be compared tightly."""
    from oracle.dcvc_oracle import state_shapes
    P = fixture.fixture_state_dict(state_shapes("dmc"))
    for pre, C in (("mv_y", 64), ("y", 96)):
        b = P[f"{pre}_prior_fusion.4.bias"].clone()
        b[C:2 * C] = 1.5                      # chunk order (q_step, scales, means): video_model.py:293,322
        P[f"{pre}_prior_fusion.4.bias"] = b
        b = P[f"{pre}_spatial_prior.4.bias"].clone()
        h = C // 2
        b[0:h] = 1.5                          # chunk order (scales_0, means_0, scales_1, means_1): common_model.py:135
        b[C:C + h] = 1.5
        P[f"{pre}_spatial_prior.4.bias"] = b
    return P


def train_inputs():
    """(x [2][4,3,64,128], ref_frame [4,3,64,128]): four anchors = four shifted crops of one moving texture."""
    fr = fixture.fixture_frames(PH, PW, N + 2, seed=5)
    return [fr[1:N + 1].contiguous(), fr[2:N + 2].contiguous()], fr[0:N].contiguous()


def train_noise(step, shapes):
    """The four U(-0.5, 0.5) tensors the reference's add_noise draws in forward_one_frame, in its order
    (y_res, mv_y_res, z, mv_z: video_model.py:547-550), after torch.manual_seed(NOISE_SEED[step])."""
    torch.manual_seed(NOISE_SEED[step])
    return [torch.nn.init.uniform_(torch.zeros(s), -0.5, 0.5) for s in shapes]


def noise_shapes(n=N, h=PH, w=PW):
    return [(n, 96, h // 16, w // 16), (n, 64, h // 16, w // 16), (n, 64, h // 64, w // 64), (n, 64, h // 64, w // 64)]


def loss_of(out):
    lam = torch.tensor(LAMBDAS)
    return torch.mean(out["bpp"] + lam * out["mse"])


def main():
    from oracle.ref_import import import_reference
    torch.set_num_threads(4)
    DMC, _, _ = import_reference()
    dmc = DMC(anchor_num=N)
    dmc.load_state_dict(train_state_dict(), strict=True)
    dmc.train()
    xs, ref = train_inputs()
    G = {"lambdas": np.array(LAMBDAS), "noise_seed": np.array(NOISE_SEED)}
    dpb = {"ref_frame": ref, "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    names = [k for k, _ in dmc.named_parameters()]
    G["param_names"] = np.array(names)
    for step, x in enumerate(xs):
        p = f"s{step}_"
        dmc.zero_grad(set_to_none=True)
        torch.manual_seed(NOISE_SEED[step])
        out = dmc.forward_one_frame(x, dpb, dmc.mv_y_q_scale, dmc.y_q_scale)
        loss = loss_of(out)
        loss.backward()
        G[p + "loss"] = np.array([loss.item()])
        for k in ("bpp_mv_y", "bpp_mv_z", "bpp_y", "bpp_z", "bpp", "mse", "me_mse"):
            G[p + k] = out[k].detach().numpy()
        norms = []
        for k, prm in dmc.named_parameters():
            norms.append(0.0 if prm.grad is None else prm.grad.double().norm().item())
        G[p + "grad_norms"] = np.array(norms)
        G[p + "grad_none"] = np.array([prm.grad is None for _, prm in dmc.named_parameters()])
        for k in FULL_GRADS:
            g = dict(dmc.named_parameters())[k].grad
            if g is not None:
                G[p + "grad:" + k] = g.numpy()
        G[p + "ref_frame"] = out["dpb"]["ref_frame"].detach().numpy()
        dpb = {k: v.detach() for k, v in out["dpb"].items()}      # core/model/dcvc_hem.py:290-291
    path = os.path.join(os.path.dirname(HERE), "tests", "golden", "train_64x128.npz")
    np.savez_compressed(path, **G)
    print("->", path, f"{os.path.getsize(path) / 1e6:.2f} MB", {k: float(G[k][0]) for k in ("s0_loss", "s1_loss")})


if __name__ == "__main__":
    main()
