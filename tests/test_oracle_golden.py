"""Pins the oracle restatement (oracle/dcvc_oracle.py + oracle/rans_oracle.c) BIT-EXACTLY to the
golden vectors generated from the real reference (oracle/gen_golden.py, tests/golden/*.npz).
CPU only.  The reference has no tests of its own."""
import os

import numpy as np
import pytest
import torch

from oracle import fixture
from oracle import dcvc_oracle as O

THREADS = 4          # gen_golden.py used 4 threads; oneDNN conv results depend on it


def _pad64(x):
    h, w = x.shape[-2:]
    return torch.nn.functional.pad(x, (0, (64 - w % 64) % 64, 0, (64 - h % 64) % 64))


@pytest.fixture(scope="module")
def models():
    torch.set_num_threads(THREADS)
    dmc = O.OracleDMC(fixture.fixture_state_dict(O.state_shapes("dmc")))
    intra = O.OracleIntra(fixture.fixture_state_dict(O.state_shapes("intra")))
    dmc.update()
    intra.update()
    return dmc, intra


def test_cdf_tables(models, golden_dir):
    dmc, intra = models
    G = np.load(os.path.join(golden_dir, "cdf_tables.npz"))
    for nm, t in (("dmc_y", dmc.y_tables), ("dmc_z", dmc.z_tables), ("dmc_z_mv", dmc.z_mv_tables),
                  ("intra_y", intra.y_tables), ("intra_z", intra.z_tables)):
        for suffix, mine in zip(("_cdf", "_len", "_off"), t):
            assert np.array_equal(G[nm + suffix], mine), nm + suffix


@pytest.mark.parametrize("case", ["tiny_64x128", "cfg1_240x416"])
def test_gop_prefix_bit_exact(models, golden_dir, case):
    dmc, intra = models
    G = np.load(os.path.join(golden_dir, case + ".npz"))
    ph, pw, qi, qmv, qy = G["meta"]
    ph, pw = int(ph), int(pw)
    fr = torch.from_numpy(G["frames_u8"])
    x0 = _pad64(fr[0:1].float() / 255.0)
    r, blob = intra.encode_decode_bytes(x0, qi, ph, pw)
    assert blob == G["i_bin"].tobytes()
    assert r["bit"] == int(G["i_bit"][0])
    assert np.array_equal(r["x_hat"].numpy(), G["i_x_hat"])
    fwd = intra.forward(x0, qi)
    assert np.array_equal(fwd["bpp_y"].numpy(), G["i_fwd_bpp_y"])
    assert np.array_equal(fwd["bpp_z"].numpy(), G["i_fwd_bpp_z"])
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    for t in (1, 2):
        p = f"p{t}_"
        x = _pad64(fr[t:t + 1].float() / 255.0)
        mvq, _ = O.get_rounded_q(qmv)
        yq, _ = O.get_rounded_q(qy)
        enc = dmc.compress(x, dpb, mvq, yq, want_symbols=True)
        assert enc["bit_stream"] == G[p + "stream"].tobytes()
        S = enc["symbols"]
        for mine, gold in (("mv_z", "sym_mv_z"), ("z", "sym_z"), ("mv_y_q_w_0", "sym_mv_y_q_w_0"),
                           ("mv_y_q_w_1", "sym_mv_y_q_w_1"), ("y_q_w_0", "sym_y_q_w_0"),
                           ("y_q_w_1", "sym_y_q_w_1"), ("mv_idx_0", "idx_mv_0"), ("mv_idx_1", "idx_mv_1"),
                           ("y_idx_0", "idx_y_0"), ("y_idx_1", "idx_y_1")):
            assert np.array_equal(S[mine].numpy(), G[p + gold]), gold
        est = dmc.forward_one_frame(x, dpb, qmv, qy)
        for k in ("bpp_mv_y", "bpp_mv_z", "bpp_y", "bpp_z", "bpp", "mse", "me_mse"):
            assert np.array_equal(est[k].numpy(), G[p + "fwd_" + k]), k
        r, blob = dmc.encode_decode_bytes(x, dpb, ph, pw, qmv, qy)
        assert blob == G[p + "bin"].tobytes()
        assert r["bit"] == int(G[p + "bit"][0])
        d = r["dpb"]
        assert np.array_equal(d["ref_frame"].numpy(), G[p + "ref_frame"])
        assert np.array_equal(d["ref_y"].numpy(), G[p + "ref_y"])
        assert np.array_equal(d["ref_mv_y"].numpy(), G[p + "ref_mv_y"])
        feat = d["ref_feature"]
        if p + "ref_feature" in G.files:
            assert np.array_equal(feat.numpy(), G[p + "ref_feature"])
        else:
            assert np.array_equal(feat[:, ::4, ::8, ::8].numpy(), G[p + "ref_feature_sub"])
        dpb = d


def _train_case():
    """Inputs / noise of tests/golden/train_64x128.npz (oracle/gen_train_golden.py), regenerated."""
    from oracle import gen_train_golden as T
    xs, ref = T.train_inputs()
    noise = [T.train_noise(s, T.noise_shapes()) for s in range(2)]
    return T, xs, ref, noise


def test_oracle_train_mode_matches_reference_autograd(golden_dir):
    """The oracle's TRAIN-mode restatement (straight-through quantiser, noisy likelihoods, LowerBound's
    custom backward) against the unmodified reference's own autograd: outputs, loss, the gradient norm of
    every parameter and a few gradients in full, two chained P-frames with batch = 4 anchors."""
    import torch
    from oracle import dcvc_oracle as O
    from oracle import fixture
    G = np.load(os.path.join(golden_dir, "train_64x128.npz"))
    T, xs, ref, noise = _train_case()
    torch.set_num_threads(4)
    names = [str(n) for n in G["param_names"]]
    P = T.train_state_dict()
    for k in names:
        P[k] = P[k].clone().requires_grad_(True)
    odmc = O.OracleDMC(P)
    dpb = {"ref_frame": ref, "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    for step, x in enumerate(xs):
        p = f"s{step}_"
        for k in names:
            P[k].grad = None
        out = odmc.forward_one_frame_train(x, dpb, P["mv_y_q_scale"], P["y_q_scale"], noise[step])
        loss = T.loss_of(out)
        loss.backward()
        np.testing.assert_allclose(loss.item(), G[p + "loss"][0], rtol=1e-5)
        for k in ("bpp_mv_y", "bpp_mv_z", "bpp_y", "bpp_z", "bpp", "mse", "me_mse"):
            np.testing.assert_allclose(out[k].detach().numpy(), G[p + k], rtol=1e-5, err_msg=k)
        norms = np.array([0.0 if P[k].grad is None else P[k].grad.double().norm().item() for k in names])
        none = np.array([P[k].grad is None for k in names])
        assert np.array_equal(none, G[p + "grad_none"]), [n for n, a, b in zip(names, none, G[p + "grad_none"]) if a != b]
        np.testing.assert_allclose(norms, G[p + "grad_norms"], rtol=2e-3, atol=1e-7)
        for key in G.files:
            if key.startswith(p + "grad:"):
                g = P[key.split(":", 1)[1]].grad.numpy()
                np.testing.assert_allclose(g, G[key], rtol=2e-3, atol=1e-6 * np.abs(G[key]).max(), err_msg=key)
        dpb = {k: v.detach() for k, v in out["dpb"].items()}
