"""CPU tests of the product's HOST logic (vcm_ts_b200/models.py, nets.py, entropy.py, arch.py,
stream_helper.py, csrc/rans_coder.cpp) with the device kernels replaced by tests/emul_backend.py.

The emulation issues the same ATen CPU ops as the oracle, so the whole wiring -- channel-slice
concats, fused epilogue flags, weight permutation of the contextual encoder, checkerboard passes,
symbol / index plane order, rANS stream order, container bytes -- must reproduce the golden vectors
generated from the real reference BIT-EXACTLY.  (The CUDA kernels themselves are checked in the
`-m gpu` tests.)
"""
import os

import numpy as np
import pytest
import torch

from oracle import dcvc_oracle as O
from oracle import fixture
from tests.emul_backend import EmulBackend
from vcm_ts_b200 import arch
from vcm_ts_b200.models import DMC, IntraNoAR

THREADS = 4


def _pad64(x):
    h, w = x.shape[-2:]
    return torch.nn.functional.pad(x, (0, (64 - w % 64) % 64, 0, (64 - h % 64) % 64))


def test_state_dict_layout_matches_reference_tree():
    for kind, spec, cls in (("dmc", arch.dmc_spec(), DMC), ("intra", arch.intra_spec(), IntraNoAR)):
        want = O.state_shapes(kind)
        assert arch.state_shapes(spec) == want
        got = {k: tuple(v.shape) for k, v in cls().state_dict().items()}
        assert got == want


@pytest.fixture(scope="module", params=["fp32", "twin"])
def models(request):
    """'twin' = the bf16-mode graph (activated twins instead of pre-activations) in fp32 math."""
    torch.set_num_threads(THREADS)
    dmc, intra = DMC(precision=request.param).eval(), IntraNoAR(precision=request.param).eval()
    dmc.load_state_dict(fixture.fixture_state_dict(O.state_shapes("dmc")), strict=True)
    intra.load_state_dict(fixture.fixture_state_dict(O.state_shapes("intra")), strict=True)
    for m in (dmc, intra):
        m._backend_factory = lambda dev, prec: EmulBackend(dev, prec)
        m.update(force=True)
    if request.param == "twin":
        assert dmc._rt.B.use_twin
    return dmc, intra


def test_cdf_tables_match_reference(models, golden_dir):
    dmc, intra = models
    G = np.load(os.path.join(golden_dir, "cdf_tables.npz"))
    for nm, t in (("dmc_y", dmc._rt.tables["y"]), ("dmc_z", dmc._rt.tables["bit_estimator_z"]),
                  ("dmc_z_mv", dmc._rt.tables["bit_estimator_z_mv"]), ("intra_y", intra._rt.tables["y"]),
                  ("intra_z", intra._rt.tables["bit_estimator_z"])):
        assert np.array_equal(G[nm + "_cdf"], t.cdf), nm
        assert np.array_equal(G[nm + "_len"], t.sizes), nm
        assert np.array_equal(G[nm + "_off"], t.offsets), nm


def test_gop_prefix_through_host_logic(models, golden_dir, tmp_path):
    dmc, intra = models
    G = np.load(os.path.join(golden_dir, "tiny_64x128.npz"))
    ph, pw, qi, qmv, qy = G["meta"]
    ph, pw = int(ph), int(pw)
    fr = torch.from_numpy(G["frames_u8"])
    x0 = _pad64(fr[0:1].float() / 255.0)
    r, blob = intra.encode_decode_bytes(x0, qi, ph, pw)
    assert blob == G["i_bin"].tobytes()
    np.testing.assert_allclose(r["x_hat"].numpy(), G["i_x_hat"], atol=2e-5)
    # file-based public call gives the same bytes
    path = str(tmp_path / "i.bin")
    r2 = intra.encode_decode(x0, qi, output_path=path, pic_width=pw, pic_height=ph)
    assert open(path, "rb").read() == blob and r2["bit"] == int(G["i_bit"][0])
    fwd = intra.forward(x0, qi)
    np.testing.assert_allclose(fwd["bpp_y"].numpy(), G["i_fwd_bpp_y"], rtol=1e-5)
    np.testing.assert_allclose(fwd["bpp_z"].numpy(), G["i_fwd_bpp_z"], rtol=1e-5)
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    for t in (1, 2):
        p = f"p{t}_"
        x = _pad64(fr[t:t + 1].float() / 255.0)
        mvq, _ = O.get_rounded_q(qmv)
        yq, _ = O.get_rounded_q(qy)
        enc = dmc.compress(x, dpb, mvq, yq)
        S = dmc._last_symbols
        for mine, gold in (("mv_z", "sym_mv_z"), ("z", "sym_z"), ("mv_y_q_w_0", "sym_mv_y_q_w_0"),
                           ("mv_y_q_w_1", "sym_mv_y_q_w_1"), ("y_q_w_0", "sym_y_q_w_0"),
                           ("y_q_w_1", "sym_y_q_w_1"), ("mv_idx_0", "idx_mv_0"), ("mv_idx_1", "idx_mv_1"),
                           ("y_idx_0", "idx_y_0"), ("y_idx_1", "idx_y_1")):
            assert np.array_equal(S[mine].numpy(), G[p + gold]), gold
        assert enc["bit_stream"] == G[p + "stream"].tobytes()
        assert set(enc) == {"dbp", "bit_stream"}
        est = dmc.forward_one_frame(x, dpb, qmv, qy)
        for k in ("bpp_mv_y", "bpp_mv_z", "bpp_y", "bpp_z", "bpp", "mse", "me_mse"):
            np.testing.assert_allclose(est[k].numpy(), G[p + "fwd_" + k], rtol=1e-5, err_msg=k)
        path = str(tmp_path / f"p{t}.bin")
        r = dmc.encode_decode(x, dpb, output_path=path, pic_width=pw, pic_height=ph,
                              mv_y_q_scale=qmv, y_q_scale=qy)
        assert open(path, "rb").read() == G[p + "bin"].tobytes()
        assert r["bit"] == int(G[p + "bit"][0])
        d = r["dpb"]
        # float tensors: the emulation feeds strided NHWC views to oneDNN, so accumulation order
        # differs from the reference's NCHW calls in the last bits; integers above are exact
        np.testing.assert_allclose(d["ref_frame"].numpy(), G[p + "ref_frame"], atol=5e-5)
        np.testing.assert_allclose(d["ref_y"].numpy(), G[p + "ref_y"], atol=1e-4)
        np.testing.assert_allclose(d["ref_mv_y"].numpy(), G[p + "ref_mv_y"], atol=1e-4)
        np.testing.assert_allclose(d["ref_feature"].numpy(), G[p + "ref_feature"], atol=1e-3, rtol=1e-4)
        dpb = d


def test_modules_pickle_before_and_after_update(models):
    import pickle
    dmc, _ = models
    clone = pickle.loads(pickle.dumps(dmc))
    assert clone._rt is None and clone.entropy_coder is None
    assert set(clone.state_dict()) == set(dmc.state_dict())


def test_no_cpu_path_in_product():
    m = DMC().eval()
    x = torch.rand(1, 3, 64, 64)
    with pytest.raises(RuntimeError):
        m.forward_one_frame(x, {"ref_frame": x, "ref_feature": None, "ref_y": None, "ref_mv_y": None}, 1.0, 1.0)


class _EagerPrograms:
    """graphs.ProgramCache stand-in for CPU: every call runs the frame body directly.  What is left
    of the CUDA-graph mode is its static, overwritten-in-place frame state (models.py `_statics`)."""

    def __init__(self, device):
        self.device, self.keys = device, []

    def invalidate(self):
        self.keys.clear()

    def run(self, key, make_gen, service):
        from vcm_ts_b200.graphs import drive_eager
        self.keys.append(key)
        return drive_eager(make_gen(), service, self.device)


def test_static_frame_state_of_graph_mode_matches_plain_mode(models, golden_dir):
    """set_cuda_graphs(True): dpb entries are static buffers read and then overwritten in place by
    every frame.  Same bitstreams and reconstructions as the plain mode, for the serial
    encode_decode chain, for the encoder-side chain of pipeline.py, and when a caller hands in
    foreign tensors."""
    dmc, intra = models
    G = np.load(os.path.join(golden_dir, "tiny_64x128.npz"))
    ph, pw, qi, qmv, qy = G["meta"]
    ph, pw = int(ph), int(pw)
    fr = torch.from_numpy(G["frames_u8"])
    xs = [_pad64(fr[t:t + 1].float() / 255.0) for t in range(fr.shape[0])]
    rec0 = intra.encode_decode_bytes(xs[0], qi, ph, pw)[0]["x_hat"]
    first = {"ref_frame": rec0, "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    mvq, _ = O.get_rounded_q(qmv)
    yq, _ = O.get_rounded_q(qy)

    from vcm_ts_b200.stream_helper import unpack_p

    def run(static):
        dmc.set_cuda_graphs(static)
        blobs, recs, enc_recs = [], [], []
        dpb, edpb = dict(first), dict(first)
        for x in xs[1:]:
            r, blob = dmc.encode_decode_bytes(x, dpb, ph, pw, qmv, qy)
            dpb = r["dpb"]
            blobs.append(blob)
            recs.append(dpb["ref_frame"].clone())
            host, edpb = dmc.compress_symbols(x, edpb, mvq, yq, want_dpb=True, clamp_ref=True)
            assert dmc.encode_symbols(host) == unpack_p(blob)[2]
            enc_recs.append(edpb["ref_frame"].clone())
        return blobs, recs, enc_recs

    try:
        want = run(False)
        dmc._rt.programs = _EagerPrograms(dmc._rt.B.device)
        got = run(True)
        keys = dmc._rt.programs.keys
        # second GOP with a foreign (cloned) ref_frame: copied into the static buffer
        got2 = run(True)
    finally:
        dmc.set_cuda_graphs(False)
    for a, b in ((want, got), (want, got2)):
        assert a[0] == b[0]
        for u, v in zip(a[1] + a[2], b[1] + b[2]):
            assert torch.equal(u, v)
    # the encoder's own reconstruction equals the decoder's (what pipeline.py relies on)
    for u, v in zip(want[1], want[2]):
        assert torch.equal(u, v)
    # first P-frame and steady-state frames select different programs; steady frames share one
    enc_keys = [k for k in keys if k[0] == "enc"]
    assert len(set(keys)) < len(keys) and len(set(enc_keys)) >= 3
