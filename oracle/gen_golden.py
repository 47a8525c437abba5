"""oracle/gen_golden.py -- TEST INFRASTRUCTURE: golden vectors from the REAL reference.

Run in the build container only (needs /root/reference and `make -C oracle ref`):

    python oracle/gen_golden.py

Imports the unmodified reference `DMC` / `IntraNoAR` (oracle/ref_import.py), loads the
deterministic fixture weights (oracle/fixture.py), and runs a 3-frame GOP prefix
(I + 2 P) through the reference's own public calls:

    IntraNoAR.encode_decode(x, q, output_path=...)      image_model.py:114-146
    DMC.compress / DMC.encode_decode(..., output_path)  video_model.py:263-352, 424-455
    DMC.forward_one_frame / IntraNoAR.forward           video_model.py:470-592, image_model.py:54-105

and stores inputs (uint8 frames), per-frame .bin bytes, symbol / scale-index planes,
reconstructions and rate numbers under tests/golden/<case>.npz.  The reference has no
golden vectors of its own, so these files are what pins the oracle
restatement (tests/test_oracle_golden.py) and, through it, the CUDA path.
"""
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import fixture  # noqa: E402
from oracle.ref_import import import_reference  # noqa: E402
from oracle.dcvc_oracle import state_shapes  # noqa: E402

CASES = {
    # name: (pic_h, pic_w, q_intra, mv_y_q, y_q, keep_full_feature)
    "tiny_64x128": (64, 128, 1.17, 1.17, 0.83, True),
    "cfg1_240x416": (240, 416, 1.0, 1.0, 1.0, False),
}
N_FRAMES = 3
THREADS = 4


def pad64(x):
    h, w = x.shape[-2:]
    ph, pw = (64 - h % 64) % 64, (64 - w % 64) % 64
    return torch.nn.functional.pad(x, (0, pw, 0, ph), mode="constant", value=0)


def frames_u8(h, w, n, seed):
    f = fixture.fixture_frames(h, w, n, seed)
    return (f * 255).round().clamp(0, 255).to(torch.uint8)


def main():
    torch.set_num_threads(THREADS)
    torch.manual_seed(0)
    DMC, IntraNoAR, _ = import_reference()
    out_dir = os.path.join(os.path.dirname(HERE), "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)

    dmc = DMC().eval()
    intra = IntraNoAR().eval()
    # pin the restated parameter tree against the real one
    for kind, mod in (("dmc", dmc), ("intra", intra)):
        ref_shapes = {k: tuple(v.shape) for k, v in mod.state_dict().items()}
        mine = {k: tuple(v) for k, v in state_shapes(kind).items()}
        assert ref_shapes == mine, (kind, set(ref_shapes) ^ set(mine))
    fixture.apply_fixture(dmc, seed=0)
    fixture.apply_fixture(intra, seed=0)
    dmc.update(force=True)
    intra.update(force=True)

    tables = {}
    for nm, helper in (("dmc_y", dmc.gaussian_encoder.cdf_helper), ("dmc_z", dmc.bit_estimator_z.cdf_helper),
                       ("dmc_z_mv", dmc.bit_estimator_z_mv.cdf_helper),
                       ("intra_y", intra.gaussian_encoder.cdf_helper),
                       ("intra_z", intra.bit_estimator_z.cdf_helper)):
        cdf, length, offset = helper.get_cdf_info()
        tables[nm + "_cdf"] = cdf.astype(np.int32)
        tables[nm + "_len"] = length.astype(np.int32)
        tables[nm + "_off"] = offset.astype(np.int32)
    np.savez_compressed(os.path.join(out_dir, "cdf_tables.npz"), **tables)

    for name, (ph, pw, qi, qmv, qy, keep_full) in CASES.items():
        G = {}
        fr = frames_u8(ph, pw, N_FRAMES, seed=1)
        G["frames_u8"] = fr.numpy()
        G["meta"] = np.array([ph, pw, qi, qmv, qy], dtype=np.float64)
        with torch.no_grad(), tempfile.TemporaryDirectory() as td:
            x0 = pad64(fr[0:1].float() / 255.0)
            # ---- I frame ------------------------------------------------------------
            binp = os.path.join(td, "i.bin")
            r = intra.encode_decode(x0, qi, binp, pic_height=ph, pic_width=pw)
            G["i_bin"] = np.frombuffer(open(binp, "rb").read(), dtype=np.uint8)
            G["i_x_hat"] = r["x_hat"].numpy()
            G["i_bit"] = np.array([r["bit"]], dtype=np.int64)
            fwd = intra.forward(x0, qi)
            G["i_fwd_bpp_y"] = fwd["bpp_y"].numpy()
            G["i_fwd_bpp_z"] = fwd["bpp_z"].numpy()
            G["i_fwd_mse"] = fwd["mse"].numpy()
            G["i_fwd_x_hat_absmax_vs_write"] = np.array(
                [(fwd["x_hat"].clamp(0, 1) - r["x_hat"]).abs().max().item()])
            dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
            # ---- P frames -----------------------------------------------------------
            for t in range(1, N_FRAMES):
                x = pad64(fr[t:t + 1].float() / 255.0)
                p = f"p{t}_"
                # symbol planes exactly as handed to the entropy coder (video_model.py:333-340)
                mvq, _ = __import__("DCVC_HEM.src.utils.stream_helper", fromlist=["x"]).get_rounded_q(qmv)
                yq, _ = __import__("DCVC_HEM.src.utils.stream_helper", fromlist=["x"]).get_rounded_q(qy)
                sym = capture_symbols(dmc, x, dpb, mvq, yq)
                for k, v in sym.items():
                    G[p + k] = v
                est = dmc.forward_one_frame(x, dpb, mv_y_q_scale=qmv, y_q_scale=qy)
                for k in ("bpp_mv_y", "bpp_mv_z", "bpp_y", "bpp_z", "bpp", "mse", "me_mse"):
                    G[p + "fwd_" + k] = est[k].numpy()
                binp = os.path.join(td, f"p{t}.bin")
                r = dmc.encode_decode(x, dpb, binp, pic_width=pw, pic_height=ph,
                                      mv_y_q_scale=qmv, y_q_scale=qy)
                G[p + "bin"] = np.frombuffer(open(binp, "rb").read(), dtype=np.uint8)
                G[p + "bit"] = np.array([r["bit"]], dtype=np.int64)
                d = r["dpb"]
                G[p + "ref_frame"] = d["ref_frame"].numpy()
                G[p + "ref_y"] = d["ref_y"].numpy()
                G[p + "ref_mv_y"] = d["ref_mv_y"].numpy()
                feat = d["ref_feature"]
                G[p + "ref_feature_stats"] = np.array(
                    [feat.double().sum().item(), feat.double().abs().sum().item(),
                     feat.abs().max().item()])
                G[p + "ref_feature" + ("" if keep_full else "_sub")] = \
                    (feat if keep_full else feat[:, ::4, ::8, ::8]).numpy()
                # the estimation path must agree bit-for-bit with the write path
                G[p + "fwd_vs_write_absmax"] = np.array(
                    [(est["dpb"]["ref_frame"].clamp(0, 1) - d["ref_frame"]).abs().max().item(),
                     (est["dpb"]["ref_feature"] - feat).abs().max().item()])
                dpb = d
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **G)
        print(name, "->", path, f"{os.path.getsize(path) / 1e6:.2f} MB",
              "bits:", [int(G[k][0]) for k in ("i_bit", "p1_bit", "p2_bit")],
              "fwd_vs_write:", G["p1_fwd_vs_write_absmax"], G["p2_fwd_vs_write_absmax"])


def capture_symbols(dmc, x, dpb, mv_q, y_q):
    """Run the reference `compress` while recording what reaches the entropy coder."""
    rec = []
    ge, bz, bzm = dmc.gaussian_encoder, dmc.bit_estimator_z, dmc.bit_estimator_z_mv
    o_ge, o_bz, o_bzm = ge.encode, bz.encode, bzm.encode

    def w_ge(sym, scales):
        rec.append(("y", sym.clone(), ge.build_indexes(scales).clone()))
        return o_ge(sym, scales)

    def w_z(tag, orig):
        def f(z):
            rec.append((tag, z.clone(), None))
            return orig(z)
        return f

    ge.encode, bz.encode, bzm.encode = w_ge, w_z("z", o_bz), w_z("z_mv", o_bzm)
    try:
        enc = dmc.compress(x, dpb, mv_q, y_q)
    finally:
        ge.encode = o_ge
        del bz.encode, bzm.encode
    assert [r[0] for r in rec] == ["z_mv", "y", "y", "z", "y", "y"]
    out = {
        "sym_mv_z": rec[0][1].to(torch.int16).numpy(), "sym_z": rec[3][1].to(torch.int16).numpy(),
        "sym_mv_y_q_w_0": rec[1][1].to(torch.int16).numpy(), "idx_mv_0": rec[1][2].to(torch.uint8).numpy(),
        "sym_mv_y_q_w_1": rec[2][1].to(torch.int16).numpy(), "idx_mv_1": rec[2][2].to(torch.uint8).numpy(),
        "sym_y_q_w_0": rec[4][1].to(torch.int16).numpy(), "idx_y_0": rec[4][2].to(torch.uint8).numpy(),
        "sym_y_q_w_1": rec[5][1].to(torch.int16).numpy(), "idx_y_1": rec[5][2].to(torch.uint8).numpy(),
        "stream": np.frombuffer(enc["bit_stream"], dtype=np.uint8),
    }
    for k in ("sym_mv_y_q_w_0", "sym_y_q_w_0"):
        assert np.abs(out[k]).max() < 30000
    return out


if __name__ == "__main__":
    main()
