"""`-m gpu` end-to-end parity: the drop-in DMC / IntraNoAR on a B200 against the golden vectors
generated from the real reference (tests/golden, oracle/gen_golden.py) and against the oracle.

Three numerics modes of the same graph: "fp32" (exact mode, fp32 FMA convs on CUDA cores), "parity"
(the tcgen05 tensor-core engine on split-bf16 operands: three bf16 planes per activation, three parts per
weight, six MMAs per product, fp32 accumulate) and "bf16" (plain bf16 tensor-core mode, what bench.py's
headline measures).  north_star's bounds: symbols / bitstreams bit-exact, bpp within 0.1 %, PSNR within
0.01 dB.

fp32 and parity modes
  * tiny_64x128: free-running I + 2 P chain -- symbols, rANS streams and `.bin` containers BIT-EXACT.
  * cfg1_240x416 (BASELINE config 1 size): every frame is teacher-forced with the oracle's dpb, so a
    last-ulp difference between GPU FMA order and oneDNN cannot cascade.  A quantiser input that
    lands within float noise of a rounding boundary may flip by one step: the test bounds the flip
    rate (<= 2e-4 of the symbols, each by +-1), and requires bits within 0.1 % and PSNR within
    0.01 dB (north_star).  Counts go to gpurun_out/e2e_<mode>_*.json.
  * cfg1_240x416 free-running I + 2 P chain against the reference goldens: symbol flips and stream
    equality are reported as first-class numbers, bits within 0.1 % and PSNR within 0.01 dB asserted.
  * one teacher-forced 1920x1088 P-frame against the oracle (parity mode; slow: the CPU oracle).
bf16 (tcgen05) mode: encoder/decoder agree with each other exactly; symbol-mismatch rates against the
fp32 goldens are bounded by stated ceilings; the north_star bpp / PSNR bounds are checked in a separate
test that is EXPECTED to fail for plain bf16 operands (xfail, non-strict) so the gap stays visible.
"""
import json
import math
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SYMS = (("mv_z", "sym_mv_z"), ("z", "sym_z"), ("mv_y_q_w_0", "sym_mv_y_q_w_0"), ("mv_y_q_w_1", "sym_mv_y_q_w_1"),
        ("y_q_w_0", "sym_y_q_w_0"), ("y_q_w_1", "sym_y_q_w_1"))
IDXS = (("mv_idx_0", "idx_mv_0"), ("mv_idx_1", "idx_mv_1"), ("y_idx_0", "idx_y_0"), ("y_idx_1", "idx_y_1"))


def _pad64(x):
    h, w = x.shape[-2:]
    return torch.nn.functional.pad(x, (0, (64 - w % 64) % 64, 0, (64 - h % 64) % 64))


def _psnr(a, b):
    mse = torch.mean((a.float() - b.float()) ** 2).item()
    return 10 * math.log10(1.0 / max(mse, 1e-12))


def _state_dicts():
    from oracle import dcvc_oracle as O
    from oracle import fixture
    return (fixture.fixture_state_dict(O.state_shapes("dmc")),
            fixture.fixture_state_dict(O.state_shapes("intra")))


def _models(precision):
    from vcm_ts_b200.models import DMC, IntraNoAR
    torch.backends.cudnn.allow_tf32 = False
    sd_p, sd_i = _state_dicts()
    dmc, intra = DMC(precision=precision).eval(), IntraNoAR(precision=precision).eval()
    dmc.load_state_dict(sd_p, strict=True)
    intra.load_state_dict(sd_i, strict=True)
    dmc, intra = dmc.to("cuda:0"), intra.to("cuda:0")
    dmc.update(force=True)
    intra.update(force=True)
    return dmc, intra


@pytest.fixture(scope="module")
def fp32_models():
    return _models("fp32")


@pytest.fixture(scope="module")
def bf16_models():
    return _models("bf16")


@pytest.fixture(scope="module")
def parity_models():
    return _models("parity")


@pytest.fixture(scope="module")
def exact_models(request, fp32_models, parity_models):
    return {"fp32": fp32_models, "parity": parity_models}


EXACT_MODES = ["fp32", "parity"]


def _report(name, obj):
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", name), "w") as f:
        json.dump(obj, f, indent=1)


@pytest.mark.parametrize("mode", EXACT_MODES)
def test_exact_modes_bit_exact_chain_tiny(exact_models, mode, golden_dir, tmp_path):
    from oracle import dcvc_oracle as O
    dmc, intra = exact_models[mode]
    G = np.load(os.path.join(golden_dir, "tiny_64x128.npz"))
    ph, pw, qi, qmv, qy = G["meta"]
    ph, pw = int(ph), int(pw)
    fr = torch.from_numpy(G["frames_u8"])
    x0 = _pad64(fr[0:1].float() / 255.0).cuda()
    r, blob = intra.encode_decode_bytes(x0, qi, ph, pw)
    assert blob == G["i_bin"].tobytes()
    assert (r["x_hat"].cpu() - torch.from_numpy(G["i_x_hat"])).abs().max() < 1e-4
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    rep = {}
    for t in (1, 2):
        p = f"p{t}_"
        x = _pad64(fr[t:t + 1].float() / 255.0).cuda()
        mvq, _ = O.get_rounded_q(qmv)
        yq, _ = O.get_rounded_q(qy)
        enc = dmc.compress(x, dpb, mvq, yq)
        S = dmc._last_symbols
        for mine, gold in SYMS:
            assert np.array_equal(S[mine].numpy(), G[p + gold]), gold
        for mine, gold in IDXS:
            rep[p + gold + "_mismatch"] = int((S[mine].numpy() != G[p + gold]).sum())
        assert enc["bit_stream"] == G[p + "stream"].tobytes()
        est = dmc.forward_one_frame(x, dpb, qmv, qy)
        for k in ("bpp_mv_y", "bpp_mv_z", "bpp_y", "bpp_z", "bpp", "mse", "me_mse"):
            np.testing.assert_allclose(est[k].cpu().numpy(), G[p + "fwd_" + k], rtol=1e-4, err_msg=k)
        path = str(tmp_path / f"p{t}.bin")
        r = dmc.encode_decode(x, dpb, output_path=path, pic_width=pw, pic_height=ph,
                              mv_y_q_scale=qmv, y_q_scale=qy)
        assert open(path, "rb").read() == G[p + "bin"].tobytes()
        assert r["bit"] == int(G[p + "bit"][0])
        d = r["dpb"]
        assert (d["ref_frame"].cpu() - torch.from_numpy(G[p + "ref_frame"])).abs().max() < 1e-4
        # decoder reproduces the encoder's reconstruction exactly
        for k in ("ref_y", "ref_mv_y", "ref_feature"):
            assert torch.equal(d[k], enc["dbp"][k]), k
        dpb = d
    _report(f"e2e_{mode}_tiny.json", rep)
    # scale indexes: identical streams prove the coded CDF rows agree; an index may still differ by
    # one where log(sigma) sits on a bin edge and both rows quantise to the same CDF
    assert sum(rep.values()) <= 2, rep


def _flip_stats(mine, gold):
    d = mine.astype(np.int32) - gold.astype(np.int32)
    return int((d != 0).sum()), int(np.abs(d).max()) if d.size else 0


@pytest.mark.parametrize("mode", EXACT_MODES)
def test_exact_modes_teacher_forced_cfg1(exact_models, mode, golden_dir):
    """416x240 (padded 448x256): GPU vs the oracle, each frame fed the oracle's own dpb."""
    from oracle import dcvc_oracle as O
    dmc, intra = exact_models[mode]
    sd_p, sd_i = _state_dicts()
    torch.set_num_threads(min(8, os.cpu_count() or 1))
    odmc, ointra = O.OracleDMC(sd_p), O.OracleIntra(sd_i)
    odmc.update()
    ointra.update()
    G = np.load(os.path.join(golden_dir, "cfg1_240x416.npz"))
    ph, pw, qi, qmv, qy = G["meta"]
    ph, pw = int(ph), int(pw)
    fr = torch.from_numpy(G["frames_u8"])
    x0 = _pad64(fr[0:1].float() / 255.0)
    rep = {}
    # ---- I frame
    oi = ointra.compress(x0, O.get_rounded_q(qi)[0], want_symbols=True)
    intra.compress(x0.cuda(), O.get_rounded_q(qi)[0])
    tot = flips = 0
    for k in ("z", "y_q_w_0", "y_q_w_1"):
        n, mx = _flip_stats(intra._last_symbols[k].numpy(), oi["symbols"][k].numpy())
        rep["i_" + k + "_flips"] = n
        assert mx <= 1, (k, mx)
        flips += n
        tot += oi["symbols"][k].numel()
    rep["i_flip_rate"] = flips / tot
    og, oblob = ointra.encode_decode_bytes(x0, qi, ph, pw)
    gg, gblob = intra.encode_decode_bytes(x0.cuda(), qi, ph, pw)
    rep["i_bits_rel"] = abs(len(gblob) - len(oblob)) / len(oblob)
    rep["i_psnr_delta_db"] = abs(_psnr(x0, gg["x_hat"].cpu()) - _psnr(x0, og["x_hat"]))
    assert oblob == G["i_bin"].tobytes()          # the oracle itself is pinned to the reference
    odpb = {"ref_frame": og["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    for t in (1, 2):
        p = f"p{t}_"
        x = _pad64(fr[t:t + 1].float() / 255.0)
        mvq, _ = O.get_rounded_q(qmv)
        yq, _ = O.get_rounded_q(qy)
        oenc = odmc.compress(x, odpb, mvq, yq, want_symbols=True)
        gdpb = {k: (v.cuda() if v is not None else None) for k, v in odpb.items()}
        genc = dmc.compress(x.cuda(), gdpb, mvq, yq)
        tot = flips = 0
        for k, _ in SYMS:
            n, mx = _flip_stats(dmc._last_symbols[k].numpy(), oenc["symbols"][k].numpy())
            rep[p + k + "_flips"] = n
            assert mx <= 1, (k, mx)
            flips += n
            tot += oenc["symbols"][k].numel()
        rep[p + "flip_rate"] = flips / tot
        rep[p + "bits_rel"] = abs(len(genc["bit_stream"]) - len(oenc["bit_stream"])) / len(oenc["bit_stream"])
        ores, _ = odmc.encode_decode_bytes(x, odpb, ph, pw, qmv, qy)
        gres, _ = dmc.encode_decode_bytes(x.cuda(), gdpb, ph, pw, qmv, qy)
        rep[p + "psnr_delta_db"] = abs(_psnr(x, gres["dpb"]["ref_frame"].cpu()) - _psnr(x, ores["dpb"]["ref_frame"]))
        rep[p + "stream_equal"] = genc["bit_stream"] == oenc["bit_stream"]
        odpb = ores["dpb"]
    _report(f"e2e_{mode}_cfg1.json", rep)
    print(mode, rep)
    for pre in ("i_", "p1_", "p2_"):
        assert rep[pre + "flip_rate"] <= 2e-4, rep
        assert rep[pre + "bits_rel"] <= 1e-3, rep
        assert rep[pre + "psnr_delta_db"] <= 0.01, rep


_FREE_RUNNING = {}


@pytest.mark.parametrize("mode", EXACT_MODES)
def test_exact_modes_free_running_cfg1(exact_models, mode, golden_dir):
    """416x240 free-running I + 2 P chain (every frame predicted from the GPU's own previous
    reconstruction) against the goldens written by the unmodified reference.  Symbol flips, scale-index
    mismatches and stream equality are reported (gpurun_out/e2e_<mode>_cfg1_free.json); bits within 0.1 %
    and PSNR within 0.01 dB of the reference are asserted (north_star), and the tensor-core parity mode must
    take exactly the decisions of the fp32-FMA mode over the whole chain."""
    from oracle import dcvc_oracle as O
    dmc, intra = exact_models[mode]
    G = np.load(os.path.join(golden_dir, "cfg1_240x416.npz"))
    ph, pw, qi, qmv, qy = G["meta"]
    ph, pw = int(ph), int(pw)
    fr = torch.from_numpy(G["frames_u8"])
    x0 = _pad64(fr[0:1].float() / 255.0).cuda()
    rep = {"mode": mode}
    r, blob = intra.encode_decode_bytes(x0, qi, ph, pw)
    flips = total = 0          # the goldens hold no I-frame symbol planes: its .bin is compared instead
    rep["i_stream_equal"] = blob == G["i_bin"].tobytes()
    rep["i_bits_rel"] = abs(r["bit"] - int(G["i_bit"][0])) / int(G["i_bit"][0])
    rep["i_psnr_delta_db"] = abs(_psnr(x0.cpu(), r["x_hat"].cpu()) - _psnr(x0.cpu(), torch.from_numpy(G["i_x_hat"])))
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    for t in (1, 2):
        p = f"p{t}_"
        x = _pad64(fr[t:t + 1].float() / 255.0).cuda()
        mvq, _ = O.get_rounded_q(qmv)
        yq, _ = O.get_rounded_q(qy)
        enc = dmc.compress(x, dpb, mvq, yq)
        S = dmc._last_symbols
        for mine, gold in SYMS:
            n, mx = _flip_stats(S[mine].numpy(), G[p + gold])
            rep[p + gold + "_flips"] = n
            flips += n
            total += G[p + gold].size
        for mine, gold in IDXS:
            rep[p + gold + "_mismatch"] = int((S[mine].numpy() != G[p + gold]).sum())
        rep[p + "stream_equal"] = enc["bit_stream"] == G[p + "stream"].tobytes()
        res, blob = dmc.encode_decode_bytes(x, dpb, ph, pw, qmv, qy)
        rep[p + "bin_equal"] = blob == G[p + "bin"].tobytes()
        rep[p + "bits_rel"] = abs(res["bit"] - int(G[p + "bit"][0])) / int(G[p + "bit"][0])
        rep[p + "psnr_delta_db"] = abs(_psnr(x.cpu(), res["dpb"]["ref_frame"].cpu()) -
                                       _psnr(x.cpu(), torch.from_numpy(G[p + "ref_frame"])))
        dpb = res["dpb"]
    rep["symbol_flips_total"] = flips
    rep["symbols_total"] = int(total)
    _report(f"e2e_{mode}_cfg1_free.json", rep)
    print(mode, rep)
    for pre in ("i_", "p1_", "p2_"):
        assert rep[pre + "bits_rel"] <= 1e-3, rep
        assert rep[pre + "psnr_delta_db"] <= 0.01, rep
    # Flips are REPORTED, not bounded, here: on this fixture ONE I-frame symbol (of 91 392) sits within float
    # noise of a rounding boundary and lands on the other side than the reference's oneDNN summation order put
    # it (the same single flip in the fp32-FMA mode and in the tensor-core parity mode); from then on the chain
    # predicts from a legitimately different reconstruction, so later symbols differ without any further
    # arithmetic disagreement (teacher-forced: 0 P-frame flips, streams byte-equal).  What must hold is that
    # both exact modes make the SAME decisions:
    # (symbol decisions; a scale index may differ by one where log(sigma) sits on a bin edge)
    _FREE_RUNNING[mode] = {k: v for k, v in rep.items() if k.endswith("_flips")}
    if len(_FREE_RUNNING) == 2:
        assert _FREE_RUNNING["fp32"] == _FREE_RUNNING["parity"], _FREE_RUNNING


@pytest.mark.slow
def test_parity_mode_teacher_forced_1080p_p_frame(parity_models):
    """BASELINE shape: one 1920x1088 P-frame (reference frame = a synthetic neighbour frame), GPU
    tensor-core parity mode against the CPU oracle on the same inputs: symbol flips reported and bounded
    (<= 2e-4, each by +-1), bits within 0.1 %.  ~1-2 minutes of CPU oracle."""
    from oracle import dcvc_oracle as O
    import bench as Bn
    dmc, _ = parity_models
    sd_p, _ = _state_dicts()
    torch.set_num_threads(os.cpu_count() or 1)
    odmc = O.OracleDMC(sd_p)
    odmc.update()
    seq = Bn.synth_sequence(2, 5)
    pad = lambda x: torch.nn.functional.pad(x, (0, Bn.WP - Bn.W, 0, Bn.HP - Bn.H))  # noqa: E731
    ref, x = pad(seq[0:1]), pad(seq[1:2])
    odpb = {"ref_frame": ref, "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    oenc = odmc.compress(x, odpb, 1.0, 1.0, want_symbols=True)
    gdpb = {"ref_frame": ref.cuda(), "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    genc = dmc.compress(x.cuda(), gdpb, 1.0, 1.0)
    rep, tot, flips = {}, 0, 0
    for k, _ in SYMS:
        n, mx = _flip_stats(dmc._last_symbols[k].numpy(), oenc["symbols"][k].numpy())
        rep[k + "_flips"] = n
        assert mx <= 1, (k, mx)
        flips += n
        tot += oenc["symbols"][k].numel()
    rep["flip_rate"] = flips / tot
    rep["symbols"] = tot
    rep["stream_equal"] = genc["bit_stream"] == oenc["bit_stream"]
    rep["bits_rel"] = abs(len(genc["bit_stream"]) - len(oenc["bit_stream"])) / len(oenc["bit_stream"])
    rep["recon_max_abs_diff"] = (genc["dbp"]["ref_frame"].cpu() - oenc["dbp"]["ref_frame"]).abs().max().item()
    _report("e2e_parity_1080p.json", rep)
    print(rep)
    assert tot == 1370880
    assert rep["flip_rate"] <= 2e-4 and rep["bits_rel"] <= 1e-3, rep


_BF16_REPORT = {}


def test_bf16_mode_self_consistent_and_close(bf16_models, golden_dir):
    """Plain-bf16 tcgen05 mode: streams decode to exactly what the encoder reconstructed (the mode is
    self-decodable ONLY: its streams are not the reference's).  Against the fp32 reference goldens the
    symbol-mismatch rate is bounded by stated ceilings (15 % of the latent symbols, 60 % of the scale
    indexes) and rate / quality by 1 % bpp / 0.1 dB; the north_star bounds (0.1 % / 0.01 dB, bit-exact
    symbols) are NOT met by this mode -- see test_bf16_mode_north_star_bounds and the parity mode."""
    from oracle import dcvc_oracle as O
    dmc, intra = bf16_models
    G = np.load(os.path.join(golden_dir, "cfg1_240x416.npz"))
    ph, pw, qi, qmv, qy = G["meta"]
    ph, pw = int(ph), int(pw)
    fr = torch.from_numpy(G["frames_u8"])
    x0 = _pad64(fr[0:1].float() / 255.0).cuda()
    rep = {}
    r, blob = intra.encode_decode_bytes(x0, qi, ph, pw)
    rep["i_bits_rel"] = abs(r["bit"] - int(G["i_bit"][0])) / int(G["i_bit"][0])
    rep["i_psnr_delta_db"] = abs(_psnr(x0.cpu(), r["x_hat"].cpu()) - _psnr(x0.cpu(), torch.from_numpy(G["i_x_hat"])))
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    for t in (1, 2):
        p = f"p{t}_"
        x = _pad64(fr[t:t + 1].float() / 255.0).cuda()
        mvq, _ = O.get_rounded_q(qmv)
        yq, _ = O.get_rounded_q(qy)
        enc = dmc.compress(x, dpb, mvq, yq)
        S = dmc._last_symbols
        for mine, gold in SYMS + IDXS:
            rep[p + gold + "_mismatch"] = int((S[mine].numpy() != G[p + gold]).sum())
            rep[p + gold + "_total"] = int(G[p + gold].size)
        dec = dmc.decompress(dpb, enc["bit_stream"], ph, pw, mvq, yq)
        for k in ("ref_y", "ref_mv_y", "ref_feature"):
            assert torch.equal(dec["dpb"][k], enc["dbp"][k]), k
        assert torch.equal(dec["dpb"]["ref_frame"], enc["dbp"]["ref_frame"].clamp(0, 1))
        res, blob = dmc.encode_decode_bytes(x, dpb, ph, pw, qmv, qy)
        rep[p + "bits_rel"] = abs(res["bit"] - int(G[p + "bit"][0])) / int(G[p + "bit"][0])
        rep[p + "psnr_delta_db"] = abs(_psnr(x.cpu(), res["dpb"]["ref_frame"].cpu()) -
                                       _psnr(x.cpu(), torch.from_numpy(G[p + "ref_frame"])))
        dpb = res["dpb"]
    for t in (1, 2):
        sm = sum(rep[f"p{t}_{g}_mismatch"] for _, g in SYMS)
        st = sum(rep[f"p{t}_{g}_total"] for _, g in SYMS)
        im = sum(rep[f"p{t}_{g}_mismatch"] for _, g in IDXS)
        it = sum(rep[f"p{t}_{g}_total"] for _, g in IDXS)
        rep[f"p{t}_symbol_mismatch_rate"] = sm / st
        rep[f"p{t}_index_mismatch_rate"] = im / it
    _report("e2e_bf16.json", rep)
    _BF16_REPORT.update(rep)
    print(rep)
    assert rep["i_bits_rel"] < 0.01 and rep["i_psnr_delta_db"] < 0.1
    for t in (1, 2):
        assert rep[f"p{t}_bits_rel"] < 0.01 and rep[f"p{t}_psnr_delta_db"] < 0.1
        assert rep[f"p{t}_symbol_mismatch_rate"] < 0.15, rep
        assert rep[f"p{t}_index_mismatch_rate"] < 0.60, rep


@pytest.mark.xfail(reason="plain bf16 operands cannot meet north_star's 0.1 % bpp / bit-exact-symbol bounds "
                          "(measured: P1 bpp off by ~0.17 %, 1-9 % of symbols differ); precision='parity' is the "
                          "tensor-core mode that does", strict=False)
def test_bf16_mode_north_star_bounds(bf16_models, golden_dir):
    if not _BF16_REPORT:
        test_bf16_mode_self_consistent_and_close(bf16_models, golden_dir)
    rep = _BF16_REPORT
    for pre in ("i_", "p1_", "p2_"):
        assert rep[pre + "bits_rel"] <= 1e-3, rep
        assert rep[pre + "psnr_delta_db"] <= 0.01, rep
    for t in (1, 2):
        assert rep[f"p{t}_symbol_mismatch_rate"] == 0.0, rep


def test_batched_estimation_path(fp32_models):
    """forward_one_frame with batch 2 and per-sample q tensors (trainer-style call in eval mode)."""
    dmc, _ = fp32_models
    g = torch.Generator().manual_seed(4)
    x = torch.rand((2, 3, 64, 128), generator=g).cuda()
    ref = (x + 0.02 * torch.randn((2, 3, 64, 128), generator=g).cuda()).clamp(0, 1)
    dpb = {"ref_frame": ref, "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    q = torch.tensor([1.0, 0.7]).reshape(2, 1, 1, 1).cuda()
    out = dmc.forward_one_frame(x, dpb, mv_y_q_scale=q, y_q_scale=q)
    assert out["bpp"].shape == (2,) and out["dpb"]["ref_frame"].shape == x.shape
    one = dmc.forward_one_frame(x[1:], {k: (v[1:] if v is not None else None) for k, v in dpb.items()},
                                mv_y_q_scale=0.7, y_q_scale=0.7)
    torch.testing.assert_close(out["bpp"][1:], one["bpp"], rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(out["dpb"]["ref_frame"][1:], one["dpb"]["ref_frame"], rtol=1e-4, atol=1e-5)


def test_1080p_round_trip_properties(bf16_models):
    """BASELINE size (1920x1080 padded to 1088), tensor-core mode, size-independent properties:
    the stream decodes to exactly the encoder's reconstruction (latents, feature, frame), decoding
    the stream again is idempotent, bits == 8 * container bytes, the coded symbols survive the host
    coder (decode(encode(s)) == s for every plane), and a second identical call is bit-identical
    (deterministic kernels)."""
    from vcm_ts_b200.stream_helper import pack_p
    dmc, intra = bf16_models
    import bench as Bn
    seq = Bn.synth_sequence(2, 3).cuda()
    pad = lambda x: torch.nn.functional.pad(x, (0, Bn.WP - Bn.W, 0, Bn.HP - Bn.H))  # noqa: E731
    r, blob_i = intra.encode_decode_bytes(pad(seq[0:1]), 1.0, Bn.H, Bn.W)
    assert r["bit"] == 8 * len(blob_i)
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    x = pad(seq[1:2])
    enc = dmc.compress(x, dpb, 1.0, 1.0)
    sym = {k: v.clone() for k, v in dmc._last_symbols.items()}
    dec = dmc.decompress(dpb, enc["bit_stream"], Bn.H, Bn.W, 1.0, 1.0)
    for k in ("ref_y", "ref_mv_y", "ref_feature"):
        assert torch.equal(dec["dpb"][k], enc["dbp"][k]), k
    assert torch.equal(dec["dpb"]["ref_frame"], enc["dbp"]["ref_frame"].clamp(0, 1))
    dec2 = dmc.decompress(dpb, enc["bit_stream"], Bn.H, Bn.W, 1.0, 1.0)
    for k in dec["dpb"]:
        assert torch.equal(dec["dpb"][k], dec2["dpb"][k]), k
    enc2 = dmc.compress(x, dpb, 1.0, 1.0)
    assert enc2["bit_stream"] == enc["bit_stream"], "compress is not deterministic"
    res, blob = dmc.encode_decode_bytes(x, dpb, Bn.H, Bn.W, 1.0, 1.0)
    assert res["bit"] == 8 * len(blob) and blob == pack_p(enc["bit_stream"], 100, 100)
    # symbols through the host coder
    rt = dmc._rt
    rt.coder.set_stream(enc["bit_stream"])
    out = torch.empty(sym["mv_z"].shape, dtype=torch.int16)
    rt.coder.decode_channels(64, out.shape[2] * out.shape[3], rt.tables["bit_estimator_z_mv"], out)
    assert torch.equal(out, sym["mv_z"])
    for s, i in (("mv_y_q_w_0", "mv_idx_0"), ("mv_y_q_w_1", "mv_idx_1")):
        o = torch.empty(sym[s].shape, dtype=torch.int16)
        rt.coder.decode_planes(sym[i], rt.tables["y"], o)
        assert torch.equal(o, sym[s]), s
    n_sym = sum(sym[k].numel() for k in ("mv_z", "z", "mv_y_q_w_0", "mv_y_q_w_1", "y_q_w_0", "y_q_w_1"))
    assert n_sym == 1370880          # symbols per 1080p P-frame
    psnr = _psnr(x[..., :Bn.H, :Bn.W].cpu(), res["dpb"]["ref_frame"][..., :Bn.H, :Bn.W].cpu())
    assert math.isfinite(psnr) and psnr > 3.0, psnr   # random-init weights: only a sanity bound
    assert float(res["dpb"]["ref_frame"].min()) >= 0.0 and float(res["dpb"]["ref_frame"].max()) <= 1.0


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_gop_pipeline_equals_serial_coding(precision):
    """pipeline.GopPipeline (encoder / host coder / decoder overlapped across frames) must produce
    byte-identical containers and bit-identical decoder reconstructions to frame-by-frame
    encode_decode: the encoder predicting from its own reconstruction changes nothing."""
    from oracle import fixture
    from vcm_ts_b200.pipeline import GopPipeline
    dev = "cuda:0"
    enc, dec = _models(precision), _models(precision)
    fr = fixture.fixture_frames(128, 192, 5, seed=9).to(dev)
    # serial reference through the public per-frame calls
    dmc, intra = dec
    want_blobs, want_rec = [], []
    r, blob = intra.encode_decode_bytes(fr[0:1], 1.0, 128, 192)
    want_blobs.append(blob)
    want_rec.append(r["x_hat"].clone())
    dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
    for t in range(1, 5):
        r, blob = dmc.encode_decode_bytes(fr[t:t + 1], dpb, 128, 192, 1.0, 1.0)
        dpb = r["dpb"]
        want_blobs.append(blob)
        want_rec.append(dpb["ref_frame"].clone())
    got_rec = {}
    pipe = GopPipeline(enc, dec, dev)
    for _ in range(2):       # twice: the pipeline object is reusable
        got_rec.clear()
        blobs = pipe.code_gop([fr[t:t + 1] for t in range(5)], 128, 192, 1.0, 1.0, 1.0,
                              on_recon=lambda t, x: got_rec.__setitem__(t, x.clone()))
        torch.cuda.synchronize()
        assert blobs == want_blobs
        for t in range(5):
            assert torch.equal(got_rec[t], want_rec[t]), t


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_cuda_graph_mode_equals_eager_coding(precision):
    """set_cuda_graphs(True): the per-frame launch sequences are captured (graphs.py) and replayed.
    Containers and reconstructions must be byte-identical to eager coding -- serial per-frame calls
    and the overlapped GopPipeline -- across three GOPs (eager pass, capture pass, pure replay)."""
    from oracle import fixture
    from vcm_ts_b200.pipeline import GopPipeline
    dev = "cuda:0"
    fr = fixture.fixture_frames(128, 192, 5, seed=11).to(dev)

    def serial(dmc, intra):
        blobs, recs = [], []
        r, blob = intra.encode_decode_bytes(fr[0:1], 1.0, 128, 192)
        blobs.append(blob)
        recs.append(r["x_hat"].clone())
        dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
        for t in range(1, 5):
            r, blob = dmc.encode_decode_bytes(fr[t:t + 1], dpb, 128, 192, 1.0, 1.0)
            dpb = r["dpb"]
            blobs.append(blob)
            recs.append(dpb["ref_frame"].clone())
        return blobs, recs

    want_blobs, want_rec = serial(*_models(precision))
    dmc, intra = _models(precision)
    dmc.set_cuda_graphs(True)
    intra.set_cuda_graphs(True)       # the I-frame too: encoder graph + decoder chain (3 segments)
    s = torch.cuda.Stream(dev)
    with torch.cuda.stream(s):
        for gop in range(3):
            blobs, recs = serial(dmc, intra)
            s.synchronize()
            assert blobs == want_blobs, gop
            for t in range(5):
                assert torch.equal(recs[t], want_rec[t]), (gop, t)
    B = dmc._rt.B
    assert B.graph_replays > 0 and len(dmc._rt.programs.progs) == 4   # enc/dec x first/steady P-frame
    assert intra._rt.B.graph_replays > 0 and len(intra._rt.programs.progs) == 2
    # also from the legacy default stream (capture runs on a side stream)
    blobs, recs = serial(dmc, intra)
    assert blobs == want_blobs and all(torch.equal(a, b) for a, b in zip(recs, want_rec))

    enc, dec = _models(precision), _models(precision)
    enc[0].set_cuda_graphs(True)
    dec[0].set_cuda_graphs(True)
    dec[1].set_cuda_graphs(True)
    pipe = GopPipeline(enc, dec, dev)
    got = {}
    for gop in range(3):
        got.clear()
        blobs = pipe.code_gop([fr[t:t + 1] for t in range(5)], 128, 192, 1.0, 1.0, 1.0,
                              on_recon=lambda t, x: got.__setitem__(t, x.clone()))
        torch.cuda.synchronize()
        assert blobs == want_blobs, gop
        for t in range(5):
            assert torch.equal(got[t], want_rec[t]), (gop, t)
    assert enc[0]._rt.B.graph_replays > 0 and dec[0]._rt.B.graph_replays > 0


def test_cuda_graphs_follow_a_weight_reload():
    """Captured graphs bake in the addresses of the packed weights: load_state_dict must drop them,
    and coding with the new weights must again equal an eager model with those weights."""
    from oracle import fixture
    dev = "cuda:0"
    fr = fixture.fixture_frames(128, 192, 4, seed=13).to(dev)
    ref_frame = fr[0:1].clone()

    def chain(dmc, n=3):
        out = []
        dpb = {"ref_frame": ref_frame, "ref_feature": None, "ref_y": None, "ref_mv_y": None}
        for t in range(1, 4):
            r, blob = dmc.encode_decode_bytes(fr[t:t + 1], dpb, 128, 192, 1.0, 1.0)
            dpb = r["dpb"]
            out.append((blob, dpb["ref_frame"].clone()))
        return out

    dmc, _ = _models("bf16")
    dmc.set_cuda_graphs(True)
    for _ in range(3):
        first = chain(dmc)
    assert dmc._rt.B.graph_replays > 0
    sd = {k: (v * 0.9 if k.endswith(".weight") and v.dim() == 4 else v) for k, v in dmc.state_dict().items()}
    dmc.load_state_dict(sd)
    assert not dmc._rt.programs.progs
    eager, _ = _models("bf16")
    eager.load_state_dict(sd)
    want = chain(eager)
    assert want[0][0] != first[0][0]
    for _ in range(3):
        got = chain(dmc)
        for (b0, r0), (b1, r1) in zip(want, got):
            assert b0 == b1 and torch.equal(r0, r1)


@pytest.mark.timeout(300)
def test_gop_pipeline_propagates_a_coder_error():
    """A failure in the host-coder stage must surface as the exception, not as a hang: the entropy thread
    drains the bounded symbol queue so the encoder thread can finish (pipeline.py)."""
    from oracle import fixture
    from vcm_ts_b200.pipeline import GopPipeline
    dev = "cuda:0"
    enc, dec = _models("bf16"), _models("bf16")
    fr = fixture.fixture_frames(128, 192, 6, seed=17).to(dev)

    def boom(host):
        raise RuntimeError("coder failed")
    enc[0].encode_symbols = boom
    pipe = GopPipeline(enc, dec, dev)
    with pytest.raises(RuntimeError, match="coder failed"):
        pipe.code_gop([fr[t:t + 1] for t in range(6)], 128, 192, 1.0, 1.0, 1.0)
