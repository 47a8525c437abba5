"""Enhancement-layer frame ops (video_coder.py:379-457, 521-530, 533-613).

CPU: the numpy oracle against the golden vectors made with the reference's own create_gradient_mask,
and the host-side box / ring preparation of vcm_ts_b200.enhancement.  GPU: the kernels, through the
C-ABI, bit-exact against the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import enhancement as OE
from vcm_ts_b200 import enhancement as E

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "enhancement.npz")


def test_oracle_matches_reference_golden():
    G = np.load(GOLD)
    for b in (0, 1, 2, 5, 9):
        for (w, h) in ((40, 23), (7, 5), (1, 1)):
            assert np.array_equal(OE.create_gradient_mask(w, h, b), G[f"mask_b{b}_{w}x{h}"]), (b, w, h)
    src, enc, enh, lp, faces = OE.golden_case()
    assert np.array_equal(OE.compute_residual(src, enc, lp, faces), G["residual"])
    for tag, (lpad, fpad) in (("p0", (0, 0)), ("p5_9", (5, 9)), ("p1_2", (1, 2))):
        assert np.array_equal(OE.fuse(src, enh, lp, faces, lpad, fpad), G["fuse_" + tag]), tag


def test_ring_model_of_the_kernel_equals_the_nested_frame_masks():
    """The kernel evaluates create_gradient_mask per pixel as ring[min(d, len - 1)], d = distance to the
    box border (csrc/enhancement.cu).  Check that closed form against the nested slice assignments."""
    for b in (0, 1, 2, 3, 5, 9, 30):
        ring = E.gradient_ring(b)
        assert ring.dtype == np.float32 and len(ring) == max(b, 1)
        for (w, h) in ((40, 23), (7, 5), (1, 1), (2, 9), (64, 64)):
            want = OE.create_gradient_mask(w, h, b)[..., 0]
            y, x = np.mgrid[0:h, 0:w]
            d = np.minimum(np.minimum(x, w - 1 - x), np.minimum(y, h - 1 - y))
            got = ring[np.minimum(d, len(ring) - 1)]
            assert np.array_equal(got, want), (b, w, h)


def test_box_normalisation_follows_numpy_slices():
    h, w = 20, 30
    boxes = [[-5, -3, 40, 25], [3, 4, 2, 1], [0, 0, 30, 20], [28, 18, -1, -1], [35, 2, 50, 9]]
    norm = E.normalise_boxes(boxes, h, w)
    for (x1, y1, x2, y2), (a, b_, c, d) in zip(boxes, norm):
        m = np.zeros((h, w), dtype=bool)
        m[y1:y2, x1:x2] = True
        n = np.zeros((h, w), dtype=bool)
        n[b_:d, a:c] = True
        assert 0 <= a <= c <= w and 0 <= b_ <= d <= h
        assert np.array_equal(m, n)


def test_no_cpu_path():
    t = torch.zeros((4, 4, 3), dtype=torch.uint8)
    with pytest.raises(RuntimeError):
        E.compute_residual(t, t)
    with pytest.raises(RuntimeError):
        E.fuse(t, t)


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to("cuda:0")


@pytest.mark.gpu
def test_kernels_match_golden_and_oracle():
    G = np.load(GOLD)
    src, enc, enh, lp, faces = OE.golden_case()
    got = E.compute_residual(_dev(src), _dev(enc), lp, faces).cpu().numpy()
    assert np.array_equal(got, G["residual"])
    for tag, (lpad, fpad) in (("p0", (0, 0)), ("p5_9", (5, 9)), ("p1_2", (1, 2))):
        got = E.fuse(_dev(src), _dev(enh), lp, faces, lpad, fpad).cpu().numpy()
        assert np.array_equal(got, G["fuse_" + tag]), tag
    # no boxes at all: residual layer empty, fuse returns the base layer
    assert not E.compute_residual(_dev(src), _dev(enc)).any()
    assert np.array_equal(E.fuse(_dev(src), _dev(enh)).cpu().numpy(), src)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1080, 1920, 3), (37, 53, 3), (64, 96, 1)])
def test_kernels_random_boxes_batched_and_ragged(shape):
    """1080p (vector path), odd sizes and C != 3 (byte path); batches with a different number of boxes
    per frame, including none, out-of-range and negative coordinates (numpy slice semantics)."""
    h, w, c = shape
    rng = np.random.default_rng(h)
    N = 3
    src = rng.integers(0, 256, (N, h, w, c), dtype=np.uint8)
    enc = rng.integers(0, 256, (N, h, w, c), dtype=np.uint8)

    def boxes(k, slack):
        out = []
        for _ in range(k):
            x1, x2 = sorted(rng.integers(-slack * w // 8, w + slack * w // 8 + 1, 2).tolist())
            y1, y2 = sorted(rng.integers(-slack * h // 8, h + slack * h // 8 + 1, 2).tolist())
            out.append([x1, y1, x2, y2])
        return out

    # residual: any coordinates (numpy slices clip / wrap them)
    lp = [boxes(4, 1), [], boxes(1, 1)]
    fc = [boxes(3, 1), [], boxes(9, 1)]
    got = E.compute_residual(_dev(src), _dev(enc), lp, fc).cpu().numpy()
    for n in range(N):
        assert np.array_equal(got[n], OE.compute_residual(src[n], enc[n], lp[n], fc[n])), n
    # fuse: boxes inside the frame (the reference's mask assignment raises otherwise, and so do we)
    lp = [boxes(4, 0), [], boxes(1, 0)]
    fc = [boxes(3, 0), [], boxes(9, 0)]
    plan = E.prepare_boxes(lp, fc, N, h, w, "cuda:0", 7, 3)
    gotf = E.fuse(_dev(src), _dev(enc), plan=plan).cpu().numpy()
    for n in range(N):
        assert np.array_equal(gotf[n], OE.fuse(src[n], enc[n], lp[n], fc[n], 7, 3)), n
    with pytest.raises(ValueError):
        E.fuse(_dev(src), _dev(enc), [[[-3, 0, 5, 5]], [], []], [], 7, 3)
    with pytest.raises(ValueError):
        OE.fuse(src[0], enc[0], [[-3, 0, 5, 5]], [], 7, 3)


@pytest.mark.gpu
def test_decoder_output_to_enhancement_layer_round_trip():
    """Residual of a frame against its 'decoded' version, fused back without borders, restores the
    source inside the boxes wherever the residual did not saturate (video_coder.py pipeline property)."""
    src, enc, _, lp, faces = OE.golden_case(seed=3, h=128, w=256)
    res = E.compute_residual(_dev(src), _dev(enc), lp, faces)
    back = E.fuse(_dev(enc), res, lp, faces, 0, 0).cpu().numpy()
    inside = np.zeros(src.shape[:2], dtype=bool)
    for x1, y1, x2, y2 in lp + faces:
        inside[y1:y2, x1:x2] = True
    ok = np.abs(src.astype(np.int32) - enc.astype(np.int32)) <= 127
    assert np.array_equal(back[inside & ok.all(-1)], src[inside & ok.all(-1)])
    assert np.array_equal(back[~inside], enc[~inside])
