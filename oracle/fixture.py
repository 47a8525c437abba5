"""oracle/fixture.py -- TEST INFRASTRUCTURE: deterministic weights and frames.

The reference ships no checkpoints (download-only, DCVC_HEM/checkpoints/download.py:9-14)
and its raw random init is numerically chaotic (hard part 1), so parity
fixtures use a construction-order-independent recipe: every state-dict entry is drawn from
a generator seeded by crc32(key), conv weights follow the reference's xavier-normal
(gain sqrt(2), common_model.py:31-36) scaled by 0.7, biases 0.01.  The same recipe is
applied to the reference modules (oracle/gen_golden.py), the oracle restatement and the
CUDA product, so all three run on identical parameters without committing 70 MB of weights.
"""
import math
import zlib

import torch
import torch.nn.functional as F

CONV_GAIN = 0.7 * math.sqrt(2.0)


def _gen(key, seed):
    g = torch.Generator(device="cpu")
    g.manual_seed((zlib.crc32(key.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
    return g


def fixture_tensor(key, shape, seed=0):
    """Deterministic fp32 tensor for state-dict entry `key` of `shape`."""
    shape = tuple(shape)
    g = _gen(key, seed)
    leaf = key.rsplit(".", 1)[-1]
    if leaf == "weight" and len(shape) == 4:
        rf = shape[2] * shape[3]
        fan_in, fan_out = shape[1] * rf, shape[0] * rf
        std = CONV_GAIN * math.sqrt(2.0 / (fan_in + fan_out))
        return torch.randn(shape, generator=g) * std
    if leaf == "bias":
        return torch.full(shape, 0.01)
    if leaf == "weight" and len(shape) == 2:          # SELayer Linear (no bias)
        bound = 1.0 / math.sqrt(shape[1])
        return (torch.rand(shape, generator=g) * 2 - 1) * bound
    if leaf in ("h", "b", "a"):                        # Bitparm (entropy_models.py:54-66)
        return torch.randn(shape, generator=g) * 0.01
    if leaf.endswith("q_basic"):                       # some entries < 0.5 exercise LowerBound
        return 0.4 + 1.2 * torch.rand(shape, generator=g)
    if leaf.endswith("q_scale"):
        n = shape[0]
        return torch.linspace(1.3, 0.6, n).reshape(shape)
    raise KeyError(f"no fixture rule for {key} {shape}")


def apply_fixture(module, seed=0):
    """Overwrite every parameter/buffer of an nn.Module with the recipe."""
    sd = module.state_dict()
    new = {k: fixture_tensor(k, v.shape, seed).to(v.dtype) for k, v in sd.items()}
    module.load_state_dict(new, strict=True)
    return module


def fixture_state_dict(shapes, seed=0):
    """shapes: {key: shape} -> {key: tensor}."""
    return {k: fixture_tensor(k, s, seed) for k, s in shapes.items()}


def fixture_frames(height, width, n_frames, seed=0):
    """Smooth moving texture in [0,1]: low-pass noise translated 1-3 px/frame + 1% noise.

    Returns float32 [n_frames, 3, height, width].
    """
    g = _gen("frames", seed)
    margin = 4 * n_frames + 8
    ch, cw = (height + margin + 15) // 16 + 2, (width + margin + 15) // 16 + 2
    coarse = torch.rand((1, 3, ch, cw), generator=g)
    mid = torch.rand((1, 3, ch * 4, cw * 4), generator=g)
    base = 0.75 * F.interpolate(coarse, scale_factor=16, mode="bicubic", align_corners=False) \
        + 0.25 * F.interpolate(mid, scale_factor=4, mode="bicubic", align_corners=False)
    base = base.clamp(0, 1)
    frames = []
    ox = oy = 0
    for t in range(n_frames):
        crop = base[0, :, oy:oy + height, ox:ox + width]
        noise = 0.01 * torch.randn(crop.shape, generator=g)
        frames.append((crop + noise).clamp(0, 1))
        ox += 1 + (t % 3)
        oy += 1 + ((t + 1) % 2)
    return torch.stack(frames).contiguous()
