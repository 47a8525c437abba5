"""oracle/dcvc_oracle.py -- TEST INFRASTRUCTURE, not product code.

CPU restatement (functional PyTorch fp32, weights in a flat {state-dict key: tensor}
map) of the DCVC-HEM P-frame / I-frame path that vcm-ts wraps.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import it.

It is pinned BIT-EXACTLY against the real reference modules run in the build container
(oracle/gen_golden.py -> tests/golden/*.npz; tests/test_oracle_golden.py), because it issues
the same ATen CPU ops in the same order as:

  DMC                 DCVC_HEM/src/models/video_model.py:131-596
  IntraNoAR           DCVC_HEM/src/models/image_model.py:16-200
  CompressionModel    DCVC_HEM/src/models/common_model.py:14-217
  building blocks     DCVC_HEM/src/models/video_net.py:14-276, DCVC_HEM/src/layers/layers.py:18-127
  entropy models      DCVC_HEM/src/entropy_models/entropy_models.py:54-288
  stream container    DCVC_HEM/src/utils/stream_helper.py:24-46,103-144

The entropy coder underneath is the plain-C restatement oracle/rans_oracle.c.
"""
import ctypes
import math
import os
import struct

import numpy as np
import torch
import torch.nn.functional as F

_HERE = os.path.dirname(os.path.abspath(__file__))


# --------------------------------------------------------------------------------------
# entropy coder (oracle/rans_oracle.c through ctypes)
# --------------------------------------------------------------------------------------
class OracleRans:
    """Same call surface as entropy_models.py:9-51 `EntropyCoder`."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            path = os.path.join(_HERE, "liboracle_rans.so")
            if not os.path.exists(path):
                raise ImportError("oracle/liboracle_rans.so missing - run `make -C oracle oracle`")
            L = ctypes.CDLL(path)
            vp, i64, i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32
            L.oracle_enc_new.restype = vp
            L.oracle_dec_new.restype = vp
            L.oracle_enc_reset.argtypes = [vp]
            L.oracle_enc_free.argtypes = [vp]
            L.oracle_dec_free.argtypes = [vp]
            L.oracle_enc_push.argtypes = [vp, vp, vp, i64, vp, i32, vp, vp]
            L.oracle_enc_flush.argtypes = [vp, ctypes.POINTER(ctypes.c_void_p)]
            L.oracle_enc_flush.restype = i64
            L.oracle_dec_set_stream.argtypes = [vp, ctypes.c_char_p, i64]
            L.oracle_dec_decode.argtypes = [vp, vp, i64, vp, i32, vp, vp, vp]
            L.oracle_pmf_to_quantized_cdf.argtypes = [vp, i32, i32, vp]
            L.oracle_pmf_to_quantized_cdf.restype = i32
            cls._lib = L
        return cls._lib

    def __init__(self):
        L = self.lib()
        self.enc = L.oracle_enc_new()
        self.dec = L.oracle_dec_new()

    def __del__(self):
        try:
            self._lib.oracle_enc_free(self.enc)
            self._lib.oracle_dec_free(self.dec)
        except Exception:
            pass

    @staticmethod
    def _i32(a):
        return np.ascontiguousarray(a, dtype=np.int32)

    def reset_encoder(self):
        self._lib.oracle_enc_reset(self.enc)

    def encode_with_indexes(self, symbols, indexes, cdf, cdf_length, offset):
        s, i, c, l, o = map(self._i32, (symbols, indexes, cdf, cdf_length, offset))
        self._lib.oracle_enc_push(self.enc, s.ctypes.data, i.ctypes.data, s.size, c.ctypes.data,
                                  c.shape[1], l.ctypes.data, o.ctypes.data)

    def flush_encoder(self):
        p = ctypes.c_void_p()
        n = self._lib.oracle_enc_flush(self.enc, ctypes.byref(p))
        return ctypes.string_at(p, n)

    def set_stream(self, stream):
        self._stream = bytes(stream)
        self._lib.oracle_dec_set_stream(self.dec, self._stream, len(self._stream))

    def decode_stream(self, indexes, cdf, cdf_length, offset):
        i, c, l, o = map(self._i32, (indexes, cdf, cdf_length, offset))
        out = np.empty(i.size, dtype=np.int32)
        self._lib.oracle_dec_decode(self.dec, i.ctypes.data, i.size, c.ctypes.data, c.shape[1],
                                    l.ctypes.data, o.ctypes.data, out.ctypes.data)
        return out


def pmf_to_quantized_cdf(pmf, precision=16):
    """ops.cpp:24-82 via the C restatement. pmf: 1-D float tensor -> int32 tensor len+1."""
    p = np.ascontiguousarray(pmf.detach().cpu().numpy(), dtype=np.float32)
    out = np.empty(p.size + 1, dtype=np.uint32)
    rc = OracleRans.lib().oracle_pmf_to_quantized_cdf(p.ctypes.data, p.size, precision, out.ctypes.data)
    if rc != 0:
        raise ValueError("pmf_to_quantized_cdf: no frequency donor")
    return torch.from_numpy(out.astype(np.int32))


def pmf_to_cdf(pmf, tail_mass, pmf_length, max_length):
    """entropy_models.py:24-32."""
    cdf = torch.zeros((len(pmf_length), int(max_length) + 2), dtype=torch.int32)
    for i, p in enumerate(pmf):
        prob = torch.cat((p[: pmf_length[i]], tail_mass[i]), dim=0)
        c = pmf_to_quantized_cdf(prob, 16)
        cdf[i, : c.size(0)] = c
    return cdf


# --------------------------------------------------------------------------------------
# stream container (stream_helper.py)
# --------------------------------------------------------------------------------------
def get_padding_size(height, width, p=64):            # stream_helper.py:24-33
    new_h = (height + p - 1) // p * p
    new_w = (width + p - 1) // p * p
    return 0, new_w - width, 0, new_h - height


def get_downsampled_shape(height, width, p):          # stream_helper.py:36-39
    new_h = (height + p - 1) // p * p
    new_w = (width + p - 1) // p * p
    return int(new_h / p + 0.5), int(new_w / p + 0.5)


def get_rounded_q(q_scale):                           # stream_helper.py:42-46
    q_scale = np.clip(q_scale, 0.01, 655.)
    q_index = int(np.round(q_scale * 100))
    return q_index / 100, q_index


def pack_p(string, mv_y_q_index, y_q_index):          # stream_helper.py:126-131
    return struct.pack(">2H", mv_y_q_index, y_q_index) + struct.pack(">I", len(string)) + string


def unpack_p(blob):                                   # stream_helper.py:134-144
    a, b = struct.unpack(">2H", blob[:4])
    (n,) = struct.unpack(">I", blob[4:8])
    return a, b, blob[8:8 + n]


def pack_i(height, width, q_index, string):           # stream_helper.py:103-110
    return struct.pack(">2I", height, width) + struct.pack(">H", q_index) + \
        struct.pack(">I", len(string)) + string


def unpack_i(blob):                                   # stream_helper.py:113-123
    h, w = struct.unpack(">2I", blob[:8])
    (q,) = struct.unpack(">H", blob[8:10])
    (n,) = struct.unpack(">I", blob[10:14])
    return h, w, q, blob[14:14 + n]


# --------------------------------------------------------------------------------------
# layers
# --------------------------------------------------------------------------------------
def conv(P, name, x, stride=1):
    w = P[name + ".weight"]
    return F.conv2d(x, w, P[name + ".bias"], stride, w.shape[-1] // 2)


class LowerBoundFn(torch.autograd.Function):          # video_net.py:14-28 (forward AND its custom backward)
    @staticmethod
    def forward(ctx, inputs, bound):
        b = torch.ones_like(inputs) * bound
        ctx.save_for_backward(inputs, b)
        return torch.max(inputs, b)

    @staticmethod
    def backward(ctx, grad_output):
        inputs, b = ctx.saved_tensors
        pass_through = (inputs >= b) | (grad_output < 0)
        return pass_through.type(grad_output.dtype) * grad_output, None


def lower_bound(x, b):                                # LowerBound.apply (same forward value as torch.max)
    return LowerBoundFn.apply(x, b)


def quant_ste(x):                                     # common_model.py:38-44 (training branch)
    return x + (torch.round(x) - x).clone().detach()


def torch_warp(feature, flow):                        # video_net.py:32-50
    N, _, H, W = flow.size()
    hor = torch.linspace(-1.0, 1.0, W, dtype=feature.dtype, device=feature.device).view(1, 1, 1, W).expand(N, -1, H, -1)
    ver = torch.linspace(-1.0, 1.0, H, dtype=feature.dtype, device=feature.device).view(1, 1, H, 1).expand(N, -1, -1, W)
    grid0 = torch.cat([hor, ver], 1)
    flow = torch.cat([flow[:, 0:1] / ((feature.size(3) - 1.0) / 2.0),
                      flow[:, 1:2] / ((feature.size(2) - 1.0) / 2.0)], 1)
    grid = grid0 + flow
    return F.grid_sample(feature, grid.permute(0, 2, 3, 1), mode="bilinear",
                         padding_mode="border", align_corners=True)


def bilinear_up2(x):                                  # video_net.py:58-63
    return F.interpolate(x, (x.size(2) * 2, x.size(3) * 2), mode="bilinear", align_corners=False)


def bilinear_down2(x):                                # video_net.py:66-71
    return F.interpolate(x, (x.size(2) // 2, x.size(3) // 2), mode="bilinear", align_corners=False)


def res_block(P, pre, x, slope=0.01, start_from_relu=True, end_with_relu=False):
    """video_net.py:74-96 `ResBlock`."""
    act = (lambda t: F.relu(t)) if slope < 0.0001 else (lambda t: F.leaky_relu(t, slope))
    out = act(x) if start_from_relu else x
    out = conv(P, pre + ".conv1", out)
    out = act(out)
    out = conv(P, pre + ".conv2", out)
    if end_with_relu:
        out = act(out)
    return x + out


def residual_block_with_stride(P, pre, x):            # layers.py:42-73
    out = conv(P, pre + ".conv1", x, 2)
    out = F.leaky_relu(out, 0.01)
    out = conv(P, pre + ".conv2", out)
    out = F.leaky_relu(out, 0.1)
    identity = conv(P, pre + ".downsample", x, 2)
    out = out + identity
    return out


def residual_block_upsample(P, pre, x):               # layers.py:76-101
    out = F.pixel_shuffle(conv(P, pre + ".subpel_conv.0", x), 2)
    out = F.leaky_relu(out, 0.01)
    out = conv(P, pre + ".conv", out)
    out = F.leaky_relu(out, 0.1)
    identity = F.pixel_shuffle(conv(P, pre + ".upsample.0", x), 2)
    out = out + identity
    return out


def residual_block(P, pre, x, slope=0.01):            # layers.py:104-127
    out = conv(P, pre + ".conv1", x)
    out = F.leaky_relu(out, slope)
    out = conv(P, pre + ".conv2", out)
    out = F.leaky_relu(out, slope)
    return out + x


def analysis(P, pre, x):                              # video_net.py:226-235 (enc)
    x = residual_block_with_stride(P, pre + ".0", x)
    x = residual_block(P, pre + ".1", x)
    x = residual_block_with_stride(P, pre + ".2", x)
    x = residual_block(P, pre + ".3", x)
    x = residual_block_with_stride(P, pre + ".4", x)
    x = residual_block(P, pre + ".5", x)
    return conv(P, pre + ".6", x, 2)


def synthesis(P, pre, x):                             # video_net.py:237-246 (dec)
    x = residual_block(P, pre + ".0", x)
    x = residual_block_upsample(P, pre + ".1", x)
    x = residual_block(P, pre + ".2", x)
    x = residual_block_upsample(P, pre + ".3", x)
    x = residual_block(P, pre + ".4", x)
    x = residual_block_upsample(P, pre + ".5", x)
    x = residual_block(P, pre + ".6", x)
    return F.pixel_shuffle(conv(P, pre + ".7.0", x), 2)


def hyper_analysis(P, pre, x):                        # video_net.py:251-262
    x = F.leaky_relu(conv(P, pre + ".0", x), 0.01)
    x = F.leaky_relu(conv(P, pre + ".2", x), 0.01)
    x = F.leaky_relu(conv(P, pre + ".4", x, 2), 0.01)
    x = F.leaky_relu(conv(P, pre + ".6", x), 0.01)
    return conv(P, pre + ".8", x, 2)


def hyper_synthesis(P, pre, x):                       # video_net.py:264-274, video_model.py:181-191
    x = F.leaky_relu(conv(P, pre + ".0", x), 0.01)
    x = F.leaky_relu(F.pixel_shuffle(conv(P, pre + ".2.0", x), 2), 0.01)
    x = F.leaky_relu(conv(P, pre + ".4", x), 0.01)
    x = F.leaky_relu(F.pixel_shuffle(conv(P, pre + ".6.0", x), 2), 0.01)
    return conv(P, pre + ".8", x)


def three_convs(P, pre, x, slope=0.2):                # prior fusion / spatial prior stacks
    x = F.leaky_relu(conv(P, pre + ".0", x), slope)
    x = F.leaky_relu(conv(P, pre + ".2", x), slope)
    return conv(P, pre + ".4", x)


def se_layer(P, pre, x):                              # video_net.py:149-162
    y = torch.mean(x, dim=(-1, -2))
    y = F.linear(y, P[pre + ".fc.0.weight"])
    y = F.relu(y)
    y = F.linear(y, P[pre + ".fc.2.weight"])
    y = torch.sigmoid(y)
    return x * y[:, :, None, None]


def conv_block_residual(P, pre, x):                   # video_net.py:165-179
    x1 = conv(P, pre + ".conv.0", x)
    x1 = F.leaky_relu(x1, 0.01)
    x1 = conv(P, pre + ".conv.2", x1)
    x1 = se_layer(P, pre + ".conv.3", x1)
    x2 = conv(P, pre + ".up_dim", x)
    return x2 + x1


def unet(P, pre, x):                                  # video_net.py:182-223
    x1 = conv_block_residual(P, pre + ".conv1", x)
    x2 = F.max_pool2d(x1, 2, 2)
    x2 = conv_block_residual(P, pre + ".conv2", x2)
    x3 = F.max_pool2d(x2, 2, 2)
    x3 = conv_block_residual(P, pre + ".conv3", x3)
    for i in range(4):
        x3 = res_block(P, f"{pre}.context_refine.{i}", x3, slope=0)
    d3 = F.pixel_shuffle(conv(P, pre + ".up3.0", x3), 2)
    d3 = conv_block_residual(P, pre + ".up_conv3", torch.cat((x2, d3), dim=1))
    d2 = F.pixel_shuffle(conv(P, pre + ".up2.0", d3), 2)
    d2 = conv_block_residual(P, pre + ".up_conv2", torch.cat((x1, d2), dim=1))
    return d2


def spynet(P, pre, im1, im2):                         # video_net.py:118-146
    L = 4
    l1, l2 = [im1], [im2]
    for lv in range(L - 1):
        l1.append(F.avg_pool2d(l1[lv], kernel_size=2, stride=2))
        l2.append(F.avg_pool2d(l2[lv], kernel_size=2, stride=2))
    sf = l2[L - 1].size()
    flow = torch.zeros([im1.size(0), 2, sf[2] // 2, sf[3] // 2], dtype=im1.dtype)
    for lv in range(L):
        flow_up = bilinear_up2(flow) * 2.0
        k = L - 1 - lv
        t = torch.cat([l1[k], torch_warp(l2[k], flow_up), flow_up], 1)
        b = f"{pre}.moduleBasic.{lv}"
        t = F.relu(conv(P, b + ".conv1", t))
        t = F.relu(conv(P, b + ".conv2", t))
        t = F.relu(conv(P, b + ".conv3", t))
        t = F.relu(conv(P, b + ".conv4", t))
        t = conv(P, b + ".conv5", t)
        flow = flow_up + t
    return flow


# --------------------------------------------------------------------------------------
# entropy models
# --------------------------------------------------------------------------------------
def bitparm(P, pre, x, final):                        # entropy_models.py:54-73
    x = x * F.softplus(P[pre + ".h"]) + P[pre + ".b"]
    if final:
        return x
    return x + torch.tanh(x) * torch.tanh(P[pre + ".a"])


def factorized_cdf(P, pre, x):                        # entropy_models.py:105-117
    x = bitparm(P, pre + ".f1", x, False)
    x = bitparm(P, pre + ".f2", x, False)
    x = bitparm(P, pre + ".f3", x, False)
    x = bitparm(P, pre + ".f4", x, True)
    return torch.sigmoid(x)


def factorized_tables(P, pre, channel):
    """entropy_models.py:119-174 `BitEstimator.update` -> (cdf, cdf_length, offset) numpy int32."""
    medians = torch.zeros((channel,))
    minima = medians + 50
    for i in range(50, 1, -1):
        samples = (torch.zeros_like(medians) - i)[None, :, None, None]
        probs = torch.squeeze(factorized_cdf(P, pre, samples))
        minima = torch.where(probs < torch.zeros_like(medians) + 0.0001,
                             torch.zeros_like(medians) + i, minima)
    maxima = medians + 50
    for i in range(50, 1, -1):
        samples = (torch.zeros_like(medians) + i)[None, :, None, None]
        probs = torch.squeeze(factorized_cdf(P, pre, samples))
        maxima = torch.where(probs > torch.zeros_like(medians) + 0.9999,
                             torch.zeros_like(medians) + i, maxima)
    minima, maxima = minima.int(), maxima.int()
    offset = -minima
    pmf_start = medians - minima
    pmf_length = maxima + minima + 1
    max_length = pmf_length.max()
    samples = torch.arange(max_length)
    samples = samples[None, :] + pmf_start[:, None, None]
    lower = factorized_cdf(P, pre, samples - 0.5).squeeze(0)
    upper = factorized_cdf(P, pre, samples + 0.5).squeeze(0)
    pmf = (upper - lower)[:, 0, :]
    tail_mass = lower[:, 0, :1] + (1.0 - upper[:, 0, -1:])
    qcdf = pmf_to_cdf(pmf, tail_mass, pmf_length, max_length)
    return (qcdf.numpy(), (pmf_length + 2).reshape(-1).int().numpy(),
            offset.reshape(-1).int().numpy())


class ScaleTable:
    """entropy_models.py:198-288 `GaussianEncoder` (tables + scale->index)."""

    def __init__(self, distribution):
        self.distribution = distribution
        self.scale_min = 0.01 if distribution == "laplace" else 0.11
        self.scale_max = 64.0
        self.scale_level = 256
        self.scale_table = torch.exp(torch.linspace(math.log(self.scale_min),
                                                    math.log(self.scale_max), self.scale_level))
        self.log_scale_min = math.log(self.scale_min)
        self.log_scale_max = math.log(self.scale_max)
        self.log_scale_step = (self.log_scale_max - self.log_scale_min) / (self.scale_level - 1)
        self.tables = None

    def _dist(self, mu, scales):
        if self.distribution == "laplace":
            return torch.distributions.laplace.Laplace(mu, scales)
        return torch.distributions.normal.Normal(mu, scales)

    def update(self):                                 # entropy_models.py:224-262
        pmf_center = torch.zeros_like(self.scale_table) + 50
        scales = torch.zeros_like(pmf_center) + self.scale_table
        d = self._dist(torch.zeros_like(scales), scales)
        for i in range(50, 1, -1):
            samples = torch.zeros_like(pmf_center) + i
            probs = torch.squeeze(d.cdf(samples))
            pmf_center = torch.where(probs > torch.zeros_like(pmf_center) + 0.9999,
                                     torch.zeros_like(pmf_center) + i, pmf_center)
        pmf_center = pmf_center.int()
        pmf_length = 2 * pmf_center + 1
        max_length = torch.max(pmf_length).item()
        samples = (torch.arange(max_length) - pmf_center[:, None]).float()
        scales = torch.zeros_like(samples) + self.scale_table[:, None]
        d = self._dist(torch.zeros_like(scales), scales)
        upper = d.cdf(samples + 0.5)
        lower = d.cdf(samples - 0.5)
        pmf = upper - lower
        tail_mass = 2 * lower[:, :1]
        qcdf = pmf_to_cdf(pmf, tail_mass, pmf_length, max_length)
        self.tables = (qcdf.numpy(), (pmf_length + 2).reshape(-1).int().numpy(),
                       (-pmf_center).reshape(-1).int().numpy())
        return self.tables

    def build_indexes(self, scales):                  # entropy_models.py:264-268
        scales = torch.maximum(scales, torch.zeros_like(scales) + 1e-5)
        idx = (torch.log(scales) - self.log_scale_min) / self.log_scale_step
        idx = idx.clamp_(0, self.scale_level - 1)
        return idx.int()


def probs_to_bits(probs):                             # common_model.py:51-55
    bits = -1.0 * torch.log(probs + 1e-5) / math.log(2.0)
    return lower_bound(bits, 0)


def laplace_bits(y, sigma):                           # common_model.py:64-69
    mu = torch.zeros_like(sigma)
    sigma = sigma.clamp(1e-5, 1e10)
    d = torch.distributions.laplace.Laplace(mu, sigma)
    return probs_to_bits(d.cdf(y + 0.5) - d.cdf(y - 0.5))


def gaussian_bits(y, sigma):                          # common_model.py:57-62
    mu = torch.zeros_like(sigma)
    sigma = sigma.clamp(0.11, 1e10)
    d = torch.distributions.normal.Normal(mu, sigma)
    return probs_to_bits(d.cdf(y + 0.5) - d.cdf(y - 0.5))


def z_bits(P, pre, z):                                # common_model.py:71-73
    return probs_to_bits(factorized_cdf(P, pre, z + 0.5) - factorized_cdf(P, pre, z - 0.5))


def checkerboard(height, width):                      # common_model.py:82-89
    micro = torch.tensor(((1, 0), (0, 1)), dtype=torch.float32)
    m0 = micro.repeat(height // 2, width // 2)[None, None]
    return m0, torch.ones_like(m0) - m0


def _masked(y, scales, means, mask, ste=False):       # common_model.py:91-102 (eval: round; train: STE)
    scales_hat = scales * mask
    means_hat = means * mask
    y_res = (y - means_hat) * mask
    y_q = quant_ste(y_res) if ste else torch.round(y_res)
    y_hat = y_q + means_hat
    return y_res, y_q, y_hat, scales_hat


def dual_prior_forward(P, y, means, scales, quant_step, prior_pre, write=False, ste=False):
    """common_model.py:104-177."""
    _, _, H, W = y.size()
    m0, m1 = checkerboard(H, W)
    quant_step = lower_bound(quant_step, 0.5)
    y = y / quant_step
    y_0, y_1 = y.chunk(2, 1)
    s0, s1 = scales.chunk(2, 1)
    u0, u1 = means.chunk(2, 1)
    r00, q00, h00, sh00 = _masked(y_0, s0, u0, m0, ste)
    r11, q11, h11, sh11 = _masked(y_1, s1, u1, m1, ste)
    params = torch.cat((h00, h11, means, scales, quant_step), dim=1)
    s0, u0, s1, u1 = three_convs(P, prior_pre, params).chunk(4, 1)
    r01, q01, h01, sh01 = _masked(y_0, s0, u0, m1, ste)
    r10, q10, h10, sh10 = _masked(y_1, s1, u1, m0, ste)
    y_res = torch.cat((r00 + r01, r11 + r10), dim=1)
    y_q = torch.cat((q00 + q01, q11 + q10), dim=1)
    y_hat = torch.cat((h00 + h01, h11 + h10), dim=1)
    scales_hat = torch.cat((sh00 + sh01, sh11 + sh10), dim=1)
    y_hat = y_hat * quant_step
    if write:
        return q00 + q11, q01 + q10, sh00 + sh11, sh01 + sh10, y_hat
    return y_res, y_q, y_hat, scales_hat


def dual_prior_decode(P, means, scales, quant_step, prior_pre, decode_fn):
    """common_model.py:182-217; decode_fn(scales_plane) -> float symbols of same shape."""
    _, _, H, W = means.size()
    m0, m1 = checkerboard(H, W)
    quant_step = torch.clamp_min(quant_step, 0.5)
    s0, s1 = scales.chunk(2, 1)
    u0, u1 = means.chunk(2, 1)
    q_r0 = decode_fn(s0 * m0 + s1 * m1)
    h00 = (q_r0 + u0) * m0
    h11 = (q_r0 + u1) * m1
    params = torch.cat((h00, h11, means, scales, quant_step), dim=1)
    s0, u0, s1, u1 = three_convs(P, prior_pre, params).chunk(4, 1)
    q_r1 = decode_fn(s0 * m1 + s1 * m0)
    h01 = (q_r1 + u0) * m1
    h10 = (q_r1 + u1) * m0
    y_hat = torch.cat((h00 + h01, h11 + h10), dim=1)
    return y_hat * quant_step


# --------------------------------------------------------------------------------------
# DMC (P-frame)
# --------------------------------------------------------------------------------------
class OracleDMC:
    """Functional twin of video_model.py:131-596 `DMC` (eval mode, batch 1 for write path)."""

    def __init__(self, P):
        self.P = P
        self.table = ScaleTable("laplace")
        self.coder = None

    def update(self):                                 # common_model.py:75-80
        self.coder = OracleRans()
        self.y_tables = self.table.update()
        self.z_tables = factorized_tables(self.P, "bit_estimator_z", 64)
        self.z_mv_tables = factorized_tables(self.P, "bit_estimator_z_mv", 64)

    # -- helpers ------------------------------------------------------------------
    def curr_q(self, basic, q_scale):                 # video_model.py:255-261
        return lower_bound(self.P[basic], 0.5) * q_scale

    def feature_pyramid(self, dpb):                   # video_model.py:17-37, 226-231
        P = self.P
        if dpb["ref_feature"] is None:
            f = conv(P, "feature_adaptor_I", dpb["ref_frame"])
        else:
            f = conv(P, "feature_adaptor_P", dpb["ref_feature"])
        l1 = res_block(P, "feature_extractor.res_block1", conv(P, "feature_extractor.conv1", f))
        l2 = res_block(P, "feature_extractor.res_block2", conv(P, "feature_extractor.conv2", l1, 2))
        l3 = res_block(P, "feature_extractor.res_block3", conv(P, "feature_extractor.conv3", l2, 2))
        return l1, l2, l3

    def context_fusion(self, c1, c2, c3):             # video_model.py:40-68
        P, n = self.P, "context_fusion_net"
        c3_up = F.pixel_shuffle(conv(P, n + ".conv3_up.0", c3), 2)
        c3_up = res_block(P, n + ".res_block3_up", c3_up)
        c3_out = res_block(P, n + ".res_block3_out", conv(P, n + ".conv3_out", c3))
        cat32 = torch.cat((c3_up, c2), dim=1)
        c2_up = F.pixel_shuffle(conv(P, n + ".conv2_up.0", cat32), 2)
        c2_up = res_block(P, n + ".res_block2_up", c2_up)
        c2_out = res_block(P, n + ".res_block2_out", conv(P, n + ".conv2_out", cat32))
        c1_out = conv(P, n + ".conv1_out", torch.cat((c2_up, c1), dim=1))
        c1_out = res_block(P, n + ".res_block1_out", c1_out)
        return c1 + c1_out, c2 + c2_out, c3 + c3_out

    def motion_compensation(self, dpb, mv):           # video_model.py:233-246
        warpframe = torch_warp(dpb["ref_frame"], mv)
        mv2 = bilinear_down2(mv) / 2
        mv3 = bilinear_down2(mv2) / 2
        f1, f2, f3 = self.feature_pyramid(dpb)
        c1, c2, c3 = torch_warp(f1, mv), torch_warp(f2, mv2), torch_warp(f3, mv3)
        c1, c2, c3 = self.context_fusion(c1, c2, c3)
        return c1, c2, c3, warpframe

    def contextual_encoder(self, x, c1, c2, c3):      # video_model.py:71-90
        P, n = self.P, "contextual_encoder"
        f = conv(P, n + ".conv1", torch.cat([x, c1], dim=1), 2)
        f = res_block(P, n + ".res1", torch.cat([f, c2], dim=1), 0.1, True, True)
        f = conv(P, n + ".conv2", f, 2)
        f = res_block(P, n + ".res2", torch.cat([f, c3], dim=1), 0.1, True, True)
        f = conv(P, n + ".conv3", f, 2)
        return conv(P, n + ".conv4", f, 2)

    def contextual_decoder(self, y_hat, c2, c3):      # video_model.py:93-112
        P, n = self.P, "contextual_decoder"
        f = F.pixel_shuffle(conv(P, n + ".up1.0", y_hat), 2)
        f = F.pixel_shuffle(conv(P, n + ".up2.0", f), 2)
        f = res_block(P, n + ".res1", torch.cat([f, c3], dim=1), 0.1, True, True)
        f = F.pixel_shuffle(conv(P, n + ".up3.0", f), 2)
        f = res_block(P, n + ".res2", torch.cat([f, c2], dim=1), 0.1, True, True)
        return F.pixel_shuffle(conv(P, n + ".up4.0", f), 2)

    def recon_generation(self, feat32, ctx64):
        """video_model.py:115-128.  NB the call site (video_model.py:330) passes
        (recon_image_feature, context1) into forward(ctx, res), so the concat order is
        [32-ch decoder feature, 64-ch context1]."""
        P, n = self.P, "recon_generation_net"
        f = conv(P, n + ".first_conv", torch.cat((feat32, ctx64), dim=1))
        f = unet(P, n + ".unet_1", f)
        f = unet(P, n + ".unet_2", f)
        return f, conv(P, n + ".recon_conv", f)

    def temporal_prior(self, c3):                     # video_model.py:193-197
        P = self.P
        t = F.leaky_relu(conv(P, "temporal_prior_encoder.0", c3, 2), 0.1)
        return conv(P, "temporal_prior_encoder.2", t, 2)

    def ctx_hyper_enc(self, y):                       # video_model.py:173-179
        P, n = self.P, "contextual_hyper_prior_encoder"
        t = F.leaky_relu(conv(P, n + ".0", y), 0.01)
        t = F.leaky_relu(conv(P, n + ".2", t, 2), 0.01)
        return conv(P, n + ".4", t, 2)

    # -- encoder-side shared trunk --------------------------------------------------
    def _analyse(self, x, dpb, mv_y_q_scale, y_q_scale, write, ste=False):
        P = self.P
        rnd = quant_ste if ste else torch.round
        q_mv = self.curr_q("mv_y_q_basic", mv_y_q_scale)
        q_y = self.curr_q("y_q_basic", y_q_scale)
        est_mv = spynet(P, "optic_flow", x, dpb["ref_frame"])
        mv_y = analysis(P, "mv_encoder", est_mv) / q_mv
        mv_z = hyper_analysis(P, "mv_hyper_prior_encoder", mv_y)
        mv_z_hat = rnd(mv_z)
        mv_params = hyper_synthesis(P, "mv_hyper_prior_decoder", mv_z_hat)
        ref_mv_y = dpb["ref_mv_y"] if dpb["ref_mv_y"] is not None else torch.zeros_like(mv_y)
        mv_q_step, mv_scales, mv_means = three_convs(
            P, "mv_y_prior_fusion", torch.cat((mv_params, ref_mv_y), dim=1)).chunk(3, 1)
        mv_dp = dual_prior_forward(P, mv_y, mv_means, mv_scales, mv_q_step,
                                   "mv_y_spatial_prior", write, ste)
        mv_y_hat = mv_dp[4 if write else 2] * q_mv
        mv_hat = synthesis(P, "mv_decoder", mv_y_hat)
        c1, c2, c3, warp_frame = self.motion_compensation(dpb, mv_hat)
        y = self.contextual_encoder(x, c1, c2, c3) / q_y
        z = self.ctx_hyper_enc(y)
        z_hat = rnd(z)
        hier = hyper_synthesis(P, "contextual_hyper_prior_decoder", z_hat)
        temporal = self.temporal_prior(c3)
        ref_y = dpb["ref_y"] if dpb["ref_y"] is not None else torch.zeros_like(y)
        q_step, scales, means = three_convs(
            P, "y_prior_fusion", torch.cat((temporal, hier, ref_y), dim=1)).chunk(3, 1)
        y_dp = dual_prior_forward(P, y, means, scales, q_step, "y_spatial_prior", write, ste)
        y_hat = y_dp[4 if write else 2] * q_y
        feat_res = self.contextual_decoder(y_hat, c2, c3)
        feature, x_hat = self.recon_generation(feat_res, c1)
        return dict(mv_z=mv_z, mv_z_hat=mv_z_hat, mv_dp=mv_dp, mv_y_hat=mv_y_hat, z=z, z_hat=z_hat,
                    y_dp=y_dp, y_hat=y_hat, feature=feature, x_hat=x_hat, warp_frame=warp_frame,
                    est_mv=est_mv, mv_hat=mv_hat)

    # -- public twins ---------------------------------------------------------------
    def _enc_z(self, z_hat, tables):                  # entropy_models.py:176-187
        C = z_hat.shape[1]
        idx = torch.arange(C).view(1, -1, 1, 1).int().repeat(1, 1, z_hat.shape[2], z_hat.shape[3])
        self.coder.encode_with_indexes(z_hat.reshape(-1).int().numpy(),
                                       idx[0].reshape(-1).int().numpy(), *tables)

    def _enc_y(self, sym, scales):                    # entropy_models.py:270-274
        idx = self.table.build_indexes(scales)
        self.coder.encode_with_indexes(sym.reshape(-1).int().numpy(),
                                       idx.reshape(-1).int().numpy(), *self.y_tables)

    def _dec_z(self, size, tables, C=64):             # entropy_models.py:189-195
        idx = torch.arange(C).view(1, -1, 1, 1).int().repeat(1, 1, size[0], size[1])
        val = self.coder.decode_stream(idx.reshape(-1).int().numpy(), *tables)
        return torch.Tensor(val).reshape(idx.shape).float()

    def _dec_y(self, scales):                         # entropy_models.py:276-281
        idx = self.table.build_indexes(scales)
        val = self.coder.decode_stream(idx.reshape(-1).int().numpy(), *self.y_tables)
        return torch.Tensor(val).reshape(scales.shape).float()

    @torch.no_grad()
    def compress(self, x, dpb, mv_y_q_scale, y_q_scale, want_symbols=False):
        """video_model.py:263-352."""
        t = self._analyse(x, dpb, mv_y_q_scale, y_q_scale, write=True)
        mq0, mq1, ms0, ms1, _ = t["mv_dp"]
        yq0, yq1, ys0, ys1, _ = t["y_dp"]
        self.coder.reset_encoder()
        self._enc_z(t["mv_z_hat"], self.z_mv_tables)
        self._enc_y(mq0, ms0)
        self._enc_y(mq1, ms1)
        self._enc_z(t["z_hat"], self.z_tables)
        self._enc_y(yq0, ys0)
        self._enc_y(yq1, ys1)
        out = {"dbp": {"ref_frame": t["x_hat"], "ref_feature": t["feature"], "ref_y": t["y_hat"],
                       "ref_mv_y": t["mv_y_hat"]},
               "bit_stream": self.coder.flush_encoder()}
        if want_symbols:
            B = self.table.build_indexes
            out["symbols"] = dict(
                mv_z=t["mv_z_hat"].to(torch.int16), z=t["z_hat"].to(torch.int16),
                mv_y_q_w_0=mq0.to(torch.int16), mv_y_q_w_1=mq1.to(torch.int16),
                y_q_w_0=yq0.to(torch.int16), y_q_w_1=yq1.to(torch.int16),
                mv_idx_0=B(ms0).to(torch.uint8), mv_idx_1=B(ms1).to(torch.uint8),
                y_idx_0=B(ys0).to(torch.uint8), y_idx_1=B(ys1).to(torch.uint8))
            out["aux"] = dict(est_mv=t["est_mv"], mv_hat=t["mv_hat"])
        return out

    @torch.no_grad()
    def decompress(self, dpb, string, height, width, mv_y_q_scale, y_q_scale):
        """video_model.py:354-422."""
        P = self.P
        q_mv = self.curr_q("mv_y_q_basic", mv_y_q_scale)
        q_y = self.curr_q("y_q_basic", y_q_scale)
        self.coder.set_stream(string)
        zs = get_downsampled_shape(height, width, 64)
        mv_z_hat = self._dec_z(zs, self.z_mv_tables)
        mv_params = hyper_synthesis(P, "mv_hyper_prior_decoder", mv_z_hat)
        ref_mv_y = dpb["ref_mv_y"]
        if ref_mv_y is None:
            _, C, H, W = mv_params.size()
            ref_mv_y = torch.zeros((1, C // 2, H, W))
        mv_q_step, mv_scales, mv_means = three_convs(
            P, "mv_y_prior_fusion", torch.cat((mv_params, ref_mv_y), dim=1)).chunk(3, 1)
        mv_y_hat = dual_prior_decode(P, mv_means, mv_scales, mv_q_step, "mv_y_spatial_prior",
                                     self._dec_y) * q_mv
        mv_hat = synthesis(P, "mv_decoder", mv_y_hat)
        c1, c2, c3, _ = self.motion_compensation(dpb, mv_hat)
        z_hat = self._dec_z(zs, self.z_tables)
        hier = hyper_synthesis(P, "contextual_hyper_prior_decoder", z_hat)
        temporal = self.temporal_prior(c3)
        ref_y = dpb["ref_y"]
        if ref_y is None:
            _, C, H, W = temporal.size()
            ref_y = torch.zeros((1, C // 2, H, W))
        q_step, scales, means = three_convs(
            P, "y_prior_fusion", torch.cat((temporal, hier, ref_y), dim=1)).chunk(3, 1)
        y_hat = dual_prior_decode(P, means, scales, q_step, "y_spatial_prior", self._dec_y) * q_y
        feat_res = self.contextual_decoder(y_hat, c2, c3)
        feature, recon = self.recon_generation(feat_res, c1)
        return {"dpb": {"ref_frame": recon.clamp(0, 1), "ref_feature": feature, "ref_y": y_hat,
                        "ref_mv_y": mv_y_hat}}

    @torch.no_grad()
    def encode_decode_bytes(self, x, dpb, pic_height, pic_width, mv_y_q_scale, y_q_scale):
        """video_model.py:424-455 with the .bin kept in memory: returns (result, blob)."""
        mv_q, mv_qi = get_rounded_q(mv_y_q_scale)
        y_q, y_qi = get_rounded_q(y_q_scale)
        enc = self.compress(x, dpb, mv_q, y_q)
        blob = pack_p(enc["bit_stream"], mv_qi, y_qi)
        a, b, string = unpack_p(blob)
        dec = self.decompress(dpb, string, pic_height, pic_width, a / 100, b / 100)
        return {"dpb": dec["dpb"], "bit": len(blob) * 8}, blob

    @torch.no_grad()
    def forward_one_frame(self, x, dpb, mv_y_q_scale, y_q_scale):
        """video_model.py:470-592 (eval mode)."""
        return self._forward_one_frame(x, dpb, mv_y_q_scale, y_q_scale, None)

    def forward_one_frame_train(self, x, dpb, mv_y_q_scale, y_q_scale, noise):
        """video_model.py:470-592 in TRAIN mode, differentiable (P entries may require grad): quant is
        the straight-through estimator (common_model.py:38-44) and the rates come from
        add_noise(y_res), add_noise(mv_y_res), add_noise(z), add_noise(mv_z) (:546-550).  `noise`: the four
        U(-0.5, 0.5) tensors in that order (the reference draws them from the global RNG in this order)."""
        return self._forward_one_frame(x, dpb, mv_y_q_scale, y_q_scale, noise)

    def _forward_one_frame(self, x, dpb, mv_y_q_scale, y_q_scale, noise):
        t = self._analyse(x, dpb, mv_y_q_scale, y_q_scale, write=False, ste=noise is not None)
        mv_y_res, mv_y_q, _, mv_scales_hat = t["mv_dp"]
        y_res, y_q, _, scales_hat = t["y_dp"]
        if noise is not None:
            y_q, mv_y_q = y_res + noise[0], mv_y_res + noise[1]
            t = dict(t, z_hat=t["z"] + noise[2], mv_z_hat=t["mv_z"] + noise[3], z_hat_q=t["z_hat"],
                     mv_z_hat_q=t["mv_z_hat"])
        B, _, H, W = x.size()
        pixel_num = H * W
        mse = torch.sum(F.mse_loss(x, t["x_hat"], reduction="none"), dim=(1, 2, 3)) / pixel_num
        me_mse = torch.sum(F.mse_loss(x, t["warp_frame"], reduction="none"), dim=(1, 2, 3)) / pixel_num
        bpp_y = torch.sum(laplace_bits(y_q, scales_hat), dim=(1, 2, 3)) / pixel_num
        bpp_mv_y = torch.sum(laplace_bits(mv_y_q, mv_scales_hat), dim=(1, 2, 3)) / pixel_num
        bpp_z = torch.sum(z_bits(self.P, "bit_estimator_z", t["z_hat"]), dim=(1, 2, 3)) / pixel_num
        bpp_mv_z = torch.sum(z_bits(self.P, "bit_estimator_z_mv", t["mv_z_hat"]), dim=(1, 2, 3)) / pixel_num
        bpp = bpp_y + bpp_z + bpp_mv_y + bpp_mv_z
        return {"bpp_mv_y": bpp_mv_y, "bpp_mv_z": bpp_mv_z, "bpp_y": bpp_y, "bpp_z": bpp_z, "bpp": bpp,
                "me_mse": me_mse, "mse": mse,
                "dpb": {"ref_frame": t["x_hat"], "ref_feature": t["feature"], "ref_y": t["y_hat"],
                        "ref_mv_y": t["mv_y_hat"]},
                "bit": torch.sum(bpp) * pixel_num, "bit_y": torch.sum(bpp_y) * pixel_num,
                "bit_z": torch.sum(bpp_z) * pixel_num, "bit_mv_y": torch.sum(bpp_mv_y) * pixel_num,
                "bit_mv_z": torch.sum(bpp_mv_z) * pixel_num}


# --------------------------------------------------------------------------------------
# IntraNoAR (I-frame)
# --------------------------------------------------------------------------------------
class OracleIntra:
    """Functional twin of image_model.py:16-200 `IntraNoAR` (eval mode)."""

    def __init__(self, P, N=192):
        self.P, self.N = P, N
        self.table = ScaleTable("gaussian")
        self.coder = None

    def update(self):
        self.coder = OracleRans()
        self.y_tables = self.table.update()
        self.z_tables = factorized_tables(self.P, "bit_estimator_z", self.N)

    def curr_q(self, q_scale):                        # image_model.py:50-52
        return lower_bound(self.P["q_basic"], 0.5) * q_scale

    def refine(self, t):                              # image_model.py:21-24
        return conv(self.P, "refine.1", unet(self.P, "refine.0", t))

    def _trunk(self, x, q_scale, write):
        P = self.P
        q = self.curr_q(q_scale)
        y = analysis(P, "enc", x) / q
        z = hyper_analysis(P, "hyper_enc", y)
        z_hat = torch.round(z)
        params = hyper_synthesis(P, "hyper_dec", z_hat)
        q_step, scales, means = three_convs(P, "y_prior_fusion", params).chunk(3, 1)
        dp = dual_prior_forward(P, y, means, scales, q_step, "y_spatial_prior", write)
        y_hat = dp[4 if write else 2] * q
        return z_hat, dp, y_hat

    @torch.no_grad()
    def forward(self, x, q_scale):                    # image_model.py:54-105
        z_hat, dp, y_hat = self._trunk(x, q_scale, False)
        _, y_q, _, scales_hat = dp
        x_hat = self.refine(synthesis(self.P, "dec", y_hat))
        _, _, H, W = x.size()
        pixel_num = H * W
        bpp_y = torch.sum(gaussian_bits(y_q, scales_hat), dim=(1, 2, 3)) / pixel_num
        bpp_z = torch.sum(z_bits(self.P, "bit_estimator_z", z_hat), dim=(1, 2, 3)) / pixel_num
        mse = torch.sum(F.mse_loss(x, x_hat, reduction="none"), dim=(1, 2, 3)) / pixel_num
        bits = torch.sum(bpp_y + bpp_z) * pixel_num
        return {"x_hat": x_hat, "mse": mse, "bit": bits.item(), "bpp": bpp_y + bpp_z,
                "bpp_y": bpp_y, "bpp_z": bpp_z}

    @torch.no_grad()
    def compress(self, x, q_scale, want_symbols=False):     # image_model.py:148-177
        z_hat, dp, _ = self._trunk(x, q_scale, True)
        q0, q1, s0, s1, _ = dp
        c = self.coder
        c.reset_encoder()
        C = self.N
        idx = torch.arange(C).view(1, -1, 1, 1).int().repeat(1, 1, z_hat.shape[2], z_hat.shape[3])
        c.encode_with_indexes(z_hat.reshape(-1).int().numpy(), idx[0].reshape(-1).int().numpy(),
                              *self.z_tables)
        for sym, sc in ((q0, s0), (q1, s1)):
            c.encode_with_indexes(sym.reshape(-1).int().numpy(),
                                  self.table.build_indexes(sc).reshape(-1).int().numpy(),
                                  *self.y_tables)
        out = {"bit_stream": c.flush_encoder()}
        if want_symbols:
            B = self.table.build_indexes
            out["symbols"] = dict(z=z_hat.to(torch.int16), y_q_w_0=q0.to(torch.int16),
                                  y_q_w_1=q1.to(torch.int16), y_idx_0=B(s0).to(torch.uint8),
                                  y_idx_1=B(s1).to(torch.uint8))
        return out

    @torch.no_grad()
    def decompress(self, bit_stream, height, width, q_scale):   # image_model.py:179-200
        P = self.P
        q = self.curr_q(q_scale)
        c = self.coder
        c.set_stream(bit_stream)
        zs = get_downsampled_shape(height, width, 64)
        idx = torch.arange(self.N).view(1, -1, 1, 1).int().repeat(1, 1, zs[0], zs[1])
        z_hat = torch.Tensor(c.decode_stream(idx.reshape(-1).int().numpy(), *self.z_tables)) \
            .reshape(idx.shape).float()
        params = hyper_synthesis(P, "hyper_dec", z_hat)
        q_step, scales, means = three_convs(P, "y_prior_fusion", params).chunk(3, 1)

        def dec(sc):
            i = self.table.build_indexes(sc)
            v = c.decode_stream(i.reshape(-1).int().numpy(), *self.y_tables)
            return torch.Tensor(v).reshape(sc.shape).float()

        y_hat = dual_prior_decode(P, means, scales, q_step, "y_spatial_prior", dec) * q
        x_hat = self.refine(synthesis(P, "dec", y_hat)).clamp_(0, 1)
        return {"x_hat": x_hat}

    @torch.no_grad()
    def encode_decode_bytes(self, x, q_scale, pic_height, pic_width):   # image_model.py:114-146
        q, qi = get_rounded_q(q_scale)
        blob = pack_i(pic_height, pic_width, qi, self.compress(x, q)["bit_stream"])
        h, w, qi2, string = unpack_i(blob)
        return {"bit": len(blob) * 8, "x_hat": self.decompress(string, h, w, qi2 / 100)["x_hat"]}, blob


def state_shapes(kind):
    """{state-dict key: shape} of the reference DMC / IntraNoAR parameter tree
    (video_model.py:144-222, image_model.py:20-45), restated so fixtures need no reference."""
    S = {}

    def c(name, cin, cout, k):
        S[name + ".weight"] = (cout, cin, k, k)
        S[name + ".bias"] = (cout,)

    def rbs(p, cin, cout):
        c(p + ".conv1", cin, cout, 3); c(p + ".conv2", cout, cout, 3); c(p + ".downsample", cin, cout, 1)

    def rb(p, ch):
        c(p + ".conv1", ch, ch, 3); c(p + ".conv2", ch, ch, 3)

    def rbu(p, ch):
        c(p + ".subpel_conv.0", ch, ch * 4, 1); c(p + ".conv", ch, ch, 3); c(p + ".upsample.0", ch, ch * 4, 1)

    def encdec(enc, dec, cin, cout, ch):
        rbs(enc + ".0", cin, ch); rb(enc + ".1", ch); rbs(enc + ".2", ch, ch); rb(enc + ".3", ch)
        rbs(enc + ".4", ch, ch); rb(enc + ".5", ch); c(enc + ".6", ch, ch, 3)
        rb(dec + ".0", ch); rbu(dec + ".1", ch); rb(dec + ".2", ch); rbu(dec + ".3", ch)
        rb(dec + ".4", ch); rbu(dec + ".5", ch); rb(dec + ".6", ch); c(dec + ".7.0", ch, cout * 4, 1)

    def hyper(enc, dec, yc, zc):
        c(enc + ".0", yc, zc, 3)
        for i in (2, 4, 6, 8):
            c(f"{enc}.{i}", zc, zc, 3)
        c(dec + ".0", zc, yc, 3); c(dec + ".2.0", yc, yc * 4, 1); c(dec + ".4", yc, yc * 3 // 2, 3)
        c(dec + ".6.0", yc * 3 // 2, yc * 3 // 2 * 4, 1); c(dec + ".8", yc * 3 // 2, yc * 2, 3)

    def resb(p, ch, bottleneck=False):
        mid = ch // 2 if bottleneck else ch
        c(p + ".conv1", ch, mid, 3); c(p + ".conv2", mid, ch, 3)

    def cbr(p, cin, cout):
        c(p + ".conv.0", cin, cout, 3); c(p + ".conv.2", cout, cout, 3)
        S[p + ".conv.3.fc.0.weight"] = (cout // 16, cout)
        S[p + ".conv.3.fc.2.weight"] = (cout, cout // 16)
        c(p + ".up_dim", cin, cout, 1)

    def unet_(p, cin, cout):
        cbr(p + ".conv1", cin, 32); cbr(p + ".conv2", 32, 64); cbr(p + ".conv3", 64, 128)
        for i in range(4):
            resb(f"{p}.context_refine.{i}", 128)
        c(p + ".up3.0", 128, 64 * 4, 1); cbr(p + ".up_conv3", 128, 64)
        c(p + ".up2.0", 64, 32 * 4, 1); cbr(p + ".up_conv2", 64, cout)

    def bitest(p, ch):
        for f in ("f1", "f2", "f3", "f4"):
            S[f"{p}.{f}.h"] = (1, ch, 1, 1); S[f"{p}.{f}.b"] = (1, ch, 1, 1)
            if f != "f4":
                S[f"{p}.{f}.a"] = (1, ch, 1, 1)

    def stack3(p, chans):
        for i, (a, b) in zip((0, 2, 4), chans):
            c(f"{p}.{i}", a, b, 3)

    if kind == "dmc":
        bitest("bit_estimator_z", 64); bitest("bit_estimator_z_mv", 64)
        for lv in range(4):
            b = f"optic_flow.moduleBasic.{lv}"
            c(b + ".conv1", 8, 32, 7); c(b + ".conv2", 32, 64, 7); c(b + ".conv3", 64, 32, 7)
            c(b + ".conv4", 32, 16, 7); c(b + ".conv5", 16, 2, 7)
        encdec("mv_encoder", "mv_decoder", 2, 2, 64)
        hyper("mv_hyper_prior_encoder", "mv_hyper_prior_decoder", 64, 64)
        stack3("mv_y_prior_fusion", ((192, 192), (192, 192), (192, 192)))
        stack3("mv_y_spatial_prior", ((256, 192), (192, 192), (192, 128)))
        c("feature_adaptor_I", 3, 64, 3); c("feature_adaptor_P", 64, 64, 1)
        for i in (1, 2, 3):
            c(f"feature_extractor.conv{i}", 64, 64, 3); resb(f"feature_extractor.res_block{i}", 64)
        n = "context_fusion_net"
        c(n + ".conv3_up.0", 64, 256, 3); resb(n + ".res_block3_up", 64); c(n + ".conv3_out", 64, 64, 3)
        resb(n + ".res_block3_out", 64); c(n + ".conv2_up.0", 128, 256, 3); resb(n + ".res_block2_up", 64)
        c(n + ".conv2_out", 128, 64, 3); resb(n + ".res_block2_out", 64); c(n + ".conv1_out", 128, 64, 3)
        resb(n + ".res_block1_out", 64)
        n = "contextual_encoder"
        c(n + ".conv1", 67, 64, 3); resb(n + ".res1", 128, True); c(n + ".conv2", 128, 64, 3)
        resb(n + ".res2", 128, True); c(n + ".conv3", 128, 64, 3); c(n + ".conv4", 64, 96, 3)
        n = "contextual_hyper_prior_encoder"
        c(n + ".0", 96, 64, 3); c(n + ".2", 64, 64, 3); c(n + ".4", 64, 64, 3)
        n = "contextual_hyper_prior_decoder"
        c(n + ".0", 64, 96, 3); c(n + ".2.0", 96, 384, 1); c(n + ".4", 96, 144, 3)
        c(n + ".6.0", 144, 576, 1); c(n + ".8", 144, 192, 3)
        c("temporal_prior_encoder.0", 64, 144, 3); c("temporal_prior_encoder.2", 144, 192, 3)
        stack3("y_prior_fusion", ((480, 384), (384, 288), (288, 288)))
        stack3("y_spatial_prior", ((384, 288), (288, 288), (288, 192)))
        n = "contextual_decoder"
        c(n + ".up1.0", 96, 256, 3); c(n + ".up2.0", 64, 256, 3); resb(n + ".res1", 128, True)
        c(n + ".up3.0", 128, 256, 3); resb(n + ".res2", 128, True); c(n + ".up4.0", 128, 128, 3)
        n = "recon_generation_net"
        c(n + ".first_conv", 96, 64, 3); unet_(n + ".unet_1", 64, 64); unet_(n + ".unet_2", 64, 64)
        c(n + ".recon_conv", 64, 3, 3)
        S["mv_y_q_basic"] = (1, 64, 1, 1); S["mv_y_q_scale"] = (4, 1, 1, 1)
        S["y_q_basic"] = (1, 96, 1, 1); S["y_q_scale"] = (4, 1, 1, 1)
    elif kind == "intra":
        N = 192
        bitest("bit_estimator_z", N)
        encdec("enc", "dec", 3, 16, N)
        unet_("refine.0", 16, 16); c("refine.1", 16, 3, 3)
        hyper("hyper_enc", "hyper_dec", N, N)
        stack3("y_prior_fusion", ((2 * N, 3 * N), (3 * N, 3 * N), (3 * N, 3 * N)))
        stack3("y_spatial_prior", ((4 * N, 3 * N), (3 * N, 3 * N), (3 * N, 2 * N)))
        S["q_basic"] = (1, N, 1, 1); S["q_scale"] = (4, 1, 1, 1)
    else:
        raise ValueError(kind)
    return S
