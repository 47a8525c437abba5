"""Entropy-model host side: CDF-table construction (once per model load, CPU, as the reference does
it) and the wrapper over the native rANS coder of libvcmts_b200.so.

Mirrors DCVC_HEM/src/entropy_models/entropy_models.py:
    EntropyCoder      :9-51     -> RansCoder (handles into csrc/rans_coder.cpp)
    BitEstimator      :94-195   -> factorized_tables()
    GaussianEncoder   :198-288  -> ScaleIndexTable
The tables must equal the reference's bit for bit (tests/golden/cdf_tables.npz), so the float math
below is issued as the same torch CPU ops in the same order.
"""
import ctypes
import math

import numpy as np
import torch

from . import _lib


def _pmf_to_cdf(pmf, tail_mass, pmf_length, max_length):
    """entropy_models.py:24-32 on top of vcm_pmf_to_quantized_cdf (ops.cpp:24-82)."""
    L = _lib.lib()
    n_tab = len(pmf_length)
    cdf = np.zeros((n_tab, int(max_length) + 2), dtype=np.int32)
    for i in range(n_tab):
        prob = torch.cat((pmf[i][: int(pmf_length[i])], tail_mass[i]), dim=0)
        p = np.ascontiguousarray(prob.detach().cpu().numpy(), dtype=np.float32)
        row = np.empty(p.size + 1, dtype=np.int32)
        rc = L.vcm_pmf_to_quantized_cdf(p.ctypes.data, p.size, 16, row.ctypes.data)
        if rc != 0:
            raise ValueError("pmf_to_quantized_cdf: " + L.vcm_last_error().decode())
        cdf[i, : row.size] = row
    return cdf


def _bitparm(h, b, a, x):
    """entropy_models.py:68-73."""
    x = x * torch.nn.functional.softplus(h) + b
    return x if a is None else x + torch.tanh(x) * torch.tanh(a)


def factorized_cdf(fp, x):
    """entropy_models.py:105-117; fp = {'f1.h': tensor[1,C,1,1], ...} on x's device."""
    for f in ("f1", "f2", "f3"):
        x = _bitparm(fp[f + ".h"], fp[f + ".b"], fp[f + ".a"], x)
    x = _bitparm(fp["f4.h"], fp["f4.b"], None, x)
    return torch.sigmoid(x)


@torch.no_grad()
def factorized_tables(fp, channel):
    """entropy_models.py:119-174 `BitEstimator.update` -> (cdf, cdf_length, offset) int32 arrays."""
    fp = {k: v.detach().float().cpu() for k, v in fp.items()}
    zero = torch.zeros((channel,))
    lo = zero + 50
    hi = zero + 50
    for i in range(50, 1, -1):
        p = torch.squeeze(factorized_cdf(fp, (zero - i)[None, :, None, None]))
        lo = torch.where(p < zero + 0.0001, zero + i, lo)
    for i in range(50, 1, -1):
        p = torch.squeeze(factorized_cdf(fp, (zero + i)[None, :, None, None]))
        hi = torch.where(p > zero + 0.9999, zero + i, hi)
    lo, hi = lo.int(), hi.int()
    pmf_start = zero - lo
    pmf_length = hi + lo + 1
    max_length = pmf_length.max()
    samples = torch.arange(max_length)[None, :] + pmf_start[:, None, None]
    lower = factorized_cdf(fp, samples - 0.5).squeeze(0)
    upper = factorized_cdf(fp, samples + 0.5).squeeze(0)
    pmf = (upper - lower)[:, 0, :]
    tail = lower[:, 0, :1] + (1.0 - upper[:, 0, -1:])
    cdf = _pmf_to_cdf(pmf, tail, pmf_length, max_length)
    return cdf, (pmf_length + 2).reshape(-1).int().numpy().copy(), (-lo).reshape(-1).int().numpy().copy()


class ScaleIndexTable:
    """entropy_models.py:198-288 `GaussianEncoder`: 256 log-spaced scales -> CDF rows."""

    def __init__(self, distribution):
        assert distribution in ("laplace", "gaussian")
        self.distribution = distribution
        self.scale_min = 0.01 if distribution == "laplace" else 0.11
        self.scale_max, self.scale_level = 64.0, 256
        self.log_scale_min = math.log(self.scale_min)
        self.log_scale_max = math.log(self.scale_max)
        self.log_scale_step = (self.log_scale_max - self.log_scale_min) / (self.scale_level - 1)

    def _cdf(self, scales, x):
        zero = torch.zeros_like(scales)
        if self.distribution == "laplace":
            return torch.distributions.laplace.Laplace(zero, scales).cdf(x)
        return torch.distributions.normal.Normal(zero, scales).cdf(x)

    @torch.no_grad()
    def update(self):
        table = torch.exp(torch.linspace(self.log_scale_min, self.log_scale_max, self.scale_level))
        center = torch.zeros_like(table) + 50
        scales = torch.zeros_like(center) + table
        for i in range(50, 1, -1):
            p = torch.squeeze(self._cdf(scales, torch.zeros_like(center) + i))
            center = torch.where(p > torch.zeros_like(center) + 0.9999, torch.zeros_like(center) + i, center)
        center = center.int()
        pmf_length = 2 * center + 1
        max_length = torch.max(pmf_length).item()
        samples = (torch.arange(max_length) - center[:, None]).float()
        scales = torch.zeros_like(samples) + table[:, None]
        upper = self._cdf(scales, samples + 0.5)
        lower = self._cdf(scales, samples - 0.5)
        cdf = _pmf_to_cdf(upper - lower, 2 * lower[:, :1], pmf_length, max_length)
        return cdf, (pmf_length + 2).reshape(-1).int().numpy().copy(), (-center).reshape(-1).int().numpy().copy()


class CdfTables:
    """Owns a native vcm_cdf_tables*."""

    def __init__(self, cdf, sizes, offsets):
        self.cdf = np.ascontiguousarray(cdf, dtype=np.int32)
        self.sizes = np.ascontiguousarray(sizes, dtype=np.int32)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int32)
        L = _lib.lib()
        self.h = L.vcm_cdf_tables_new(self.cdf.ctypes.data, self.sizes.ctypes.data, self.offsets.ctypes.data,
                                      self.cdf.shape[0], self.cdf.shape[1])
        if not self.h:
            raise _lib.VcmError("vcm_cdf_tables_new: " + L.vcm_last_error().decode())

    def __del__(self):
        try:
            if self.h:
                _lib.lib().vcm_cdf_tables_free(self.h)
        except Exception:
            pass


class RansCoder:
    """EntropyCoder twin (entropy_models.py:9-51) over int16 symbols / uint8 table indexes held in
    host (pinned) torch tensors.  ctypes releases the GIL for the duration of each native call."""

    def __init__(self):
        L = _lib.lib()
        self.L = L
        self.enc = L.vcm_rans_enc_new()
        self.dec = L.vcm_rans_dec_new()
        self._stream = None
        self._live = []        # tables pushed from since the last flush: the encoder points into them

    def __del__(self):
        try:
            self.L.vcm_rans_enc_free(self.enc)
            self.L.vcm_rans_dec_free(self.dec)
        except Exception:
            pass

    def reset_encoder(self):
        self.L.vcm_rans_enc_reset(self.enc)
        self._live.clear()

    def encode_planes(self, sym, idx, tables):
        """sym int16 / idx uint8 host tensors of equal numel (GaussianEncoder.encode, :270-274)."""
        assert sym.dtype == torch.int16 and idx.dtype == torch.uint8 and sym.numel() == idx.numel()
        assert sym.device.type == "cpu" and sym.is_contiguous() and idx.is_contiguous()
        self._live.append(tables)
        _lib.check(self.L.vcm_rans_enc_push(self.enc, sym.data_ptr(), idx.data_ptr(), sym.numel(), tables.h),
                   "vcm_rans_enc_push")

    def encode_channels(self, sym, tables):
        """sym int16 host tensor [1, C, h, w]; table index = channel (BitEstimator.encode, :183-187)."""
        assert sym.dtype == torch.int16 and sym.device.type == "cpu" and sym.is_contiguous()
        assert sym.dim() == 4 and sym.shape[0] == 1, "the bitstream path is batch 1 (entropy_models.py:186)"
        C, hw = sym.shape[1], sym.shape[2] * sym.shape[3]
        self._live.append(tables)
        _lib.check(self.L.vcm_rans_enc_push_channels(self.enc, sym.data_ptr(), C, hw, tables.h),
                   "vcm_rans_enc_push_channels")

    def flush_encoder(self):
        p = ctypes.c_void_p()
        n = self.L.vcm_rans_enc_flush(self.enc, ctypes.byref(p))
        self._live.clear()
        if n < 0:
            raise _lib.VcmError("vcm_rans_enc_flush: " + self.L.vcm_last_error().decode())
        return ctypes.string_at(p, n)

    def set_stream(self, stream):
        self._stream = bytes(stream)
        _lib.check(self.L.vcm_rans_dec_set_stream(self.dec, self._stream, len(self._stream)),
                   "vcm_rans_dec_set_stream")

    def decode_planes(self, idx, tables, out):
        assert idx.dtype == torch.uint8 and out.dtype == torch.int16 and idx.numel() == out.numel()
        assert idx.device.type == "cpu" and out.device.type == "cpu"
        _lib.check(self.L.vcm_rans_dec_decode(self.dec, idx.data_ptr(), idx.numel(), tables.h, out.data_ptr()),
                   "vcm_rans_dec_decode")
        return out

    def decode_channels(self, C, hw, tables, out):
        assert out.dtype == torch.int16 and out.numel() == C * hw and out.device.type == "cpu"
        _lib.check(self.L.vcm_rans_dec_decode_channels(self.dec, C, hw, tables.h, out.data_ptr()),
                   "vcm_rans_dec_decode_channels")
        return out
