"""Device backend: NHWC tensor views + one method per C-ABI kernel of libvcmts_b200.so.

`View` is how the reference's torch.cat / chunk / split (video_model.py:59-63,84-87,108-110,124,
290,318; common_model.py:133,157-160) disappear: producers write into channel slices of one NHWC
buffer and consumers read slices.

`CudaBackend` is the product path.  It raises if the library or a B200 is missing; there is no
CPU path in this package (tests/ inject their own emulation backend to check host logic).
"""
import ctypes

import torch

from . import _lib
from ._lib import BF16, F32, ConvDesc, DpDesc, check


class View:
    """Channel slice [c0, c0+C) of a contiguous NHWC tensor `t` of shape [N, H, W, Ct]."""
    __slots__ = ("t", "c0", "C")

    def __init__(self, t, c0=0, C=None):
        assert t.dim() == 4 and t.is_contiguous()
        self.t, self.c0 = t, c0
        self.C = t.shape[3] - c0 if C is None else C
        assert 0 <= c0 and c0 + self.C <= t.shape[3]

    N = property(lambda s: s.t.shape[0])
    H = property(lambda s: s.t.shape[1])
    W = property(lambda s: s.t.shape[2])
    cs = property(lambda s: s.t.shape[3])
    dtype = property(lambda s: s.t.dtype)

    @property
    def ptr(self):
        return self.t.data_ptr() + self.c0 * self.t.element_size()

    def sl(self, a, b):
        assert 0 <= a < b <= self.C
        return View(self.t, self.c0 + a, b - a)

    def dense(self):
        """torch tensor [N,H,W,C] (a strided alias, for tests / debugging)."""
        return self.t[..., self.c0:self.c0 + self.C]


def _code(dtype):
    if dtype == torch.float32:
        return F32
    if dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"unsupported activation dtype {dtype}")


class PackedConv:
    """Device-side packed weights of one nn.Conv2d (fp32 [tap][Cin_p][Cout] + optional UMMA bf16)."""
    __slots__ = ("w", "bias", "K", "Cin", "Cout", "w_umma", "bias_tc", "ps", "cin_tc", "fresh")

    def __init__(self, w, bias, K, Cin, Cout, w_umma=None):
        self.w, self.bias, self.K, self.Cin, self.Cout, self.w_umma = w, bias, K, Cin, Cout, w_umma
        self.bias_tc, self.ps, self.cin_tc = bias, False, Cin
        self.fresh = True      # packed by a kernel that may still be running: first launch without PDL


PRECISIONS = ("fp32", "bf16", "parity")


class CudaBackend:
    """precision: 'fp32' (exact mode, fp32 FMA convs on CUDA cores), 'bf16' (tcgen05 tensor-core convs,
    bf16 activations) or 'parity' (tcgen05 convs on split-bf16 operands: fp32 activations in HBM, every
    conv input split into three bf16 planes and every weight into three bf16 parts, six MMAs per
    product -> fp32-class accuracy on the tensor pipe; the graph is the exact mode's)."""

    def __init__(self, device, precision="fp32"):
        self.L = _lib.lib()
        self.device = torch.device(device)
        if self.device.type != "cuda" or not torch.cuda.is_available():
            raise RuntimeError("vcm_ts_b200 needs a CUDA device (B200, sm_100a); there is no CPU path")
        with torch.cuda.device(self.device):
            if not self.L.vcm_device_ok():
                raise RuntimeError("libvcmts_b200.so: no compute-capability-10.x device visible")
        if precision not in PRECISIONS:
            raise ValueError("precision must be one of %s" % (PRECISIONS,))
        self.precision = precision
        self.act_dtype = torch.bfloat16 if precision == "bf16" else torch.float32
        # bf16 mode: producers also store the activated twin a following ResBlock reads, so the
        # tensor-core engine never needs a pre-activation on TMA-loaded tiles
        self.use_twin = precision == "bf16"
        self.launches = 0          # kernels launched (graph replays add their kernel count)
        self.graph_replays = 0
        # programmatic dependent launch of the tensor-core convs (VCM_CONV_PDL in include/vcmts_b200.h): the
        # prologue + weight loads of conv n+1 overlap the tail of conv n.  Opt-in (VCM_B200_PDL=1): measured, it
        # changes neither the 8-sequence throughput (other streams' kernels already fill the SMs a finishing
        # conv leaves) nor the single-sequence rate (43.4 / 28.9 frames/s with, 43.9 / 28.9 without).
        import os
        self.pdl = os.environ.get("VCM_B200_PDL", "0") == "1"
        # SELayer means out of the producing conv's epilogue (bf16 tensor-core mode; VCM_B200_FUSED_SE=0: off)
        self.fused_se_sums = precision == "bf16" and os.environ.get("VCM_B200_FUSED_SE", "1") == "1"

    # -- plumbing -------------------------------------------------------------------------
    @property
    def stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def new(self, N, H, W, C, dtype=None, zero=False):
        f = torch.zeros if zero else torch.empty
        return View(f((N, H, W, C), dtype=dtype or self.act_dtype, device=self.device))

    def empty(self, shape, dtype):
        return torch.empty(shape, dtype=dtype, device=self.device)

    def pack_conv(self, weight, bias, in_map=None, cin_p=None, ps=False):
        """ps: the conv is stored through PixelShuffle(2).  The tensor-core packing then orders the
        GEMM columns position-major (c' = pos * Cout/4 + oc <- reference channel 4*oc + pos) so the
        epilogue's channel blocks are contiguous output channels (tc_epilogue.cuh)."""
        Cout, Cin, K, _ = weight.shape
        w = weight.detach().to(self.device, torch.float32).contiguous()
        cin_p = cin_p or (len(in_map) if in_map is not None else Cin)
        m = None
        if in_map is not None:
            m = torch.tensor(list(in_map), dtype=torch.int32, device=self.device)
        out = torch.empty((K * K, cin_p, Cout), dtype=torch.float32, device=self.device)
        check(self.L.vcm_pack_conv_weight(w.data_ptr(), out.data_ptr(), Cout, Cin, K, cin_p,
                                          m.data_ptr() if m is not None else None, self.stream),
              "vcm_pack_conv_weight")
        pc = PackedConv(out, bias.detach().to(self.device, torch.float32).contiguous(), K, cin_p, Cout)
        if self.precision == "parity":
            # three bf16 parts of the fp32 weights, input channels padded to 16 (zero rows)
            cp = (cin_p + 15) // 16 * 16
            nbytes = self.L.vcm_conv2d_tc_weight_bytes(Cout, cp, K)
            if nbytes > 0:
                full = list(in_map) if in_map is not None else list(range(Cin))
                full = full + [-1] * (cp - len(full))
                m3 = torch.tensor(full, dtype=torch.int32, device=self.device)
                wu = torch.empty((3 * nbytes,), dtype=torch.uint8, device=self.device)
                if ps:
                    assert Cout % 4 == 0
                    perm = torch.arange(Cout, device=self.device).reshape(Cout // 4, 4).t().reshape(-1)
                    w = w[perm].contiguous()
                    pc.bias_tc = pc.bias[perm].contiguous()
                    pc.ps = True
                check(self.L.vcm_pack_conv_weight_umma_split3(w.data_ptr(), wu.data_ptr(), Cout, Cin, K, cp,
                                                              m3.data_ptr(), self.stream),
                      "vcm_pack_conv_weight_umma_split3")
                pc.w_umma, pc.cin_tc = wu, cp
        if self.precision == "bf16" and cin_p % 16 == 0:
            nbytes = self.L.vcm_conv2d_tc_weight_bytes(Cout, cin_p, K)
            if nbytes > 0:
                wu = torch.empty((nbytes,), dtype=torch.uint8, device=self.device)
                if ps:
                    assert Cout % 4 == 0
                    perm = torch.arange(Cout, device=self.device).reshape(Cout // 4, 4).t().reshape(-1)
                    w = w[perm].contiguous()
                    pc.bias_tc = pc.bias[perm].contiguous()
                    pc.ps = True
                check(self.L.vcm_pack_conv_weight_umma(w.data_ptr(), wu.data_ptr(), Cout, Cin, K, cin_p,
                                                       m.data_ptr() if m is not None else None,
                                                       self.stream), "vcm_pack_conv_weight_umma")
                pc.w_umma = wu
        self.launches += 1
        return pc

    # -- convolution ------------------------------------------------------------------------
    def conv(self, x, pc, out=None, stride=1, pre=None, post=None, res=None, res_scale=None,
             res2=None, ps=False, out_dtype=None, slope2=None, out2=None, se_sums=False):
        """Fused conv.  With slope2 set, also stores leaky_relu(y, slope2) to `out2` and returns
        (y, y_act).  se_sums=True (bf16 tensor-core mode): returns (y, sums) where sums is a float
        [rows, Cout] tensor of channel sums of y for `se_scale_from_sums`, or None where the engine cannot
        produce them for this launch (then use `se_scale(y, ...)`)."""
        K = pc.K
        assert x.C == pc.Cin, f"conv input has {x.C} channels, weights expect {pc.Cin}"
        pad = K // 2
        Ho = (x.H + 2 * pad - K) // stride + 1
        Wo = (x.W + 2 * pad - K) // stride + 1
        if ps:
            oh, ow, oc = 2 * Ho, 2 * Wo, pc.Cout // 4
        else:
            oh, ow, oc = Ho, Wo, pc.Cout
        if out is None:
            out = self.new(x.N, oh, ow, oc, dtype=out_dtype)
        assert (out.N, out.H, out.W, out.C) == (x.N, oh, ow, oc), "conv output view mismatch"
        flags = 0
        if self.pdl and not pc.fresh:
            flags |= _lib.CONV_PDL
        pc.fresh = False
        if pre is not None:
            flags |= _lib.CONV_PRE_ACT
        if post is not None:
            flags |= _lib.CONV_POST_ACT
        if ps:
            flags |= _lib.CONV_PIXEL_SHUFFLE
        if res is not None:
            flags |= _lib.CONV_RES_SCALE if res_scale is not None else _lib.CONV_RES
            assert (res.N, res.H, res.W, res.C) == (out.N, out.H, out.W, out.C) and res.dtype == out.dtype
        if res2 is not None:
            flags |= _lib.CONV_RES2
            assert (res2.N, res2.H, res2.W, res2.C) == (out.N, out.H, out.W, out.C) and res2.dtype == out.dtype
        if slope2 is not None:
            flags |= _lib.CONV_OUT2
            if out2 is None:
                out2 = self.new(out.N, out.H, out.W, out.C, dtype=out.dtype)
            assert (out2.N, out2.H, out2.W, out2.C) == (out.N, out.H, out.W, out.C) and out2.dtype == out.dtype
        d = ConvDesc(x.N, x.H, x.W, pc.Cin, pc.Cout, K, stride, x.cs, out.cs,
                     res.cs if res is not None else 0, res2.cs if res2 is not None else 0,
                     out2.cs if slope2 is not None else 0, flags,
                     float(pre or 0.0), float(post or 0.0), float(slope2 or 0.0),
                     _code(x.dtype), _code(out.dtype))
        args = [ctypes.byref(d), x.ptr, None, pc.bias.data_ptr(), res.ptr if res is not None else None,
                res_scale.data_ptr() if res_scale is not None else None,
                res2.ptr if res2 is not None else None, out.ptr,
                out2.ptr if slope2 is not None else None, self.stream]
        if self.precision == "parity" and pc.w_umma is not None and x.dtype == torch.float32:
            # split the fp32 input (pre-activation applied on the way) into three bf16 planes and run
            # the tensor-core engine on them; the epilogue works in fp32 as in the exact mode
            cp = pc.cin_tc
            d3 = ConvDesc(x.N, x.H, x.W, cp, pc.Cout, K, stride, 3 * cp, out.cs,
                          res.cs if res is not None else 0, res2.cs if res2 is not None else 0,
                          out2.cs if slope2 is not None else 0,
                          (flags & ~_lib.CONV_PRE_ACT) | _lib.CONV_SPLIT3,
                          0.0, float(post or 0.0), float(slope2 or 0.0), BF16, _code(out.dtype), cp)
            if self.L.vcm_conv2d_tc_supported(ctypes.byref(d3)) == 1:
                if pc.ps != bool(ps):
                    raise RuntimeError("conv weights were packed for ps=%s but the conv is launched with ps=%s"
                                       % (pc.ps, bool(ps)))
                xs = torch.empty((x.N, x.H, x.W, 3 * cp), dtype=torch.bfloat16, device=self.device)
                check(self.L.vcm_split_bf16x3(x.ptr, xs.data_ptr(), x.N * x.H * x.W, x.C, cp, x.cs, 3 * cp, cp,
                                              1 if pre is not None else 0, float(pre or 0.0), self.stream),
                      "vcm_split_bf16x3")
                args[0], args[1] = ctypes.byref(d3), xs.data_ptr()
                args[2], args[3] = pc.w_umma.data_ptr(), pc.bias_tc.data_ptr()
                check(self.L.vcm_conv2d_tc_fwd(*args), "vcm_conv2d_tc_fwd")
                self.launches += 2
                return (out, out2) if slope2 is not None else out
        sums = None
        if self.precision == "bf16" and pc.w_umma is not None and \
                self.L.vcm_conv2d_tc_supported(ctypes.byref(d)) == 1:
            if pc.ps != bool(ps):
                raise RuntimeError("conv weights were packed for ps=%s but the conv is launched with ps=%s"
                                   % (pc.ps, bool(ps)))
            args[2], args[3] = pc.w_umma.data_ptr(), pc.bias_tc.data_ptr()
            if se_sums and slope2 is None and (x.ptr % 16 == 0) and (out.ptr % 16 == 0) and \
                    (res is None or res.ptr % 16 == 0):
                rows = self.L.vcm_conv2d_se_sum_rows(ctypes.byref(d))
                if rows > 0:
                    sums = self.empty((rows, pc.Cout), torch.float32)
                    d.flags |= _lib.CONV_SE_SUM
                    args[8] = sums.data_ptr()
            rc = self.L.vcm_conv2d_tc_fwd(*args)
            if rc == 2 and not pc.ps:     # e.g. a view whose base is not 16-byte aligned: fp32-FMA engine
                args[2], args[3] = pc.w.data_ptr(), pc.bias.data_ptr()
                rc = self.L.vcm_conv2d_fwd(*args)
            check(rc, "vcm_conv2d_tc_fwd")
        else:
            args[2] = pc.w.data_ptr()
            check(self.L.vcm_conv2d_fwd(*args), "vcm_conv2d_fwd")
        self.launches += 1
        if se_sums:
            return out, sums
        return (out, out2) if slope2 is not None else out

    # -- layout / elementwise ---------------------------------------------------------------
    def from_nchw(self, t, out=None, dtype=None):
        t = t.detach().to(self.device, torch.float32).contiguous()
        N, C, H, W = t.shape
        if out is None:
            out = self.new(N, H, W, C, dtype=dtype)
        assert (out.N, out.H, out.W, out.C) == (N, H, W, C)
        check(self.L.vcm_nchw_to_nhwc(t.data_ptr(), out.ptr, N, C, H, W, out.cs, _code(out.dtype),
                                      self.stream), "vcm_nchw_to_nhwc")
        self.launches += 1
        return out

    def to_nchw(self, v, clamp=None, out=None):
        if out is None:
            out = torch.empty((v.N, v.C, v.H, v.W), dtype=torch.float32, device=self.device)
        assert tuple(out.shape) == (v.N, v.C, v.H, v.W) and out.dtype == torch.float32 and out.is_contiguous()
        lo, hi = clamp if clamp is not None else (0.0, 0.0)
        check(self.L.vcm_nhwc_to_nchw(v.ptr, out.data_ptr(), v.N, v.C, v.H, v.W, v.cs, _code(v.dtype),
                                      1 if clamp is not None else 0, lo, hi, self.stream),
              "vcm_nhwc_to_nchw")
        self.launches += 1
        return out

    def copy(self, src, dst):
        assert (src.N, src.H, src.W, src.C) == (dst.N, dst.H, dst.W, dst.C)
        check(self.L.vcm_copy_channels(src.ptr, dst.ptr, src.N * src.H * src.W, src.C, src.cs, dst.cs,
                                       _code(src.dtype), _code(dst.dtype), self.stream), "vcm_copy_channels")
        self.launches += 1
        return dst

    def scale(self, src, vec, op, out=None):
        """op: 'div' | 'mul' by vec [N, C] fp32."""
        if out is None:
            out = self.new(src.N, src.H, src.W, src.C, dtype=src.dtype)
        assert vec.shape == (src.N, src.C) and vec.dtype == torch.float32 and vec.is_contiguous()
        check(self.L.vcm_scale_channels(src.ptr, out.ptr, vec.data_ptr(), src.N, src.H * src.W, src.C,
                                        src.cs, out.cs, 0 if op == "div" else 1, _code(src.dtype),
                                        self.stream), "vcm_scale_channels")
        self.launches += 1
        return out

    def maxpool(self, x, out=None):
        if out is None:
            out = self.new(x.N, x.H // 2, x.W // 2, x.C, dtype=x.dtype)
        check(self.L.vcm_maxpool2(x.ptr, out.ptr, x.N, x.H, x.W, x.C, x.cs, out.cs, _code(x.dtype),
                                  self.stream), "vcm_maxpool2")
        self.launches += 1
        return out

    def se_scale(self, x, w1, w2):
        """SELayer scale [N, C] fp32 of view x; w1 [C/16, C], w2 [C, C/16] fp32 device tensors."""
        hw = x.H * x.W
        nblk = self.L.vcm_se_partial_count(hw)
        partial = self.empty((x.N * nblk * x.C,), torch.float32)
        scale = self.empty((x.N, x.C), torch.float32)
        check(self.L.vcm_se_scale(x.ptr, partial.data_ptr(), w1.data_ptr(), w2.data_ptr(),
                                  scale.data_ptr(), x.N, hw, x.C, w1.shape[0], x.cs, _code(x.dtype),
                                  self.stream), "vcm_se_scale")
        self.launches += 2
        return scale

    def se_scale_from_sums(self, sums, x, w1, w2):
        """SELayer scale [1, C] from the channel sums a `conv(..., se_sums=True)` wrote for its output x."""
        assert x.N == 1 and sums.shape[1] == x.C
        scale = self.empty((1, x.C), torch.float32)
        check(self.L.vcm_se_scale_from_sums(sums.data_ptr(), sums.shape[0], w1.data_ptr(), w2.data_ptr(),
                                            scale.data_ptr(), x.H * x.W, x.C, w1.shape[0], self.stream),
              "vcm_se_scale_from_sums")
        self.launches += 1
        return scale

    # -- motion glue (fp32) -----------------------------------------------------------------
    def avgpool(self, x, out=None):
        assert x.dtype == torch.float32
        if out is None:
            out = self.new(x.N, x.H // 2, x.W // 2, x.C, dtype=torch.float32)
        check(self.L.vcm_avgpool2(x.ptr, out.ptr, x.N, x.H, x.W, x.C, x.cs, out.cs, self.stream),
              "vcm_avgpool2")
        self.launches += 1
        return out

    def flow_up2(self, flow, out=None, gain=2.0):
        assert flow.dtype == torch.float32 and flow.C == 2
        if out is None:
            out = self.new(flow.N, flow.H * 2, flow.W * 2, 2, dtype=torch.float32)
        check(self.L.vcm_flow_up2(flow.ptr, out.ptr, flow.N, flow.H, flow.W, flow.cs, out.cs, gain,
                                  self.stream), "vcm_flow_up2")
        self.launches += 1
        return out

    def flow_down2(self, flow, out=None, gain=0.5):
        assert flow.dtype == torch.float32 and flow.C == 2
        if out is None:
            out = self.new(flow.N, flow.H // 2, flow.W // 2, 2, dtype=torch.float32)
        check(self.L.vcm_flow_down2(flow.ptr, out.ptr, flow.N, flow.H, flow.W, flow.cs, out.cs, gain,
                                    self.stream), "vcm_flow_down2")
        self.launches += 1
        return out

    def spynet_pyramid(self, x_nchw, ref_nchw, lvl, im2):
        """Both frames -> four pyramid levels each, one launch.  lvl: four [N,h,w,8] fp32 views (channels 0..2
        receive im1), im2: four dense [N,h,w,3] fp32 views."""
        x = x_nchw.detach().to(self.device, torch.float32).contiguous()
        r = ref_nchw.detach().to(self.device, torch.float32).contiguous()
        N, _, H, W = x.shape
        assert all(v.cs == 8 and v.c0 == 0 for v in lvl) and all(v.cs == 3 and v.c0 == 0 for v in im2)
        pl = (ctypes.c_void_p * 4)(*[v.ptr for v in lvl])
        pi = (ctypes.c_void_p * 4)(*[v.ptr for v in im2])
        check(self.L.vcm_spynet_pyramid(x.data_ptr(), r.data_ptr(), pl, pi, N, H, W, self.stream),
              "vcm_spynet_pyramid")
        self.launches += 1

    def spynet_level_input(self, flow_prev, im2, lvl, buf16=None):
        assert flow_prev.dtype == torch.float32 and flow_prev.C == 2 and lvl.cs == 8 and im2.cs == 3
        check(self.L.vcm_spynet_level_input(flow_prev.ptr, im2.ptr, lvl.ptr, buf16.ptr if buf16 is not None else None,
                                            lvl.N, lvl.H, lvl.W, flow_prev.cs, self.stream),
              "vcm_spynet_level_input")
        self.launches += 1

    def warp(self, feat, flow, out=None):
        assert flow.dtype == torch.float32 and flow.C == 2
        assert (flow.N, flow.H, flow.W) == (feat.N, feat.H, feat.W)
        if out is None:
            out = self.new(feat.N, feat.H, feat.W, feat.C, dtype=feat.dtype)
        assert out.dtype == feat.dtype
        check(self.L.vcm_flow_warp(feat.ptr, flow.ptr, out.ptr, feat.N, feat.H, feat.W, feat.C, feat.cs,
                                   flow.cs, out.cs, _code(feat.dtype), self.stream), "vcm_flow_warp")
        self.launches += 1
        return out

    # -- training path: backward kernels (training.py) ----------------------------------------
    def wgrad(self, x, dy, K, stride, Cin, Cout):
        """dW (OIHW fp32) of a conv with input view x and output-gradient view dy."""
        assert x.C == Cin and dy.C == Cout and x.N == dy.N
        d = ConvDesc(x.N, x.H, x.W, Cin, Cout, K, stride, x.cs, dy.cs, 0, 0, 0, 0, 0.0, 0.0, 0.0,
                     _code(x.dtype), _code(dy.dtype), 0)
        n = self.L.vcm_conv2d_wgrad_scratch_floats(ctypes.byref(d))
        scratch = self.empty((max(int(n), 1),), torch.float32)
        dw = self.empty((Cout, Cin, K, K), torch.float32)
        check(self.L.vcm_conv2d_wgrad(ctypes.byref(d), x.ptr, dy.ptr, dw.data_ptr(), scratch.data_ptr(),
                                      self.stream), "vcm_conv2d_wgrad")
        self.launches += 2
        return dw

    def warp_bwd(self, feat, flow, gout):
        """Backward of warp(): (gfeat fp32 [N,H,W,C], gflow fp32 [N,H,W,2]), dense NHWC tensors."""
        assert flow.dtype == torch.float32 and flow.C == 2 and gout.dtype == feat.dtype
        gfeat = torch.zeros((feat.N, feat.H, feat.W, feat.C), dtype=torch.float32, device=self.device)
        gflow = torch.zeros((feat.N, feat.H, feat.W, 2), dtype=torch.float32, device=self.device)
        check(self.L.vcm_flow_warp_bwd(feat.ptr, flow.ptr, gout.ptr, gfeat.data_ptr(), gflow.data_ptr(), feat.N,
                                       feat.H, feat.W, feat.C, feat.cs, flow.cs, gout.cs, _code(feat.dtype),
                                       self.stream), "vcm_flow_warp_bwd")
        self.launches += 1
        return gfeat, gflow

    # -- entropy epilogues ------------------------------------------------------------------
    def _dp_desc(self, fusion, C, mode, table, y=None, prior=None, params=None, yhat=None):
        return DpDesc(fusion.N, fusion.H, fusion.W, C, y.cs if y is not None else 0, fusion.cs,
                      prior.cs if prior is not None else 0, params.cs if params is not None else 0,
                      yhat.cs if yhat is not None else 0, _code(fusion.dtype), mode,
                      table.log_scale_min, table.log_scale_step)

    def dp_pass0(self, mode, table, fusion, params, C, y=None, yq=None, sym=None, idx=None,
                 y_q_full=None, scales_full=None):
        d = self._dp_desc(fusion, C, mode, table, y=y, params=params)
        p = lambda t: t.data_ptr() if t is not None else None  # noqa: E731
        check(self.L.vcm_dual_prior_pass0(ctypes.byref(d), y.ptr if y is not None else None, fusion.ptr,
                                          params.ptr, p(yq), p(sym), p(idx), p(y_q_full),
                                          p(scales_full), self.stream), "vcm_dual_prior_pass0")
        self.launches += 1

    def dp_pass1(self, mode, table, fusion, prior, params, C, q_outer=None, y_hat=None, yq=None,
                 sym=None, idx=None, y_q_full=None, scales_full=None):
        d = self._dp_desc(fusion, C, mode, table, prior=prior, params=params, yhat=y_hat)
        p = lambda t: t.data_ptr() if t is not None else None  # noqa: E731
        check(self.L.vcm_dual_prior_pass1(ctypes.byref(d), p(yq), fusion.ptr, prior.ptr, params.ptr,
                                          p(q_outer), y_hat.ptr if y_hat is not None else None, p(sym),
                                          p(idx), p(y_q_full), p(scales_full), self.stream),
              "vcm_dual_prior_pass1")
        self.launches += 1

    def round_latent(self, z, z_hat=None, sym=None):
        if z_hat is None:
            z_hat = self.new(z.N, z.H, z.W, z.C, dtype=z.dtype)
        check(self.L.vcm_round_latent(z.ptr, z_hat.ptr, sym.data_ptr() if sym is not None else None,
                                      z.N, z.H, z.W, z.C, z.cs, z_hat.cs, _code(z.dtype), self.stream),
              "vcm_round_latent")
        self.launches += 1
        return z_hat

    def symbols_to_latent(self, sym, N, H, W, C, out=None):
        if out is None:
            out = self.new(N, H, W, C)
        check(self.L.vcm_symbols_to_latent(sym.data_ptr(), out.ptr, N, H, W, C, out.cs,
                                           _code(out.dtype), self.stream), "vcm_symbols_to_latent")
        self.launches += 1
        return out

    def latent_bits(self, y_q_full, scales_full, N, dist):
        per = y_q_full.numel() // N
        nblk = self.L.vcm_bits_partial_count(per)
        partial = self.empty((N * nblk,), torch.float32)
        out = self.empty((N,), torch.float32)
        check(self.L.vcm_latent_bits(y_q_full.data_ptr(), scales_full.data_ptr(), partial.data_ptr(),
                                     out.data_ptr(), N, per, 0 if dist == "laplace" else 1, self.stream),
              "vcm_latent_bits")
        self.launches += 2
        return out

    def factorized_bits(self, z_hat, fparams):
        hw = z_hat.H * z_hat.W
        nblk = self.L.vcm_bits_partial_count(hw * z_hat.C)
        partial = self.empty((z_hat.N * nblk,), torch.float32)
        out = self.empty((z_hat.N,), torch.float32)
        check(self.L.vcm_factorized_bits(z_hat.ptr, fparams.data_ptr(), partial.data_ptr(),
                                         out.data_ptr(), z_hat.N, hw, z_hat.C, z_hat.cs,
                                         _code(z_hat.dtype), self.stream), "vcm_factorized_bits")
        self.launches += 2
        return out

    def sq_err_sum(self, a, b):
        a = a.detach().to(self.device, torch.float32).contiguous()
        b = b.detach().to(self.device, torch.float32).contiguous()
        N = a.shape[0]
        per = a.numel() // N
        nblk = self.L.vcm_bits_partial_count(per)
        partial = self.empty((N * nblk,), torch.float32)
        out = self.empty((N,), torch.float32)
        check(self.L.vcm_sq_err_sum(a.data_ptr(), b.data_ptr(), partial.data_ptr(), out.data_ptr(), N,
                                    per, self.stream), "vcm_sq_err_sum")
        self.launches += 2
        return out
