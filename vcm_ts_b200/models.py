"""Drop-in `DMC` and `IntraNoAR` modules: the reference's operator API
(DCVC_HEM/src/models/video_model.py:131-596, image_model.py:16-200, common_model.py:14-217) on top
of the sm_100a kernels of libvcmts_b200.so.

Same constructor arguments, state-dict keys, method signatures and result dictionaries (including
the reference's `"dbp"` key in `DMC.compress`), so `video_coder.py`, `benchmark_videos_decoding.py`
and `DCVC_HEM/test_video.py` can import these classes instead of the reference's.  Callers hand in
and get back ordinary NCHW fp32 tensors; internally everything is NHWC views (backend.py) and every
nn.Conv2d / activation / PixelShuffle / concat is a fused device op (nets.py).

`precision` (constructor kwarg or VCM_B200_PRECISION env): "fp32" = exact mode (fp32 FMA convs,
meant to reproduce the reference's quantised symbols), "bf16" = tcgen05 tensor-core mode.
There is no CPU path: calling a compute method without a B200 raises.
"""
import math
import os
import time

import numpy as np
import torch
from torch import nn

from . import arch
from .backend import CudaBackend, View
from .graphs import ProgramCache, drive_eager
from .entropy import CdfTables, RansCoder, ScaleIndexTable, factorized_tables
from .nets import Graph
from .stream_helper import (decode_i, decode_p, encode_i, encode_p, filesize, get_downsampled_shape,
                            get_rounded_q, get_state_dict, pack_i, pack_p, unpack_i, unpack_p)

F32 = torch.float32


class _Node(nn.Module):
    """Anonymous container so parameter names match the reference's module tree."""


class _Staging:
    """Pinned host buffers for the symbol / scale-index hand-off with the host entropy coder
    (entropy_models.py:183-195, 270-281 move int32 tensors with blocking pageable copies).  Device -> host:
    every tensor of a hand-off is copied asynchronously into its pinned buffer and ONE event is waited for;
    host -> device copies of decoded symbols are asynchronous from pinned memory (the next launch on the
    stream orders after them).  Buffers are reused round-robin (`depth` generations per tag), so a consumer
    on another thread (pipeline.py) may hold the previous generations while the next frame is staged."""

    def __init__(self, device, depth=4):
        self.device, self.depth = device, depth
        self.bufs, self.gen = {}, {}
        self.pin = device.type == "cuda"
        # blocking event waits: the waiting thread sleeps instead of spinning, which matters when 8 ranks x 8
        # coding threads share 32 logical CPUs with the rANS coder (VCM_B200_BLOCKING_SYNC=0: spin)
        self.blocking = os.environ.get("VCM_B200_BLOCKING_SYNC", "1") == "1"

    def _buf(self, key, shape, dtype):
        b = self.bufs.get(key)
        if b is None:
            b = torch.empty(shape, dtype=dtype, pin_memory=self.pin)
            self.bufs[key] = b
        return b

    def to_host(self, tag, *tensors):
        g = self.gen[tag] = (self.gen.get(tag, -1) + 1) % self.depth
        outs = []
        for i, t in enumerate(tensors):
            b = self._buf((tag, g, i, tuple(t.shape), t.dtype), tuple(t.shape), t.dtype)
            b.copy_(t, non_blocking=True)
            outs.append(b)
        if self.pin:
            ev = torch.cuda.Event(blocking=self.blocking)
            ev.record(torch.cuda.current_stream(self.device))
            ev.synchronize()
        return outs

    def host_buf(self, tag, shape, dtype):
        return self._buf((tag, tuple(shape), dtype), tuple(shape), dtype)


class _Runtime:
    """Per-process, per-device state that must not be pickled: backend, packed weights, coder."""

    def __init__(self, backend):
        self.B = backend
        self.staging = _Staging(backend.device)
        self.packed = {}
        self.lin = {}
        self.persist = {}
        self.coder = None
        self.tables = {}
        self.qcache = {}
        self.statics = {}
        self.static_ids = {}
        self.programs = ProgramCache(backend)


class CompressionModel(nn.Module):
    """common_model.py:14-49,75-80 twin: parameter tree from a layer table, lazy runtime, update()."""

    def __init__(self, spec, y_distribution, precision=None):
        super().__init__()
        self.y_distribution = y_distribution
        self._precision = precision or os.environ.get("VCM_B200_PRECISION", "fp32")
        self._rt = None
        self._graphs = os.environ.get("VCM_B200_CUDA_GRAPHS", "0") == "1"
        self._backend_factory = None          # tests inject an emulation backend here
        self.entropy_coder = None             # reference attribute; set by update()
        self._scale_table = ScaleIndexTable(y_distribution)
        self._fparam_names = []
        for kind, name, meta in spec.items:
            if kind == arch.CONV:
                cin, cout, k = meta
                node = self._node(name)
                w = torch.empty(cout, cin, k, k)
                nn.init.xavier_normal_(w, math.sqrt(2))          # common_model.py:31-36
                node.weight = nn.Parameter(w)
                node.bias = nn.Parameter(torch.full((cout,), 0.01))
            elif kind == arch.LINEAR:
                cin, cout = meta
                node = self._node(name)
                w = torch.empty(cout, cin)
                nn.init.kaiming_uniform_(w, a=math.sqrt(5))      # nn.Linear default
                node.weight = nn.Parameter(w)
            elif kind == arch.BITPARM:
                (ch,) = meta
                self._fparam_names.append(name)
                for f in ("f1", "f2", "f3", "f4"):
                    node = self._node(f"{name}.{f}")
                    node.h = nn.Parameter(torch.nn.init.normal_(torch.empty(1, ch, 1, 1), 0, 0.01))
                    node.b = nn.Parameter(torch.nn.init.normal_(torch.empty(1, ch, 1, 1), 0, 0.01))
                    if f != "f4":
                        node.a = nn.Parameter(torch.nn.init.normal_(torch.empty(1, ch, 1, 1), 0, 0.01))
            else:
                self.register_parameter(name, nn.Parameter(torch.ones(meta)))

    def _node(self, dotted):
        m = self
        for part in dotted.split("."):
            if not hasattr(m, part):
                m.add_module(part, _Node())
            m = getattr(m, part)
        return m

    # ---- pickling (benchmark_videos_decoding.py:248-254 pickles modules into a spawn worker)
    def __getstate__(self):
        st = dict(self.__dict__)
        st["_rt"] = None
        st["_backend_factory"] = None
        st["entropy_coder"] = None
        return st

    # ---- runtime ------------------------------------------------------------------------
    def set_precision(self, precision):
        if precision != self._precision:
            self._precision, self._rt = precision, None
        return self

    def load_state_dict(self, *args, **kwargs):
        out = super().load_state_dict(*args, **kwargs)
        self.reset_cuda_graphs()        # captured graphs point at the previously packed weights
        return out

    def set_cuda_graphs(self, enabled=True):
        """Opt-in: capture the per-frame launch sequences of compress / decompress into CUDA graphs
        (graphs.py).  In this mode the returned dpb is a set of STATIC buffers owned by the model
        (ref_frame an NCHW fp32 tensor as usual, ref_feature / ref_y / ref_mv_y internal NHWC views):
        the next call overwrites them in place, so use them (or copy them) before coding the next
        frame.  Symbols, bitstreams and reconstructions are identical to the eager mode.  One model
        instance holds one chain per side and must be driven from one thread / stream at a time;
        `load_state_dict` drops the captured graphs (in-place edits of parameters behind the model's
        back are not detected while replaying -- call `reset_cuda_graphs()` after them)."""
        self._graphs = bool(enabled)
        return self

    def reset_cuda_graphs(self):
        """Drop the captured graphs; the next frames run eagerly once and are captured again."""
        if self._rt is not None:
            self._rt.programs.invalidate()
            self._rt.qcache.clear()

    @property
    def device(self):
        return next(self.parameters()).device

    def _runtime(self):
        dev = self.device
        if self._rt is None or self._rt.B.device != dev:
            if self._backend_factory is not None:
                backend = self._backend_factory(dev, self._precision)
            else:
                backend = CudaBackend(dev, self._precision)      # raises without a B200
            rt = _Runtime(backend)
            rt.graph = Graph(backend, lambda n, m=None, ps=False: self._packed(rt, n, m, ps),
                             lambda n: self._lin(rt, n))
            self._rt = rt
        return self._rt

    def _packed(self, rt, name, in_map, ps=False):
        w = self.get_parameter(name + ".weight")
        b = self.get_parameter(name + ".bias")
        key = (name, tuple(in_map) if in_map is not None else None, ps)
        tag = (w._version, w.data_ptr(), b._version, b.data_ptr())
        hit = rt.packed.get(key)
        if hit is None or hit[0] != tag:
            hit = (tag, rt.B.pack_conv(w, b, in_map, ps=ps))
            rt.packed[key] = hit
        return hit[1]

    def _lin(self, rt, name):
        w = self.get_parameter(name)
        tag = (w._version, w.data_ptr())
        hit = rt.lin.get(name)
        if hit is None or hit[0] != tag:
            hit = (tag, w.detach().to(rt.B.device, F32).contiguous())
            rt.lin[name] = hit
        return hit[1]

    def _persist(self, rt, key, N, H, W, C, dtype):
        """Zero-initialised buffer that survives across frames (pad channels stay 0)."""
        k = (key, N, H, W, C, dtype)
        v = rt.persist.get(k)
        if v is None:
            v = rt.B.new(N, H, W, C, dtype=dtype, zero=True)
            rt.persist[k] = v
        return v

    def _fparams(self, rt, name):
        """[11, C] fp32: f1.h f1.b f1.a f2.h f2.b f2.a f3.h f3.b f3.a f4.h f4.b (vcm_factorized_bits)."""
        rows = []
        for f in ("f1", "f2", "f3", "f4"):
            for p in ("h", "b", "a"):
                if f == "f4" and p == "a":
                    continue
                rows.append(self.get_parameter(f"{name}.{f}.{p}").detach().reshape(-1))
        return torch.stack(rows).to(rt.B.device, F32).contiguous()

    def _fdict(self, name):
        out = {}
        for f in ("f1", "f2", "f3", "f4"):
            for p in ("h", "b", "a"):
                if f == "f4" and p == "a":
                    continue
                out[f"{f}.{p}"] = self.get_parameter(f"{name}.{f}.{p}")
        return out

    def update(self, force=False):
        """common_model.py:75-80: build the CDF tables and the native coder (CPU, once)."""
        rt = self._runtime()
        if rt.coder is not None and not force:
            return
        rt.coder = RansCoder()
        self.entropy_coder = rt.coder
        rt.tables = {"y": CdfTables(*self._scale_table.update())}
        for name in self._fparam_names:
            ch = self.get_parameter(name + ".f1.h").shape[1]
            rt.tables[name] = CdfTables(*factorized_tables(self._fdict(name), ch))

    def _need_coder(self):
        rt = self._runtime()
        if rt.coder is None:
            raise RuntimeError("call update(force=True) before compress/decompress (as in the reference)")
        return rt

    # ---- q helpers ------------------------------------------------------------------------
    def _curr_q(self, rt, basic, q_scale, N):
        """LowerBound(q_basic, .5) * q_scale -> contiguous [N, C] fp32 on the device."""
        if not torch.is_tensor(q_scale):
            key = (id(basic), basic._version, basic.data_ptr(), float(q_scale), N)
            hit = rt.qcache.get(key)
            if hit is None:
                hit = rt.qcache[key] = self._curr_q(rt, basic, torch.tensor(float(q_scale)), N)
            return hit
        qb = torch.max(basic.detach(), torch.ones_like(basic) * 0.5)
        if torch.is_tensor(q_scale):
            q = qb * q_scale.detach().to(qb.device, qb.dtype)
        else:
            q = qb * float(q_scale)
        q = q.reshape(q.shape[0], -1)
        if q.shape[0] != N:
            q = q.expand(N, -1)
        return q.to(rt.B.device, F32).contiguous()

    # ---- NCHW fp32 tensors handed to callers keep their internal NHWC twin -------------------
    @staticmethod
    def _export(B, view, clamp=None):
        """NHWC view -> the NCHW fp32 tensor callers see.  The view rides along (only when the values
        are identical, i.e. no clamp) so that feeding the tensor back next frame skips the layout
        round trip; an in-place edit by the caller bumps `_version` and invalidates it."""
        t = B.to_nchw(view, clamp=clamp)
        if clamp is None:
            t._vcm_view = (view, t._version)
        return t

    @staticmethod
    def _import(B, t, dtype=None):
        """NCHW tensor from a caller -> NHWC view (cached twin if still valid).  Internal dpb entries
        (set_cuda_graphs) already are views."""
        if isinstance(t, View):
            return t
        hit = getattr(t, "_vcm_view", None)
        want = dtype or B.act_dtype
        if hit is not None and hit[1] == t._version and hit[0].dtype == want and \
                hit[0].t.device == B.device and (hit[0].N, hit[0].C, hit[0].H, hit[0].W) == tuple(t.shape):
            return hit[0]
        return B.from_nchw(t, dtype=dtype)

    @staticmethod
    def _dpb_into(B, entry, out):
        """Previous-frame latent (NCHW tensor from a caller, or an internal view) -> channel slice."""
        if isinstance(entry, View):
            return B.copy(entry, out)
        return B.from_nchw(entry, out=out)

    # ---- host <-> device symbol traffic -----------------------------------------------------
    def _to_host(self, *tensors):
        return self._runtime().staging.to_host("enc", *tensors)

    def _decode_plane_fn(self, rt):
        def decode_plane(idx_dev):
            (idx,) = rt.staging.to_host("dec_idx", idx_dev)
            out = rt.staging.host_buf("dec_sym", idx.shape, torch.int16)
            rt.coder.decode_planes(idx, rt.tables["y"], out)
            return out.to(rt.B.device, non_blocking=True)
        return decode_plane

    def _decode_z(self, rt, table_name, C, zs):
        out = rt.staging.host_buf("dec_z", (1, C, zs[0], zs[1]), torch.int16)
        rt.coder.decode_channels(C, zs[0] * zs[1], rt.tables[table_name], out)
        return rt.B.symbols_to_latent(out.to(rt.B.device, non_blocking=True), 1, zs[0], zs[1], C)


# ==========================================================================================
# P-frame model
# ==========================================================================================
class DMC(CompressionModel):
    """video_model.py:131-596."""

    def __init__(self, anchor_num=4, precision=None):
        super().__init__(arch.dmc_spec(anchor_num), "laplace", precision)
        self.DMC_version = "1.19"
        self.channel_mv, self.channel_N, self.channel_M = 64, 64, 96
        self.anchor_num = int(anchor_num)

    @staticmethod
    def get_q_scales_from_ckpt(ckpt_path):
        ckpt = get_state_dict(ckpt_path)
        return ckpt["y_q_scale"].reshape(-1), ckpt["mv_y_q_scale"].reshape(-1)

    def get_curr_mv_y_q(self, q_scale):
        return torch.max(self.mv_y_q_basic, torch.ones_like(self.mv_y_q_basic) * 0.5) * q_scale

    def get_curr_y_q(self, q_scale):
        return torch.max(self.y_q_basic, torch.ones_like(self.y_q_basic) * 0.5) * q_scale

    # ---- shared graph pieces ------------------------------------------------------------------
    def _ctx_buffer(self, rt, N, H, W):
        """[feat32 (32) | context1 (64) | x (3) | zero pad] : recon reads [0,96) in the reference's
        cat order (video_model.py:124,330), the contextual encoder reads [32, 32+Cin_p)."""
        pad = self._ENC_PAD if rt.B.precision == "bf16" else 1
        return self._persist(rt, "ctx", N, H, W, 99 + pad, rt.B.act_dtype)

    # bf16 mode: zero channels after x so that the stride-2 conv1 of the contextual encoder sees 96 = 3 x 32
    # input channels (context1 64 + x 3 + 29 zeros).  With 13 (80 = 5 x 16 channels) its per-tap engine moved
    # 45 boxes of 32-byte rows per tile and TMA's per-row cost made it 320 us at 1080p; 27 boxes of 64-byte rows
    # cost less even with the extra zero bytes.
    _ENC_PAD = int(os.environ.get("VCM_B200_ENC_PAD", "29"))

    def _enc_in_map(self, rt):
        pad = self._ENC_PAD if rt.B.precision == "bf16" else 0
        # buffer order (context1 0..63, x 64..66, pad) -> reference order cat([x, context1])
        return [3 + j for j in range(64)] + [0, 1, 2] + [-1] * pad

    def multi_scale_feature_extractor(self, dpb):
        """video_model.py:226-231, :17-37 -> three NHWC views."""
        rt = self._runtime()
        g, B = rt.graph, rt.B
        if dpb["ref_feature"] is None:
            xin, m = g.frame_input(dpb["ref_frame"])
            f = g.conv("feature_adaptor_I", xin, in_map=m)
        else:
            f = g.conv("feature_adaptor_P", self._import(B, dpb["ref_feature"]))
        t, ta = g.conv_t("feature_extractor.conv1", f, 0.01)
        l1 = g.res_block("feature_extractor.res_block1", t, x_act=ta)
        t, ta = g.conv_t("feature_extractor.conv2", l1, 0.01, stride=2)
        l2 = g.res_block("feature_extractor.res_block2", t, x_act=ta)
        t, ta = g.conv_t("feature_extractor.conv3", l2, 0.01, stride=2)
        l3 = g.res_block("feature_extractor.res_block3", t, x_act=ta)
        return l1, l2, l3

    def motion_compensation(self, dpb, mv, want_warp_frame=False):
        """video_model.py:233-246 + MultiScaleContextFusion :40-68.  mv: fp32 flow view.
        Returns (ctx buffer holding context1 at [32,96), E2 = [f | context2], E3 = [f | context3],
        warp_frame view or None).  E2 / E3 are (plain, activated-twin-or-None) pairs: the twins
        (LeakyReLU 0.1) feed the bottleneck ResBlocks of the contextual encoder / decoder."""
        rt = self._runtime()
        g, B = rt.graph, rt.B
        N, H, W = mv.N, mv.H, mv.W
        act = B.act_dtype
        warp_frame = None
        if want_warp_frame:
            warp_frame = B.warp(B.from_nchw(dpb["ref_frame"], dtype=F32), mv)
        mv2 = B.flow_down2(mv, gain=0.5)
        mv3 = B.flow_down2(mv2, gain=0.5)
        f1, f2, f3 = self.multi_scale_feature_extractor(dpb)
        cat_a = B.new(N, H // 2, W // 2, 128, dtype=act)    # (context3_up | context2)
        cat_b = B.new(N, H, W, 128, dtype=act)              # (context2_up | context1)
        c1 = B.warp(f1, mv, out=cat_b.sl(64, 128))
        c2 = B.warp(f2, mv2, out=cat_a.sl(64, 128))
        c3 = B.warp(f3, mv3)
        ctx = self._ctx_buffer(rt, N, H, W)
        e2 = B.new(N, H // 2, W // 2, 128, dtype=act)
        e3 = B.new(N, H // 4, W // 4, 128, dtype=act)
        e2a = B.new(N, H // 2, W // 2, 128, dtype=act) if B.use_twin else None
        e3a = B.new(N, H // 4, W // 4, 128, dtype=act) if B.use_twin else None
        tw = 0.1 if B.use_twin else None
        n = "context_fusion_net"
        t, ta = g.conv_t(n + ".conv3_up.0", c3, 0.01, ps=True)
        g.res_block(n + ".res_block3_up", t, x_act=ta, out=cat_a.sl(0, 64))
        t, ta = g.conv_t(n + ".conv3_out", c3, 0.01)
        g.res_block(n + ".res_block3_out", t, x_act=ta, res2=c3, out=e3.sl(64, 128), twin_out=tw,
                    out2=e3a.sl(64, 128) if e3a else None)
        t, ta = g.conv_t(n + ".conv2_up.0", cat_a, 0.01, ps=True)
        g.res_block(n + ".res_block2_up", t, x_act=ta, out=cat_b.sl(0, 64))
        t, ta = g.conv_t(n + ".conv2_out", cat_a, 0.01)
        g.res_block(n + ".res_block2_out", t, x_act=ta, res2=c2, out=e2.sl(64, 128), twin_out=tw,
                    out2=e2a.sl(64, 128) if e2a else None)
        t, ta = g.conv_t(n + ".conv1_out", cat_b, 0.01)
        g.res_block(n + ".res_block1_out", t, x_act=ta, res2=c1, out=ctx.sl(32, 96))
        return ctx, (e2, e2a), (e3, e3a), warp_frame

    def _mv_prior(self, rt, mv_z_hat, dpb):
        """mv hyper decoder + prior fusion (video_model.py:285-293 / :371-380)."""
        g, B = rt.graph, rt.B
        N, h, w = mv_z_hat.N, mv_z_hat.H * 4, mv_z_hat.W * 4
        fin = B.new(N, h, w, 192, zero=dpb["ref_mv_y"] is None)
        g.hyper_synthesis("mv_hyper_prior_decoder", mv_z_hat, out=fin.sl(0, 128))
        if dpb["ref_mv_y"] is not None:
            self._dpb_into(B, dpb["ref_mv_y"], fin.sl(128, 192))
        return g.three_convs("mv_y_prior_fusion", fin)

    def _y_prior(self, rt, z_hat, e3, dpb):
        """hierarchical + temporal prior -> fusion (video_model.py:312-322 / :394-405)."""
        g, B = rt.graph, rt.B
        N, h, w = z_hat.N, z_hat.H * 4, z_hat.W * 4
        fin = B.new(N, h, w, 480, zero=dpb["ref_y"] is None)
        t = g.conv("temporal_prior_encoder.0", e3[0].sl(64, 128), stride=2, post=0.1)
        g.conv("temporal_prior_encoder.2", t, stride=2, out=fin.sl(0, 192))
        g.hyper_synthesis("contextual_hyper_prior_decoder", z_hat, out=fin.sl(192, 384))
        if dpb["ref_y"] is not None:
            self._dpb_into(B, dpb["ref_y"], fin.sl(384, 480))
        return g.three_convs("y_prior_fusion", fin)

    def _contextual_encoder(self, rt, ctx, e2, e3):
        """video_model.py:71-90; x already sits in ctx[96:99)."""
        g = rt.graph
        n = "contextual_encoder"
        m = self._enc_in_map(rt)
        (e2, e2a), (e3, e3a) = e2, e3
        g.conv_t(n + ".conv1", ctx.sl(32, 32 + len(m)), 0.1, in_map=m, stride=2, out=e2.sl(0, 64),
                 out2=e2a.sl(0, 64) if e2a else None)
        f = g.res_block(n + ".res1", e2, 0.1, True, True, x_act=e2a)
        g.conv_t(n + ".conv2", f, 0.1, stride=2, out=e3.sl(0, 64), out2=e3a.sl(0, 64) if e3a else None)
        f = g.res_block(n + ".res2", e3, 0.1, True, True, x_act=e3a)
        f = g.conv(n + ".conv3", f, stride=2)
        return g.conv(n + ".conv4", f, stride=2)

    def _reconstruct(self, rt, y_hat, ctx, e2, e3, feature_out=None):
        """ContextualDecoder (video_model.py:93-112) + ReconGeneration (:115-128)."""
        g = rt.graph
        n = "contextual_decoder"
        (e2, e2a), (e3, e3a) = e2, e3
        f = g.conv(n + ".up1.0", y_hat, ps=True)
        g.conv_t(n + ".up2.0", f, 0.1, ps=True, out=e3.sl(0, 64), out2=e3a.sl(0, 64) if e3a else None)
        f = g.res_block(n + ".res1", e3, 0.1, True, True, x_act=e3a)
        g.conv_t(n + ".up3.0", f, 0.1, ps=True, out=e2.sl(0, 64), out2=e2a.sl(0, 64) if e2a else None)
        f = g.res_block(n + ".res2", e2, 0.1, True, True, x_act=e2a)
        g.conv(n + ".up4.0", f, ps=True, out=ctx.sl(0, 32))
        n = "recon_generation_net"
        f = g.conv(n + ".first_conv", ctx.sl(0, 96))
        f = g.unet(n + ".unet_1", f)
        feature = g.unet(n + ".unet_2", f, out=feature_out)
        x_hat = g.conv(n + ".recon_conv", feature, out_dtype=F32)
        return feature, x_hat

    def _analyse(self, x, dpb, mv_y_q_scale, y_q_scale, want_full, outs=None):
        """Encoder-side trunk shared by compress() and forward_one_frame() (video_model.py:263-331 /
        :470-538).  outs: static next-frame buffers (`_statics`) that mv_y_hat / y_hat / feature are
        written to in place of fresh tensors (each is read, as the previous frame's, before the
        kernel that overwrites it)."""
        rt = self._runtime()
        g, B = rt.graph, rt.B
        N, _, H, W = x.shape
        q_mv = self._curr_q(rt, self.mv_y_q_basic, mv_y_q_scale, N)
        q_y = self._curr_q(rt, self.y_q_basic, y_q_scale, N)
        table = self._scale_table
        outs = outs or {}

        est_mv = g.spynet("optic_flow", x, dpb["ref_frame"])
        if B.use_twin:
            # tensor-core mode: the 2-channel flow as bf16 padded to 16 channels (zero weights on the
            # pad), so the stride-2 first layer of the mv encoder runs on tcgen05 like every other conv
            mv_in = self._persist(rt, "mv_in", N, H, W, 16, B.act_dtype)
            B.copy(est_mv, mv_in.sl(0, 2))
            mv_y = g.analysis("mv_encoder", mv_in, [0, 1] + [-1] * 14)
        else:
            mv_y = g.analysis("mv_encoder", est_mv)
        mv_y = B.scale(mv_y, q_mv, "div")
        mv_z = g.hyper_analysis("mv_hyper_prior_encoder", mv_y)
        zs = (N, 64, mv_z.H, mv_z.W)
        mv_z_sym = B.empty(zs, torch.int16)
        mv_z_hat = B.round_latent(mv_z, sym=mv_z_sym)
        mv_fusion = self._mv_prior(rt, mv_z_hat, dpb)
        mv_dp = g.dual_prior_encode(table, mv_y, mv_fusion, "mv_y_spatial_prior", q_mv, want_full,
                                    y_hat=outs.get("ref_mv_y"))
        mv_y_hat = mv_dp["y_hat"]
        mv_hat = g.synthesis("mv_decoder", mv_y_hat, out_dtype=F32)

        ctx, e2, e3, warp_frame = self.motion_compensation(dpb, mv_hat, want_warp_frame=want_full)
        B.from_nchw(x, out=ctx.sl(96, 99))
        y = B.scale(self._contextual_encoder(rt, ctx, e2, e3), q_y, "div")
        n = "contextual_hyper_prior_encoder"
        t = g.conv(n + ".0", y, post=0.01)
        t = g.conv(n + ".2", t, stride=2, post=0.01)
        z = g.conv(n + ".4", t, stride=2)
        z_sym = B.empty(zs, torch.int16)
        z_hat = B.round_latent(z, sym=z_sym)
        y_fusion = self._y_prior(rt, z_hat, e3, dpb)
        y_dp = g.dual_prior_encode(table, y, y_fusion, "y_spatial_prior", q_y, want_full,
                                   y_hat=outs.get("ref_y"))
        y_hat = y_dp["y_hat"]
        feature, x_hat = self._reconstruct(rt, y_hat, ctx, e2, e3, outs.get("ref_feature"))
        return dict(mv_z_sym=mv_z_sym, z_sym=z_sym, mv_z_hat=mv_z_hat, z_hat=z_hat, mv_dp=mv_dp, y_dp=y_dp,
                    mv_y_hat=mv_y_hat, y_hat=y_hat, feature=feature, x_hat=x_hat, warp_frame=warp_frame,
                    est_mv=est_mv, mv_hat=mv_hat)

    # ---- static next-frame state + captured programs (set_cuda_graphs) ---------------------------
    def _statics(self, rt, H, W, side):
        """Fixed-address frame state for the captured programs of one side ("enc" / "dec"): the
        current frame / first latent, and the four dpb entries, which every coded frame reads (as the
        previous frame's) and then overwrites.  The two sides keep separate chains, so
        compress(dpb) followed by decompress(dpb) on one model still sees the same `dpb`."""
        st = rt.statics.get((H, W, side))
        if st is None:
            B = rt.B
            st = {"ref_frame": torch.empty((1, 3, H, W), dtype=F32, device=B.device),
                  "ref_feature": B.new(1, H, W, 64),
                  "ref_y": B.new(1, H // 16, W // 16, 96),
                  "ref_mv_y": B.new(1, H // 16, W // 16, 64)}
            for k, v in st.items():
                rt.static_ids[id(v)] = (side, k)
            if side == "enc":
                st["x"] = torch.empty((1, 3, H, W), dtype=F32, device=B.device)
            else:
                st["mv_z_sym"] = B.empty((1, 64, H // 64, W // 64), torch.int16)
            rt.statics[(H, W, side)] = st
        return st

    def _load_statics(self, rt, st, dpb, side):
        """dpb from the caller -> the dpb the frame body reads.  Entries that are static buffers of
        this model (the dpb it returned for the previous frame, either side) are read where they
        lie; anything else is copied into `st`.  Also returns the pattern that selects the program."""
        B = rt.B
        inner, pattern = {}, []
        for k in ("ref_frame", "ref_feature", "ref_y", "ref_mv_y"):
            v = dpb[k]
            if v is None:
                inner[k] = None
                pattern.append(None)
                continue
            owner = rt.static_ids.get(id(v))
            if owner is not None and owner[1] == k:
                inner[k] = v
                pattern.append(owner[0])
                continue
            if k == "ref_frame":
                st[k].copy_(v.detach().to(B.device, F32))
            else:
                B.copy(self._import(B, v), st[k])
            inner[k] = st[k]
            pattern.append(side)
        return inner, tuple(pattern)

    def _enc_body(self, rt, x, dpb, q_mv, q_y, outs, clamp_ref):
        """Generator with no host hand-off (graphs.py): one captured graph."""
        t = self._analyse(x, dpb, q_mv, q_y, False, outs)
        if outs is not None:
            rt.B.to_nchw(t["x_hat"], clamp=(0.0, 1.0) if clamp_ref else None, out=outs["ref_frame"])
        return (t["mv_z_sym"], *t["mv_dp"]["sym"], *t["mv_dp"]["idx"],
                t["z_sym"], *t["y_dp"]["sym"], *t["y_dp"]["idx"])
        yield   # noqa: unreachable -- makes this a generator

    # ---- public API ------------------------------------------------------------------------------
    @torch.no_grad()
    def compress_symbols(self, x, dpb, mv_y_q_scale, y_q_scale, want_dpb=True, clamp_ref=False):
        """Device half of compress(): analysis nets + quantisation, symbols / scale indexes copied to
        the host.  Returns (host symbol tuple in stream order, dpb or None).  clamp_ref: ref_frame
        clamped to [0,1] like the decoder's (video_model.py:413), for encoder-side prediction chains
        (pipeline.py)."""
        rt = self._need_coder()
        assert x.shape[0] == 1, "the bitstream path is batch 1 (entropy_models.py:186)"
        B = rt.B
        if self._graphs:
            H, W = x.shape[2:]
            st = self._statics(rt, H, W, "enc")
            inner, pattern = self._load_statics(rt, st, dpb, "enc")
            st["x"].copy_(x.detach().to(B.device, F32))
            outs = st if want_dpb else None
            key = ("enc", H, W, pattern, float(mv_y_q_scale), float(y_q_scale), want_dpb, clamp_ref)
            syms = rt.programs.run(key, lambda: self._enc_body(rt, st["x"], inner, float(mv_y_q_scale),
                                                                float(y_q_scale), outs, clamp_ref), None)
            out = {k: st[k] for k in ("ref_frame", "ref_feature", "ref_y", "ref_mv_y")} if want_dpb else None
            return tuple(self._to_host(*syms)), out
        t = self._analyse(x, dpb, mv_y_q_scale, y_q_scale, want_full=False)
        host = tuple(self._to_host(t["mv_z_sym"], *t["mv_dp"]["sym"], *t["mv_dp"]["idx"],
                                   t["z_sym"], *t["y_dp"]["sym"], *t["y_dp"]["idx"]))
        out = None
        if want_dpb:
            out = {
                "ref_frame": self._export(B, t["x_hat"], clamp=(0.0, 1.0) if clamp_ref else None),
                "ref_feature": self._export(B, t["feature"]),
                "ref_y": self._export(B, t["y_hat"]),
                "ref_mv_y": self._export(B, t["mv_y_hat"]),
            }
        return host, out

    def encode_symbols(self, host):
        """Host half of compress(): the sequential rANS coder (GIL released inside).  Stream order:
        z_mv, mv_y pass 0/1, z, y pass 0/1 (video_model.py:333-340)."""
        rt = self._need_coder()
        mv_z, ms0, ms1, mi0, mi1, z, ys0, ys1, yi0, yi1 = host
        c = rt.coder
        c.reset_encoder()
        c.encode_channels(mv_z, rt.tables["bit_estimator_z_mv"])
        c.encode_planes(ms0, mi0, rt.tables["y"])
        c.encode_planes(ms1, mi1, rt.tables["y"])
        c.encode_channels(z, rt.tables["bit_estimator_z"])
        c.encode_planes(ys0, yi0, rt.tables["y"])
        c.encode_planes(ys1, yi1, rt.tables["y"])
        return c.flush_encoder()

    @torch.no_grad()
    def compress(self, x, dpb, mv_y_q_scale, y_q_scale, _want_dpb=True):
        """video_model.py:263-352."""
        host, dbp = self.compress_symbols(x, dpb, mv_y_q_scale, y_q_scale, want_dpb=_want_dpb)
        bit_stream = self.encode_symbols(host)
        mv_z, ms0, ms1, mi0, mi1, z, ys0, ys1, yi0, yi1 = host
        self._last_symbols = dict(mv_z=mv_z, mv_y_q_w_0=ms0, mv_y_q_w_1=ms1, mv_idx_0=mi0, mv_idx_1=mi1,
                                  z=z, y_q_w_0=ys0, y_q_w_1=ys1, y_idx_0=yi0, y_idx_1=yi1)
        return {"dbp": dbp, "bit_stream": bit_stream}   # dbp is None for encode_decode() (decoder's is used)

    def _dec_body(self, rt, dpb, q_mv_scale, q_y_scale, zs, outs):
        """Decoder frame body (video_model.py:354-422) as a generator: yields the host-coder requests
        of graphs.py (two latents, four index planes), returns (x_hat, feature, y_hat, mv_y_hat)."""
        g, B = rt.graph, rt.B
        q_mv = self._curr_q(rt, self.mv_y_q_basic, q_mv_scale, 1)
        q_y = self._curr_q(rt, self.y_q_basic, q_y_scale, 1)
        table = self._scale_table
        outs = outs or {}

        def planes(gen):
            try:
                idx = next(gen)
                while True:
                    sym = yield ("plane", idx)
                    idx = gen.send(sym)
            except StopIteration as e:
                return e.value

        mv_z_hat = B.symbols_to_latent(outs["mv_z_sym"], 1, zs[0], zs[1], 64)
        mv_fusion = self._mv_prior(rt, mv_z_hat, dpb)
        mv_y_hat = yield from planes(g.dual_prior_decode_gen(table, mv_fusion, 64, "mv_y_spatial_prior",
                                                             q_mv, y_hat=outs.get("ref_mv_y")))
        mv_hat = g.synthesis("mv_decoder", mv_y_hat, out_dtype=F32)
        ctx, e2, e3, _ = self.motion_compensation(dpb, mv_hat)
        z_sym = yield ("z", "bit_estimator_z", 64, zs)
        z_hat = B.symbols_to_latent(z_sym, 1, zs[0], zs[1], 64)
        y_fusion = self._y_prior(rt, z_hat, e3, dpb)
        y_hat = yield from planes(g.dual_prior_decode_gen(table, y_fusion, 96, "y_spatial_prior", q_y,
                                                          y_hat=outs.get("ref_y")))
        feature, x_hat = self._reconstruct(rt, y_hat, ctx, e2, e3, outs.get("ref_feature"))
        if "ref_frame" in outs:
            B.to_nchw(x_hat, clamp=(0.0, 1.0), out=outs["ref_frame"])
        return x_hat, feature, y_hat, mv_y_hat

    def _dec_service(self, rt):
        """Host side of the decoder requests: the sequential rANS decoder (GIL released inside)."""
        def service(req):
            if req[0] == "plane":
                (idx,) = rt.staging.to_host("dec_idx", req[1])
                out = rt.staging.host_buf("dec_sym", idx.shape, torch.int16)
                rt.coder.decode_planes(idx, rt.tables["y"], out)
                return out
            _, name, C, zs = req
            out = rt.staging.host_buf("dec_z:" + name, (1, C, zs[0], zs[1]), torch.int16)
            rt.coder.decode_channels(C, zs[0] * zs[1], rt.tables[name], out)
            return out
        return service

    @torch.no_grad()
    def decompress(self, dpb, string, height, width, mv_y_q_scale, y_q_scale):
        """video_model.py:354-422."""
        rt = self._need_coder()
        B = rt.B
        rt.coder.set_stream(string)
        zs = get_downsampled_shape(height, width, 64)
        service = self._dec_service(rt)
        mv_z = service(("z", "bit_estimator_z_mv", 64, zs))      # first in the stream, no device input
        if self._graphs:
            H, W = zs[0] * 64, zs[1] * 64
            st = self._statics(rt, H, W, "dec")
            inner, pattern = self._load_statics(rt, st, dpb, "dec")
            st["mv_z_sym"].copy_(mv_z)
            key = ("dec", H, W, pattern, float(mv_y_q_scale), float(y_q_scale))
            rt.programs.run(key, lambda: self._dec_body(rt, inner, float(mv_y_q_scale), float(y_q_scale),
                                                        zs, st), service)
            return {"dpb": {k: st[k] for k in ("ref_frame", "ref_feature", "ref_y", "ref_mv_y")}}
        gen = self._dec_body(rt, dpb, mv_y_q_scale, y_q_scale, zs, {"mv_z_sym": mv_z.to(B.device)})
        x_hat, feature, y_hat, mv_y_hat = drive_eager(gen, service, B.device)
        return {
            "dpb": {
                "ref_frame": self._export(B, x_hat, clamp=(0.0, 1.0)),
                "ref_feature": self._export(B, feature),
                "ref_y": self._export(B, y_hat),
                "ref_mv_y": self._export(B, mv_y_hat),
            }
        }

    def encode_decode_bytes(self, x, dpb, pic_height, pic_width, mv_y_q_scale, y_q_scale):
        """encode_decode(write) with the `.bin` kept in memory: returns (result, blob)."""
        mv_y_q_scale, mv_y_q_index = get_rounded_q(mv_y_q_scale)
        y_q_scale, y_q_index = get_rounded_q(y_q_scale)
        t0 = time.time()
        encoded = self.compress(x, dpb, mv_y_q_scale, y_q_scale, _want_dpb=False)
        blob = pack_p(encoded["bit_stream"], mv_y_q_index, y_q_index)
        t1 = time.time()
        a, b, string = unpack_p(blob)
        decoded = self.decompress(dpb, string, pic_height, pic_width, a / 100, b / 100)
        t2 = time.time()
        return {"dpb": decoded["dpb"], "bit": len(blob) * 8, "encoding_time": t1 - t0,
                "decoding_time": t2 - t1}, blob

    def encode_decode(self, x, dpb, output_path=None, pic_width=None, pic_height=None,
                      mv_y_q_scale=None, y_q_scale=None):
        """video_model.py:424-468."""
        if output_path is not None:
            mv_y_q_scale, mv_y_q_index = get_rounded_q(mv_y_q_scale)
            y_q_scale, y_q_index = get_rounded_q(y_q_scale)
            t0 = time.time()
            encoded = self.compress(x, dpb, mv_y_q_scale, y_q_scale, _want_dpb=False)
            encode_p(encoded["bit_stream"], mv_y_q_index, y_q_index, output_path)
            bits = filesize(output_path) * 8
            t1 = time.time()
            mv_y_q_index, y_q_index, string = decode_p(output_path)
            decoded = self.decompress(dpb, string, pic_height, pic_width,
                                      mv_y_q_index / 100, y_q_index / 100)
            t2 = time.time()
            return {"dpb": decoded["dpb"], "bit": bits, "encoding_time": t1 - t0,
                    "decoding_time": t2 - t1}
        encoded = self.forward_one_frame(x, dpb, mv_y_q_scale=mv_y_q_scale, y_q_scale=y_q_scale)
        return {
            "dpb": encoded["dpb"],
            "bit_y": encoded["bit_y"].item(),
            "bit_z": encoded["bit_z"].item(),
            "bit_mv_y": encoded["bit_mv_y"].item(),
            "bit_mv_z": encoded["bit_mv_z"].item(),
            "bit": encoded["bit"].item(),
            "decoding_time": 0,
        }

    def forward_one_frame(self, x, dpb, mv_y_q_scale=None, y_q_scale=None):
        """video_model.py:470-592.  In train() mode under autograd the differentiable graph of training.py
        runs: straight-through quantiser, noisy likelihoods, gradients to every parameter through the
        device conv / warp kernels.  In eval() mode (or under torch.no_grad()) the fused inference graph
        below runs and the result carries no gradients."""
        if self.training and torch.is_grad_enabled():
            from .training import TrainGraph, default_noise
            rt = self._runtime()
            if getattr(rt, "train_graph", None) is None:
                rt.train_graph = TrainGraph(self, rt)
            noise = getattr(self, "_noise_fn", None) or default_noise
            return rt.train_graph.forward_one_frame(x, dpb, mv_y_q_scale, y_q_scale, noise)
        with torch.no_grad():
            rt = self._runtime()
            B = rt.B
            t = self._analyse(x, dpb, mv_y_q_scale, y_q_scale, want_full=True)
            N, _, H, W = x.shape
            pixel_num = H * W
            x_hat = B.to_nchw(t["x_hat"])
            mse = B.sq_err_sum(x, x_hat) / pixel_num
            me_mse = B.sq_err_sum(x, B.to_nchw(t["warp_frame"])) / pixel_num
            bpp_y = B.latent_bits(t["y_dp"]["y_q"], t["y_dp"]["scales_hat"], N, "laplace") / pixel_num
            bpp_mv_y = B.latent_bits(t["mv_dp"]["y_q"], t["mv_dp"]["scales_hat"], N, "laplace") / pixel_num
            bpp_z = B.factorized_bits(t["z_hat"], self._fparams(rt, "bit_estimator_z")) / pixel_num
            bpp_mv_z = B.factorized_bits(t["mv_z_hat"], self._fparams(rt, "bit_estimator_z_mv")) / pixel_num
            bpp = bpp_y + bpp_z + bpp_mv_y + bpp_mv_z
            return {
                "bpp_mv_y": bpp_mv_y, "bpp_mv_z": bpp_mv_z, "bpp_y": bpp_y, "bpp_z": bpp_z, "bpp": bpp,
                "me_mse": me_mse, "mse": mse,
                "dpb": {"ref_frame": x_hat, "ref_feature": B.to_nchw(t["feature"]),
                        "ref_y": B.to_nchw(t["y_hat"]), "ref_mv_y": B.to_nchw(t["mv_y_hat"])},
                "bit": torch.sum(bpp) * pixel_num, "bit_y": torch.sum(bpp_y) * pixel_num,
                "bit_z": torch.sum(bpp_z) * pixel_num, "bit_mv_y": torch.sum(bpp_mv_y) * pixel_num,
                "bit_mv_z": torch.sum(bpp_mv_z) * pixel_num,
            }

    def forward(self, x, dpb, mv_y_q_scale=None, y_q_scale=None):
        return self.forward_one_frame(x, dpb, mv_y_q_scale=mv_y_q_scale, y_q_scale=y_q_scale)


# ==========================================================================================
# I-frame model
# ==========================================================================================
class IntraNoAR(CompressionModel):
    """image_model.py:16-200."""

    def __init__(self, N=192, anchor_num=4, precision=None):
        super().__init__(arch.intra_spec(N, anchor_num), "gaussian", precision)
        self.N = int(N)
        self.anchor_num = int(anchor_num)

    @staticmethod
    def get_q_scales_from_ckpt(ckpt_path):
        return get_state_dict(ckpt_path)["q_scale"].reshape(-1)

    def get_curr_q(self, q_scale):
        return torch.max(self.q_basic, torch.ones_like(self.q_basic) * 0.5) * q_scale

    def _trunk(self, rt, x, q, want_full):
        g, B = rt.graph, rt.B
        xin, m = g.frame_input(x)
        y = B.scale(g.analysis("enc", xin, m), q, "div")
        z = g.hyper_analysis("hyper_enc", y)
        z_sym = B.empty((x.shape[0], self.N, z.H, z.W), torch.int16)
        z_hat = B.round_latent(z, sym=z_sym)
        fusion = g.three_convs("y_prior_fusion", g.hyper_synthesis("hyper_dec", z_hat))
        dp = g.dual_prior_encode(self._scale_table, y, fusion, "y_spatial_prior", q, want_full)
        return z_sym, z_hat, dp

    def _synth(self, rt, y_hat):
        g = rt.graph
        t = g.synthesis("dec", y_hat)
        return g.conv("refine.1", g.unet("refine.0", t), out_dtype=F32)

    def forward(self, x, q_scale=None):
        """image_model.py:54-105.  In train() mode under autograd: the differentiable graph of training.py
        (straight-through quantiser, noisy likelihoods, gradients through the device conv kernels); otherwise
        the fused inference graph."""
        if self.training and torch.is_grad_enabled():
            from .training import TrainGraph, default_noise
            rt = self._runtime()
            if getattr(rt, "train_graph", None) is None:
                rt.train_graph = TrainGraph(self, rt)
            return rt.train_graph.intra_forward(x, q_scale, getattr(self, "_noise_fn", None) or default_noise)
        with torch.no_grad():
            rt = self._runtime()
            B = rt.B
            N, _, H, W = x.shape
            q = self._curr_q(rt, self.q_basic, q_scale, N)
            _, z_hat, dp = self._trunk(rt, x, q, True)
            x_hat = B.to_nchw(self._synth(rt, dp["y_hat"]))
            pixel_num = H * W
            bpp_y = B.latent_bits(dp["y_q"], dp["scales_hat"], N, "gaussian") / pixel_num
            bpp_z = B.factorized_bits(z_hat, self._fparams(rt, "bit_estimator_z")) / pixel_num
            mse = B.sq_err_sum(x, x_hat) / pixel_num
            bits = torch.sum(bpp_y + bpp_z) * pixel_num
            return {"x_hat": x_hat, "mse": mse, "bit": bits.item(), "bpp": bpp_y + bpp_z,
                    "bpp_y": bpp_y, "bpp_z": bpp_z}

    # ---- captured programs (set_cuda_graphs): static input frame / output reconstruction per resolution
    def _intra_statics(self, rt, H, W):
        st = rt.statics.get((H, W, "intra"))
        if st is None:
            dev = rt.B.device
            st = {"x": torch.empty((1, 3, H, W), dtype=F32, device=dev),
                  "x_hat": torch.empty((1, 3, H, W), dtype=F32, device=dev),
                  "z_sym": rt.B.empty((1, self.N, H // 64, W // 64), torch.int16)}
            rt.statics[(H, W, "intra")] = st
        return st

    def _enc_body(self, rt, x, q_scale):
        q = self._curr_q(rt, self.q_basic, q_scale, 1)
        z_sym, _, dp = self._trunk(rt, x, q, False)
        return (z_sym, *dp["sym"], *dp["idx"])
        yield   # noqa: unreachable -- makes this a generator (graphs.py)

    def _dec_body(self, rt, z_sym, zs, q_scale, out):
        g, B = rt.graph, rt.B
        q = self._curr_q(rt, self.q_basic, q_scale, 1)
        z_hat = B.symbols_to_latent(z_sym, 1, zs[0], zs[1], self.N)
        fusion = g.three_convs("y_prior_fusion", g.hyper_synthesis("hyper_dec", z_hat))
        gen = g.dual_prior_decode_gen(self._scale_table, fusion, self.N, "y_spatial_prior", q)
        try:
            idx = next(gen)
            while True:
                sym = yield ("plane", idx)
                idx = gen.send(sym)
        except StopIteration as e:
            y_hat = e.value
        return B.to_nchw(self._synth(rt, y_hat), clamp=(0.0, 1.0), out=out)

    @torch.no_grad()
    def compress(self, x, q_scale):
        """image_model.py:148-177.  Stream order: z, y pass 0, y pass 1."""
        rt = self._need_coder()
        assert x.shape[0] == 1, "the bitstream path is batch 1 (entropy_models.py:186)"
        if self._graphs:
            H, W = x.shape[2:]
            st = self._intra_statics(rt, H, W)
            st["x"].copy_(x.detach().to(rt.B.device, F32))
            syms = rt.programs.run(("ienc", H, W, float(q_scale)),
                                   lambda: self._enc_body(rt, st["x"], float(q_scale)), None)
            z, s0, s1, i0, i1 = self._to_host(*syms)
        else:
            q = self._curr_q(rt, self.q_basic, q_scale, 1)
            z_sym, _, dp = self._trunk(rt, x, q, False)
            z, s0, s1, i0, i1 = self._to_host(z_sym, *dp["sym"], *dp["idx"])
        c = rt.coder
        c.reset_encoder()
        c.encode_channels(z, rt.tables["bit_estimator_z"])
        c.encode_planes(s0, i0, rt.tables["y"])
        c.encode_planes(s1, i1, rt.tables["y"])
        self._last_symbols = dict(z=z, y_q_w_0=s0, y_q_w_1=s1, y_idx_0=i0, y_idx_1=i1)
        return {"bit_stream": c.flush_encoder()}

    @torch.no_grad()
    def decompress(self, bit_stream, height, width, q_scale):
        """image_model.py:179-200.  With set_cuda_graphs(True) the returned x_hat is a static buffer of the
        model that the next decompress() of the same resolution overwrites (use or copy it first)."""
        rt = self._need_coder()
        g, B = rt.graph, rt.B
        rt.coder.set_stream(bit_stream)
        zs = get_downsampled_shape(height, width, 64)
        if self._graphs:
            H, W = zs[0] * 64, zs[1] * 64
            st = self._intra_statics(rt, H, W)
            z = rt.staging.host_buf("dec_z", (1, self.N, zs[0], zs[1]), torch.int16)
            rt.coder.decode_channels(self.N, zs[0] * zs[1], rt.tables["bit_estimator_z"], z)
            st["z_sym"].copy_(z, non_blocking=True)

            def service(req):
                (idx,) = rt.staging.to_host("dec_idx", req[1])
                out = rt.staging.host_buf("dec_sym", idx.shape, torch.int16)
                rt.coder.decode_planes(idx, rt.tables["y"], out)
                return out
            rt.programs.run(("idec", H, W, float(q_scale)),
                            lambda: self._dec_body(rt, st["z_sym"], zs, float(q_scale), st["x_hat"]), service)
            return {"x_hat": st["x_hat"]}
        q = self._curr_q(rt, self.q_basic, q_scale, 1)
        z_hat = self._decode_z(rt, "bit_estimator_z", self.N, zs)
        fusion = g.three_convs("y_prior_fusion", g.hyper_synthesis("hyper_dec", z_hat))
        y_hat = g.dual_prior_decode(self._scale_table, fusion, self.N, "y_spatial_prior", q,
                                    self._decode_plane_fn(rt))
        return {"x_hat": B.to_nchw(self._synth(rt, y_hat), clamp=(0.0, 1.0))}

    def encode_decode_bytes(self, x, q_scale, pic_height, pic_width):
        t0 = time.time()
        q_scale, q_index = get_rounded_q(q_scale)
        blob = pack_i(pic_height, pic_width, q_index, self.compress(x, q_scale)["bit_stream"])
        t1 = time.time()
        height, width, q_index, bit_stream = unpack_i(blob)
        x_hat = self.decompress(bit_stream, height, width, q_index / 100)["x_hat"]
        t2 = time.time()
        return {"bit": len(blob) * 8, "x_hat": x_hat, "encoding_time": t1 - t0,
                "decoding_time": t2 - t1}, blob

    def encode_decode(self, x, q_scale, output_path=None, pic_width=None, pic_height=None):
        """image_model.py:114-146."""
        if output_path is None:
            return self.forward(x, q_scale)
        assert pic_height is not None
        assert pic_width is not None
        t0 = time.time()
        q_scale, q_index = get_rounded_q(q_scale)
        bit_stream = self.compress(x, q_scale)["bit_stream"]
        encode_i(pic_height, pic_width, q_index, bit_stream, output_path)
        bit = filesize(output_path) * 8
        t1 = time.time()
        height, width, q_index, bit_stream = decode_i(output_path)
        x_hat = self.decompress(bit_stream, height, width, q_index / 100)["x_hat"]
        t2 = time.time()
        return {"bit": bit, "x_hat": x_hat, "encoding_time": t1 - t0, "decoding_time": t2 - t1}
