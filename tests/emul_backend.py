"""TEST INFRASTRUCTURE: a torch-CPU emulation of vcm_ts_b200.backend.CudaBackend.

Each method restates the *contract* of one C-ABI kernel (include/vcmts_b200.h) with plain torch
ops, so the host-side graphs in vcm_ts_b200/nets.py + models.py (channel-slice wiring, fused
epilogue flags, weight permutations, stream order) can be checked against the oracle on a machine
without a GPU.  It is injected through `model._backend_factory` by tests only; the product never
imports it and has no CPU path.
"""
import math

import torch
import torch.nn.functional as F

from vcm_ts_b200.backend import PackedConv, View


class EmulBackend:
    precision = "fp32"
    act_dtype = torch.float32

    def __init__(self, device="cpu", precision="fp32"):
        self.device = torch.device(device)
        self.launches = 0
        # "twin": fp32 math but the bf16-mode graph (activated twins instead of pre-activations)
        self.use_twin = precision != "fp32"
        # the bf16-mode graph also takes the SELayer means out of the producing conv (VCM_CONV_SE_SUM)
        self.fused_se_sums = self.use_twin
        self.tc_like = []      # (name-less) record of convs that asked for a pre-activation

    def new(self, N, H, W, C, dtype=None, zero=False):
        f = torch.zeros if zero else (lambda s, dtype: torch.full(s, float("nan"), dtype=dtype))
        return View(f((N, H, W, C), dtype=dtype or self.act_dtype))

    def empty(self, shape, dtype):
        return torch.zeros(shape, dtype=dtype)

    def pack_conv(self, weight, bias, in_map=None, cin_p=None, ps=False):
        w = weight.detach().float()
        Cout, Cin, K, _ = w.shape
        if in_map is not None:
            w2 = torch.zeros(Cout, len(in_map), K, K)
            for j, src in enumerate(in_map):
                if src >= 0:
                    w2[:, j] = w[:, src]
            w = w2
        pc = PackedConv(w, bias.detach().float(), K, w.shape[1], Cout)
        return pc

    @staticmethod
    def _nchw(v):
        return v.dense().permute(0, 3, 1, 2).float()

    @staticmethod
    def _store(v, t_nchw):
        v.dense().copy_(t_nchw.permute(0, 2, 3, 1).to(v.dtype))

    def conv(self, x, pc, out=None, stride=1, pre=None, post=None, res=None, res_scale=None,
             res2=None, ps=False, out_dtype=None, slope2=None, out2=None, se_sums=False):
        if se_sums:
            # contract of VCM_CONV_SE_SUM: rows of per-"warp" channel sums of the output that add up to the
            # channel totals, or None where the engine has no fused sums (here: as the engines, above 64 channels)
            assert slope2 is None
            y = self.conv(x, pc, out=out, stride=stride, pre=pre, post=post, res=res, res_scale=res_scale,
                          res2=res2, ps=ps, out_dtype=out_dtype)
            if y.N != 1 or y.C > 64:
                return y, None
            t = self._nchw(y)[0].reshape(y.C, -1)
            parts = torch.stack([c.sum(dim=1) for c in torch.chunk(t, 8, dim=1)])
            return y, parts.contiguous()
        assert x.C == pc.Cin
        if pre is not None:
            self.tc_like.append((pc.K, pc.Cin, pc.Cout))
        t = self._nchw(x)
        assert not torch.isnan(t).any(), "conv reads an unwritten buffer"
        if pre is not None:
            t = F.leaky_relu(t, pre)
        y = F.conv2d(t, pc.w, pc.bias, stride, pc.K // 2)
        if post is not None:
            y = F.leaky_relu(y, post)
        if ps:
            y = F.pixel_shuffle(y, 2)
        if res is not None and res_scale is None:
            y = self._nchw(res) + y
        if res is not None and res_scale is not None:
            y = y + self._nchw(res) * res_scale[:, :, None, None]
        if res2 is not None:
            y = self._nchw(res2) + y
        if out is None:
            out = self.new(y.shape[0], y.shape[2], y.shape[3], y.shape[1], dtype=out_dtype)
        assert (out.N, out.C, out.H, out.W) == tuple(y.shape)
        self._store(out, y)
        self.launches += 1
        if slope2 is not None:
            if out2 is None:
                out2 = self.new(y.shape[0], y.shape[2], y.shape[3], y.shape[1], dtype=out.dtype)
            self._store(out2, F.leaky_relu(y, slope2))
            return out, out2
        return out

    def from_nchw(self, t, out=None, dtype=None):
        t = t.detach().float()
        N, C, H, W = t.shape
        if out is None:
            out = self.new(N, H, W, C, dtype=dtype)
        self._store(out, t)
        return out

    def to_nchw(self, v, clamp=None, out=None):
        t = self._nchw(v).contiguous()
        t = t.clamp(*clamp) if clamp is not None else t
        if out is not None:
            out.copy_(t)
            return out
        return t

    def copy(self, src, dst):
        dst.dense().copy_(src.dense().to(dst.dtype))
        return dst

    def scale(self, src, vec, op, out=None):
        if out is None:
            out = self.new(src.N, src.H, src.W, src.C, dtype=src.dtype)
        t = self._nchw(src)
        v = vec[:, :, None, None]
        self._store(out, t / v if op == "div" else t * v)
        return out

    def maxpool(self, x, out=None):
        y = F.max_pool2d(self._nchw(x), 2, 2)
        if out is None:
            out = self.new(x.N, x.H // 2, x.W // 2, x.C, dtype=x.dtype)
        self._store(out, y)
        return out

    def se_scale(self, x, w1, w2):
        m = torch.mean(self._nchw(x), dim=(-1, -2))
        return torch.sigmoid(F.linear(F.relu(F.linear(m, w1)), w2)).contiguous()

    def se_scale_from_sums(self, sums, x, w1, w2):
        m = (sums.sum(dim=0) / float(x.H * x.W))[None]
        return torch.sigmoid(F.linear(F.relu(F.linear(m, w1)), w2)).contiguous()

    def avgpool(self, x, out=None):
        y = F.avg_pool2d(self._nchw(x), 2, 2)
        if out is None:
            out = self.new(x.N, x.H // 2, x.W // 2, x.C, dtype=torch.float32)
        self._store(out, y)
        return out

    def flow_up2(self, flow, out=None, gain=2.0):
        t = self._nchw(flow)
        y = F.interpolate(t, (t.shape[2] * 2, t.shape[3] * 2), mode="bilinear", align_corners=False) * gain
        if out is None:
            out = self.new(flow.N, flow.H * 2, flow.W * 2, 2, dtype=torch.float32)
        self._store(out, y)
        return out

    def flow_down2(self, flow, out=None, gain=0.5):
        t = self._nchw(flow)
        y = F.interpolate(t, (t.shape[2] // 2, t.shape[3] // 2), mode="bilinear", align_corners=False) * gain
        if out is None:
            out = self.new(flow.N, flow.H // 2, flow.W // 2, 2, dtype=torch.float32)
        self._store(out, y)
        return out

    def warp(self, feat, flow, out=None):
        f, fl = self._nchw(feat), self._nchw(flow)
        N, _, H, W = fl.shape
        hor = torch.linspace(-1.0, 1.0, W).view(1, 1, 1, W).expand(N, -1, H, -1)
        ver = torch.linspace(-1.0, 1.0, H).view(1, 1, H, 1).expand(N, -1, -1, W)
        grid = torch.cat([hor, ver], 1) + torch.cat([fl[:, 0:1] / ((W - 1.0) / 2.0),
                                                     fl[:, 1:2] / ((H - 1.0) / 2.0)], 1)
        y = F.grid_sample(f, grid.permute(0, 2, 3, 1), mode="bilinear", padding_mode="border",
                          align_corners=True)
        if out is None:
            out = self.new(feat.N, feat.H, feat.W, feat.C, dtype=feat.dtype)
        self._store(out, y)
        return out

    # ---- training path: contracts of vcm_conv2d_wgrad / vcm_flow_warp_bwd -------------------------
    def wgrad(self, x, dy, K, stride, Cin, Cout):
        with torch.enable_grad():
            w = torch.zeros((Cout, Cin, K, K), requires_grad=True)
            y = F.conv2d(self._nchw(x).contiguous(), w, None, stride, K // 2)
            (gw,) = torch.autograd.grad(y, w, self._nchw(dy).contiguous())
        return gw

    def warp_bwd(self, feat, flow, gout):
        with torch.enable_grad():
            f = self._nchw(feat).detach().requires_grad_(True)
            fl = self._nchw(flow).detach().requires_grad_(True)
            N, _, H, W = fl.shape
            hor = torch.linspace(-1.0, 1.0, W).view(1, 1, 1, W).expand(N, -1, H, -1)
            ver = torch.linspace(-1.0, 1.0, H).view(1, 1, H, 1).expand(N, -1, -1, W)
            grid = torch.cat([hor, ver], 1) + torch.cat([fl[:, 0:1] / ((W - 1.0) / 2.0),
                                                         fl[:, 1:2] / ((H - 1.0) / 2.0)], 1)
            y = F.grid_sample(f, grid.permute(0, 2, 3, 1), mode="bilinear", padding_mode="border",
                              align_corners=True)
            gf, gfl = torch.autograd.grad(y, (f, fl), self._nchw(gout))
        return gf.permute(0, 2, 3, 1).contiguous(), gfl.permute(0, 2, 3, 1).contiguous()

    # ---- entropy epilogues (contract of vcm_dual_prior_pass0/1) ------------------------------
    @staticmethod
    def _index(table, sc):
        s = torch.maximum(sc, torch.zeros_like(sc) + 1e-5)
        i = (torch.log(s) - table.log_scale_min) / table.log_scale_step
        return i.clamp_(0, 255).int().to(torch.uint8)

    @staticmethod
    def _masks(h, w):
        yy, xx = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
        m0 = ((yy + xx) % 2 == 0)[None, None]
        return m0, ~m0

    def dp_pass0(self, mode, table, fusion, params, C, y=None, yq=None, sym=None, idx=None,
                 y_q_full=None, scales_full=None):
        f = self._nchw(fusion)
        qs = torch.clamp_min(f[:, :C], 0.5)
        sc, mu = f[:, C:2 * C], f[:, 2 * C:3 * C]
        m0, m1 = self._masks(f.shape[2], f.shape[3])
        act = torch.cat([m0.expand(-1, C // 2, -1, -1), m1.expand(-1, C // 2, -1, -1)], 1)
        P = self._nchw(params)
        P[:, C:2 * C], P[:, 2 * C:3 * C], P[:, 3 * C:4 * C] = mu, sc, qs
        h = C // 2
        if mode == 0:
            v = self._nchw(y) / qs
            yq.copy_(v.permute(0, 2, 3, 1))
            q = torch.round(v - mu) * act
            sym.copy_((q[:, :h] + q[:, h:]).to(torch.int16))
            idx.copy_(self._index(table, sc[:, :h] * m0 + sc[:, h:] * m1))
        elif mode == 1:
            idx.copy_(self._index(table, sc[:, :h] * m0 + sc[:, h:] * m1))
        else:
            s = sym.float()
            q = torch.cat([s, s], 1) * act
        if mode != 1:
            P[:, :C] = (q + mu) * act
            if y_q_full is not None:
                y_q_full.copy_(q.permute(0, 2, 3, 1))
                scales_full.copy_((sc * act).permute(0, 2, 3, 1))
        self._store(params, P)

    def dp_pass1(self, mode, table, fusion, prior, params, C, q_outer=None, y_hat=None, yq=None,
                 sym=None, idx=None, y_q_full=None, scales_full=None):
        p = self._nchw(prior)
        h = C // 2
        sc = torch.cat([p[:, 0:h], p[:, C:C + h]], 1)
        mu = torch.cat([p[:, h:C], p[:, C + h:2 * C]], 1)
        m0, m1 = self._masks(p.shape[2], p.shape[3])
        act = torch.cat([m1.expand(-1, h, -1, -1), m0.expand(-1, h, -1, -1)], 1)
        P = self._nchw(params)
        if mode == 0:
            v = yq.permute(0, 3, 1, 2)
            q = torch.round(v - mu) * act
            sym.copy_((q[:, :h] + q[:, h:]).to(torch.int16))
            idx.copy_(self._index(table, sc[:, :h] * m1 + sc[:, h:] * m0))
        elif mode == 1:
            idx.copy_(self._index(table, sc[:, :h] * m1 + sc[:, h:] * m0))
            return
        else:
            s = sym.float()
            q = torch.cat([s, s], 1) * act
        part = (q + mu) * act
        out = (P[:, :C] + part) * P[:, 3 * C:4 * C] * q_outer[:, :, None, None]
        self._store(y_hat, out)
        if y_q_full is not None:
            a = act.expand_as(q).permute(0, 2, 3, 1)
            y_q_full[a] = q.permute(0, 2, 3, 1)[a]
            scales_full[a] = sc.permute(0, 2, 3, 1)[a]

    def round_latent(self, z, z_hat=None, sym=None):
        q = torch.round(self._nchw(z))
        if z_hat is None:
            z_hat = self.new(z.N, z.H, z.W, z.C, dtype=z.dtype)
        self._store(z_hat, q)
        if sym is not None:
            sym.copy_(q.to(torch.int16))
        return z_hat

    def symbols_to_latent(self, sym, N, H, W, C, out=None):
        if out is None:
            out = self.new(N, H, W, C)
        self._store(out, sym.float())
        return out

    @staticmethod
    def _bits(p):
        return torch.clamp_min(-1.0 * torch.log(p + 1e-5) / math.log(2.0), 0)

    def latent_bits(self, y_q_full, scales_full, N, dist):
        if dist == "laplace":
            d = torch.distributions.laplace.Laplace(torch.zeros_like(scales_full), scales_full.clamp(1e-5, 1e10))
        else:
            d = torch.distributions.normal.Normal(torch.zeros_like(scales_full), scales_full.clamp(0.11, 1e10))
        p = d.cdf(y_q_full + 0.5) - d.cdf(y_q_full - 0.5)
        return self._bits(p).reshape(N, -1).sum(1)

    def factorized_bits(self, z_hat, fparams):
        z = self._nchw(z_hat)
        r = lambda i: fparams[i][None, :, None, None]  # noqa: E731

        def cdf(x):
            for l in range(3):
                x = x * F.softplus(r(3 * l)) + r(3 * l + 1)
                x = x + torch.tanh(x) * torch.tanh(r(3 * l + 2))
            return torch.sigmoid(x * F.softplus(r(9)) + r(10))

        return self._bits(cdf(z + 0.5) - cdf(z - 0.5)).reshape(z.shape[0], -1).sum(1)

    def sq_err_sum(self, a, b):
        return ((a.float() - b.float()) ** 2).reshape(a.shape[0], -1).sum(1)
