"""Training path of the drop-in `DMC`: train-mode
`forward_one_frame` with gradients to every parameter, as the reference's trainers call it
(core/model/dcvc_hem.py:254-322 -> `self.dmc.forward_one_frame(input, dpb, self.dmc.mv_y_q_scale,
self.dmc.y_q_scale)`, `loss.backward()` in core/engine/train_multi.py:203-268, DDP wrap at :174-191).

What runs where
  * every nn.Conv2d -- forward, data gradient and weight gradient -- is a device kernel of
    libvcmts_b200.so behind `torch.autograd.Function`s: forward = the fused conv engines (tcgen05 in
    "bf16" / "parity" precision), data gradient = the SAME engine on the spatially flipped, transposed
    weights (stride-2 convs: on the zero-interleaved output gradient), weight gradient =
    `vcm_conv2d_wgrad`; `flow_warp` forward / backward are `vcm_flow_warp` / `vcm_flow_warp_bwd`;
  * the element-wise glue of the training graph (activations, residual adds, PixelShuffle, concat,
    pooling, the straight-through quantiser, uniform noise, likelihoods -> bits, MSE) are plain torch
    ops on the same device tensors: autograd supplies their backward.  Activations are channels-last
    (NHWC memory), so conv inputs / outputs alias the kernels' NHWC views without copies.

Numerics follow the reference's train mode (common_model.py:38-49: quant = x + (round(x) - x).detach(),
add_noise = x + U(-0.5, 0.5); video_model.py:548-557: bits from the noisy residuals; LowerBound's
custom backward video_net.py:14-28).  There is no CPU path: the conv Functions call the CUDA library.
"""
import math

import torch
import torch.nn.functional as F

from .backend import View

CL = torch.channels_last
F32 = torch.float32


def _view(t):
    """NCHW-shaped tensor -> NHWC View of the same memory (made channels-last first if needed)."""
    t = t.contiguous(memory_format=CL)
    return View(t.permute(0, 2, 3, 1)), t


def _nchw(v):
    """NHWC View over its own buffer -> NCHW-shaped channels-last tensor."""
    return v.dense().permute(0, 3, 1, 2)


class LowerBoundFn(torch.autograd.Function):
    """video_net.py:14-28."""

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


class ConvFn(torch.autograd.Function):
    """y = leaky_relu(conv2d(x, W, b, stride, pad=K//2), post) on the device engines."""

    @staticmethod
    def forward(ctx, x, weight, bias, g, name, stride, post, out_dtype):
        B = g.B
        xv, xc = _view(x)
        pc = g.packed(name, weight, bias)
        y = B.conv(xv, pc, stride=stride, post=post, out_dtype=out_dtype)
        out = _nchw(y)
        ctx.g, ctx.name, ctx.stride, ctx.post = g, name, stride, post
        ctx.save_for_backward(xc, weight, out if post is not None else None)
        return out

    @staticmethod
    def backward(ctx, gy):
        xc, weight, out = ctx.saved_tensors
        g, B, stride = ctx.g, ctx.g.B, ctx.stride
        Cout, Cin, K, _ = weight.shape
        gy = gy.contiguous(memory_format=CL)
        if ctx.post is not None:
            gy = torch.where(out > 0, gy, gy * ctx.post)
        gyv, gyc = _view(gy)
        gx = gw = gb = None
        if ctx.needs_input_grad[2]:
            gb = gyc.sum(dim=(0, 2, 3), dtype=F32).to(weight.dtype)
        if ctx.needs_input_grad[1]:
            xw, gyw, cin_w = xc, gyc, Cin
            if xc.dtype == torch.bfloat16:
                # the tcgen05 weight-gradient kernel wants bf16 operands and 16-aligned input channels: cast
                # an fp32 output gradient, zero-pad a 2/3/8/67-channel input (the pad columns are dropped)
                if gyw.dtype != torch.bfloat16:
                    gyw = gyw.to(torch.bfloat16)
                if Cin % 16:
                    cin_w = (Cin + 15) // 16 * 16
                    xw = F.pad(xc, (0, 0, 0, 0, 0, cin_w - Cin)).contiguous(memory_format=CL)
            xv, _ = _view(xw)
            gwv, _ = _view(gyw)
            gw = B.wgrad(xv, gwv, K, stride, cin_w, Cout)
            if cin_w != Cin:
                gw = gw[:, :Cin]
            gw = gw.to(weight.dtype)
        if ctx.needs_input_grad[0]:
            pcT = g.packed_dgrad(ctx.name, weight)
            src = gyv
            if stride == 2:
                # transposed conv = stride-1 conv of the zero-interleaved gradient (even sizes, pad K//2)
                N, _, H, W = xc.shape
                up = torch.zeros((N, Cout, H, W), dtype=gyc.dtype, device=gyc.device).contiguous(memory_format=CL)
                up[:, :, ::2, ::2] = gyc
                src, _ = _view(up)
            gx = _nchw(B.conv(src, pcT, stride=1, out_dtype=xc.dtype))
        return gx, gw, gb, None, None, None, None, None


class WarpFn(torch.autograd.Function):
    """video_net.py:32-55 flow_warp: bilinear, border, align_corners=True."""

    @staticmethod
    def forward(ctx, feat, flow, g):
        fv, fc = _view(feat)
        lv, lc = _view(flow.to(F32))
        ctx.g = g
        ctx.save_for_backward(fc, lc)
        return _nchw(g.B.warp(fv, lv))

    @staticmethod
    def backward(ctx, gout):
        fc, lc = ctx.saved_tensors
        fv, _ = _view(fc)
        lv, _ = _view(lc)
        gv, _ = _view(gout.to(fc.dtype))
        gfeat, gflow = ctx.g.B.warp_bwd(fv, lv, gv)
        return gfeat.permute(0, 3, 1, 2).to(fc.dtype), gflow.permute(0, 3, 1, 2), None


class TrainGraph:
    """The DMC network of video_model.py / video_net.py / layers.py as a differentiable graph over
    NCHW-shaped channels-last tensors.  `model`: the DMC module (parameters by reference name)."""

    def __init__(self, model, rt):
        self.m, self.rt, self.B = model, rt, rt.B
        self.act = rt.B.act_dtype

    # ---- packed weights, keyed by parameter version (the optimizer step bumps it)
    def packed(self, name, weight, bias):
        key = (name, "train")
        tag = (weight._version, weight.data_ptr(), bias._version, bias.data_ptr())
        hit = self.rt.packed.get(key)
        if hit is None or hit[0] != tag:
            hit = (tag, self.B.pack_conv(weight, bias))
            self.rt.packed[key] = hit
        return hit[1]

    def packed_dgrad(self, name, weight):
        key = (name, "dgrad")
        tag = (weight._version, weight.data_ptr())
        hit = self.rt.packed.get(key)
        if hit is None or hit[0] != tag:
            wt = weight.detach().flip(2, 3).transpose(0, 1).contiguous()     # [Cin, Cout, K, K]
            zero = torch.zeros((wt.shape[0],), dtype=F32, device=weight.device)
            hit = (tag, self.B.pack_conv(wt, zero))
            self.rt.packed[key] = hit
        return hit[1]

    # ---- primitives
    def conv(self, name, x, stride=1, post=None, ps=False, out_dtype=None):
        w = self.m.get_parameter(name + ".weight")
        b = self.m.get_parameter(name + ".bias")
        if x.dtype != self.act:
            x = x.to(self.act)
        y = ConvFn.apply(x, w, b, self, name, stride, post, out_dtype)
        return F.pixel_shuffle(y, 2) if ps else y

    def warp(self, feat, flow):
        return WarpFn.apply(feat, flow, self)

    @staticmethod
    def act_(x, slope):
        return F.relu(x) if slope < 0.0001 else F.leaky_relu(x, slope)

    # ---- video_net.py:74-96
    def res_block(self, pre, x, slope=0.01, start_from_relu=True, end_with_relu=False):
        s = 0.0 if slope < 0.0001 else slope
        t = self.act_(x, s) if start_from_relu else x
        t = self.conv(pre + ".conv1", t, post=s)
        t = self.conv(pre + ".conv2", t, post=s if end_with_relu else None)
        return x + t

    # ---- layers.py:42-127
    def residual_block_with_stride(self, pre, x):
        idn = self.conv(pre + ".downsample", x, stride=2)
        t = self.conv(pre + ".conv1", x, stride=2, post=0.01)
        return self.conv(pre + ".conv2", t, post=0.1) + idn

    def residual_block_upsample(self, pre, x):
        idn = self.conv(pre + ".upsample.0", x, ps=True)
        t = self.conv(pre + ".subpel_conv.0", x, post=0.01, ps=True)
        return self.conv(pre + ".conv", t, post=0.1) + idn

    def residual_block(self, pre, x, slope=0.01):
        t = self.conv(pre + ".conv1", x, post=slope)
        return self.conv(pre + ".conv2", t, post=slope) + x

    # ---- video_net.py:226-276
    def analysis(self, pre, x, out_dtype=None):
        x = self.residual_block_with_stride(pre + ".0", x)
        x = self.residual_block(pre + ".1", x)
        x = self.residual_block_with_stride(pre + ".2", x)
        x = self.residual_block(pre + ".3", x)
        x = self.residual_block_with_stride(pre + ".4", x)
        x = self.residual_block(pre + ".5", x)
        return self.conv(pre + ".6", x, stride=2, out_dtype=out_dtype)

    def synthesis(self, pre, x, out_dtype=None):
        x = self.residual_block(pre + ".0", x)
        x = self.residual_block_upsample(pre + ".1", x)
        x = self.residual_block(pre + ".2", x)
        x = self.residual_block_upsample(pre + ".3", x)
        x = self.residual_block(pre + ".4", x)
        x = self.residual_block_upsample(pre + ".5", x)
        x = self.residual_block(pre + ".6", x)
        return self.conv(pre + ".7.0", x, ps=True, out_dtype=out_dtype)

    def hyper_analysis(self, pre, x):
        x = self.conv(pre + ".0", x, post=0.01)
        x = self.conv(pre + ".2", x, post=0.01)
        x = self.conv(pre + ".4", x, stride=2, post=0.01)
        x = self.conv(pre + ".6", x, post=0.01)
        return self.conv(pre + ".8", x, stride=2, out_dtype=F32)

    def hyper_synthesis(self, pre, x):
        x = self.conv(pre + ".0", x, post=0.01)
        x = self.conv(pre + ".2.0", x, post=0.01, ps=True)
        x = self.conv(pre + ".4", x, post=0.01)
        x = self.conv(pre + ".6.0", x, post=0.01, ps=True)
        return self.conv(pre + ".8", x, out_dtype=F32)

    def three_convs(self, pre, x, slope=0.2):
        x = self.conv(pre + ".0", x, post=slope)
        x = self.conv(pre + ".2", x, post=slope)
        return self.conv(pre + ".4", x, out_dtype=F32)

    # ---- video_net.py:149-223
    def conv_block_residual(self, pre, x):
        t = self.conv(pre + ".conv.0", x, post=0.01)
        x1 = self.conv(pre + ".conv.2", t)
        w1 = self.m.get_parameter(pre + ".conv.3.fc.0.weight")
        w2 = self.m.get_parameter(pre + ".conv.3.fc.2.weight")
        s = torch.sigmoid(F.linear(F.relu(F.linear(x1.float().mean(dim=(-1, -2)), w1)), w2))
        return self.conv(pre + ".up_dim", x) + x1 * s[:, :, None, None].to(x1.dtype)

    def unet(self, pre, x):
        x1 = self.conv_block_residual(pre + ".conv1", x)
        x2 = self.conv_block_residual(pre + ".conv2", F.max_pool2d(x1, 2, 2))
        x3 = self.conv_block_residual(pre + ".conv3", F.max_pool2d(x2, 2, 2))
        for i in range(4):
            x3 = self.res_block(f"{pre}.context_refine.{i}", x3, slope=0)
        d3 = self.conv(pre + ".up3.0", x3, ps=True)
        d3 = self.conv_block_residual(pre + ".up_conv3", torch.cat((x2, d3), dim=1))
        d2 = self.conv(pre + ".up2.0", d3, ps=True)
        return self.conv_block_residual(pre + ".up_conv2", torch.cat((x1, d2), dim=1))

    # ---- video_net.py:99-146
    def spynet(self, pre, im1, im2):
        L = 4
        l1, l2 = [im1], [im2]
        for k in range(L - 1):
            l1.append(F.avg_pool2d(l1[k], 2, 2))
            l2.append(F.avg_pool2d(l2[k], 2, 2))
        N, _, h, w = l2[L - 1].shape
        flow = torch.zeros((N, 2, h // 2, w // 2), dtype=F32, device=im1.device)
        for lv in range(L):
            up = F.interpolate(flow, (flow.shape[2] * 2, flow.shape[3] * 2), mode="bilinear", align_corners=False) * 2.0
            k = L - 1 - lv
            t = torch.cat([l1[k], self.warp(l2[k], up), up], 1)
            b = f"{pre}.moduleBasic.{lv}"
            t = self.conv(b + ".conv1", t, post=0.0)
            t = self.conv(b + ".conv2", t, post=0.0)
            t = self.conv(b + ".conv3", t, post=0.0)
            t = self.conv(b + ".conv4", t, post=0.0)
            flow = up + self.conv(b + ".conv5", t, out_dtype=F32)
        return flow

    # ---- common_model.py:38-49, 82-177
    def quant(self, x):
        return x + (torch.round(x) - x).detach()

    def _masked(self, y, scales, means, mask):
        scales_hat = scales * mask
        means_hat = means * mask
        y_res = (y - means_hat) * mask
        y_q = self.quant(y_res)
        return y_res, y_q, y_q + means_hat, scales_hat

    def dual_prior(self, y, means, scales, quant_step, prior_pre):
        _, _, H, W = y.shape
        micro = torch.tensor(((1, 0), (0, 1)), dtype=F32, device=y.device)
        m0 = micro.repeat(H // 2, W // 2)[None, None]
        m1 = torch.ones_like(m0) - m0
        quant_step = LowerBoundFn.apply(quant_step, 0.5)
        y = y / quant_step
        y_0, y_1 = y.chunk(2, 1)
        s0, s1 = scales.chunk(2, 1)
        u0, u1 = means.chunk(2, 1)
        r00, q00, h00, sh00 = self._masked(y_0, s0, u0, m0)
        r11, q11, h11, sh11 = self._masked(y_1, s1, u1, m1)
        params = torch.cat((h00, h11, means, scales, quant_step), dim=1)
        s0, u0, s1, u1 = self.three_convs(prior_pre, params).chunk(4, 1)
        r01, q01, h01, sh01 = self._masked(y_0, s0, u0, m1)
        r10, q10, h10, sh10 = self._masked(y_1, s1, u1, m0)
        y_res = torch.cat((r00 + r01, r11 + r10), dim=1)
        y_q = torch.cat((q00 + q01, q11 + q10), dim=1)
        y_hat = torch.cat((h00 + h01, h11 + h10), dim=1) * quant_step
        scales_hat = torch.cat((sh00 + sh01, sh11 + sh10), dim=1)
        return y_res, y_q, y_hat, scales_hat

    # ---- common_model.py:51-73, entropy_models.py:54-117
    @staticmethod
    def probs_to_bits(probs):
        bits = -1.0 * torch.log(probs + 1e-5) / math.log(2.0)
        return LowerBoundFn.apply(bits, 0)

    def laplace_bits(self, y, sigma):
        sigma = sigma.clamp(1e-5, 1e10)
        d = torch.distributions.laplace.Laplace(torch.zeros_like(sigma), sigma)
        return self.probs_to_bits(d.cdf(y + 0.5) - d.cdf(y - 0.5))

    def gaussian_bits(self, y, sigma):
        sigma = sigma.clamp(0.11, 1e10)
        d = torch.distributions.normal.Normal(torch.zeros_like(sigma), sigma)
        return self.probs_to_bits(d.cdf(y + 0.5) - d.cdf(y - 0.5))

    def z_bits(self, z, name):
        def cdf(x):
            for f in ("f1", "f2", "f3", "f4"):
                h = self.m.get_parameter(f"{name}.{f}.h")
                b = self.m.get_parameter(f"{name}.{f}.b")
                x = x * F.softplus(h) + b
                if f != "f4":
                    x = x + torch.tanh(x) * torch.tanh(self.m.get_parameter(f"{name}.{f}.a"))
            return torch.sigmoid(x)
        return self.probs_to_bits(cdf(z + 0.5) - cdf(z - 0.5))

    # ---- video_model.py:17-68, 226-246
    def feature_pyramid(self, dpb):
        if dpb["ref_feature"] is None:
            f = self.conv("feature_adaptor_I", dpb["ref_frame"])
        else:
            f = self.conv("feature_adaptor_P", dpb["ref_feature"])
        n = "feature_extractor"
        l1 = self.res_block(n + ".res_block1", self.conv(n + ".conv1", f))
        l2 = self.res_block(n + ".res_block2", self.conv(n + ".conv2", l1, stride=2))
        l3 = self.res_block(n + ".res_block3", self.conv(n + ".conv3", l2, stride=2))
        return l1, l2, l3

    def context_fusion(self, c1, c2, c3):
        n = "context_fusion_net"
        c3_up = self.res_block(n + ".res_block3_up", self.conv(n + ".conv3_up.0", c3, ps=True))
        c3_out = self.res_block(n + ".res_block3_out", self.conv(n + ".conv3_out", c3))
        cat2 = torch.cat((c3_up, c2), dim=1)
        c2_up = self.res_block(n + ".res_block2_up", self.conv(n + ".conv2_up.0", cat2, ps=True))
        c2_out = self.res_block(n + ".res_block2_out", self.conv(n + ".conv2_out", cat2))
        c1_out = self.res_block(n + ".res_block1_out", self.conv(n + ".conv1_out", torch.cat((c2_up, c1), dim=1)))
        return c1 + c1_out, c2 + c2_out, c3 + c3_out

    def motion_compensation(self, dpb, mv):
        warp_frame = self.warp(dpb["ref_frame"].to(F32), mv)
        down = lambda t: F.interpolate(t, (t.shape[2] // 2, t.shape[3] // 2), mode="bilinear",   # noqa: E731
                                       align_corners=False)
        mv2 = down(mv) / 2
        mv3 = down(mv2) / 2
        f1, f2, f3 = self.feature_pyramid(dpb)
        c1, c2, c3 = self.warp(f1, mv), self.warp(f2, mv2), self.warp(f3, mv3)
        return (*self.context_fusion(c1, c2, c3), warp_frame)

    # ---- video_model.py:71-128
    def contextual_encoder(self, x, c1, c2, c3):
        n = "contextual_encoder"
        f = self.conv(n + ".conv1", torch.cat([x.to(c1.dtype), c1], dim=1), stride=2)
        f = self.res_block(n + ".res1", torch.cat([f, c2], dim=1), 0.1, True, True)
        f = self.conv(n + ".conv2", f, stride=2)
        f = self.res_block(n + ".res2", torch.cat([f, c3], dim=1), 0.1, True, True)
        f = self.conv(n + ".conv3", f, stride=2)
        return self.conv(n + ".conv4", f, stride=2, out_dtype=F32)

    def contextual_decoder(self, y_hat, c2, c3):
        n = "contextual_decoder"
        f = self.conv(n + ".up1.0", y_hat, ps=True)
        f = self.conv(n + ".up2.0", f, ps=True)
        f = self.res_block(n + ".res1", torch.cat([f, c3], dim=1), 0.1, True, True)
        f = self.conv(n + ".up3.0", f, ps=True)
        f = self.res_block(n + ".res2", torch.cat([f, c2], dim=1), 0.1, True, True)
        return self.conv(n + ".up4.0", f, ps=True)

    def recon_generation(self, res, ctx):
        # video_model.py:330 calls recon_generation_net(recon_image_feature, context1); ReconGeneration
        # concatenates its first argument first (:124): (32 decoder channels | 64 context channels)
        n = "recon_generation_net"
        f = self.conv(n + ".first_conv", torch.cat((res, ctx), dim=1))
        f = self.unet(n + ".unet_1", f)
        feature = self.unet(n + ".unet_2", f)
        return feature, self.conv(n + ".recon_conv", feature, out_dtype=F32)

    # ---- video_model.py:470-592 (training branch)
    def forward_one_frame(self, x, dpb, mv_y_q_scale, y_q_scale, noise_fn):
        m = self.m
        x = x.to(F32)
        curr_mv_y_q = LowerBoundFn.apply(m.mv_y_q_basic, 0.5) * mv_y_q_scale
        curr_y_q = LowerBoundFn.apply(m.y_q_basic, 0.5) * y_q_scale
        est_mv = self.spynet("optic_flow", x, dpb["ref_frame"].to(F32))
        mv_y = self.analysis("mv_encoder", est_mv, out_dtype=F32) / curr_mv_y_q
        mv_z = self.hyper_analysis("mv_hyper_prior_encoder", mv_y)
        mv_z_hat = self.quant(mv_z)
        mv_params = self.hyper_synthesis("mv_hyper_prior_decoder", mv_z_hat)
        ref_mv_y = dpb["ref_mv_y"] if dpb["ref_mv_y"] is not None else torch.zeros_like(mv_y)
        mv_q_step, mv_scales, mv_means = self.three_convs(
            "mv_y_prior_fusion", torch.cat((mv_params, ref_mv_y.to(F32)), dim=1)).chunk(3, 1)
        mv_y_res, mv_y_q, mv_y_hat, mv_scales_hat = self.dual_prior(mv_y, mv_means, mv_scales, mv_q_step,
                                                                    "mv_y_spatial_prior")
        mv_y_hat = mv_y_hat * curr_mv_y_q
        mv_hat = self.synthesis("mv_decoder", mv_y_hat, out_dtype=F32)
        c1, c2, c3, warp_frame = self.motion_compensation(dpb, mv_hat)
        y = self.contextual_encoder(x, c1, c2, c3) / curr_y_q
        n = "contextual_hyper_prior_encoder"
        t = self.conv(n + ".0", y, post=0.01)
        t = self.conv(n + ".2", t, stride=2, post=0.01)
        z = self.conv(n + ".4", t, stride=2, out_dtype=F32)
        z_hat = self.quant(z)
        hier = self.hyper_synthesis("contextual_hyper_prior_decoder", z_hat)
        t = self.conv("temporal_prior_encoder.0", c3, stride=2, post=0.1)
        temporal = self.conv("temporal_prior_encoder.2", t, stride=2, out_dtype=F32)
        ref_y = dpb["ref_y"] if dpb["ref_y"] is not None else torch.zeros_like(y)
        q_step, scales, means = self.three_convs(
            "y_prior_fusion", torch.cat((temporal, hier, ref_y.to(F32)), dim=1)).chunk(3, 1)
        y_res, y_q, y_hat, scales_hat = self.dual_prior(y, means, scales, q_step, "y_spatial_prior")
        y_hat = y_hat * curr_y_q
        res = self.contextual_decoder(y_hat, c2, c3)
        feature, recon = self.recon_generation(res, c1)

        B_, _, H, W = x.shape
        pixel_num = H * W
        mse = torch.sum((x - recon) ** 2, dim=(1, 2, 3)) / pixel_num
        me_mse = torch.sum((x - warp_frame.to(F32)) ** 2, dim=(1, 2, 3)) / pixel_num
        if m.training:
            y_for_bit = y_res + noise_fn(y_res)
            mv_y_for_bit = mv_y_res + noise_fn(mv_y_res)
            z_for_bit = z + noise_fn(z)
            mv_z_for_bit = mv_z + noise_fn(mv_z)
        else:
            y_for_bit, mv_y_for_bit, z_for_bit, mv_z_for_bit = y_q, mv_y_q, z_hat, mv_z_hat
        bpp_y = torch.sum(self.laplace_bits(y_for_bit, scales_hat), dim=(1, 2, 3)) / pixel_num
        bpp_mv_y = torch.sum(self.laplace_bits(mv_y_for_bit, mv_scales_hat), dim=(1, 2, 3)) / pixel_num
        bpp_z = torch.sum(self.z_bits(z_for_bit, "bit_estimator_z"), dim=(1, 2, 3)) / pixel_num
        bpp_mv_z = torch.sum(self.z_bits(mv_z_for_bit, "bit_estimator_z_mv"), dim=(1, 2, 3)) / pixel_num
        bpp = bpp_y + bpp_z + bpp_mv_y + bpp_mv_z
        return {"bpp_mv_y": bpp_mv_y, "bpp_mv_z": bpp_mv_z, "bpp_y": bpp_y, "bpp_z": bpp_z, "bpp": bpp,
                "me_mse": me_mse, "mse": mse,
                "dpb": {"ref_frame": recon, "ref_feature": feature.to(F32), "ref_y": y_hat, "ref_mv_y": mv_y_hat},
                "bit": torch.sum(bpp) * pixel_num, "bit_y": torch.sum(bpp_y) * pixel_num,
                "bit_z": torch.sum(bpp_z) * pixel_num, "bit_mv_y": torch.sum(bpp_mv_y) * pixel_num,
                "bit_mv_z": torch.sum(bpp_mv_z) * pixel_num}


    # ---- image_model.py:54-105 (IntraNoAR.forward, training branch)
    def intra_forward(self, x, q_scale, noise_fn):
        m = self.m
        x = x.to(F32)
        curr_q = LowerBoundFn.apply(m.q_basic, 0.5) * q_scale
        y = self.analysis("enc", x, out_dtype=F32) / curr_q
        z = self.hyper_analysis("hyper_enc", y)
        z_hat = self.quant(z)
        params = self.hyper_synthesis("hyper_dec", z_hat)
        q_step, scales, means = self.three_convs("y_prior_fusion", params).chunk(3, 1)
        y_res, y_q, y_hat, scales_hat = self.dual_prior(y, means, scales, q_step, "y_spatial_prior")
        y_hat = y_hat * curr_q
        t = self.synthesis("dec", y_hat)
        x_hat = self.conv("refine.1", self.unet("refine.0", t), out_dtype=F32)
        if m.training:
            y_for_bit, z_for_bit = y_res + noise_fn(y_res), z + noise_fn(z)
        else:
            y_for_bit, z_for_bit = y_q, z_hat
        _, _, H, W = x.shape
        pixel_num = H * W
        bpp_y = torch.sum(self.gaussian_bits(y_for_bit, scales_hat), dim=(1, 2, 3)) / pixel_num
        bpp_z = torch.sum(self.z_bits(z_for_bit, "bit_estimator_z"), dim=(1, 2, 3)) / pixel_num
        mse = torch.sum((x - x_hat) ** 2, dim=(1, 2, 3)) / pixel_num
        bits = torch.sum(bpp_y + bpp_z) * pixel_num
        return {"x_hat": x_hat, "mse": mse, "bit": bits.item(), "bpp": bpp_y + bpp_z, "bpp_y": bpp_y, "bpp_z": bpp_z}


def default_noise(x):
    """common_model.py:46-49 add_noise: U(-0.5, 0.5), detached."""
    return torch.empty_like(x).uniform_(-0.5, 0.5)
