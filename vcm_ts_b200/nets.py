"""Host-side network graphs of the DCVC-HEM path, expressed as sequences of fused device ops on
NHWC views (backend.py).  This mirrors the *structure* of the reference modules

    video_net.py:74-96 ResBlock, :99-146 MEBasic/ME_Spynet, :149-223 SELayer/ConvBlockResidual/UNet,
    :226-276 get_enc_dec_models / get_hyper_enc_dec_models,
    layers.py:42-127 ResidualBlockWithStride / ResidualBlockUpsample / ResidualBlock,
    video_model.py:17-128 FeatureExtractor / MultiScaleContextFusion / ContextualEncoder /
    ContextualDecoder / ReconGeneration, common_model.py:104-217 dual spatial prior,

but every activation, residual add, PixelShuffle, concat and chunk of those modules is folded into
the epilogue / addressing of the producing convolution.
"""
import torch

F32 = torch.float32


class Graph:
    """Shared building blocks.  `B` is a backend, `pc(name, in_map, ps)` returns packed conv weights,
    `lin(name)` a device fp32 tensor of a Linear weight."""

    def __init__(self, B, pc, lin):
        self.B, self.pc, self.lin = B, pc, lin

    def conv(self, name, x, in_map=None, **kw):
        return self.B.conv(x, self.pc(name, in_map, bool(kw.get("ps", False))), **kw)

    def conv_t(self, name, x, slope2, out=None, out2=None, **kw):
        """Conv whose output feeds a ResBlock: returns (y, y_act).  y_act = leaky_relu(y, slope2)
        is stored by the same epilogue when the backend runs the twin graph (bf16 mode), else None."""
        if not self.B.use_twin:
            return self.conv(name, x, out=out, **kw), None
        return self.conv(name, x, out=out, out2=out2, slope2=slope2, **kw)

    # ---- video_net.py:74-96
    def res_block(self, pre, x, slope=0.01, start_from_relu=True, end_with_relu=False, out=None,
                  res2=None, x_act=None, twin_out=None, out2=None):
        """x + conv2(act(conv1(act(x)))).  x_act: pre-activated twin of x (then conv1 needs no
        pre-activation).  twin_out: slope of an activated twin of the result to emit as well
        (returns (y, y_act))."""
        s = 0.0 if slope < 0.0001 else slope
        if x_act is not None and start_from_relu:
            t = self.conv(pre + ".conv1", x_act, post=s)
        else:
            t = self.conv(pre + ".conv1", x, pre=s if start_from_relu else None, post=s)
        kw = dict(post=s if end_with_relu else None, res=x, res2=res2, out=out)
        if twin_out is not None:
            return self.conv_t(pre + ".conv2", t, twin_out, out2=out2, **kw)
        return self.conv(pre + ".conv2", t, **kw)

    # ---- layers.py:42-73
    def residual_block_with_stride(self, pre, x, in_map=None):
        idn = self.conv(pre + ".downsample", x, in_map=in_map, stride=2)
        t = self.conv(pre + ".conv1", x, in_map=in_map, stride=2, post=0.01)
        return self.conv(pre + ".conv2", t, post=0.1, res=idn)

    def frame_input(self, x_nchw):
        """NCHW fp32 frame -> (NHWC view, in_map).  Tensor-core mode pads the 3 channels to 16 (zero
        weights on the pad) so first-layer convs run on tcgen05; exact mode keeps 3 channels."""
        B = self.B
        if not B.use_twin:
            return B.from_nchw(x_nchw), None
        N, C, H, W = x_nchw.shape
        buf = B.new(N, H, W, 16, zero=True)
        B.from_nchw(x_nchw, out=buf.sl(0, C))
        return buf, list(range(C)) + [-1] * (16 - C)

    # ---- layers.py:76-101
    def residual_block_upsample(self, pre, x):
        idn = self.conv(pre + ".upsample.0", x, ps=True)
        t = self.conv(pre + ".subpel_conv.0", x, ps=True, post=0.01)
        return self.conv(pre + ".conv", t, post=0.1, res=idn)

    # ---- layers.py:104-127
    def residual_block(self, pre, x, slope=0.01):
        t = self.conv(pre + ".conv1", x, post=slope)
        return self.conv(pre + ".conv2", t, post=slope, res=x)

    # ---- video_net.py:226-235
    def analysis(self, pre, x, in_map=None):
        x = self.residual_block_with_stride(pre + ".0", x, in_map)
        x = self.residual_block(pre + ".1", x)
        x = self.residual_block_with_stride(pre + ".2", x)
        x = self.residual_block(pre + ".3", x)
        x = self.residual_block_with_stride(pre + ".4", x)
        x = self.residual_block(pre + ".5", x)
        return self.conv(pre + ".6", x, stride=2)

    # ---- video_net.py:237-246
    def synthesis(self, pre, x, out=None, out_dtype=None):
        x = self.residual_block(pre + ".0", x)
        x = self.residual_block_upsample(pre + ".1", x)
        x = self.residual_block(pre + ".2", x)
        x = self.residual_block_upsample(pre + ".3", x)
        x = self.residual_block(pre + ".4", x)
        x = self.residual_block_upsample(pre + ".5", x)
        x = self.residual_block(pre + ".6", x)
        return self.conv(pre + ".7.0", x, ps=True, out=out, out_dtype=out_dtype)

    # ---- video_net.py:251-262
    def hyper_analysis(self, pre, x):
        x = self.conv(pre + ".0", x, post=0.01)
        x = self.conv(pre + ".2", x, post=0.01)
        x = self.conv(pre + ".4", x, stride=2, post=0.01)
        x = self.conv(pre + ".6", x, post=0.01)
        return self.conv(pre + ".8", x, stride=2)

    # ---- video_net.py:264-274, video_model.py:181-191
    def hyper_synthesis(self, pre, x, out=None):
        x = self.conv(pre + ".0", x, post=0.01)
        x = self.conv(pre + ".2.0", x, ps=True, post=0.01)
        x = self.conv(pre + ".4", x, post=0.01)
        x = self.conv(pre + ".6.0", x, ps=True, post=0.01)
        return self.conv(pre + ".8", x, out=out)

    # ---- prior fusion / spatial prior stacks (video_model.py:150-164,199-213)
    def three_convs(self, pre, x, slope=0.2):
        x = self.conv(pre + ".0", x, post=slope)
        x = self.conv(pre + ".2", x, post=slope)
        return self.conv(pre + ".4", x)

    # ---- video_net.py:149-179: up_dim(x) + conv(x) * se(conv(x))
    def conv_block_residual(self, pre, x, out=None, twin_out=None):
        t = self.conv(pre + ".conv.0", x, post=0.01)
        w1, w2 = self.lin(pre + ".conv.3.fc.0.weight"), self.lin(pre + ".conv.3.fc.2.weight")
        if getattr(self.B, "fused_se_sums", False):
            # tensor-core mode: the SE mean comes out of the epilogue of the conv that produces x1 where the
            # engine can do it (one image, whole tiles); no extra pass over x1
            x1, sums = self.conv(pre + ".conv.2", t, se_sums=True)
            s = self.B.se_scale_from_sums(sums, x1, w1, w2) if sums is not None else self.B.se_scale(x1, w1, w2)
        else:
            x1 = self.conv(pre + ".conv.2", t)
            s = self.B.se_scale(x1, w1, w2)
        if twin_out is not None:
            return self.conv_t(pre + ".up_dim", x, twin_out, res=x1, res_scale=s, out=out)
        return self.conv(pre + ".up_dim", x, res=x1, res_scale=s, out=out)

    # ---- video_net.py:182-223
    def unet(self, pre, x, out=None):
        B = self.B
        N, H, W = x.N, x.H, x.W
        cat2 = B.new(N, H, W, 64, dtype=x.dtype)             # (x1 | d2)
        cat3 = B.new(N, H // 2, W // 2, 128, dtype=x.dtype)  # (x2 | d3)
        x1 = self.conv_block_residual(pre + ".conv1", x, out=cat2.sl(0, 32))
        x2 = self.conv_block_residual(pre + ".conv2", B.maxpool(x1), out=cat3.sl(0, 64))
        x3, x3a = self.conv_block_residual(pre + ".conv3", B.maxpool(x2), twin_out=0.0)
        for i in range(4):
            if i < 3:
                x3, x3a = self.res_block(f"{pre}.context_refine.{i}", x3, slope=0, x_act=x3a, twin_out=0.0)
            else:
                x3 = self.res_block(f"{pre}.context_refine.{i}", x3, slope=0, x_act=x3a)
        self.conv(pre + ".up3.0", x3, ps=True, out=cat3.sl(64, 128))
        d3 = self.conv_block_residual(pre + ".up_conv3", cat3)
        self.conv(pre + ".up2.0", d3, ps=True, out=cat2.sl(32, 64))
        return self.conv_block_residual(pre + ".up_conv2", cat2, out=out)

    # ---- video_net.py:118-146 (ME_Spynet) with :99-115 (MEBasic)
    def spynet(self, pre, x_nchw, ref_nchw):
        """x, ref: NCHW fp32 frames -> flow view [N,H,W,2] fp32."""
        B = self.B
        N, _, H, W = x_nchw.shape
        L = 4
        if hasattr(B, "spynet_pyramid") and H % 8 == 0 and W % 8 == 0:
            return self._spynet_fused(pre, x_nchw, ref_nchw)
        # level buffers hold cat[im1_l, warp(im2_l, flow_up), flow_up] (8 ch, fp32)
        bufs = [B.new(N, H >> k, W >> k, 8, dtype=F32) for k in range(L)]
        im2 = [B.from_nchw(ref_nchw, dtype=F32)]
        B.from_nchw(x_nchw, out=bufs[0].sl(0, 3))
        for k in range(1, L):
            B.avgpool(bufs[k - 1].sl(0, 3), out=bufs[k].sl(0, 3))
            im2.append(B.avgpool(im2[k - 1]))
        flow = B.new(N, (H >> (L - 1)) // 2, (W >> (L - 1)) // 2, 2, dtype=F32, zero=True)
        for lv in range(L):
            k = L - 1 - lv
            buf = bufs[k]
            flow_up = B.flow_up2(flow, out=buf.sl(6, 8), gain=2.0)
            B.warp(im2[k], flow_up, out=buf.sl(3, 6))
            b = f"{pre}.moduleBasic.{lv}"
            if B.use_twin:
                # tensor-core mode: bf16 copy of the level input padded to 16 channels (zero weights
                # on the pad) so the 7x7 8->32 conv runs on tcgen05 as well
                buf16 = B.new(N, buf.H, buf.W, 16, dtype=B.act_dtype, zero=True)
                B.copy(buf, buf16.sl(0, 8))
                t = self.conv(b + ".conv1", buf16, in_map=list(range(8)) + [-1] * 8, post=0.0)
            else:
                t = self.conv(b + ".conv1", buf, post=0.0, out_dtype=B.act_dtype)
            t = self.conv(b + ".conv2", t, post=0.0)
            t = self.conv(b + ".conv3", t, post=0.0)
            t = self.conv(b + ".conv4", t, post=0.0)
            flow = self.conv(b + ".conv5", t, res=flow_up, out_dtype=F32)
        return flow

    def _spynet_fused(self, pre, x_nchw, ref_nchw):
        """ME_Spynet with the fused glue kernels: ONE pyramid launch for both images and all levels, ONE launch
        per level for flow_up*2 + warp + concat (+ the bf16 16-channel copy the tensor-core conv1 reads):
        5 glue launches instead of 20, bit-identical values."""
        B = self.B
        N, _, H, W = x_nchw.shape
        L = 4
        bufs = [B.new(N, H >> k, W >> k, 8, dtype=F32) for k in range(L)]
        im2 = [B.new(N, H >> k, W >> k, 3, dtype=F32) for k in range(L)]
        B.spynet_pyramid(x_nchw, ref_nchw, bufs, im2)
        flow = B.new(N, (H >> (L - 1)) // 2, (W >> (L - 1)) // 2, 2, dtype=F32, zero=True)
        for lv in range(L):
            k = L - 1 - lv
            buf = bufs[k]
            b = f"{pre}.moduleBasic.{lv}"
            if B.use_twin:
                buf16 = B.new(N, buf.H, buf.W, 16, dtype=B.act_dtype)
                B.spynet_level_input(flow, im2[k], buf, buf16)
                t = self.conv(b + ".conv1", buf16, in_map=list(range(8)) + [-1] * 8, post=0.0)
            else:
                B.spynet_level_input(flow, im2[k], buf)
                t = self.conv(b + ".conv1", buf, post=0.0, out_dtype=B.act_dtype)
            t = self.conv(b + ".conv2", t, post=0.0)
            t = self.conv(b + ".conv3", t, post=0.0)
            t = self.conv(b + ".conv4", t, post=0.0)
            flow = self.conv(b + ".conv5", t, res=buf.sl(6, 8), out_dtype=F32)
        return flow

    # ---- common_model.py:104-180 (forward_dual_prior / compress_dual_prior), both passes
    def dual_prior_encode(self, table, y, fusion, prior_pre, q_outer, want_full=False, y_hat=None):
        """y: latent / outer q; fusion: (q_step, scales, means) view.  Returns dict with the two
        symbol planes, two index planes, y_hat (already * q_step * q_outer; written to `y_hat` when
        given) and, if want_full, y_q / scales_hat for the likelihood kernel."""
        B = self.B
        N, h, w, C = y.N, y.H, y.W, y.C
        params = B.new(N, h, w, 4 * C, dtype=y.dtype)
        yq = B.empty((N, h, w, C), F32)
        plane = (N, C // 2, h, w)
        sym0, sym1 = B.empty(plane, torch.int16), B.empty(plane, torch.int16)
        idx0, idx1 = B.empty(plane, torch.uint8), B.empty(plane, torch.uint8)
        yqf = B.empty((N, h, w, C), F32) if want_full else None
        scf = B.empty((N, h, w, C), F32) if want_full else None
        B.dp_pass0(0, table, fusion, params, C, y=y, yq=yq, sym=sym0, idx=idx0, y_q_full=yqf,
                   scales_full=scf)
        prior = self.three_convs(prior_pre, params)
        if y_hat is None:
            y_hat = B.new(N, h, w, C, dtype=y.dtype)
        B.dp_pass1(0, table, fusion, prior, params, C, q_outer=q_outer, y_hat=y_hat, yq=yq, sym=sym1,
                   idx=idx1, y_q_full=yqf, scales_full=scf)
        return dict(sym=(sym0, sym1), idx=(idx0, idx1), y_hat=y_hat, y_q=yqf, scales_hat=scf)

    # ---- common_model.py:182-217 (decompress_dual_prior)
    def dual_prior_decode_gen(self, table, fusion, C, prior_pre, q_outer, y_hat=None):
        """Generator form: yields each uint8 index plane [N,C/2,h,w] (device) and is sent back the
        int16 symbol plane (device) the host coder decoded for it; returns y_hat.  The two yields are
        the only host hand-offs, so the device work between them can be captured in CUDA graphs
        (graphs.py)."""
        B = self.B
        N, h, w = fusion.N, fusion.H, fusion.W
        params = B.new(N, h, w, 4 * C, dtype=fusion.dtype)
        plane = (N, C // 2, h, w)
        idx = B.empty(plane, torch.uint8)
        B.dp_pass0(1, table, fusion, params, C, idx=idx)
        sym = yield idx
        B.dp_pass0(2, table, fusion, params, C, sym=sym)
        prior = self.three_convs(prior_pre, params)
        idx = B.empty(plane, torch.uint8)
        B.dp_pass1(1, table, fusion, prior, params, C, idx=idx)
        sym = yield idx
        if y_hat is None:
            y_hat = B.new(N, h, w, C, dtype=fusion.dtype)
        B.dp_pass1(2, table, fusion, prior, params, C, q_outer=q_outer, y_hat=y_hat, sym=sym)
        return y_hat

    def dual_prior_decode(self, table, fusion, C, prior_pre, q_outer, decode_plane):
        """decode_plane(idx uint8 device tensor [N,C/2,h,w]) -> int16 device tensor of symbols."""
        gen = self.dual_prior_decode_gen(table, fusion, C, prior_pre, q_outer)
        try:
            idx = next(gen)
            while True:
                idx = gen.send(decode_plane(idx))
        except StopIteration as e:
            return e.value
