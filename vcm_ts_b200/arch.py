"""Parameter layout of the drop-in modules: the state-dict keys and shapes of the reference's
`DMC` (video_model.py:144-222) and `IntraNoAR` (image_model.py:20-45) module trees, generated from a
compact layer table so checkpoints written for the reference load unchanged.
"""

CONV, LINEAR, BITPARM, QPARAM = "conv", "linear", "bitparm", "q"


class Spec:
    """Ordered list of (kind, key-prefix, meta)."""

    def __init__(self):
        self.items = []

    def conv(self, name, cin, cout, k):
        self.items.append((CONV, name, (cin, cout, k)))

    def linear(self, name, cin, cout):
        self.items.append((LINEAR, name, (cin, cout)))

    def bit_estimator(self, name, ch):
        self.items.append((BITPARM, name, (ch,)))

    def q(self, name, shape):
        self.items.append((QPARAM, name, tuple(shape)))

    # ---- composite blocks, named as the reference's submodules are indexed
    def res_stride(self, p, cin, cout):              # layers.py:42-73
        self.conv(p + ".conv1", cin, cout, 3)
        self.conv(p + ".conv2", cout, cout, 3)
        self.conv(p + ".downsample", cin, cout, 1)

    def res_plain(self, p, ch):                      # layers.py:104-127
        self.conv(p + ".conv1", ch, ch, 3)
        self.conv(p + ".conv2", ch, ch, 3)

    def res_up(self, p, ch):                         # layers.py:76-101
        self.conv(p + ".subpel_conv.0", ch, ch * 4, 1)
        self.conv(p + ".conv", ch, ch, 3)
        self.conv(p + ".upsample.0", ch, ch * 4, 1)

    def enc_dec(self, enc, dec, cin, cout, ch):      # video_net.py:226-248
        for i, kind in enumerate("sPsPsP"):
            if kind == "s":
                self.res_stride(f"{enc}.{i}", cin if i == 0 else ch, ch)
            else:
                self.res_plain(f"{enc}.{i}", ch)
        self.conv(enc + ".6", ch, ch, 3)
        for i, kind in enumerate("PuPuPuP"):
            (self.res_plain if kind == "P" else self.res_up)(f"{dec}.{i}", ch)
        self.conv(dec + ".7.0", ch, cout * 4, 1)

    def hyper(self, enc, dec, yc, zc):               # video_net.py:251-276
        self.conv(enc + ".0", yc, zc, 3)
        for i in (2, 4, 6, 8):
            self.conv(f"{enc}.{i}", zc, zc, 3)
        m = yc * 3 // 2
        self.conv(dec + ".0", zc, yc, 3)
        self.conv(dec + ".2.0", yc, yc * 4, 1)
        self.conv(dec + ".4", yc, m, 3)
        self.conv(dec + ".6.0", m, m * 4, 1)
        self.conv(dec + ".8", m, yc * 2, 3)

    def res_block(self, p, ch, bottleneck=False):    # video_net.py:74-96
        mid = ch // 2 if bottleneck else ch
        self.conv(p + ".conv1", ch, mid, 3)
        self.conv(p + ".conv2", mid, ch, 3)

    def conv_block_residual(self, p, cin, cout):     # video_net.py:149-179
        self.conv(p + ".conv.0", cin, cout, 3)
        self.conv(p + ".conv.2", cout, cout, 3)
        self.linear(p + ".conv.3.fc.0", cout, cout // 16)
        self.linear(p + ".conv.3.fc.2", cout // 16, cout)
        self.conv(p + ".up_dim", cin, cout, 1)

    def unet(self, p, cin, cout):                    # video_net.py:182-223
        self.conv_block_residual(p + ".conv1", cin, 32)
        self.conv_block_residual(p + ".conv2", 32, 64)
        self.conv_block_residual(p + ".conv3", 64, 128)
        for i in range(4):
            self.res_block(f"{p}.context_refine.{i}", 128)
        self.conv(p + ".up3.0", 128, 256, 1)
        self.conv_block_residual(p + ".up_conv3", 128, 64)
        self.conv(p + ".up2.0", 64, 128, 1)
        self.conv_block_residual(p + ".up_conv2", 64, cout)

    def stack3(self, p, chans):
        for i, (a, b) in zip((0, 2, 4), chans):
            self.conv(f"{p}.{i}", a, b, 3)


def dmc_spec(anchor_num=4):
    """video_model.py:131-224."""
    s = Spec()
    s.bit_estimator("bit_estimator_z", 64)
    s.bit_estimator("bit_estimator_z_mv", 64)
    for lv in range(4):                              # video_net.py:99-127
        b = f"optic_flow.moduleBasic.{lv}"
        for i, (a, c) in enumerate(((8, 32), (32, 64), (64, 32), (32, 16), (16, 2)), 1):
            s.conv(f"{b}.conv{i}", a, c, 7)
    s.enc_dec("mv_encoder", "mv_decoder", 2, 2, 64)
    s.hyper("mv_hyper_prior_encoder", "mv_hyper_prior_decoder", 64, 64)
    s.stack3("mv_y_prior_fusion", ((192, 192), (192, 192), (192, 192)))
    s.stack3("mv_y_spatial_prior", ((256, 192), (192, 192), (192, 128)))
    s.conv("feature_adaptor_I", 3, 64, 3)
    s.conv("feature_adaptor_P", 64, 64, 1)
    for i in (1, 2, 3):                              # video_model.py:17-37
        s.conv(f"feature_extractor.conv{i}", 64, 64, 3)
        s.res_block(f"feature_extractor.res_block{i}", 64)
    n = "context_fusion_net"                         # video_model.py:40-68
    s.conv(n + ".conv3_up.0", 64, 256, 3)
    s.res_block(n + ".res_block3_up", 64)
    s.conv(n + ".conv3_out", 64, 64, 3)
    s.res_block(n + ".res_block3_out", 64)
    s.conv(n + ".conv2_up.0", 128, 256, 3)
    s.res_block(n + ".res_block2_up", 64)
    s.conv(n + ".conv2_out", 128, 64, 3)
    s.res_block(n + ".res_block2_out", 64)
    s.conv(n + ".conv1_out", 128, 64, 3)
    s.res_block(n + ".res_block1_out", 64)
    n = "contextual_encoder"                         # video_model.py:71-90
    s.conv(n + ".conv1", 67, 64, 3)
    s.res_block(n + ".res1", 128, True)
    s.conv(n + ".conv2", 128, 64, 3)
    s.res_block(n + ".res2", 128, True)
    s.conv(n + ".conv3", 128, 64, 3)
    s.conv(n + ".conv4", 64, 96, 3)
    n = "contextual_hyper_prior_encoder"             # video_model.py:173-179
    s.conv(n + ".0", 96, 64, 3)
    s.conv(n + ".2", 64, 64, 3)
    s.conv(n + ".4", 64, 64, 3)
    n = "contextual_hyper_prior_decoder"             # video_model.py:181-191
    s.conv(n + ".0", 64, 96, 3)
    s.conv(n + ".2.0", 96, 384, 1)
    s.conv(n + ".4", 96, 144, 3)
    s.conv(n + ".6.0", 144, 576, 1)
    s.conv(n + ".8", 144, 192, 3)
    s.conv("temporal_prior_encoder.0", 64, 144, 3)   # video_model.py:193-197
    s.conv("temporal_prior_encoder.2", 144, 192, 3)
    s.stack3("y_prior_fusion", ((480, 384), (384, 288), (288, 288)))
    s.stack3("y_spatial_prior", ((384, 288), (288, 288), (288, 192)))
    n = "contextual_decoder"                         # video_model.py:93-112
    s.conv(n + ".up1.0", 96, 256, 3)
    s.conv(n + ".up2.0", 64, 256, 3)
    s.res_block(n + ".res1", 128, True)
    s.conv(n + ".up3.0", 128, 256, 3)
    s.res_block(n + ".res2", 128, True)
    s.conv(n + ".up4.0", 128, 128, 3)
    n = "recon_generation_net"                       # video_model.py:115-128
    s.conv(n + ".first_conv", 96, 64, 3)
    s.unet(n + ".unet_1", 64, 64)
    s.unet(n + ".unet_2", 64, 64)
    s.conv(n + ".recon_conv", 64, 3, 3)
    s.q("mv_y_q_basic", (1, 64, 1, 1))
    s.q("mv_y_q_scale", (anchor_num, 1, 1, 1))
    s.q("y_q_basic", (1, 96, 1, 1))
    s.q("y_q_scale", (anchor_num, 1, 1, 1))
    return s


def intra_spec(N=192, anchor_num=4):
    """image_model.py:16-48."""
    s = Spec()
    s.bit_estimator("bit_estimator_z", N)
    s.enc_dec("enc", "dec", 3, 16, N)
    s.unet("refine.0", 16, 16)
    s.conv("refine.1", 16, 3, 3)
    s.hyper("hyper_enc", "hyper_dec", N, N)
    s.stack3("y_prior_fusion", ((2 * N, 3 * N), (3 * N, 3 * N), (3 * N, 3 * N)))
    s.stack3("y_spatial_prior", ((4 * N, 3 * N), (3 * N, 3 * N), (3 * N, 2 * N)))
    s.q("q_basic", (1, N, 1, 1))
    s.q("q_scale", (anchor_num, 1, 1, 1))
    return s


def state_shapes(spec):
    """{state-dict key: shape}."""
    out = {}
    for kind, name, meta in spec.items:
        if kind == CONV:
            cin, cout, k = meta
            out[name + ".weight"] = (cout, cin, k, k)
            out[name + ".bias"] = (cout,)
        elif kind == LINEAR:
            cin, cout = meta
            out[name + ".weight"] = (cout, cin)
        elif kind == BITPARM:
            (ch,) = meta
            for f in ("f1", "f2", "f3", "f4"):
                out[f"{name}.{f}.h"] = (1, ch, 1, 1)
                out[f"{name}.{f}.b"] = (1, ch, 1, 1)
                if f != "f4":
                    out[f"{name}.{f}.a"] = (1, ch, 1, 1)
        else:
            out[name] = meta
    return out
