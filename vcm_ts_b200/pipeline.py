"""GOP coding with the three stages of `encode_decode` overlapped ACROSS frames of one sequence:

    device encoder (analysis nets, stream A)  ->  host rANS encoder  ->  device+host decoder (stream B)

`DMC.encode_decode` (video_model.py:424-468) runs them back to back per frame, so the GPU idles while
the sequential entropy coder works and vice versa.  Here frame t+1 is analysed while frame t is in
the host coder and frame t-1 is being decoded.  That is legal because the encoder's own
reconstruction is bit-identical to the decoder's (same deterministic kernels; asserted by
tests/test_gpu_e2e.py), so the encoder can predict from it without waiting for the decoder -- what
every practical video encoder does.  Every frame still goes through a real bitstream, and the
reconstructions handed back are the DECODER's.  The I-frame is coded serially (its reconstruction
only exists on the decoder side, image_model.py:148-200).

Two model instances are needed (encoder side / decoder side: each owns per-frame scratch buffers).
"""
import queue
import threading

import torch

from .stream_helper import get_rounded_q, pack_p, unpack_p


class GopPipeline:
    def __init__(self, enc_models, dec_models, device):
        """enc_models / dec_models: (DMC, IntraNoAR) pairs with identical weights, on `device`,
        `update()`d."""
        self.enc_dmc, self.enc_intra = enc_models
        self.dec_dmc, self.dec_intra = dec_models
        self.device = torch.device(device)
        # the decoder chain (host rANS hand-offs in series with device work) is the critical path of
        # a single sequence: its kernels take SMs ahead of the encoder's, which runs ahead anyway
        self.s_enc = torch.cuda.Stream(self.device)
        self.s_dec = torch.cuda.Stream(self.device, priority=-1)

    def code_gop(self, frames, pic_height, pic_width, q_i, q_mv, q_y, on_recon=None):
        """frames: sequence of callables or tensors giving the padded [1,3,Hp,Wp] frame t (called /
        used on the encoder thread, on its stream).  on_recon(t, x_hat_nchw) runs on the decoder thread
        (its stream) for every decoded frame.  Returns the list of container blobs (bytes)."""
        n = len(frames)
        blobs = [None] * n
        errs = []
        q_sym = queue.Queue(maxsize=2)
        q_blob = queue.Queue(maxsize=2)
        mvq, mvq_i = get_rounded_q(q_mv)
        yq, yq_i = get_rounded_q(q_y)
        get = lambda t: frames[t]() if callable(frames[t]) else frames[t]   # noqa: E731

        # ---- I-frame, serial (on the decoder stream: its reconstruction seeds both chains)
        with torch.cuda.stream(self.s_dec):
            r, blob0 = self.dec_intra.encode_decode_bytes(get(0), q_i, pic_height, pic_width)
            blobs[0] = blob0
            rec0 = r["x_hat"]
            if on_recon is not None:
                on_recon(0, rec0)
            self.s_dec.synchronize()
        if n == 1:
            return blobs
        first = {"ref_frame": rec0, "ref_feature": None, "ref_y": None, "ref_mv_y": None}

        def enc_body():
            try:
                torch.cuda.set_device(self.device)
                dpb = dict(first)
                with torch.cuda.stream(self.s_enc):
                    for t in range(1, n):
                        host, dpb = self.enc_dmc.compress_symbols(get(t), dpb, mvq, yq, want_dpb=True, clamp_ref=True)
                        q_sym.put((t, host))
            except BaseException as e:   # noqa: BLE001
                errs.append(e)
            finally:
                q_sym.put(None)

        def ent_body():
            try:
                while True:
                    item = q_sym.get()
                    if item is None:
                        break
                    t, host = item
                    q_blob.put((t, pack_p(self.enc_dmc.encode_symbols(host), mvq_i, yq_i)))
            except BaseException as e:   # noqa: BLE001
                errs.append(e)
                while q_sym.get() is not None:    # drain so the encoder thread can finish (bounded queue)
                    pass
            finally:
                q_blob.put(None)

        def dec_body():
            try:
                torch.cuda.set_device(self.device)
                dpb = dict(first)
                with torch.cuda.stream(self.s_dec):
                    while True:
                        item = q_blob.get()
                        if item is None:
                            break
                        t, blob = item
                        a, b, string = unpack_p(blob)
                        dpb = self.dec_dmc.decompress(dpb, string, pic_height, pic_width, a / 100, b / 100)["dpb"]
                        blobs[t] = blob
                        if on_recon is not None:
                            on_recon(t, dpb["ref_frame"])
                    self.s_dec.synchronize()
            except BaseException as e:   # noqa: BLE001
                errs.append(e)
                while q_blob.get() is not None:   # drain so the producers can finish
                    pass

        ts = [threading.Thread(target=f) for f in (enc_body, ent_body, dec_body)]
        for th in ts:
            th.start()
        for th in ts:
            th.join()
        if errs:
            raise errs[0]
        return blobs
