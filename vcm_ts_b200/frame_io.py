"""Frame I/O around the codec path on the device: the uint8 frames a PNG decoder
yields go to the GPU as uint8 (4x fewer H2D bytes than the reference's float32 tensors,
`DCVC_HEM/src/utils/png_reader.py:37-39`, `test_video.py:62-65,107-125`) and are normalised and
padded there; reconstructions come back as uint8 HWC ready for the PNG encoder
(`save_torch_image`, `stream_helper.py:151-154`).  Values are bit-identical to the reference's
numpy formulas.  No CPU path: raises without the CUDA library / a device tensor."""
import torch

from . import _lib
from ._lib import check
from .stream_helper import get_padding_size


def frames_from_u8(u8_hwc, device=None, padded=True):
    """uint8 [N,H,W,3] or [H,W,3] (host, ideally pinned, or device) -> float32 NCHW on the GPU,
    zero-padded right/bottom to multiples of 64 like the reference's callers do."""
    L = _lib.lib()
    t = u8_hwc if u8_hwc.dim() == 4 else u8_hwc[None]
    if t.dtype != torch.uint8 or t.shape[-1] != 3:
        raise TypeError("expected uint8 [N,H,W,3]")
    if not t.is_cuda:
        if device is None:
            raise RuntimeError("vcm_ts_b200.frame_io needs a CUDA device; there is no CPU path")
        t = t.to(device, non_blocking=True)
    t = t.contiguous()
    N, H, W, _ = t.shape
    Hp, Wp = H, W
    if padded:
        _, pr, _, pb = get_padding_size(H, W)
        Hp, Wp = H + pb, W + pr
    out = torch.empty((N, 3, Hp, Wp), dtype=torch.float32, device=t.device)
    check(L.vcm_u8hwc_to_frame(t.data_ptr(), out.data_ptr(), N, H, W, Hp, Wp,
                               torch.cuda.current_stream(t.device).cuda_stream), "vcm_u8hwc_to_frame")
    return out


def frames_to_u8(x_nchw, height, width, out=None):
    """float32 NCHW reconstruction [N,3,Hs,Ws] on the GPU -> uint8 [N,height,width,3] (device tensor, or
    copied into `out`, e.g. a pinned host buffer, on the current stream)."""
    L = _lib.lib()
    if not x_nchw.is_cuda:
        raise RuntimeError("vcm_ts_b200.frame_io needs a CUDA tensor; there is no CPU path")
    x = x_nchw.detach().to(torch.float32).contiguous()
    N, C, Hs, Ws = x.shape
    if C != 3 or height > Hs or width > Ws:
        raise ValueError("expected [N,3,Hs>=height,Ws>=width]")
    dev = torch.empty((N, height, width, 3), dtype=torch.uint8, device=x.device)
    check(L.vcm_frame_to_u8hwc(x.data_ptr(), dev.data_ptr(), N, height, width, Hs, Ws,
                               torch.cuda.current_stream(x.device).cuda_stream), "vcm_frame_to_u8hwc")
    if out is not None:
        out.copy_(dev, non_blocking=True)
        return out
    return dev
