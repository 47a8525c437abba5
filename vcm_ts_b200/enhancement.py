"""Enhancement-layer frame ops on the device: what `compute_residuals`
(video_coder.py:379-457) and `fuse_layers` (video_coder.py:533-613) do per frame in numpy between
PNG round trips, done on uint8 HWC frames that already sit on the GPU next to the decoder output
(`vcm_ts_b200.frame_io.frames_to_u8`).  Same argument meaning as the reference: box lists are the
detectors' `[x1, y1, x2, y2]` lists, licence plates first, faces second (later boxes overwrite
earlier ones), `*_padding` is the `border_size` of `create_gradient_mask` (video_coder.py:521-530).
Results are bit-identical to the numpy statements.  No CPU path: raises without the CUDA library /
a device tensor."""
import numpy as np
import torch

from . import _lib
from ._lib import check

MAX_BOXES = 256


def normalise_boxes(boxes, height, width):
    """Resolve the numpy slice semantics of `mask[y1:y2, x1:x2]` (negative indexes count from the
    end, out-of-range ones are clipped, empty slices select nothing) -> list of in-range boxes."""
    out = []
    for x1, y1, x2, y2 in boxes:
        xa, xb, _ = slice(int(x1), int(x2)).indices(width)
        ya, yb, _ = slice(int(y1), int(y2)).indices(height)
        out.append([xa, ya, max(xa, xb), max(ya, yb)])
    return out


def gradient_ring(border_size):
    """The distinct values of create_gradient_mask (video_coder.py:521-530), outermost frame first:
    float32(1 - linspace(0.9, 0, border)); {1} when there is no border."""
    if border_size > 0:
        return (1 - np.linspace(0.9, 0.0, int(border_size))).astype(np.float32)
    return np.ones(1, dtype=np.float32)


def _frames(*ts):
    out = []
    for t in ts:
        if not torch.is_tensor(t) or not t.is_cuda:
            raise RuntimeError("vcm_ts_b200.enhancement needs CUDA tensors; there is no CPU path")
        if t.dtype != torch.uint8:
            raise TypeError("expected uint8 HWC frames")
        out.append((t if t.dim() == 4 else t[None]).contiguous())
    if any(o.shape != out[0].shape for o in out):
        raise ValueError("frames must have the same shape")
    return out


def _box_tensors(per_frame, H, W, device, borders=None):
    """per_frame: one list of boxes per frame.  -> (boxes int32 [T,4], start int32 [N+1], ring_start, ring, max)."""
    flat, start, rings, rstart = [], [0], [], [0]
    for f, boxes in enumerate(per_frame):
        nb = normalise_boxes(boxes, H, W)
        if len(nb) > MAX_BOXES:
            raise ValueError("at most %d boxes per frame" % MAX_BOXES)
        flat += nb
        start.append(len(flat))
        if borders is not None:
            for b in borders[f]:
                r = gradient_ring(b)
                rings.append(r)
                rstart.append(rstart[-1] + len(r))
    bt = torch.tensor(flat, dtype=torch.int32).reshape(-1, 4).to(device)
    st = torch.tensor(start, dtype=torch.int32).to(device)
    mx = max([start[i + 1] - start[i] for i in range(len(per_frame))] or [0])
    if borders is None:
        return bt, st, None, None, mx
    rv = torch.from_numpy(np.concatenate(rings) if rings else np.zeros(0, np.float32)).to(device)
    rs = torch.tensor(rstart, dtype=torch.int32).to(device)
    return bt, st, rs, rv, mx


def _per_frame(boxes, N):
    """A flat list of boxes (one frame) or a list of per-frame lists."""
    if N == 1 and (len(boxes) == 0 or not isinstance(boxes[0][0], (list, tuple, np.ndarray))):
        return [list(boxes)]
    if len(boxes) == 0:
        return [[] for _ in range(N)]
    if len(boxes) != N:
        raise ValueError("need one box list per frame")
    return [list(b) for b in boxes]


def prepare_boxes(lp_bboxes, face_bboxes, N, H, W, device, liplates_padding=None, faces_padding=None):
    """Upload the box lists of N frames once (they change per frame in the reference's loops, but a
    caller coding many frames can prepare them ahead of the launch).  With paddings: a plan for
    `fuse` (boxes must lie inside the frame, as the reference's mask assignment requires); without:
    a plan for `compute_residual` (numpy slice semantics for out-of-range boxes)."""
    lp, fc = _per_frame(lp_bboxes, N), _per_frame(face_bboxes, N)
    per = [a + b for a, b in zip(lp, fc)]
    if liplates_padding is None:
        return _box_tensors(per, H, W, device)
    for boxes in per:
        for x1, y1, x2, y2 in boxes:
            if not (0 <= x1 <= x2 <= W and 0 <= y1 <= y2 <= H):
                # the reference's `mask[y1:y2, x1:x2] = create_gradient_mask(x2 - x1, y2 - y1, ...)` cannot
                # broadcast for such a box
                raise ValueError("fuse: box %r does not lie inside the %dx%d frame" % ([x1, y1, x2, y2], W, H))
    borders = [[liplates_padding] * len(a) + [faces_padding] * len(b) for a, b in zip(lp, fc)]
    return _box_tensors(per, H, W, device, borders)


def compute_residual(source_u8, encoded_u8, lp_bboxes=(), face_bboxes=(), out=None, plan=None):
    """video_coder.py:415-447 per frame: uint8(clip(source - encoded + 128, 0, 255) * mask), mask = 1
    inside any licence-plate / face box.  Frames: uint8 [H,W,C] or [N,H,W,C] on the GPU."""
    L = _lib.lib()
    src, enc = _frames(source_u8, encoded_u8)
    N, H, W, C = src.shape
    bt, st, _, _, mx = plan or prepare_boxes(lp_bboxes, face_bboxes, N, H, W, src.device)
    if out is None:
        out = torch.empty_like(src)
    check(L.vcm_enh_residual(src.data_ptr(), enc.data_ptr(), out.data_ptr(), N, H, W, C,
                             bt.data_ptr() if bt.numel() else None, st.data_ptr(), mx,
                             torch.cuda.current_stream(src.device).cuda_stream), "vcm_enh_residual")
    return out if source_u8.dim() == 4 else out[0]


def fuse(base_u8, enhancement_u8, lp_bboxes=(), face_bboxes=(), liplates_padding=0, faces_padding=0, out=None,
         plan=None):
    """video_coder.py:570-600 per frame: uint8(clip(base + mask * (enhancement - 128), 0, 255)) with
    the gradient masks of create_gradient_mask (licence plates first, faces overwrite them)."""
    L = _lib.lib()
    base, enh = _frames(base_u8, enhancement_u8)
    N, H, W, C = base.shape
    bt, st, rs, rv, mx = plan or prepare_boxes(lp_bboxes, face_bboxes, N, H, W, base.device,
                                               liplates_padding, faces_padding)
    if out is None:
        out = torch.empty_like(base)
    has = bt.numel() > 0
    check(L.vcm_enh_fuse(base.data_ptr(), enh.data_ptr(), out.data_ptr(), N, H, W, C,
                         bt.data_ptr() if has else None, st.data_ptr(), rs.data_ptr() if has else None,
                         rv.data_ptr() if has else None, mx,
                         torch.cuda.current_stream(base.device).cuda_stream), "vcm_enh_fuse")
    return out if base_u8.dim() == 4 else out[0]
