"""Generates tests/golden/enhancement.npz (build container only: reads /root/reference).

`video_coder.py` cannot be imported here (cv2 / ffmpeg / detector packages are absent), so the
reference's own `create_gradient_mask` is taken from its source with `ast` and executed unmodified;
the residual / fuse results are then produced by the reference's loop-body statements
(video_coder.py:419-447, 575-600), written out below verbatim in meaning, using THAT mask function.

    python -m oracle.gen_enhancement_golden
"""
import ast
import os

import numpy as np

from oracle import enhancement as E
from oracle.ref_import import REF_ROOT


def reference_gradient_mask():
    src = open(os.path.join(REF_ROOT, "video_coder.py")).read()
    fn = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "create_gradient_mask")
    ns = {"np": np}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "video_coder.py", "exec"), ns)
    return ns["create_gradient_mask"]


def main():
    ref_mask = reference_gradient_mask()
    src, enc, enh, lp, faces = E.golden_case()
    G = {}
    for b in (0, 1, 2, 5, 9):
        for (w, h) in ((40, 23), (7, 5), (1, 1)):
            G[f"mask_b{b}_{w}x{h}"] = ref_mask(w=w, h=h, border_size=b)
    # compute_residuals body (video_coder.py:419-447)
    source_frame, encoded_frame = src.astype(np.float32), enc.astype(np.float32)
    h, w, c = source_frame.shape
    residual = source_frame - encoded_frame
    residual = np.clip(residual + 128, 0.0, 255.0)
    mask = np.zeros(shape=(h, w, 1), dtype=np.float32)
    for [x1, y1, x2, y2] in lp:
        mask[y1:y2, x1:x2] = 1.0
    for [x1, y1, x2, y2] in faces:
        mask[y1:y2, x1:x2] = 1.0
    result_frame = residual.astype(np.float32) * mask
    G["residual"] = result_frame.astype(np.uint8)
    # fuse_layers body (video_coder.py:575-600) for two padding settings
    for tag, (lpad, fpad) in (("p0", (0, 0)), ("p5_9", (5, 9)), ("p1_2", (1, 2))):
        dcvc_hem_frame = src.astype(np.float32)
        enhancement_frame = enh.astype(np.float32) - 128
        mask = np.zeros(shape=(h, w, 1), dtype=np.float32)
        for [x1, y1, x2, y2] in lp:
            mask[y1:y2, x1:x2] = ref_mask(w=x2 - x1, h=y2 - y1, border_size=lpad)
        for [x1, y1, x2, y2] in faces:
            mask[y1:y2, x1:x2] = ref_mask(w=x2 - x1, h=y2 - y1, border_size=fpad)
        result_frame = dcvc_hem_frame
        result_frame += mask * enhancement_frame
        result_frame = np.clip(result_frame, 0, 255)
        G["fuse_" + tag] = result_frame.astype(np.uint8)
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                       "enhancement.npz")
    np.savez_compressed(out, **G)
    print(out, {k: v.shape for k, v in G.items()})


if __name__ == "__main__":
    main()
