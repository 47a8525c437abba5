"""Frame container + padding helpers, byte-compatible with DCVC_HEM/src/utils/stream_helper.py
(:24-46 padding / z shape / q rounding, :49-56 checkpoint key stripping, :59-62 filesize,
:103-144 big-endian I / P frame files).  In-memory `pack_*` / `unpack_*` twins are added so the
hot loop does not have to round-trip through the file system.
"""
import os
import struct

import numpy as np
import torch


def get_padding_size(height, width, p=64):
    new_h = (height + p - 1) // p * p
    new_w = (width + p - 1) // p * p
    return 0, new_w - width, 0, new_h - height          # left, right, up, down


def get_downsampled_shape(height, width, p):
    new_h = (height + p - 1) // p * p
    new_w = (width + p - 1) // p * p
    return int(new_h / p + 0.5), int(new_w / p + 0.5)


def get_rounded_q(q_scale):
    q_scale = np.clip(q_scale, 0.01, 655.)
    q_index = int(np.round(q_scale * 100))
    return q_index / 100, q_index


def get_state_dict(ckpt_path):
    ckpt = torch.load(ckpt_path, map_location=torch.device("cpu"))
    for key in ("state_dict", "net"):
        if key in ckpt:
            ckpt = ckpt[key]
    return {k.replace("module.", ""): v for k, v in ckpt.items()}


def filesize(filepath):
    if not os.path.isfile(filepath):
        raise ValueError(f'Invalid file "{filepath}".')
    return os.stat(filepath).st_size


def pack_i(height, width, q_index, bit_stream):
    return struct.pack(">2I", height, width) + struct.pack(">H", q_index) + \
        struct.pack(">I", len(bit_stream)) + bit_stream


def unpack_i(blob):
    height, width = struct.unpack(">2I", blob[:8])
    (q_index,) = struct.unpack(">H", blob[8:10])
    (n,) = struct.unpack(">I", blob[10:14])
    return height, width, q_index, blob[14:14 + n]


def pack_p(string, mv_y_q_index, y_q_index):
    return struct.pack(">2H", mv_y_q_index, y_q_index) + struct.pack(">I", len(string)) + string


def unpack_p(blob):
    mv_y_q_index, y_q_index = struct.unpack(">2H", blob[:4])
    (n,) = struct.unpack(">I", blob[4:8])
    return mv_y_q_index, y_q_index, blob[8:8 + n]


def encode_i(height, width, q_index, bit_stream, output):
    with open(output, "wb") as f:
        f.write(pack_i(height, width, q_index, bit_stream))


def decode_i(inputpath):
    with open(inputpath, "rb") as f:
        return unpack_i(f.read())


def encode_p(string, mv_y_q_index, y_q_index, output):
    with open(output, "wb") as f:
        f.write(pack_p(string, mv_y_q_index, y_q_index))


def decode_p(inputpath):
    with open(inputpath, "rb") as f:
        return unpack_p(f.read())
