"""Multi-GPU inference = independent coding units (sequences / GOPs) dealt to ranks, no data-path
collective.  Mirrors the reference's job distribution (`DCVC_HEM/test_video.py:240-261`:
`process_idx % gpu_num`) and the fact that every I-frame resets the decoded-picture buffer
(`video_coder.py:122-130`), so a GOP is the smallest unit that can move to another GPU.

Only the merge of per-unit results (bits, PSNR, timings -- a few numbers per frame) crosses ranks,
through `torch.distributed.all_gather_object` (gloo on CPU in the tests, NCCL's side channel on the
GPU box); the frames and latents never leave their GPU.
"""
import torch.distributed as dist


def gop_units(n_frames, gop):
    """[(first_frame, n_frames_in_gop)] of one sequence (`video_coder.py:119-121`: frame % gop == 0
    is an I-frame)."""
    return [(s, min(gop, n_frames - s)) for s in range(0, n_frames, gop)]


def assign_units(n_units, world_size, rank):
    """Static round-robin: unit u goes to rank u % world_size."""
    if not 0 <= rank < world_size:
        raise ValueError("rank out of range")
    return list(range(rank, n_units, world_size))


def code_units(units, rank, world_size, code_fn):
    """Runs code_fn(unit_index, unit) on this rank's share; returns {unit_index: result}."""
    mine = assign_units(len(units), world_size, rank)
    return {u: code_fn(u, units[u]) for u in mine}


def merge_results(local, n_units):
    """Every rank ends up with the list of all unit results in unit order.  `local`: {unit: result}
    (picklable, small).  Without an initialised process group this is the single-process identity."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        parts = [None] * dist.get_world_size()
        dist.all_gather_object(parts, local)
    else:
        parts = [local]
    merged = {}
    for p in parts:
        for u, r in p.items():
            if u in merged:
                raise RuntimeError(f"unit {u} coded twice")
            merged[u] = r
    missing = [u for u in range(n_units) if u not in merged]
    if missing:
        raise RuntimeError(f"units {missing} were not coded by any rank")
    return [merged[u] for u in range(n_units)]
