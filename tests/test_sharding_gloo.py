"""N>1 path on CPU: two gloo ranks code disjoint GOPs of one sequence through the product's host
logic (device kernels emulated, tests/emul_backend.py) and merge; the merged per-frame bitstreams
must equal a single-process run byte for byte (the path has no data-path collective, so sharding
must not change a single bit)."""
import hashlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vcm_ts_b200 import sharding


def test_assignment_is_a_partition():
    for n, w in ((0, 2), (1, 2), (5, 2), (8, 8), (9, 4)):
        seen = sorted(u for r in range(w) for u in sharding.assign_units(n, w, r))
        assert seen == list(range(n))
    assert sharding.gop_units(7, 3) == [(0, 3), (3, 3), (6, 1)]
    with pytest.raises(ValueError):
        sharding.assign_units(4, 2, 2)


def _models():
    from oracle import dcvc_oracle as O
    from oracle import fixture
    from tests.emul_backend import EmulBackend
    from vcm_ts_b200.models import DMC, IntraNoAR
    dmc, intra = DMC().eval(), IntraNoAR().eval()
    dmc.load_state_dict(fixture.fixture_state_dict(O.state_shapes("dmc")))
    intra.load_state_dict(fixture.fixture_state_dict(O.state_shapes("intra")))
    for m in (dmc, intra):
        m._backend_factory = lambda dev, precision: EmulBackend(dev, precision)
        m.update(force=True)
    return dmc, intra


def _code_gop(models, frames, unit):
    dmc, intra = models
    first, n = unit
    out, dpb = [], None
    for t in range(first, first + n):
        x = frames[t:t + 1]
        if t == first:
            r, blob = intra.encode_decode_bytes(x, 1.0, 64, 64)
            dpb = {"ref_frame": r["x_hat"], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
        else:
            r, blob = dmc.encode_decode_bytes(x, dpb, 64, 64, 1.0, 1.0)
            dpb = r["dpb"]
        out.append((t, r["bit"], hashlib.sha256(blob).hexdigest()))
    return out


def _frames():
    from oracle import fixture
    return fixture.fixture_frames(64, 64, 4, seed=5)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        frames = _frames()
        units = sharding.gop_units(frames.shape[0], 2)
        models = _models()
        local = sharding.code_units(units, rank, world, lambda u, unit: _code_gop(models, frames, unit))
        merged = sharding.merge_results(local, len(units))
        q.put((rank, sorted(local), merged))
    finally:
        dist.destroy_process_group()


def test_two_gloo_ranks_reproduce_single_process(tmp_path):
    torch.set_num_threads(2)
    frames = _frames()
    units = sharding.gop_units(frames.shape[0], 2)
    models = _models()
    want = sharding.merge_results(sharding.code_units(units, 0, 1, lambda u, unit: _code_gop(models, frames, unit)),
                                  len(units))
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    shares = {r: mine for r, mine, _ in got}
    assert shares == {0: [0], 1: [1]}
    for _, _, merged in got:
        assert merged == want, "sharded coding changed a bitstream"
