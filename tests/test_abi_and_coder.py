"""CPU tests: the C-ABI library loads and exports every symbol include/vcmts_b200.h declares, and
the product's host entropy coder (csrc/rans_coder.cpp) is byte-identical to the oracle C coder --
and, where oracle/_ref was built (build container), to the reference's own unmodified C++."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import dcvc_oracle as O
from vcm_ts_b200 import _lib
from vcm_ts_b200.entropy import CdfTables, RansCoder, ScaleIndexTable

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "vcmts_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(vcm_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) > 40
    L = ctypes.CDLL(_lib.SO) if os.path.exists(_lib.SO) else _lib.lib()
    for name in sorted(declared):
        assert hasattr(L, name), f"{name} declared in the header but not exported"
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    assert _lib.lib().vcm_version() >= 100


def _random_case(seed, n=20000):
    rng = np.random.default_rng(seed)
    table = ScaleIndexTable("laplace").update()
    idx = rng.integers(0, 256, n).astype(np.uint8)
    sym = np.rint(rng.laplace(0, 1, n) * (1 + idx / 16.0)).astype(np.int16)
    sym[::97] = rng.integers(-3000, 3000, sym[::97].size)          # force bypass escapes
    return table, sym, idx


@pytest.mark.parametrize("seed", [0, 1])
def test_coder_bytes_equal_oracle_and_round_trip(seed):
    table, sym, idx = _random_case(seed)
    T = CdfTables(*table)
    c = RansCoder()
    c.reset_encoder()
    s_t, i_t = torch.from_numpy(sym), torch.from_numpy(idx)
    c.encode_planes(s_t, i_t, T)
    mine = c.flush_encoder()
    o = O.OracleRans()
    o.reset_encoder()
    o.encode_with_indexes(sym.astype(np.int32), idx.astype(np.int32), *table)
    assert mine == o.flush_encoder()
    c.set_stream(mine)
    out = torch.empty(sym.size, dtype=torch.int16)
    c.decode_planes(i_t, T, out)
    assert np.array_equal(out.numpy(), sym)


def test_coder_channel_mode_and_mixed_stream():
    rng = np.random.default_rng(5)
    ytab = ScaleIndexTable("gaussian").update()
    P = {k: torch.randn(s) * 0.3 for k, s in O.state_shapes("dmc").items() if k.startswith("bit_estimator_z.")}
    ztab = O.factorized_tables(P, "bit_estimator_z", 64)
    z = rng.integers(-6, 7, (1, 64, 3, 5)).astype(np.int16)
    idx = rng.integers(0, 256, 4000).astype(np.uint8)
    sym = rng.integers(-4, 5, 4000).astype(np.int16)
    TY, TZ = CdfTables(*ytab), CdfTables(*ztab)
    c = RansCoder()
    c.reset_encoder()
    c.encode_channels(torch.from_numpy(z), TZ)
    c.encode_planes(torch.from_numpy(sym), torch.from_numpy(idx), TY)
    mine = c.flush_encoder()
    o = O.OracleRans()
    o.reset_encoder()
    zi = np.repeat(np.arange(64), 15).astype(np.int32)
    o.encode_with_indexes(z.reshape(-1).astype(np.int32), zi, *ztab)
    o.encode_with_indexes(sym.astype(np.int32), idx.astype(np.int32), *ytab)
    assert mine == o.flush_encoder()
    c.set_stream(mine)
    zo = torch.empty((1, 64, 3, 5), dtype=torch.int16)
    c.decode_channels(64, 15, TZ, zo)
    so = torch.empty(4000, dtype=torch.int16)
    c.decode_planes(torch.from_numpy(idx), TY, so)
    assert np.array_equal(zo.numpy(), z) and np.array_equal(so.numpy(), sym)


def test_coder_bytes_equal_reference_cpp():
    """oracle/_ref = the reference's rans_interface.cpp compiled unmodified (build container)."""
    from oracle import ref_import
    try:
        ref = ref_import.load_ref_native("MLCodec_rans")
    except ImportError:
        pytest.skip("oracle/_ref not built here")
    table, sym, idx = _random_case(7, 5000)
    cdf, sizes, offs = table
    enc = ref.BufferedRansEncoder()
    enc.encode_with_indexes(sym.astype(np.int32).tolist(), idx.astype(np.int32).tolist(), cdf.tolist(),
                            sizes.tolist(), offs.tolist())
    want = enc.flush()
    c = RansCoder()
    c.reset_encoder()
    c.encode_planes(torch.from_numpy(sym), torch.from_numpy(idx), CdfTables(*table))
    assert c.flush_encoder() == want


def test_coder_edge_frequencies_division_free_encode_and_mode_fast_path():
    """The encoder replaces x / freq by a round-up reciprocal and the decoder tries the most probable
    value first (rans_coder.cpp).  Tables built to hit every special case: freq 1, 2, powers of two
    and their neighbours, the largest legal freq, a dominant symbol, an escape-heavy table.  Bytes
    must equal the division-based oracle port (and the reference's own C++ when it is built), and
    decode back."""
    rng = np.random.default_rng(11)
    rows = [
        [1, 1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 255, 256, 257, 1023, 1024, 1025, 4096, 32768],
        [65535 - 6, 1, 1, 1, 1, 1, 1],             # dominant symbol (decoder fast path), freq-1 tail
        [1, 1, 1, 1, 1, 1, 65535 - 6],             # dominant symbol is the escape: fast path must not take it
        [32768, 16384, 8192, 4096, 2048, 1024, 512, 256, 128, 64, 32, 16, 8, 4, 2, 1],
        [3] * 200,
    ]
    stride = max(len(r) for r in rows) + 1
    cdf = np.zeros((len(rows), stride), dtype=np.int32)
    sizes = np.zeros(len(rows), dtype=np.int32)
    offs = np.array([-3, 0, -2, -8, -100], dtype=np.int32)
    for i, r in enumerate(rows):
        f = np.array(r, dtype=np.int64)
        f[np.argmax(f)] += 65536 - f.sum()         # make the row sum to 2^16 exactly
        assert f.min() >= 1 and f.sum() == 65536
        cdf[i, 1:len(r) + 1] = np.cumsum(f)
        sizes[i] = len(r) + 1
    n = 60000
    idx = rng.integers(0, len(rows), n).astype(np.uint8)
    sym = np.empty(n, dtype=np.int16)
    for i in range(len(rows)):
        m = idx == i
        k = int(m.sum())
        f = np.diff(cdf[i, :sizes[i]]).astype(np.float64)
        v = rng.choice(len(f), size=k, p=f / f.sum())          # in-table values by their own probability
        far = rng.random(k) < 0.05                             # plus out-of-table values on both sides
        v = np.where(far, rng.integers(-40, len(f) + 40, k), v)
        sym[m] = (v + offs[i]).astype(np.int16)
    table = (torch.from_numpy(cdf), torch.from_numpy(sizes), torch.from_numpy(offs))
    T = CdfTables(*table)
    c = RansCoder()
    c.reset_encoder()
    c.encode_planes(torch.from_numpy(sym), torch.from_numpy(idx), T)
    mine = c.flush_encoder()
    o = O.OracleRans()
    o.reset_encoder()
    o.encode_with_indexes(sym.astype(np.int32), idx.astype(np.int32), *table)
    assert mine == o.flush_encoder()
    from oracle import ref_import
    try:
        ref = ref_import.load_ref_native("MLCodec_rans")
        enc = ref.BufferedRansEncoder()
        enc.encode_with_indexes(sym.astype(np.int32).tolist(), idx.astype(np.int32).tolist(), cdf.tolist(),
                                sizes.tolist(), offs.tolist())
        assert mine == enc.flush()
    except ImportError:
        pass
    c.set_stream(mine)
    out = torch.empty(n, dtype=torch.int16)
    c.decode_planes(torch.from_numpy(idx), T, out)
    assert np.array_equal(out.numpy(), sym)


def test_pmf_to_quantized_cdf_matches_oracle():
    rng = np.random.default_rng(3)
    L = _lib.lib()
    for n in (2, 7, 33, 101):
        p = rng.random(n).astype(np.float32) ** 4
        p /= p.sum()
        mine = np.empty(n + 1, dtype=np.int32)
        assert L.vcm_pmf_to_quantized_cdf(p.ctypes.data, n, 16, mine.ctypes.data) == 0
        want = O.pmf_to_quantized_cdf(torch.from_numpy(p)).numpy()
        assert np.array_equal(mine, want)
