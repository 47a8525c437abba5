"""vcm_sq_err_sum_noalloc: the allocation-free variant of vcm_sq_err_sum (include/vcmts_b200.h).

No counterpart in the reference; it exists for callers that cannot call cudaMalloc.  Its partial
sums live in a `static __device__` array inside the library instead of caller scratch.  The file
name makes pytest collect it last.
"""
import pytest
import torch

from vcm_ts_b200 import _lib

pytestmark = pytest.mark.gpu

RED_CHUNK = 8192
MAX_PARTIALS = 8192


def _both(L, a, b):
    """(existing entry point with caller scratch, new entry point) on the same inputs."""
    N = a.shape[0]
    per = a.numel() // N
    stream = torch.cuda.current_stream().cuda_stream
    nblk = L.vcm_bits_partial_count(per)
    partial = torch.empty((N * nblk,), device=a.device, dtype=torch.float32)
    want = torch.empty((N,), device=a.device, dtype=torch.float32)
    got = torch.full((N,), float("nan"), device=a.device, dtype=torch.float32)
    _lib.check(L.vcm_sq_err_sum(a.data_ptr(), b.data_ptr(), partial.data_ptr(), want.data_ptr(), N, per,
                                stream), "vcm_sq_err_sum")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    _lib.check(L.vcm_sq_err_sum_noalloc(a.data_ptr(), b.data_ptr(), got.data_ptr(), N, per, stream),
               "vcm_sq_err_sum_noalloc")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return want.cpu(), got.cpu(), free0 - free1


def test_sq_err_sum_noalloc_matches_scratch_variant_twice():
    L = _lib.lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    g = torch.Generator().manual_seed(7)
    # call 1: two padded 1080p RGB frames (766 partial sums per sample); call 2: a smaller, ragged
    # shape, so slots the first call filled but the second does not own must not leak into its sums
    shapes = [(2, 3, 1088, 1920), (3, 3, 67, 129)]
    grown = []
    for shape in shapes:
        a = torch.rand(shape, generator=g).to(dev)
        b = (a.cpu() + 0.05 * torch.randn(shape, generator=g)).to(dev)
        want, got, d_free = _both(L, a, b)
        grown.append(d_free)
        # same kernels, same fixed summation order: bit-identical, not merely close
        assert torch.equal(got, want), (got, want)
        ref = ((a.double() - b.double()) ** 2).reshape(shape[0], -1).sum(1).float().cpu()
        torch.testing.assert_close(got, ref, rtol=1e-5, atol=0)
    # the second call finds the scratch address cached: no device memory may be taken by it
    assert grown[1] <= 0, grown


def test_sq_err_sum_noalloc_rejects_sizes_beyond_the_scratch():
    L = _lib.lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    a = torch.zeros((1, 16), device=dev)
    out = torch.zeros((1,), device=dev)
    per = (MAX_PARTIALS + 1) * RED_CHUNK          # one partial sum too many; the pointers are never read
    rc = L.vcm_sq_err_sum_noalloc(a.data_ptr(), a.data_ptr(), out.data_ptr(), 1, per, None)
    assert rc != 0
    msg = L.vcm_last_error().decode()
    assert "exceed" in msg and str(MAX_PARTIALS) in msg, msg
    # exactly at the limit is accepted by the size check (N samples x partials each)
    N = 4
    per_ok = (MAX_PARTIALS // N) * RED_CHUNK
    big_a = torch.rand((N, per_ok), device=dev)
    big_b = torch.rand((N, per_ok), device=dev)
    want, got, _ = _both(L, big_a, big_b)
    assert torch.equal(got, want)
