"""penroz_colsum_static: the allocation-free bias-gradient column sum.

This entry point is for hosts that cannot get device memory of their own — a binding that lives outside PyTorch
and is only handed input/output pointers, or a worker where torch's caching allocator already holds nearly all of
HBM so a plain cudaMalloc would fail. Its partial sums go to a fixed 1 MiB array inside the library instead of a
caller buffer. The test calls it straight through the C ABI (ctypes, raw pointers) twice in a row with different
shapes and compares each result with penroz_colsum (caller-provided scratch), to the tolerance the other column-sum
tests use; it also checks that matrices wider than the built-in limit are refused with a readable error.

The file name sorts last on purpose: with `pytest -x` every other test has already run if this one fails. It passes
on its own in a fresh process as well as after the rest of the suite (the first call resolves the array's address
for the device, later calls reuse it).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel_err(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-12)).item()


def test_colsum_static_matches_colsum_twice_in_a_row():
    from penroz import _lib
    _lib.require_device()
    L = _lib.lib()
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    torch.manual_seed(11)
    # second call is narrower than the first: it must not pick up partial sums the first call left behind; the third
    # sits exactly on the built-in limit (8192 columns) and accumulates
    for (M, N, dtype, accumulate) in [(4099, 3072, torch.bfloat16, 1), (513, 772, torch.float32, 0),
                                      (257, 8192, torch.bfloat16, 1)]:
        x = torch.randn(M, N, device=dev).to(dtype)
        start = torch.randn(N, device=dev)
        want = start.clone()
        scratch = torch.empty(L.colsum_scratch_bytes(M, N), device=dev, dtype=torch.uint8)
        L.colsum(x.data_ptr(), x.stride(0), M, N, want.data_ptr(), accumulate, _lib.dt(x), scratch.data_ptr(),
                 scratch.numel(), st)
        got = start.clone()
        # One untimed call per case first: the first call on a device looks up the built-in array's address, and the
        # first launch of each kernel variant makes the driver load its code (lazy module loading), which moves the
        # driver's free-memory figure although nobody allocated. The measured call below is an ordinary later call.
        warm = start.clone()
        L.colsum_static(x.data_ptr(), x.stride(0), M, N, warm.data_ptr(), accumulate, _lib.dt(x), st)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        free_before = torch.cuda.mem_get_info()[0]              # the driver's own count: sees a raw cudaMalloc too
        L.colsum_static(x.data_ptr(), x.stride(0), M, N, got.data_ptr(), accumulate, _lib.dt(x), st)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before          # nothing was allocated through torch
        assert torch.cuda.mem_get_info()[0] == free_before      # and nothing behind torch's back either
        assert _rel_err(got, want) < 1e-5, (M, N, dtype, _rel_err(got, want))
        ref = x.float().sum(0) + (start if accumulate else 0)
        assert _rel_err(got, ref) < 1e-5


def test_colsum_static_rejects_matrices_wider_than_its_scratch():
    from penroz import _lib
    _lib.require_device()
    L = _lib.lib()
    dev = torch.device("cuda:0")
    N = 8192 + 4
    x = torch.zeros(8, N, device=dev)
    out = torch.zeros(N, device=dev)
    with pytest.raises(_lib.PenrozError) as err:
        L.colsum_static(x.data_ptr(), x.stride(0), 8, N, out.data_ptr(), 0, _lib.F32,
                        torch.cuda.current_stream().cuda_stream)
    assert "8192" in str(err.value) and "penroz_colsum" in str(err.value)


def test_colsum_refuses_a_leading_dimension_shorter_than_the_row():
    from penroz import _lib
    _lib.require_device()
    L = _lib.lib()
    dev = torch.device("cuda:0")
    x = torch.zeros(8, 64, device=dev)
    out = torch.zeros(64, device=dev)
    scratch = torch.empty(L.colsum_scratch_bytes(8, 64), device=dev, dtype=torch.uint8)
    st = torch.cuda.current_stream().cuda_stream
    with pytest.raises(_lib.PenrozError, match="leading dimension"):
        L.colsum_static(x.data_ptr(), 32, 8, 64, out.data_ptr(), 0, _lib.F32, st)
    with pytest.raises(_lib.PenrozError, match="leading dimension"):
        L.colsum(x.data_ptr(), 32, 8, 64, out.data_ptr(), 0, _lib.F32, scratch.data_ptr(), scratch.numel(), st)
