"""The allocation-free statistics entry point (`scb_bn_stats_noalloc`), called through the C ABI.

It exists for callers that may not allocate device memory at the call site (for instance a host
framework whose pool already owns the GPU): its scratch is a fixed array inside the library.  The
check: on two calls in a row -- the second one finds the array as the first one left it -- the
(mean, M2, count) message equals the one of the regular entry point, which gets caller-provided
scratch, to the 1e-5 the other batch-norm tests use.  Collected last on purpose (file name)."""
import ctypes

import pytest
import torch

from simclr_pytorch_distributed_b200 import _lib

pytestmark = pytest.mark.gpu


def _stats(lib, fn_noalloc, x, channels_last):
    C = x.shape[1]
    if channels_last:
        x = x.contiguous(memory_format=torch.channels_last)
        geom = (1, x.numel() // C, C, 1)
    else:
        x = x.contiguous()
        geom = (0, x.shape[0], C, x.numel() // (x.shape[0] * C))
    out = torch.full((2 * C + 1,), float('nan'), device=x.device)
    stream = torch.cuda.current_stream().cuda_stream
    if fn_noalloc:
        rc = lib.scb_bn_stats_noalloc(x.data_ptr(), _lib.dtype_code(x.dtype), geom[0], geom[1], geom[2], geom[3],
                                      out.data_ptr(), stream)
    else:
        ws = torch.zeros(lib.scb_bn_workspace_bytes(C), dtype=torch.uint8, device=x.device)
        rc = lib.scb_bn_stats(x.data_ptr(), _lib.dtype_code(x.dtype), geom[0], geom[1], geom[2], geom[3],
                              out.data_ptr(), ws.data_ptr(), ws.numel(), stream)
    _lib.check(rc, 'bn_stats')
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_noalloc_matches_regular_entry_point_twice(dtype):
    lib = _lib.load()
    torch.manual_seed(21)
    cases = [((64, 256, 8, 8), True), ((32, 48, 5, 7), False)]
    for call in range(2):                       # two calls in a row share the built-in array
        for shape, cl in cases:
            x = (torch.randn(shape, device='cuda') * 1.7 + 0.6 * (call + 1)).to(dtype)
            ref = _stats(lib, False, x, cl)
            got = _stats(lib, True, x, cl)
            C = shape[1]
            assert torch.equal(got[2 * C], ref[2 * C])                       # element count
            assert torch.allclose(got[:C], ref[:C], rtol=1e-5, atol=1e-5)    # mean
            assert torch.allclose(got[C:2 * C], ref[C:2 * C], rtol=1e-5, atol=1e-5 * float(ref[C:2 * C].max()))


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch():
    lib = _lib.load()
    limit = _lib.BN_NOALLOC_MAX_CHANNELS
    x = torch.zeros(2, limit + 8, device='cuda')
    out = torch.empty(2 * (limit + 8) + 1, device='cuda')
    rc = lib.scb_bn_stats_noalloc(x.data_ptr(), _lib.F32, 1, 2, limit + 8, 1, out.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    msg = lib.scb_last_error().decode()
    assert 'exceed the built-in scratch' in msg and str(limit) in msg
