"""vfi_psnr_u8_noalloc (no counterpart in the reference): the allocation-free
variant of vfi_psnr_u8, its accumulators in a static __device__ array of the library.  Called through the C ABI twice in a row
and compared with the existing entry point -- integer SSE bit-exact, PSNR equal (the tolerance tests/test_gpu_reduce.py uses for
vfi_psnr_u8 itself is exact SSE).  Kept in a file that collects last."""
import ctypes as C

import numpy as np
import pytest
import torch

import vfi_b200  # noqa: F401
from vfi_b200 import _abi
from vfi_b200.synth import noisy_copy

pytestmark = pytest.mark.gpu


def _pair(n, H, W, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    b = noisy_copy(a, seed=seed + 1, amp=6)
    return torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()


def test_noalloc_psnr_equals_vfi_psnr_u8_two_calls_in_a_row(engine):
    lib = _abi.lib()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # second call: FEWER pairs and different data, so sums left in the shared array by the first call would show
    for n, H, W, seed in ((5, 360, 640, 1), (3, 271, 333, 7)):
        a, b = _pair(n, H, W, seed)
        b[0] = a[0]                                                   # an identical pair: SSE 0, PSNR inf
        sse = torch.full((n,), -1, dtype=torch.int64, device="cuda")
        ps = torch.full((n,), -1.0, dtype=torch.float64, device="cuda")
        rc = lib.vfi_psnr_u8_noalloc(engine._h, C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), n, a[0].numel(),
                                     C.c_void_p(sse.data_ptr()), C.c_void_p(ps.data_ptr()), stream)
        assert rc == 0, lib.vfi_last_error(engine._h)
        want_sse, want_ps = engine.psnr_sse(a, b)                     # the existing entry point (vfi_psnr_u8)
        assert sse.tolist() == want_sse.tolist()
        assert sse.tolist() == [int(((a[i].cpu().numpy().astype(np.int64) - b[i].cpu().numpy()) ** 2).sum()) for i in range(n)]
        assert torch.equal(ps, want_ps) and float(ps[0]) == float("inf")
    # the host wrapper is the same call; either output may be omitted
    a, b = _pair(2, 45, 71, 3)
    sse = torch.empty(2, dtype=torch.int64, device="cuda")
    engine.psnr_sse_noalloc(a, b, sse)
    assert sse.tolist() == engine.psnr_sse(a, b)[0].tolist()


def test_noalloc_rejects_more_pairs_than_the_static_scratch_holds(engine):
    lib = _abi.lib()
    a, b = _pair(1, 8, 8, 0)
    out = torch.empty(1, dtype=torch.int64, device="cuda")
    args = (engine._h, C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()))
    rc = lib.vfi_psnr_u8_noalloc(*args, _abi.NOALLOC_MAX_PAIRS + 1, 1, C.c_void_p(out.data_ptr()), C.c_void_p(0), C.c_void_p(0))
    assert rc == _abi.E_ARG and b"exceed the static scratch" in lib.vfi_last_error(engine._h)   # checked before any launch
    assert lib.vfi_psnr_u8_noalloc(*args, 1, a[0].numel(), C.c_void_p(0), C.c_void_p(0), C.c_void_p(0)) == _abi.E_ARG
