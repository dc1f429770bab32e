"""Guard-band test of every C-ABI entry point that writes a caller-supplied buffer: each result tensor the Engine hands to the
library is carved out of the middle of a larger allocation whose 4 KiB on either side hold a known pattern, at ragged frame
sizes (odd widths, heights below one tile, sizes that are no multiple of any kernel's tile); after the call the pattern must be
intact.  compute-sanitizer is not available on the GPU pool, so this is the repo's own check for writes past either end of an
output (parity of the values is the other tests' job; here only the bytes AROUND them matter)."""
import numpy as np
import pytest
import torch

import vfi_b200  # noqa: F401
from vfi_b200 import _abi
from vfi_b200 import engine as engine_mod

pytestmark = pytest.mark.gpu
GUARD = 4096                      # bytes on either side; a multiple of torch's 512-byte allocation alignment
PATTERN = 0xA5


class _GuardedTorch:
    """Stands in for the `torch` module inside vfi_b200.engine: empty / empty_like / zeros return the middle of a padded buffer."""

    def __init__(self):
        self.buffers = []

    def __getattr__(self, name):
        return getattr(torch, name)

    def _carve(self, shape, dtype, device, fill):
        shape = tuple(int(s) for s in (shape if isinstance(shape, (tuple, list, torch.Size)) else (shape,)))
        nbytes = int(np.prod(shape, dtype=np.int64)) * torch.empty((), dtype=dtype).element_size()
        body = (nbytes + 511) // 512 * 512
        raw = torch.full((GUARD + body + GUARD,), PATTERN, dtype=torch.uint8, device=device)
        self.buffers.append((raw, nbytes))
        view = raw[GUARD:GUARD + nbytes].view(dtype).view(shape)
        if fill is not None:
            view.fill_(fill)
        return view

    def empty(self, *shape, dtype=torch.float32, device=None, **kw):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list, torch.Size)):
            shape = shape[0]
        return self._carve(shape, dtype, device, None)

    def zeros(self, *shape, dtype=torch.float32, device=None, **kw):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list, torch.Size)):
            shape = shape[0]
        return self._carve(shape, dtype, device, 0)

    def empty_like(self, t, **kw):
        return self._carve(tuple(t.shape), t.dtype, t.device, None)

    def check(self, what):
        torch.cuda.synchronize()
        assert self.buffers, f"{what}: no result buffer went through the guard allocator"
        for raw, nbytes in self.buffers:
            head = raw[:GUARD]
            tail = raw[GUARD + nbytes:]                       # includes the round-up slack after the tensor's last byte
            assert bool((head == PATTERN).all()), f"{what}: wrote BEFORE a {nbytes}-byte result"
            assert bool((tail == PATTERN).all()), f"{what}: wrote PAST the end of a {nbytes}-byte result"
        self.buffers.clear()


@pytest.fixture()
def guarded(monkeypatch):
    g = _GuardedTorch()
    monkeypatch.setattr(engine_mod, "torch", g)
    return g


def _frames(n, H, W, seed, ch=3):
    rng = np.random.default_rng(seed)
    shape = (n, H, W, ch) if ch else (n, H, W)
    return torch.from_numpy(rng.integers(0, 256, shape, dtype=np.uint8)).cuda()


RAGGED = [(45, 71), (33, 47), (129, 250), (7, 300), (270, 482)]


@pytest.mark.parametrize("H,W", RAGGED)
def test_metric_and_statistic_outputs_stay_in_bounds(engine, guarded, H, W):
    a, b = _frames(3, H, W, 1), _frames(3, H, W, 2)
    engine.psnr_sse(a, b); guarded.check("vfi_psnr_u8")
    engine.moments(a, b); guarded.check("vfi_moments_u8")
    if H >= 7 and W >= 7:
        engine.ssim(a, b, 1); guarded.check("vfi_ssim_gray_u8(1)")
        engine.psnr_ssim(a, b); guarded.check("vfi_psnr_ssim_u8")
    if H // 4 >= 7 and W // 4 >= 7:
        engine.ssim(a, b, 4); guarded.check("vfi_ssim_gray_u8(4)")
    engine.laplacian_var(a); guarded.check("vfi_laplacian_var")
    flow = torch.randn(3, H, W, 2, device="cuda") * 5
    engine.flow_stats(flow); guarded.check("vfi_flow_stats")
    engine.add_weighted(a, 0.3, b, 0.7); guarded.check("vfi_add_weighted_u8")
    engine.resize_gray(a, (W // 2 + 1, H // 2 + 1)); guarded.check("vfi_resize_gray_u8 (down)")
    engine.resize_gray(a[..., 0].contiguous(), (W + 13, H + 5)); guarded.check("vfi_resize_gray_u8 (up, gray)")
    engine.canny(a[..., 1].contiguous()); guarded.check("vfi_canny_u8")
    engine.remap_flow(a[0], flow[0], 0.5); guarded.check("vfi_remap_flow_u8")
    He, We = H & ~1, W & ~1
    if He >= 2 and We >= 2:
        engine.to_yuv420(a[:, :He, :We].contiguous()); guarded.check("vfi_rgb_to_yuv420_u8")


@pytest.mark.parametrize("H,W", [(100, 100), (129, 250), (270, 482)])
def test_farneback_and_hud_outputs_stay_in_bounds(engine, guarded, H, W):
    a, b = _frames(2, H, W, 3), _frames(2, H, W, 4)
    engine.farneback(a, b, want_flow=True, want_stats=True); guarded.check("vfi_farneback (rgb)")
    engine.farneback(a[..., 0].contiguous(), b[..., 0].contiguous(), want_flow=False, want_stats=True)
    guarded.check("vfi_farneback (gray, stats only)")
    hud = engine_mod.HudState(engine)                         # its thumbnail ring comes from the guard allocator too
    for i in range(3):
        hud.update(a[i % 2])
    guarded.check("vfi_hud_update")


@pytest.mark.parametrize("H,W", [(45, 71), (129, 250), (270, 482)])
@pytest.mark.parametrize("precision", [_abi.PRECISION_BF16, _abi.PRECISION_FP32])
def test_ifnet_outputs_stay_in_bounds(engine, guarded, H, W, precision):
    a, b = _frames(2, H, W, 5), _frames(2, H, W, 6)
    engine.infer(a, b, [0.25, 0.5, 0.75], 1.0, precision, want_float=True, want_flow=True)
    guarded.check("vfi_ifnet_infer_ex (frames, float frames, flow)")
    mask = (torch.rand(2, H, W, device="cuda") < 0.1).to(torch.uint8)
    # reflect padding needs a frame larger than the pad (torch's F.pad raises the same way): 45 x 71 cannot be padded to the
    # 128-pixel multiple that scale 0.5 and the runner's convention ask for, so that case stays at scale 1 / base.py's padding
    small = min(H, W) < 128
    engine.infer(a, b, [0.5], 1.0 if small else 0.5, precision, hud_mask=mask,
                 flags=_abi.FLAG_BGR | (0 if small else _abi.FLAG_PAD_REFLECT_EDGE))
    guarded.check("vfi_ifnet_infer_ex (HUD composite, BGR, scale 0.5 + runner padding where the frame allows)")
    flow = torch.randn(2, H, W, 4, device="cuda") * 3
    m = torch.randn(2, H, W, device="cuda")
    engine.warp_blend(a, b, flow, m, want_float=True); guarded.check("vfi_warp_blend")


def test_the_guard_itself_catches_a_one_byte_overrun(guarded):
    t = guarded.empty((3, 5), dtype=torch.uint8, device="cuda")
    raw, nbytes = guarded.buffers[-1]
    raw[GUARD + nbytes] = 0                                  # the byte right after the tensor
    with pytest.raises(AssertionError, match="PAST the end"):
        guarded.check("self-test")
    guarded.buffers.clear()
    t = guarded.zeros(7, dtype=torch.float64, device="cuda")
    raw, nbytes = guarded.buffers[-1]
    raw[GUARD - 1] = 0                                       # the byte right before it
    with pytest.raises(AssertionError, match="BEFORE"):
        guarded.check("self-test")
    guarded.buffers.clear()
    assert t.shape == (7,) and float(t.sum()) == 0.0


# ---- the library's OWN intermediate buffers: vfi_set_option("ws_guard", 1) (csrc/abi.cu ws_guard_check) ------------------------
# Every buffer vfi_ifnet_infer_ex carves from its workspace is followed by a 4 KiB pattern gap that a device kernel checks after
# the blend; the workspace is pattern-filled first, so equal frames with and without the guard also show that no kernel reads
# workspace bytes it (or an earlier kernel of the same call) did not write.
KERNEL_VARIANTS = [{}, {"gc_persistent": 0}, {"fuse_gather3": 1}, {"fuse_gather": 0}, {"v2_tma": 0}, {"v2_pm": 0}, {"conv_v2": 0}, {"t_fast": 0}]
OPTION_DEFAULTS = {"gc_persistent": 1, "fuse_gather3": 0, "fuse_gather": 1, "v2_tma": 1, "v2_pm": 1, "conv_v2": 1, "t_fast": 1}


def _infer_guarded(engine, a, b, ts, scale, precision, **kw):
    want = engine.infer(a, b, ts, scale, precision, **kw).clone()
    engine.set_option("ws_guard", 1)
    try:
        got = engine.infer(a, b, ts, scale, precision, **kw)          # raises VFIError if a gap byte changed
    finally:
        engine.set_option("ws_guard", 0)
    assert torch.equal(got, want), "frames depend on workspace bytes no kernel of the call wrote"


@pytest.mark.parametrize("variant", KERNEL_VARIANTS, ids=lambda v: "-".join(f"{k}{x}" for k, x in v.items()) or "default")
@pytest.mark.parametrize("H,W", [(45, 71), (129, 250), (270, 482)])
def test_ifnet_workspace_buffers_stay_in_bounds_bf16(engine, H, W, variant):
    a, b = _frames(2, H, W, 7), _frames(2, H, W, 8)
    for k, v in variant.items():
        engine.set_option(k, v)
    try:
        _infer_guarded(engine, a, b, [0.25, 0.5, 0.75], 1.0, _abi.PRECISION_BF16)
        if min(H, W) >= 128:
            _infer_guarded(engine, a, b, [0.5], 0.5, _abi.PRECISION_BF16, flags=_abi.FLAG_PAD_REFLECT_EDGE)
        if min(H, W) >= 260:
            _infer_guarded(engine, a, b, [0.5], 0.25, _abi.PRECISION_BF16)
    finally:
        for k in variant:
            engine.set_option(k, OPTION_DEFAULTS[k])


@pytest.mark.parametrize("H,W", [(45, 71), (270, 482)])
def test_ifnet_workspace_buffers_stay_in_bounds_fp32_and_clip_mode(engine, H, W):
    clip = _frames(4, H, W, 9)
    _infer_guarded(engine, clip[:-1], clip[1:], [0.5], 1.0, _abi.PRECISION_FP32)       # consecutive pairs: shared Head images
    _infer_guarded(engine, clip[:-1], clip[1:], [0.25, 0.5, 0.75], 1.0, _abi.PRECISION_BF16)
    if min(H, W) >= 128:
        _infer_guarded(engine, clip[:1], clip[1:2], [0.5], 0.5, _abi.PRECISION_FP32)


def test_workspace_guard_self_test_fires(engine):
    a, b = _frames(1, 129, 250, 10), _frames(1, 129, 250, 11)
    engine.set_option("ws_guard", 2)                                   # the library overwrites one gap byte itself
    try:
        with pytest.raises(_abi.VFIError, match="workspace guard"):
            engine.infer(a, b, [0.5], 1.0, _abi.PRECISION_BF16)
    finally:
        engine.set_option("ws_guard", 0)
    engine.infer(a, b, [0.5], 1.0, _abi.PRECISION_BF16)                # and the handle is usable afterwards


@pytest.mark.parametrize("H,W", [(100, 100), (129, 250), (270, 482), (360, 641)])
def test_farneback_workspace_buffers_stay_in_bounds(engine, H, W):
    a, b = _frames(3, H, W, 12), _frames(3, H, W, 13)
    for kw in ({}, {"levels": 5, "iterations": 2}, {"poly_n": 7, "poly_sigma": 1.5, "winsize": 21}):
        want_flow, want_stats = engine.farneback(a, b, want_flow=True, want_stats=True, **kw)
        want_flow = want_flow.clone()
        engine.set_option("ws_guard", 1)
        try:
            flow, stats = engine.farneback(a, b, want_flow=True, want_stats=True, **kw)   # raises VFIError if a gap byte changed
            stats_only = engine.farneback(a, b, want_flow=False, want_stats=True, **kw)   # final flow lands in the workspace
        finally:
            engine.set_option("ws_guard", 0)
        # the pattern-filled workspace changes nothing: no kernel reads bytes the call did not write
        assert torch.equal(flow, want_flow) and torch.equal(stats, want_stats) and torch.equal(stats_only, want_stats)
