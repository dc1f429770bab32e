"""Thin host wrapper over the C ABI: one vfi_handle per CUDA device, torch tensors for device memory
and streams only.  Every method launches the sm_100a kernels; nothing here computes on the CPU."""
import ctypes as C

import numpy as np
import torch

from . import _abi
from . import ifnet_params

_ENGINES = {}


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def library_device_bytes():
    """Device memory held by every live library handle of this process (workspaces + packed weights)."""
    return sum(e.device_bytes() for e in _ENGINES.values())


class Engine:
    def __init__(self, device=0):
        if not torch.cuda.is_available():
            raise RuntimeError("vfi_b200 needs an sm_100 CUDA device; there is no CPU fallback")
        self.device = torch.device("cuda", device)
        self._lib = _abi.lib()
        h = C.c_void_p()
        rc = self._lib.vfi_create(device, C.byref(h))
        if rc != 0:
            raise _abi.VFIError(f"vfi_create(device={device}) failed ({rc}): needs an sm_100 (B200) GPU")
        self._h = h
        self.weights_loaded = False
        self._graph_launches = 0      # kernels launched by CUDA-graph replays (the library counts them at capture only)
        self._ts_last = None          # timesteps the handle's device copy holds (set by the last eager infer)

    @staticmethod
    def get(device=None):
        """The per-device engine singleton; None / 'cuda' = the CURRENT CUDA device (the rank's GPU under torchrun)."""
        if isinstance(device, str):
            device = torch.device(device)
        if isinstance(device, torch.device):
            device = device.index
        if device is None:
            if not torch.cuda.is_available():
                raise RuntimeError("vfi_b200 needs an sm_100 CUDA device; there is no CPU fallback")
            device = torch.cuda.current_device()
        if device not in _ENGINES:
            _ENGINES[device] = Engine(device)
        return _ENGINES[device]

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.vfi_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @property
    def launches(self):
        return int(self._lib.vfi_launch_count(self._h)) + self._graph_launches

    def fence(self):
        """Order the current stream into the handle's call sequence (vfi_fence): before and after a graph replay."""
        self._check(self._lib.vfi_fence(self._h, _stream()), "vfi_fence")

    def device_bytes(self):
        """Device memory owned by the library handle (workspace + scratch + packed weights): invisible to torch's stats."""
        return int(self._lib.vfi_device_bytes(self._h))

    def set_option(self, key, value):
        self._check(self._lib.vfi_set_option(self._h, key.encode(), int(value)), "vfi_set_option")

    def profile(self, on):
        self._check(self._lib.vfi_profile_enable(self._h, 1 if on else 0), "vfi_profile_enable")

    def profile_read(self):
        """{class: (ms, launches)} since the last read; synchronises the device."""
        n = len(_abi.PROFILE_CLASSES)
        ms = (C.c_double * n)()
        cnt = (C.c_longlong * n)()
        self._check(self._lib.vfi_profile_read(self._h, C.cast(ms, C.c_void_p), C.cast(cnt, C.c_void_p), n),
                    "vfi_profile_read")
        return {k: (ms[i], int(cnt[i])) for i, k in enumerate(_abi.PROFILE_CLASSES)}

    def _check(self, rc, what):
        _abi.check(self._h, rc, what)

    def _u8(self, x):
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(np.ascontiguousarray(x))
        if x.dtype != torch.uint8:
            raise TypeError("expected uint8 frames")
        return x.to(self.device, non_blocking=True).contiguous()

    # ---- IFNet ---------------------------------------------------------------------------
    def load_weights(self, state_dict, lrelu_slope=None):
        flat = ifnet_params.flatten(state_dict)
        slope = C.byref(C.c_float(lrelu_slope)) if lrelu_slope is not None else None
        with torch.cuda.device(self.device):
            rc = self._lib.vfi_ifnet_load_weights(self._h, flat.ctypes.data_as(C.c_void_p), flat.size,
                                                  C.cast(slope, C.c_void_p) if slope is not None else C.c_void_p(0))
        self._check(rc, "vfi_ifnet_load_weights")
        self.weights_loaded = True

    def infer(self, img0, img1, timesteps, scale=1.0, precision=_abi.PRECISION_BF16, hud_mask=None,
              want_float=False, want_flow=False, out=None, flags=0):
        """img0/img1: (B,H,W,3) or (H,W,3) uint8 (numpy or torch).  Returns uint8 (B,T,H,W,3) on the GPU
        (plus float frames / final flow when asked).  flags: _abi.FLAG_BGR | _abi.FLAG_PAD_REFLECT_EDGE (the
        conventions of scripts/run_experiments.py:365-386)."""
        img0, img1 = self._u8(img0), self._u8(img1)
        if img0.dim() == 3:
            img0, img1 = img0[None], img1[None]
        B, H, W, _ = img0.shape
        ts = (C.c_float * len(timesteps))(*[float(t) for t in timesteps])
        T = len(timesteps)
        with torch.cuda.device(self.device):
            if out is None:
                out = torch.empty((B, T, H, W, 3), dtype=torch.uint8, device=self.device)
            of = torch.empty((B, T, H, W, 3), dtype=torch.float32, device=self.device) if want_float else None
            fl = torch.empty((B, T, H, W, 4), dtype=torch.float32, device=self.device) if want_flow else None
            hm = self._u8(hud_mask) if hud_mask is not None else None
            rc = self._lib.vfi_ifnet_infer_ex(self._h, _ptr(img0), _ptr(img1), B, H, W, C.cast(ts, C.c_void_p), T,
                                              float(scale), int(precision), int(flags), _ptr(hm), _ptr(out), _ptr(of),
                                              _ptr(fl), _stream())
        self._check(rc, "vfi_ifnet_infer_ex")
        self._ts_last = tuple(float(t) for t in timesteps)
        res = (out,)
        if want_float:
            res += (of,)
        if want_flow:
            res += (fl,)
        return res if len(res) > 1 else out

    def warp_blend(self, img0, img1, flow, mask, want_float=False):
        img0, img1 = self._u8(img0), self._u8(img1)
        N, H, W, _ = img0.shape
        flow = flow.to(self.device, torch.float32).contiguous()
        mask = mask.to(self.device, torch.float32).contiguous()
        with torch.cuda.device(self.device):
            out = torch.empty((N, H, W, 3), dtype=torch.uint8, device=self.device)
            of = torch.empty((N, H, W, 3), dtype=torch.float32, device=self.device) if want_float else None
            rc = self._lib.vfi_warp_blend(self._h, _ptr(img0), _ptr(img1), _ptr(flow), _ptr(mask), N, H, W,
                                          _ptr(out), _ptr(of), _stream())
        self._check(rc, "vfi_warp_blend")
        return (out, of) if want_float else out

    def debug_conv(self, layer, precision, x, residual=None):
        x = x.to(self.device, torch.float32).contiguous()
        N, Hin, Win, _ = x.shape
        from .ifnet_layers import layer_out_shape
        oshape = layer_out_shape(layer, N, Hin, Win)
        with torch.cuda.device(self.device):
            y = torch.zeros(oshape, dtype=torch.float32, device=self.device)
            r = residual.to(self.device, torch.float32).contiguous() if residual is not None else None
            rc = self._lib.vfi_debug_conv(self._h, layer, precision, _ptr(x), N, Hin, Win, _ptr(r), _ptr(y), _stream())
        self._check(rc, "vfi_debug_conv")
        return y

    # ---- metrics / routing reductions ----------------------------------------------------
    def psnr_sse(self, a, b):
        """Exact integer SSE per frame pair.  a, b: (n, ...) uint8, same shape."""
        a, b = self._u8(a), self._u8(b)
        n = a.shape[0]
        elems = a[0].numel()
        with torch.cuda.device(self.device):
            sse = torch.empty(n, dtype=torch.int64, device=self.device)
            ps = torch.empty(n, dtype=torch.float64, device=self.device)
            rc = self._lib.vfi_psnr_u8(self._h, _ptr(a), _ptr(b), n, elems, _ptr(sse), _ptr(ps), _stream())
        self._check(rc, "vfi_psnr_u8")
        return sse, ps

    def psnr_sse_noalloc(self, a, b, sse_out, psnr_out=None):
        """vfi_psnr_u8_noalloc: vfi_psnr_u8 with the accumulators in the library's static device array -- nothing is
        allocated, here or in the library.  a, b: contiguous uint8 CUDA tensors (n, ...) on this device; sse_out (n,) int64 and
        psnr_out (n,) float64 (optional) are the caller's CUDA tensors.  n <= _abi.NOALLOC_MAX_PAIRS; overlapping calls race."""
        for t in (a, b, sse_out) + ((psnr_out,) if psnr_out is not None else ()):
            if not (isinstance(t, torch.Tensor) and t.is_cuda and t.is_contiguous()):
                raise TypeError("psnr_sse_noalloc takes contiguous CUDA tensors (it does not copy or allocate)")
        rc = self._lib.vfi_psnr_u8_noalloc(self._h, _ptr(a), _ptr(b), a.shape[0], a[0].numel(), _ptr(sse_out), _ptr(psnr_out),
                                           _stream())
        self._check(rc, "vfi_psnr_u8_noalloc")

    def moments(self, a, b):
        """Exact global moments per pair: (n,5) int64 = sum a, sum b, sum a^2, sum b^2, sum ab over every byte."""
        a, b = self._u8(a), self._u8(b)
        n = a.shape[0]
        with torch.cuda.device(self.device):
            out = torch.empty((n, 5), dtype=torch.int64, device=self.device)
            rc = self._lib.vfi_moments_u8(self._h, _ptr(a), _ptr(b), n, a[0].numel(), _ptr(out), _stream())
        self._check(rc, "vfi_moments_u8")
        return out

    def ssim(self, a, b, downscale=1):
        a, b = self._u8(a), self._u8(b)
        n, H, W, _ = a.shape
        with torch.cuda.device(self.device):
            out = torch.empty(n, dtype=torch.float64, device=self.device)
            rc = self._lib.vfi_ssim_gray_u8(self._h, _ptr(a), _ptr(b), n, H, W, downscale, _ptr(out), _stream())
        self._check(rc, "vfi_ssim_gray_u8")
        return out

    def psnr_ssim(self, a, b):
        """PSNR (exact integer SSE) and full-resolution SSIM of n frame pairs from one pass over the frames.
        a, b: (n,H,W,3) uint8.  Returns (sse int64, psnr float64, ssim float64), each (n,), on the device."""
        a, b = self._u8(a), self._u8(b)
        n, H, W, _ = a.shape
        with torch.cuda.device(self.device):
            sse = torch.empty(n, dtype=torch.int64, device=self.device)
            ps = torch.empty(n, dtype=torch.float64, device=self.device)
            ss = torch.empty(n, dtype=torch.float64, device=self.device)
            rc = self._lib.vfi_psnr_ssim_u8(self._h, _ptr(a), _ptr(b), n, H, W, _ptr(sse), _ptr(ps), _ptr(ss), _stream())
        self._check(rc, "vfi_psnr_ssim_u8")
        return sse, ps, ss

    def flow_stats(self, flow):
        """flow: (n,H,W,2) float32 -> (n,3) float64 [mean, max, std]."""
        if isinstance(flow, np.ndarray):
            flow = torch.from_numpy(np.ascontiguousarray(flow))
        flow = flow.to(self.device, torch.float32, non_blocking=True).contiguous()
        n, H, W, _ = flow.shape
        with torch.cuda.device(self.device):
            out = torch.empty((n, 3), dtype=torch.float64, device=self.device)
            rc = self._lib.vfi_flow_stats(self._h, _ptr(flow), n, H, W, _ptr(out), _stream())
        self._check(rc, "vfi_flow_stats")
        return out

    def farneback(self, img0, img1, pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2,
                  want_flow=True, want_stats=False):
        """Dense Farneback flow of n frame pairs on the device (defaults = adaptive_pipeline.py:119-128).
        img0/img1: (n,H,W,3) RGB or (n,H,W) gray uint8.  Returns flow (n,H,W,2) f32 and/or stats (n,3) f64."""
        img0, img1 = self._u8(img0), self._u8(img1)
        ch = 3 if img0.dim() == 4 else 1
        n, H, W = img0.shape[:3]
        with torch.cuda.device(self.device):
            flow = torch.empty((n, H, W, 2), dtype=torch.float32, device=self.device) if want_flow else None
            st = torch.empty((n, 3), dtype=torch.float64, device=self.device) if want_stats else None
            rc = self._lib.vfi_farneback(self._h, _ptr(img0), _ptr(img1), n, H, W, ch, float(pyr_scale), int(levels),
                                         int(winsize), int(iterations), int(poly_n), float(poly_sigma), _ptr(flow),
                                         _ptr(st), _stream())
        self._check(rc, "vfi_farneback")
        if want_flow and want_stats:
            return flow, st
        return flow if want_flow else st

    def resize_gray(self, img, dsize, bgr=False):
        """cv2.resize(cv2.cvtColor(img, *2GRAY), dsize) on the device.  img: (n,H,W,3) or (n,H,W) uint8; dsize = (dw, dh)."""
        img = self._u8(img)
        ch = 3 if img.dim() == 4 else 1
        n, H, W = img.shape[:3]
        dw, dh = int(dsize[0]), int(dsize[1])
        with torch.cuda.device(self.device):
            out = torch.empty((n, dh, dw), dtype=torch.uint8, device=self.device)
            rc = self._lib.vfi_resize_gray_u8(self._h, _ptr(img), n, H, W, ch, 1 if bgr else 0, dh, dw, _ptr(out), _stream())
        self._check(rc, "vfi_resize_gray_u8")
        return out

    def add_weighted(self, a, alpha, b, beta, out=None):
        """cv2.addWeighted(a, alpha, b, beta, 0) for uint8 arrays of one shape, on the device."""
        a, b = self._u8(a), self._u8(b)
        with torch.cuda.device(self.device):
            if out is None:
                out = torch.empty_like(a)
            rc = self._lib.vfi_add_weighted_u8(self._h, _ptr(a), _ptr(b), a.numel(), float(alpha), float(beta), _ptr(out), _stream())
        self._check(rc, "vfi_add_weighted_u8")
        return out

    def canny(self, gray, low=100, high=200, want_edges=True):
        """cv2.Canny(gray, low, high) on the device.  gray: (n,H,W) uint8 -> (edges (n,H,W) uint8 | None, density (n,) f64)."""
        gray = self._u8(gray)
        n, H, W = gray.shape
        with torch.cuda.device(self.device):
            edges = torch.empty((n, H, W), dtype=torch.uint8, device=self.device) if want_edges else None
            dens = torch.empty(n, dtype=torch.float64, device=self.device)
            rc = self._lib.vfi_canny_u8(self._h, _ptr(gray), n, H, W, float(low), float(high), _ptr(edges), _ptr(dens), _stream())
        self._check(rc, "vfi_canny_u8")
        return edges, dens

    def canny_density(self, gray, low=100, high=200):
        return self.canny(gray, low, high, want_edges=False)[1]

    def remap_flow(self, img, flow, t):
        """cv2.remap(img, x + flow[...,0]*t, y + flow[...,1]*t, INTER_LINEAR) on the device.  img (H,W,C) uint8, flow (H,W,2) f32."""
        img = self._u8(img)
        flow = flow.to(self.device, torch.float32).contiguous() if isinstance(flow, torch.Tensor) else \
            torch.from_numpy(np.ascontiguousarray(flow, dtype=np.float32)).to(self.device)
        H, W = img.shape[:2]
        ch = img.shape[2] if img.dim() == 3 else 1
        with torch.cuda.device(self.device):
            out = torch.empty_like(img)
            rc = self._lib.vfi_remap_flow_u8(self._h, _ptr(img), _ptr(flow), float(t), H, W, ch, _ptr(out), _stream())
        self._check(rc, "vfi_remap_flow_u8")
        return out

    def to_yuv420(self, frames, bgr=False, out=None):
        """cv2.cvtColor(f, COLOR_RGB2YUV_I420) of every frame on the device.  frames: (..., H, W, 3) uint8 -> (..., H*3//2, W)."""
        frames = self._u8(frames)
        H, W = frames.shape[-3], frames.shape[-2]
        n = frames.numel() // (H * W * 3)
        with torch.cuda.device(self.device):
            if out is None:
                out = torch.empty(tuple(frames.shape[:-3]) + (H * 3 // 2, W), dtype=torch.uint8, device=self.device)
            rc = self._lib.vfi_rgb_to_yuv420_u8(self._h, _ptr(frames), n, H, W, 1 if bgr else 0, _ptr(out), _stream())
        self._check(rc, "vfi_rgb_to_yuv420_u8")
        return out

    def laplacian_var(self, rgb):
        rgb = self._u8(rgb)
        n, H, W, _ = rgb.shape
        with torch.cuda.device(self.device):
            out = torch.empty(n, dtype=torch.float64, device=self.device)
            rc = self._lib.vfi_laplacian_var(self._h, _ptr(rgb), n, H, W, _ptr(out), _stream())
        self._check(rc, "vfi_laplacian_var")
        return out


class HudState:
    """Device-resident ring of 180x320 gray thumbnails + the current mask (AdaptiveRouter.frame_history)."""

    def __init__(self, engine, hist_len=10, var_threshold=10.0):
        self.e = engine
        self.hist_len = hist_len
        self.var_threshold = float(var_threshold)
        self.count = C.c_longlong(0)
        self.ring = torch.zeros((hist_len, 180, 320), dtype=torch.uint8, device=engine.device)
        self.mask = None
        self.coverage = None

    def __len__(self):
        return min(int(self.count.value), self.hist_len)

    def update(self, rgb, mask_out=None, cov_out=None):
        """Push one frame.  mask_out (H,W) uint8 / cov_out (1,) float64 device tensors receive the results when given
        (the batched clip path keeps one mask per pair of a batch)."""
        rgb = self.e._u8(rgb)
        H, W, _ = rgb.shape
        with torch.cuda.device(self.e.device):
            mask = mask_out if mask_out is not None else torch.empty((H, W), dtype=torch.uint8, device=self.e.device)
            cov = cov_out if cov_out is not None else torch.empty(1, dtype=torch.float64, device=self.e.device)
            rc = self.e._lib.vfi_hud_update(self.e._h, _ptr(rgb), H, W, _ptr(self.ring), self.hist_len,
                                            C.byref(self.count), self.var_threshold, _ptr(mask), _ptr(cov), _stream())
        self.e._check(rc, "vfi_hud_update")
        self.mask, self.coverage = mask, cov
        return mask, cov
