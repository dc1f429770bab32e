"""RIFE v4.25 behind the reference's RIFEModel interface, running on the B200 kernels.

Drop-in for the reference's models/sota/rife_wrapper.py: same constructor (:34-40), model-directory
search and FileNotFoundError behaviour (:44-72), `info` (:74-85), `load` (:87-108), `interpolate`
(:110-154), Lanczos `upscale` (:156-164), `interpolate_recursive` (:166-203), `RIFELiteModel`
(:206-224) and `get_rife_model` (:227-245).  What changed underneath: to_tensor, reflect padding, the
IFNet pass for ALL timesteps (Head features computed once per pair), unpad and the *255/clip/uint8
conversion run as one C-ABI call (vfi_ifnet_infer) on the GPU; only uint8 frames cross PCIe.
Deviations: pad multiple is max(32, 64/scale) so 1440p works.  A train_log directory
without flownet.pkl raises FileNotFoundError at load() like the reference does when its weights are missing --
unless the caller opts into seeded random-init weights (allow_random_weights=True or VFI_B200_RANDOM_WEIGHTS=1:
the synthetic benchmark configs, tests and smoke; there is no network to fetch the release weights).
"""
import os
import sys
from pathlib import Path
from typing import List, Optional

import cv2
import numpy as np
import torch

from ..base import BaseModel, ModelInfo
from ... import _abi, ifnet_params
from ...engine import Engine

_REPO_ROOT = Path(__file__).resolve().parents[3]


class RIFEModel(BaseModel):
    """RIFE v4.25/4.26 - Real-time Intermediate Flow Estimation (arbitrary timesteps)."""

    def __init__(self, model_dir: Optional[str] = None, version: str = 'v4.25', device: str = 'cuda',
                 uhd_mode: bool = False, *, precision: str = 'bf16', weight_seed: int = 1234,
                 allow_random_weights: bool = False):
        super().__init__(device)
        if model_dir is None:
            for cand in (Path('external/Practical-RIFE'), Path('../external/Practical-RIFE'),
                         Path.home() / 'gaming-vfisr' / 'external' / 'Practical-RIFE',
                         _REPO_ROOT / 'external' / 'Practical-RIFE'):
                if cand.exists():
                    model_dir = str(cand)
                    break
            if model_dir is None:
                raise FileNotFoundError("Could not find Practical-RIFE directory. "
                                        "Please clone it to external/Practical-RIFE")
        self.model_dir = Path(model_dir)
        self.version = version
        self.uhd_mode = uhd_mode
        self.weights_dir = self.model_dir / 'train_log'
        if not self.weights_dir.exists():
            raise FileNotFoundError(f"RIFE weights not found at {self.weights_dir}. "
                                    "Download from: https://github.com/hzwer/Practical-RIFE/releases")
        if precision not in ('bf16', 'fp32'):
            raise ValueError("precision must be 'bf16' or 'fp32'")
        self.precision = precision
        self.weight_seed = weight_seed
        self.allow_random_weights = allow_random_weights or os.environ.get('VFI_B200_RANDOM_WEIGHTS') == '1'
        self._engine = None
        self._state_dict = None
        self._pinned = {}
        self._copier = None
        self._graphs = {}
        self.use_graph = False if os.environ.get('VFI_B200_NO_GRAPH') == '1' else \
            ('always' if os.environ.get('VFI_B200_GRAPH') == '1' else True)

    @property
    def info(self) -> ModelInfo:
        return ModelInfo(name=f'RIFE_{self.version}', type='sota', supports_vfi=True, supports_sr=False,
                         supports_joint=False, parameters=9_800_000, requires_gpu=True,
                         description="Real-time VFI, supports arbitrary timesteps, ~400fps @ 1080p")

    # the reference keeps the torch module in self._model; here it is the engine handle
    def state_dict(self):
        if self._state_dict is None:
            if (self.weights_dir / 'flownet.pkl').exists():
                self._state_dict = ifnet_params.load_pkl(self.weights_dir)
            elif self.allow_random_weights:
                print(f"WARNING: {self.weights_dir / 'flownet.pkl'} not found -- using RANDOM-INIT IFNet weights (seed "
                      f"{self.weight_seed}); the interpolated frames are meaningless outside the synthetic benchmarks",
                      file=sys.stderr)
                self._state_dict = ifnet_params.random_init(self.weight_seed)
            else:
                raise FileNotFoundError(f"RIFE weights not found at {self.weights_dir / 'flownet.pkl'}. "
                                        "Download from: https://github.com/hzwer/Practical-RIFE/releases "
                                        "(or pass allow_random_weights=True for the synthetic benchmark configs)")
        return self._state_dict

    def load(self) -> None:
        print(f"Loading RIFE {self.version}...")
        if self.device.type != 'cuda':
            raise RuntimeError("RIFE (b200) needs a CUDA sm_100 device; there is no CPU fallback")
        self._engine = Engine.get(self.device)
        self._engine.load_weights(self.state_dict())
        self._model = self._engine
        self._loaded = True
        print(f"RIFE {self.version} loaded successfully")

    def _pin(self, frame: np.ndarray, slot: int) -> torch.Tensor:
        key = (slot, frame.shape)
        buf = self._pinned.get(key)
        if buf is None:
            buf = torch.empty(frame.shape, dtype=torch.uint8).pin_memory()
            self._pinned[key] = buf
        return buf

    def _stage(self, frame: np.ndarray, slot: int) -> torch.Tensor:
        """uint8 frame -> pinned host buffer -> async H2D copy (3 B/px cross PCIe, not 12)."""
        buf = self._pin(frame, slot)
        buf.numpy()[...] = frame
        return buf.to(self.device, non_blocking=True)

    def _stage_pair(self, frame0: np.ndarray, frame1: np.ndarray):
        """Both frames of a pair: the second frame's copy into its pinned buffer runs on a helper thread (numpy releases the
        GIL for the copy) while this thread stages the first and starts its upload: the two 6 MB host copies of a 1080p pair
        (0.5 ms each) overlap each other and the first H2D transfer."""
        if frame0.nbytes < (1 << 20):
            return self._stage(frame0, 0), self._stage(frame1, 1)
        if self._copier is None:
            from concurrent.futures import ThreadPoolExecutor
            self._copier = ThreadPoolExecutor(1)
        b1 = self._pin(frame1, 1)
        fut = self._copier.submit(np.copyto, b1.numpy(), frame1)
        d0 = self._stage(frame0, 0)
        fut.result()
        return d0, b1.to(self.device, non_blocking=True)

    def interpolate(self, frame0: np.ndarray, frame1: np.ndarray, num_frames: int = 3,
                    timestamps: Optional[List[float]] = None) -> List[np.ndarray]:
        self.ensure_loaded()
        if timestamps is None:
            timestamps = self.get_default_timestamps(num_frames)
        return self._interpolate_masked(frame0, frame1, timestamps, None)

    def _interpolate_masked(self, frame0, frame1, timestamps, hud_mask_dev):
        """interpolate() with the adaptive pipeline's HUD composite fused into the blend kernel's store.
        One pair per call is the reference's call shape (scripts/run_benchmarks.py:142-149): the ~70 launches of a pair are
        replayed from a CUDA graph (captured per frame size / timestep list), the frames go up from pinned staging buffers
        and the result comes down into pinned memory (the returned arrays are views of it: no pageable bounce copy)."""
        if len(timestamps) == 0:
            return []
        ts = tuple(float(t) for t in timestamps)
        H, W = frame0.shape[:2]
        # measured on B200 (profiles/r02_latency.txt): the replay wins 11 % at 360p (launch-bound), nothing at 720p, and loses
        # 1 % at 1080p (GPU-bound; the fixed staging buffers cost more than the launches) -> graphs up to 720p only
        small = H * W <= 1280 * 720 or self.use_graph == 'always'
        if not (self.use_graph and small) or frame0.shape != frame1.shape or frame0.dtype != np.uint8 or frame0.ndim != 3:
            return self._interpolate_eager(frame0, frame1, list(ts), hud_mask_dev)
        key = (H, W, ts, self.uhd_mode, self.precision)
        g = self._graphs.get(key)
        if g is None:
            if len(self._graphs) >= 4:                       # a handful of shapes at most: drop the oldest graph's buffers
                self._graphs.pop(next(iter(self._graphs)))
            g = self._graphs[key] = _PairGraph(self, H, W, ts)
        return g.run(frame0, frame1, hud_mask_dev)

    def _interpolate_eager(self, frame0, frame1, timestamps, hud_mask_dev):
        prec = _abi.PRECISION_BF16 if self.precision == 'bf16' else _abi.PRECISION_FP32
        scale = 0.5 if self.uhd_mode else 1.0
        hm = hud_mask_dev[None] if hud_mask_dev is not None else None
        d0, d1 = self._stage_pair(frame0, frame1)
        out = self._engine.infer(d0[None], d1[None], timestamps, scale, prec, hud_mask=hm)
        host = torch.empty(out.shape[1:], dtype=torch.uint8, pin_memory=True)   # pinned, recycled by torch's host allocator
        host.copy_(out[0], non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()   # synchronises, as the reference's .cpu() does (base.py:200)
        host = host.numpy()
        return [host[i] for i in range(len(timestamps))]

    def interpolate_batch(self, frames0, frames1, timestamps, hud_masks=None, device_out=False):
        """Extra (not in the reference): B pairs x T timesteps in one launch sequence.
        frames*: (B,H,W,3) uint8 numpy or CUDA tensors.  Returns (B,T,H,W,3) uint8."""
        self.ensure_loaded()
        prec = _abi.PRECISION_BF16 if self.precision == 'bf16' else _abi.PRECISION_FP32
        scale = 0.5 if self.uhd_mode else 1.0
        out = self._engine.infer(frames0, frames1, list(timestamps), scale, prec, hud_mask=hud_masks)
        return out if device_out else out.cpu().numpy()

    def clip_streamer(self, pairs: int, H: int, W: int, timestamps, depth: int = 2, output: str = 'rgb') -> "ClipStreamer":
        """Extra (not in the reference): pipelined clip interpolation, see ClipStreamer.  output='yuv420': the results come
        down as planar I420 frames (cv2 COLOR_RGB2YUV_I420, (H*3/2, W) uint8 each) -- half the device -> host bytes."""
        self.ensure_loaded()
        return ClipStreamer(self, pairs, H, W, list(timestamps), depth, output)

    def upscale(self, frame: np.ndarray, scale: float = 1.333) -> np.ndarray:
        """RIFE does no SR: Lanczos on the host exactly as the reference (rife_wrapper.py:156-164)."""
        h, w = frame.shape[:2]
        return cv2.resize(frame, (int(w * scale), int(h * scale)), interpolation=cv2.INTER_LANCZOS4)

    def interpolate_recursive(self, frame0: np.ndarray, frame1: np.ndarray, num_iterations: int = 2) -> List[np.ndarray]:
        """2^n frame multiplication by repeated midpoint interpolation (rife_wrapper.py:166-203)."""
        self.ensure_loaded()
        if num_iterations < 1:
            return []
        seq = [frame0, frame1]
        for _ in range(num_iterations):
            nxt = [seq[0]]
            for a, b in zip(seq[:-1], seq[1:]):
                nxt += [self.interpolate(a, b, num_frames=1)[0], b]
            seq = nxt
        return seq[1:-1]


class _PairGraph:
    """The launch sequence of ONE pair (B = 1, fixed timesteps / size) captured once and replayed per call."""

    def __init__(self, model: "RIFEModel", H: int, W: int, ts):
        self.m, self.ts, self.T = model, ts, len(ts)
        dev, eng = model.device, model._engine
        self.prec = _abi.PRECISION_BF16 if model.precision == 'bf16' else _abi.PRECISION_FP32
        self.scale = 0.5 if model.uhd_mode else 1.0
        with torch.cuda.device(dev):
            self.stream = torch.cuda.Stream(dev)
            self.f0 = torch.empty((1, H, W, 3), dtype=torch.uint8, device=dev)
            self.f1 = torch.empty((1, H, W, 3), dtype=torch.uint8, device=dev)
            self.mask = torch.zeros((1, H, W), dtype=torch.uint8, device=dev)
            self.out = torch.empty((1, self.T, H, W, 3), dtype=torch.uint8, device=dev)
            self.pin = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
            self.mask_zero = True
            self.f0.zero_()
            self.f1.zero_()
            self.stream.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(self.stream):
                for _ in range(2):      # warm-up: sizes the workspace, uploads the timesteps, sets kernel attributes
                    eng.infer(self.f0, self.f1, list(ts), self.scale, self.prec, hud_mask=self.mask, out=self.out)
                self.stream.synchronize()
                l0 = eng.launches
                self.graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.graph, stream=self.stream):
                    eng.infer(self.f0, self.f1, list(ts), self.scale, self.prec, hud_mask=self.mask, out=self.out)
                self.n_launches = eng.launches - l0
            torch.cuda.current_stream(dev).wait_stream(self.stream)

    def run(self, frame0, frame1, hud_mask_dev):
        eng, dev = self.m._engine, self.m.device
        if eng._ts_last != self.ts:          # another call changed the handle's device copy of the timesteps: go eager once
            return self.m._interpolate_eager(frame0, frame1, list(self.ts), hud_mask_dev)
        self.pin[0].numpy()[...] = frame0
        self.pin[1].numpy()[...] = frame1
        cur = torch.cuda.current_stream(dev)
        self.stream.wait_stream(cur)         # the HUD mask (and anything else the caller queued) is ready
        with torch.cuda.stream(self.stream):
            self.f0[0].copy_(self.pin[0], non_blocking=True)
            self.f1[0].copy_(self.pin[1], non_blocking=True)
            if hud_mask_dev is not None:
                self.mask[0].copy_(hud_mask_dev, non_blocking=True)
                self.mask_zero = False
            elif not self.mask_zero:
                self.mask.zero_()
                self.mask_zero = True
            eng.fence()
            self.graph.replay()
            eng._graph_launches += self.n_launches
            eng.fence()
            host = torch.empty(self.out.shape[1:], dtype=torch.uint8, pin_memory=True)
            host.copy_(self.out[0], non_blocking=True)
        self.stream.synchronize()            # the reference's .cpu() synchronises here (base.py:200)
        host = host.numpy()
        return [host[i] for i in range(self.T)]


class ClipStreamer:
    """Streams a clip through RIFE in batches of `pairs` consecutive frame pairs with the host<->device copies
    overlapped with the kernels: three CUDA streams (upload, compute, download) and `depth` buffer slots.
    A batch is the pairs+1 consecutive frames clip[s : s+pairs+1] (frame k and k+1 form pair k), so every frame crosses
    PCIe once, as uint8.  The reference's hot loop (scripts/run_benchmarks.py:142) does upload, compute and
    download serially per pair.

        st = model.clip_streamer(pairs=8, H=1080, W=1920, timestamps=[0.25, 0.5, 0.75])
        slot = st.submit(frames_pinned)      # (pairs+1, H, W, 3) uint8 pinned host tensor; returns immediately
        out = st.result(slot)                # (pairs, T, H, W, 3) uint8 pinned host tensor (valid until the slot is reused)
    """

    def __init__(self, model: "RIFEModel", pairs: int, H: int, W: int, timestamps, depth: int = 2, output: str = 'rgb'):
        if output not in ('rgb', 'yuv420'):
            raise ValueError("output must be 'rgb' or 'yuv420'")
        if output == 'yuv420' and (H % 2 or W % 2):
            raise ValueError("yuv420 output needs even frame sizes")
        self.m, self.B, self.ts, self.depth, self.output = model, pairs, timestamps, depth, output
        dev = model.device
        T = len(timestamps)
        self.prec = _abi.PRECISION_BF16 if model.precision == 'bf16' else _abi.PRECISION_FP32
        self.scale = 0.5 if model.uhd_mode else 1.0
        with torch.cuda.device(dev):
            self.s_in, self.s_cmp, self.s_out = (torch.cuda.Stream(dev) for _ in range(3))
            self.in_dev = [torch.empty((pairs + 1, H, W, 3), dtype=torch.uint8, device=dev) for _ in range(depth)]
            self.out_dev = [torch.empty((pairs, T, H, W, 3), dtype=torch.uint8, device=dev) for _ in range(depth)]
            res_shape = (pairs, T, H, W, 3) if output == 'rgb' else (pairs, T, H * 3 // 2, W)
            self.yuv_dev = [torch.empty(res_shape, dtype=torch.uint8, device=dev) for _ in range(depth)] if output == 'yuv420' else None
            self.out_host = [torch.empty(res_shape, dtype=torch.uint8).pin_memory() for _ in range(depth)]
            self.ev_in = [torch.cuda.Event() for _ in range(depth)]
            self.ev_cmp = [torch.cuda.Event() for _ in range(depth)]
            self.ev_out = [torch.cuda.Event() for _ in range(depth)]
        self.k = 0
        self.h2d_bytes = (pairs + 1) * H * W * 3
        self.d2h_bytes = pairs * T * H * W * 3 // (1 if output == 'rgb' else 2)

    def submit(self, frames_host: torch.Tensor) -> int:
        i = self.k % self.depth
        first = self.k < self.depth
        cur = torch.cuda.current_stream(self.m.device)
        if self.k == 0:
            for s in (self.s_in, self.s_cmp, self.s_out):
                s.wait_stream(cur)
        with torch.cuda.stream(self.s_in):
            if not first:
                self.s_in.wait_event(self.ev_cmp[i])        # the kernels that read this input slot are done
            self.in_dev[i].copy_(frames_host, non_blocking=True)
            self.ev_in[i].record(self.s_in)
        with torch.cuda.stream(self.s_cmp):
            self.s_cmp.wait_event(self.ev_in[i])
            if not first:
                self.s_cmp.wait_event(self.ev_out[i])       # the previous result in this output slot has left
            self.m._engine.infer(self.in_dev[i][:-1], self.in_dev[i][1:], self.ts, self.scale, self.prec,
                                 out=self.out_dev[i])
            if self.yuv_dev is not None:
                self.m._engine.to_yuv420(self.out_dev[i], out=self.yuv_dev[i])
            self.ev_cmp[i].record(self.s_cmp)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_cmp[i])
            self.out_host[i].copy_(self.out_dev[i] if self.yuv_dev is None else self.yuv_dev[i], non_blocking=True)
            self.ev_out[i].record(self.s_out)
        self.k += 1
        return i

    def result(self, slot: int) -> torch.Tensor:
        self.ev_out[slot].synchronize()
        return self.out_host[slot]

    def join(self) -> None:
        """Make the current stream wait for everything submitted so far (for event timing / ordering)."""
        cur = torch.cuda.current_stream(self.m.device)
        for s in (self.s_in, self.s_cmp, self.s_out):
            cur.wait_stream(s)


class RIFELiteModel(RIFEModel):
    """RIFE Lite: same interface, different `info` (rife_wrapper.py:206-224)."""

    @property
    def info(self) -> ModelInfo:
        return ModelInfo(name=f'RIFE_Lite_{self.version}', type='sota', supports_vfi=True, supports_sr=False,
                         supports_joint=False, parameters=4_500_000, requires_gpu=True,
                         description="RIFE Lite - faster than full RIFE, slight quality tradeoff")


def get_rife_model(version: str = 'v4.25', lite: bool = False, **kwargs) -> RIFEModel:
    return (RIFELiteModel if lite else RIFEModel)(version=version, **kwargs)
