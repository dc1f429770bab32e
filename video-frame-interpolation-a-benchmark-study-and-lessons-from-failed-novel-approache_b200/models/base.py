"""Model interface, API-compatible with the reference's models/base.py.

Same public surface (reference models/base.py): ModelInfo :17-27, InferenceResult :30-41,
BaseModel :44-233 (info / load / interpolate / upscale / process_pair / ensure_loaded and the
to_tensor / to_numpy / pad_to_multiple / unpad / get_default_timestamps helpers), JointModel :236-288,
TwoStageModel :291-364.  Frames are (H, W, 3) uint8 RGB numpy arrays in and out.
"""
import abc
import dataclasses
import time
from typing import List, Optional, Tuple

import numpy as np
import torch


@dataclasses.dataclass
class ModelInfo:
    name: str
    type: str                      # 'traditional' | 'sota' | 'novel'
    supports_vfi: bool
    supports_sr: bool
    supports_joint: bool
    parameters: Optional[int] = None
    requires_gpu: bool = True
    description: str = ""


@dataclasses.dataclass
class InferenceResult:
    frames: List[np.ndarray]
    inference_time_ms: float
    vram_peak_mb: float
    model_used: str = ""
    extra_info: dict = None

    def __post_init__(self):
        self.extra_info = {} if self.extra_info is None else self.extra_info


def _timed(fn):
    """Wall-clock + peak-VRAM bracket used by every process_pair (base.py:140-158)."""
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.reset_peak_memory_stats()
    t0 = time.perf_counter()
    frames = fn()
    ms = (time.perf_counter() - t0) * 1000.0
    peak = 0
    if cuda:  # torch's allocator statistics + what the library handle cudaMalloc'ed itself (workspace, packed weights)
        from ..engine import library_device_bytes
        peak = (torch.cuda.max_memory_allocated() + library_device_bytes()) / 1e6
    return frames, ms, peak


class BaseModel(abc.ABC):
    def __init__(self, device: str = 'cuda'):
        self.device = torch.device(device if torch.cuda.is_available() else 'cpu')
        self._model = None
        self._loaded = False

    @property
    @abc.abstractmethod
    def info(self) -> ModelInfo:
        ...

    @abc.abstractmethod
    def load(self) -> None:
        ...

    @abc.abstractmethod
    def interpolate(self, frame0: np.ndarray, frame1: np.ndarray, num_frames: int = 3,
                    timestamps: Optional[List[float]] = None) -> List[np.ndarray]:
        ...

    @abc.abstractmethod
    def upscale(self, frame: np.ndarray, scale: float = 1.333) -> np.ndarray:
        ...

    def process_pair(self, frame0: np.ndarray, frame1: np.ndarray, num_intermediate: int = 3,
                     target_scale: float = 1.333) -> InferenceResult:
        """[frame0, interp_1..n, frame1], each upscaled; timed end to end (base.py:114-165)."""
        if not self._loaded:
            raise RuntimeError(f"Model {self.info.name} not loaded. Call load() first.")

        def run():
            mids = self.interpolate(frame0, frame1, num_intermediate)
            return [self.upscale(f, target_scale) for f in [frame0, *mids, frame1]]

        frames, ms, peak = _timed(run)
        return InferenceResult(frames=frames, inference_time_ms=ms, vram_peak_mb=peak, model_used=self.info.name)

    def ensure_loaded(self) -> None:
        if not self._loaded:
            self.load()
            self._loaded = True

    # ---- helpers kept for callers that use them directly ----------------------------------
    def to_tensor(self, frame: np.ndarray) -> torch.Tensor:
        """(H,W,C) uint8 -> (1,C,H,W) float32 in [0,1] on self.device."""
        t = torch.from_numpy(frame).permute(2, 0, 1).float() / 255.0
        return t.unsqueeze(0).to(self.device)

    def to_numpy(self, tensor: torch.Tensor) -> np.ndarray:
        """(1,C,H,W)/(C,H,W) float -> (H,W,C) uint8 with the reference's truncating cast."""
        if tensor.dim() == 4:
            tensor = tensor.squeeze(0)
        arr = tensor.permute(1, 2, 0).cpu().numpy()
        return (arr * 255).clip(0, 255).astype(np.uint8)

    def pad_to_multiple(self, tensor: torch.Tensor, multiple: int = 32) -> Tuple[torch.Tensor, Tuple[int, int, int, int]]:
        h, w = tensor.shape[-2:]
        ph, pw = (-h) % multiple, (-w) % multiple
        if ph == 0 and pw == 0:
            return tensor, (0, 0, 0, 0)
        pad = (0, pw, 0, ph)
        return torch.nn.functional.pad(tensor, pad, mode='reflect'), pad

    def unpad(self, tensor: torch.Tensor, original_h: int, original_w: int) -> torch.Tensor:
        return tensor[:, :, :original_h, :original_w]

    def get_default_timestamps(self, num_frames: int) -> List[float]:
        return [(i + 1) / (num_frames + 1) for i in range(num_frames)]


class JointModel(BaseModel):
    """Joint VFI+SR models: one pass produces all frames at target resolution (base.py:236-288)."""

    @abc.abstractmethod
    def joint_process(self, frame0: np.ndarray, frame1: np.ndarray, num_intermediate: int = 3,
                      target_scale: float = 1.333) -> List[np.ndarray]:
        ...

    def process_pair(self, frame0, frame1, num_intermediate: int = 3, target_scale: float = 1.333) -> InferenceResult:
        if not self._loaded:
            raise RuntimeError(f"Model {self.info.name} not loaded. Call load() first.")
        frames, ms, peak = _timed(lambda: self.joint_process(frame0, frame1, num_intermediate, target_scale))
        return InferenceResult(frames=frames, inference_time_ms=ms, vram_peak_mb=peak, model_used=self.info.name)


class TwoStageModel(BaseModel):
    """A VFI model followed by an SR model (base.py:291-364)."""

    def __init__(self, vfi_model: BaseModel, sr_model: BaseModel):
        self.vfi, self.sr = vfi_model, sr_model
        self._loaded = False

    @property
    def info(self) -> ModelInfo:
        a, b = self.vfi.info, self.sr.info
        return ModelInfo(name=f'{a.name}+{b.name}', type='sota', supports_vfi=True, supports_sr=True,
                         supports_joint=False, parameters=(a.parameters or 0) + (b.parameters or 0),
                         requires_gpu=a.requires_gpu or b.requires_gpu,
                         description=f"Two-stage: {a.name} for VFI, {b.name} for SR")

    def load(self) -> None:
        print(f"Loading two-stage model: {self.info.name}")
        self.vfi.load()
        self.sr.load()
        self._loaded = True

    def interpolate(self, frame0, frame1, num_frames=3, timestamps=None):
        return self.vfi.interpolate(frame0, frame1, num_frames, timestamps)

    def upscale(self, frame, scale=1.333):
        return self.sr.upscale(frame, scale)

    def process_pair(self, frame0, frame1, num_intermediate=3, target_scale=1.333) -> InferenceResult:
        if not self._loaded:
            raise RuntimeError("Model not loaded")

        def run():
            mids = self.vfi.interpolate(frame0, frame1, num_intermediate)
            return [self.sr.upscale(f, target_scale) for f in [frame0, *mids, frame1]]

        frames, ms, peak = _timed(run)
        return InferenceResult(frames=frames, inference_time_ms=ms, vram_peak_mb=peak, model_used=self.info.name,
                               extra_info={'vfi_model': self.vfi.info.name, 'sr_model': self.sr.info.name})
