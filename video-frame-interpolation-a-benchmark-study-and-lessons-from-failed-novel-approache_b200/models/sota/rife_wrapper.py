"""RIFE v4.25 behind the reference's RIFEModel interface, running on the B200 kernels.

Drop-in for the reference's models/sota/rife_wrapper.py: same constructor (:34-40), model-directory
search and FileNotFoundError behaviour (:44-72), `info` (:74-85), `load` (:87-108), `interpolate`
(:110-154), Lanczos `upscale` (:156-164), `interpolate_recursive` (:166-203), `RIFELiteModel`
(:206-224) and `get_rife_model` (:227-245).  What changed underneath: to_tensor, reflect padding, the
IFNet pass for ALL timesteps (Head features computed once per pair), unpad and the *255/clip/uint8
conversion run as one C-ABI call (vfi_ifnet_infer) on the GPU; only uint8 frames cross PCIe.
Deviations: pad multiple is max(32, 64/scale) so 1440p works; a train_log
directory without flownet.pkl means seeded random-init weights (there is no network to fetch them).
"""
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
                 uhd_mode: bool = False, *, precision: str = 'bf16', weight_seed: int = 1234):
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
        self._engine = None
        self._state_dict = None
        self._pinned = {}

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
            else:
                self._state_dict = ifnet_params.random_init(self.weight_seed)
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

    def _stage(self, frame: np.ndarray, slot: int) -> torch.Tensor:
        """uint8 frame -> pinned host buffer -> async H2D copy (3 B/px cross PCIe, not 12)."""
        key = (slot, frame.shape)
        buf = self._pinned.get(key)
        if buf is None:
            buf = torch.empty(frame.shape, dtype=torch.uint8).pin_memory()
            self._pinned[key] = buf
        buf.numpy()[...] = frame
        return buf.to(self.device, non_blocking=True)

    def interpolate(self, frame0: np.ndarray, frame1: np.ndarray, num_frames: int = 3,
                    timestamps: Optional[List[float]] = None) -> List[np.ndarray]:
        self.ensure_loaded()
        if timestamps is None:
            timestamps = self.get_default_timestamps(num_frames)
        return self._interpolate_masked(frame0, frame1, timestamps, None)

    def _interpolate_masked(self, frame0, frame1, timestamps, hud_mask_dev):
        """interpolate() with the adaptive pipeline's HUD composite fused into the blend kernel's store."""
        if len(timestamps) == 0:
            return []
        prec = _abi.PRECISION_BF16 if self.precision == 'bf16' else _abi.PRECISION_FP32
        scale = 0.5 if self.uhd_mode else 1.0
        hm = hud_mask_dev[None] if hud_mask_dev is not None else None
        out = self._engine.infer(self._stage(frame0, 0)[None], self._stage(frame1, 1)[None], timestamps, scale, prec,
                                 hud_mask=hm)
        host = out[0].cpu().numpy()          # synchronises, as the reference's .cpu() does (base.py:200)
        return [host[i] for i in range(len(timestamps))]

    def interpolate_batch(self, frames0, frames1, timestamps, hud_masks=None, device_out=False):
        """Extra (not in the reference): B pairs x T timesteps in one launch sequence.
        frames*: (B,H,W,3) uint8 numpy or CUDA tensors.  Returns (B,T,H,W,3) uint8."""
        self.ensure_loaded()
        prec = _abi.PRECISION_BF16 if self.precision == 'bf16' else _abi.PRECISION_FP32
        scale = 0.5 if self.uhd_mode else 1.0
        out = self._engine.infer(frames0, frames1, list(timestamps), scale, prec, hud_mask=hud_masks)
        return out if device_out else out.cpu().numpy()

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


class RIFELiteModel(RIFEModel):
    """RIFE Lite: same interface, different `info` (rife_wrapper.py:206-224)."""

    @property
    def info(self) -> ModelInfo:
        return ModelInfo(name=f'RIFE_Lite_{self.version}', type='sota', supports_vfi=True, supports_sr=False,
                         supports_joint=False, parameters=4_500_000, requires_gpu=True,
                         description="RIFE Lite - faster than full RIFE, slight quality tradeoff")


def get_rife_model(version: str = 'v4.25', lite: bool = False, **kwargs) -> RIFEModel:
    return (RIFELiteModel if lite else RIFEModel)(version=version, **kwargs)
