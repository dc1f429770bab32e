"""AdaptiveVFI routing + RIFE fast path behind the reference's AdaptivePipeline interface.

Drop-in for the reference's models/novel/adaptive_pipeline.py: ContentAnalysis :30-40, RoutingStats
:43-71, AdaptiveRouter :74-286 (same six constructor thresholds :81-96, same public methods and
state), AdaptivePipeline :289-483.  The per-pair statistics run as reduction kernels through the
C ABI: scene-cut SSIM (vfi_ssim_gray_u8, downscale 4), dense Farneback flow with the reference's
parameters (:119-128) and its |flow| mean/max/std (vfi_farneback -> vfi_flow_stats, the flow field never
leaves the GPU), Laplacian variance (vfi_laplacian_var), HUD temporal-variance mask (vfi_hud_update);
the decision tree is vfi_route_decide.  `flow_fn=cv2_farneback_flow` swaps OpenCV's own flow back in
(host, uploaded) -- used by the tests that compare every float64 field bit for bit with the reference.
Fixes kept minimal: the RIFE / VFIMamba imports point at models/sota.
"""
import dataclasses
import time
from collections import deque
from typing import List, Optional, Tuple

import cv2
import numpy as np
import torch

from ..base import BaseModel, InferenceResult, ModelInfo
from ... import _abi
from ...engine import Engine, HudState


@dataclasses.dataclass
class ContentAnalysis:
    motion_mean: float
    motion_max: float
    motion_std: float
    has_particles: bool
    is_scene_change: bool
    hud_coverage: float
    recommended_model: str
    confidence: float


@dataclasses.dataclass
class RoutingStats:
    total: int = 0
    rife_count: int = 0
    vfimamba_count: int = 0
    scene_change_count: int = 0

    def add(self, model: str):
        self.total += 1
        field = {'rife': 'rife_count', 'vfimamba': 'vfimamba_count', 'scene_change': 'scene_change_count'}.get(model)
        if field:
            setattr(self, field, getattr(self, field) + 1)

    def to_dict(self) -> dict:
        if not self.total:
            return {'total': 0}
        pct = lambda n: n / self.total * 100
        return {'total': self.total, 'rife': self.rife_count, 'rife_pct': pct(self.rife_count),
                'vfimamba': self.vfimamba_count, 'vfimamba_pct': pct(self.vfimamba_count),
                'scene_change': self.scene_change_count, 'scene_change_pct': pct(self.scene_change_count)}


FARNEBACK = dict(pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2)   # reference :119-128


def cv2_farneback_flow(frame0: np.ndarray, frame1: np.ndarray) -> np.ndarray:
    """OpenCV's own Farneback on the host (reference parameters adaptive_pipeline.py:119-128): optional flow_fn."""
    g0 = cv2.cvtColor(frame0, cv2.COLOR_RGB2GRAY)
    g1 = cv2.cvtColor(frame1, cv2.COLOR_RGB2GRAY)
    return cv2.calcOpticalFlowFarneback(g0, g1, None, pyr_scale=0.5, levels=3, winsize=15, iterations=3,
                                        poly_n=5, poly_sigma=1.2, flags=0)


class AdaptiveRouter:
    """Analyses a frame pair and recommends a model; statistics are computed on the GPU."""

    def __init__(self, motion_threshold_low: float = 5.0, motion_threshold_high: float = 25.0,
                 scene_change_threshold: float = 0.65, particle_threshold: float = 0.4,
                 hud_variance_threshold: float = 10.0, hud_history_frames: int = 10, *, device=None, flow_fn=None):
        self.motion_threshold_low = motion_threshold_low
        self.motion_threshold_high = motion_threshold_high
        self.scene_change_threshold = scene_change_threshold
        self.particle_threshold = particle_threshold
        self.hud_variance_threshold = hud_variance_threshold
        self.hud_history_frames = hud_history_frames
        self.hud_mask: Optional[np.ndarray] = None
        self._device = device
        self._flow_fn = flow_fn          # None: Farneback on the GPU (vfi_farneback); callable(f0, f1) -> host flow (H,W,2)
        self._engine = None
        self._hud = None
        self._hud_mask_dev = None
        self._last = {}

    # -- plumbing ------------------------------------------------------------------------
    @property
    def engine(self) -> Engine:
        if self._engine is None:
            self._engine = Engine.get(self._device)      # None -> the current CUDA device
        return self._engine

    def reset(self) -> None:
        """Forget the HUD history and mask (a fresh router with the same thresholds): a new clip starts."""
        self._hud = None
        self._hud_mask_dev = None
        self.hud_mask = None
        self._last = {}

    def _hud_state(self) -> HudState:
        if self._hud is None:
            self._hud = HudState(self.engine, self.hud_history_frames, self.hud_variance_threshold)
        return self._hud

    @property
    def frame_history(self) -> deque:
        """The last <= hud_history_frames gray 320x180 thumbnails, oldest first (reference :105)."""
        d = deque(maxlen=self.hud_history_frames)
        if self._hud is not None:
            n, c, L = len(self._hud), int(self._hud.count.value), self.hud_history_frames
            ring = self._hud.ring.cpu().numpy()
            for k in range(c - n, c):
                d.append(ring[k % L])
        return d

    def _dev(self, frame):
        return self.engine._u8(frame)

    # -- the reference's public methods ----------------------------------------------------
    def compute_motion(self, frame0: np.ndarray, frame1: np.ndarray) -> Tuple[float, float, float, np.ndarray]:
        """(mean, max, std, |flow| map) -- reference :108-138."""
        if self._flow_fn is None:
            flow, st = self.engine.farneback(self._dev(frame0)[None], self._dev(frame1)[None], want_stats=True, **FARNEBACK)
            flow = flow[0]
        else:
            flow = torch.from_numpy(self._flow_fn(frame0, frame1)).to(self.engine.device)
            st = self.engine.flow_stats(flow[None])
        st = st[0].tolist()
        # the reference hands the |flow| map back (:131,138); callers on the hot path (analyze) never ask for it
        mag = torch.sqrt(flow[..., 0] ** 2 + flow[..., 1] ** 2).cpu().numpy()
        return st[0], st[1], st[2], mag

    def detect_scene_change(self, frame0: np.ndarray, frame1: np.ndarray) -> Tuple[bool, float]:
        """SSIM of the quarter-resolution gray pair < threshold -- reference :140-160."""
        score = float(self.engine.ssim(self._dev(frame0)[None], self._dev(frame1)[None], downscale=4)[0])
        return score < self.scene_change_threshold, score

    def detect_particles(self, frame0: np.ndarray, flow_magnitude, flow_std: float) -> Tuple[bool, float]:
        """sqrt(min(std/20,1) * min(var(Laplacian)/500,1)) > threshold -- reference :162-191."""
        lv = float(self.engine.laplacian_var(self._dev(frame0)[None])[0])
        self._last['laplacian_var'] = lv
        _, _, hp, ps = _abi.route_decide(1.0, 0.0, 0.0, flow_std, lv, 0.0, self.particle_threshold, 1.0, 1e300)
        return hp, ps

    def detect_hud(self, frame: np.ndarray) -> Tuple[np.ndarray, float]:
        """Static-region mask from the variance of the last five thumbnails -- reference :193-235."""
        mask, cov = self._hud_state().update(self._dev(frame))
        cov = float(cov[0])
        if len(self._hud) < 5:
            # the reference returns zeros here WITHOUT touching self.hud_mask (:207-209)
            return np.zeros(frame.shape[:2], dtype=bool), 0.0
        self._hud_mask_dev = mask
        self.hud_mask = mask.cpu().numpy().astype(bool)
        return self.hud_mask, cov

    def analyze(self, frame0: np.ndarray, frame1: np.ndarray) -> ContentAnalysis:
        """Scene cut (early return) -> motion -> particles -> HUD -> decision tree (reference :237-286)."""
        e = self.engine
        d0, d1 = self._dev(frame0), self._dev(frame1)
        score = float(e.ssim(d0[None], d1[None], downscale=4)[0])
        if score < self.scene_change_threshold:
            return ContentAnalysis(0, 0, 0, False, True, 0, 'scene_change', 1.0 - score)
        if self._flow_fn is None:
            st = e.farneback(d0[None], d1[None], want_flow=False, want_stats=True, **FARNEBACK)
        else:
            st = e.flow_stats(torch.from_numpy(self._flow_fn(frame0, frame1))[None])
        lv = e.laplacian_var(d0[None])
        mean, mx, std = st[0].tolist()
        lv = float(lv[0])
        _, hud_cov = self._detect_hud_dev(d0, frame0.shape[:2])
        route, conf, hp, ps = _abi.route_decide(score, mean, mx, std, lv, self.scene_change_threshold,
                                                self.particle_threshold, self.motion_threshold_low,
                                                self.motion_threshold_high)
        self._last = {'ssim': score, 'laplacian_var': lv, 'particle_score': ps}
        return ContentAnalysis(motion_mean=mean, motion_max=mx, motion_std=std, has_particles=hp,
                               is_scene_change=False, hud_coverage=hud_cov,
                               recommended_model=_abi.ROUTE_NAMES[route], confidence=conf)

    def _detect_hud_dev(self, dframe, shape):
        mask, cov = self._hud_state().update(dframe)
        if len(self._hud) < 5:
            return None, 0.0
        self._hud_mask_dev = mask
        self.hud_mask = None            # host copy made lazily by hud_mask_host()
        return mask, float(cov[0])

    def hud_mask_host(self):
        if self.hud_mask is None and self._hud_mask_dev is not None:
            self.hud_mask = self._hud_mask_dev.cpu().numpy().astype(bool)
        return self.hud_mask


class AdaptivePipeline(BaseModel):
    """Routes pairs by content; RIFE fast path on the B200 kernels; HUD pixels copied from the nearer input."""

    def __init__(self, device: str = 'cuda', motion_threshold_low: float = 5.0, motion_threshold_high: float = 25.0,
                 enable_vfimamba: bool = True, sr_model_name: str = 'lanczos', *, allow_random_weights: bool = False,
                 flow_fn=None):
        super().__init__(device)
        self._allow_random_weights = allow_random_weights
        self.enable_vfimamba = enable_vfimamba
        self.sr_model_name = sr_model_name
        self.router = AdaptiveRouter(motion_threshold_low=motion_threshold_low,
                                     motion_threshold_high=motion_threshold_high,
                                     device=self.device if self.device.type == 'cuda' else None, flow_fn=flow_fn)
        self.stats = RoutingStats()
        self._rife = None
        self._vfimamba = None
        self._sr = None

    @property
    def info(self) -> ModelInfo:
        return ModelInfo(name='AdaptivePipeline', type='novel', supports_vfi=True, supports_sr=True,
                         supports_joint=False, parameters=27_700_000, requires_gpu=True,
                         description="Novel adaptive routing: fast RIFE for easy content, quality VFIMamba for hard content")

    def load(self) -> None:
        print("Loading Adaptive Pipeline...")
        from ..sota.rife_wrapper import RIFEModel
        self._rife = RIFEModel(device=str(self.device), allow_random_weights=self._allow_random_weights)
        self._rife.load()
        if self.enable_vfimamba:
            try:
                from ..sota.vfimamba_wrapper import VFIMambaModel  # not part of this build: falls back below
                self._vfimamba = VFIMambaModel(device=str(self.device))
                self._vfimamba.load()
            except Exception as e:  # same fallback as the reference (:358-361)
                print(f"Warning: Could not load VFIMamba: {e}")
                print("Will use RIFE for all frames")
                self.enable_vfimamba = False
        self._loaded = True
        print("Adaptive Pipeline ready")

    def interpolate(self, frame0: np.ndarray, frame1: np.ndarray, num_frames: int = 3,
                    timestamps: Optional[List[float]] = None) -> List[np.ndarray]:
        return self._interpolate_routed(frame0, frame1, num_frames, timestamps)[0]

    def _interpolate_routed(self, frame0, frame1, num_frames=3, timestamps=None):
        """interpolate() that also hands back the ContentAnalysis it routed on."""
        self.ensure_loaded()
        analysis = self.router.analyze(frame0, frame1)
        return self._route(analysis, frame0, frame1, num_frames, timestamps), analysis

    def _route(self, analysis, frame0, frame1, num_frames, timestamps):
        if analysis.is_scene_change:
            self.stats.add('scene_change')
            return [frame0.copy() for _ in range(num_frames)]
        use_hud = analysis.hud_coverage > 0.01 and self.router._hud_mask_dev is not None
        if analysis.recommended_model == 'vfimamba' and self.enable_vfimamba:
            self.stats.add('vfimamba')
            out = self._vfimamba.interpolate(frame0, frame1, num_frames, timestamps)
            if use_hud:  # foreign model: composite on its host frames (reference :399-409)
                ts = timestamps if timestamps is not None else self._rife.get_default_timestamps(num_frames)
                m = self.router.hud_mask_host()
                for fr, t in zip(out, ts):
                    fr[m] = (frame0 if t < 0.5 else frame1)[m]
            return out
        self.stats.add('rife')
        ts = timestamps if timestamps is not None else self._rife.get_default_timestamps(num_frames)
        return self._rife._interpolate_masked(frame0, frame1, ts, self.router._hud_mask_dev if use_hud else None)

    def upscale(self, frame: np.ndarray, scale: float = 1.333) -> np.ndarray:
        h, w = frame.shape[:2]
        return cv2.resize(frame, (int(w * scale), int(h * scale)), interpolation=cv2.INTER_LANCZOS4)

    def process_pair(self, frame0: np.ndarray, frame1: np.ndarray, num_intermediate: int = 3,
                     target_scale: float = 1.333) -> InferenceResult:
        """Analyse (for extra_info) + adaptive interpolate (analyses again, as the reference :443,:382) + upscale."""
        self.ensure_loaded()
        if torch.cuda.is_available():
            torch.cuda.reset_peak_memory_stats()
        t0 = time.perf_counter()
        analysis = self.router.analyze(frame0, frame1)
        mids = self.interpolate(frame0, frame1, num_intermediate)
        frames = [self.upscale(f, target_scale) for f in [frame0, *mids, frame1]]
        ms = (time.perf_counter() - t0) * 1000
        return InferenceResult(frames=frames, inference_time_ms=ms, vram_peak_mb=self._vram_mb(), model_used=self.info.name,
                               extra_info={'analysis': {'motion_mean': analysis.motion_mean,
                                                        'motion_max': analysis.motion_max,
                                                        'has_particles': analysis.has_particles,
                                                        'is_scene_change': analysis.is_scene_change,
                                                        'hud_coverage': analysis.hud_coverage,
                                                        'recommended_model': analysis.recommended_model},
                                           'routing_stats': self.stats.to_dict()})

    def process_pair_fast(self, frame0: np.ndarray, frame1: np.ndarray, num_intermediate: int = 3,
                          target_scale: float = 1.333) -> InferenceResult:
        """Extra: process_pair with ONE analysis per pair (the reference analyses twice, :443 + :382,
        pushing the HUD history twice).  Same outputs for a clip whose HUD history is pushed once per pair."""
        self.ensure_loaded()
        t0 = time.perf_counter()
        mids, analysis = self._interpolate_routed(frame0, frame1, num_intermediate)
        frames = [self.upscale(f, target_scale) for f in [frame0, *mids, frame1]]
        ms = (time.perf_counter() - t0) * 1000
        return InferenceResult(frames=frames, inference_time_ms=ms, vram_peak_mb=self._vram_mb(), model_used=self.info.name,
                               extra_info={'analysis': dataclasses.asdict(analysis), 'routing_stats': self.stats.to_dict()})

    def _vram_mb(self) -> float:
        from ...engine import library_device_bytes
        return (torch.cuda.max_memory_allocated() + library_device_bytes()) / 1e6

    def process_clip(self, frames, num_frames: int = 3, rank: int = 0, world: int = 1, group=None,
                     on_pair=None) -> dict:
        """Extra (not in the reference): interpolate a whole clip sharded by frame-pair index over `world`
        ranks.  Each rank replays the HUD pushes it needs from the pairs before its range,
        runs interpolate() on its own pairs, and the route counters are summed with ONE all-reduce.
        frames: (n_pairs+1, H, W, 3) uint8.  on_pair(j, mids, analysis) receives each pair's result.
        Returns {'range', 'routes' (clip-order int8, -1 where unknown), 'stats' (whole clip)}."""
        from ... import sharding
        self.ensure_loaded()
        n_pairs = len(frames) - 1
        cuts = {}

        def is_cut(j):
            if j not in cuts:
                cuts[j] = self.router.detect_scene_change(frames[j], frames[j + 1])[0]
            return cuts[j]

        code = {'rife': _abi.ROUTE_RIFE, 'vfimamba': _abi.ROUTE_VFIMAMBA, 'scene_change': _abi.ROUTE_SCENE_CHANGE}
        local = []

        def process(j):
            mids, a = self._interpolate_routed(frames[j], frames[j + 1], num_frames)
            local.append(code[a.recommended_model])
            if on_pair is not None:
                on_pair(j, mids, a)

        (lo, hi), _ = sharding.run_sharded(n_pairs, rank, world, is_cut,
                                           lambda j: self.router._detect_hud_dev(self.router._dev(frames[j]), None),
                                           process)
        dev = self.device if self.device.type == 'cuda' else torch.device('cpu')
        st = self.stats
        vec = torch.tensor([st.total, st.rife_count, st.vfimamba_count, st.scene_change_count], dtype=torch.int64,
                           device=dev)
        vec = sharding.reduce_counters(vec, group).tolist()
        routes = sharding.gather_flags(torch.tensor(local, dtype=torch.int8, device=dev), n_pairs, rank, world, group)
        whole = RoutingStats(*vec)
        return {'range': (lo, hi), 'routes': routes.cpu().numpy(), 'stats': whole.to_dict()}

    def process_clip_batched(self, frames, num_frames: int = 3, rank: int = 0, world: int = 1, group=None,
                             batch: int = 8, out=None, on_batch=None, n_pairs: Optional[int] = None, base: int = 0,
                             collect: bool = True) -> dict:
        """Extra (not in the reference): process_clip at device speed.  Same per-pair semantics as interpolate()
        (:369-411: cut -> copies of frame0; else RIFE with the HUD composite when coverage > 0.01; the route
        recommendation is computed and counted either way), evaluated `batch` pairs at a time on a DEVICE-RESIDENT
        clip: one scene-cut SSIM launch for the shard, then per batch one Farneback + flow-statistics pass, one
        Laplacian pass, the (sequential) HUD pushes of the non-cut pairs each writing its own mask slice, ONE
        device->host read of the 8 x {mean, max, std, laplacian, coverage} scalars for the decision tree, and one
        IFNet call over the batch with the per-pair masks.  Sharding, HUD warm-up replay and the end-of-clip
        all-reduce / all-gather are those of process_clip.
        frames: (n_pairs+1, H, W, 3) uint8 CUDA tensor (a numpy / host tensor is uploaded whole).  A rank that holds only
        its part of the clip passes frames = clip[base : ...] with `base` and the clip's `n_pairs` (it needs the frames from
        16 pairs before its range, for the HUD warm-up replay across cuts, to one past its last pair).
        out: optional (hi-lo, T, H, W, 3) uint8 CUDA tensor for the shard's frames; on_batch(j0, frames_dev, analyses).
        collect=False skips the end-of-clip collectives (this rank's counters and flags only, padded with -1).
        Returns {'range', 'routes', 'cuts', 'hud_coverage' (clip order), 'stats', 'analyses' (this shard)}."""
        from ... import sharding
        self.ensure_loaded()
        if self.enable_vfimamba:
            raise RuntimeError("process_clip_batched runs the RIFE fast path only; use process_clip with VFIMamba loaded")
        r, e = self.router, self.router.engine
        dev = e.device
        if not (isinstance(frames, torch.Tensor) and frames.is_cuda):
            frames = e._u8(frames)
        if n_pairs is None:
            n_pairs = base + frames.shape[0] - 1
        H, W = int(frames.shape[1]), int(frames.shape[2])
        F = lambda a, b: frames[a - base:b - base]          # clip frames [a, b)
        lo, hi = sharding.shard_range(n_pairs, rank, world)
        ts = self._rife.get_default_timestamps(num_frames)
        T = len(ts)
        prec = _abi.PRECISION_BF16 if self._rife.precision == 'bf16' else _abi.PRECISION_FP32
        scale = 0.5 if self._rife.uhd_mode else 1.0
        code = {'rife': _abi.ROUTE_RIFE, 'vfimamba': _abi.ROUTE_VFIMAMBA, 'scene_change': _abi.ROUTE_SCENE_CHANGE}
        scores = {}

        def score_range(a, b):                      # scene-cut SSIM of pairs [a, b), one launch per <= 64 pairs
            for s0 in range(a, b, 64):
                s1 = min(b, s0 + 64)
                sc = e.ssim(F(s0, s1), F(s0 + 1, s1 + 1), downscale=4).tolist()
                scores.update(zip(range(s0, s1), sc))

        def is_cut(j):
            if j not in scores:
                score_range(max(base, j - 15), j + 1)
            return scores[j] < r.scene_change_threshold

        analyses, local_routes, local_cov = [], [], []
        if lo < hi:
            score_range(lo, hi)
            _, replay = sharding.warmup_start(lo, is_cut)
            for j in replay:
                r._detect_hud_dev(frames[j - base], None)
            hud = r._hud_state()
            masks = torch.zeros((batch, H, W), dtype=torch.uint8, device=dev)
            covs = torch.zeros((batch,), dtype=torch.float64, device=dev)
            for j0 in range(lo, hi, batch):
                nb = min(batch, hi - j0)
                f0, f1 = F(j0, j0 + nb), F(j0 + 1, j0 + nb + 1)
                cut = [is_cut(j) for j in range(j0, j0 + nb)]
                st = lv = None
                if not all(cut):
                    st = e.farneback(f0, f1, want_flow=False, want_stats=True, **FARNEBACK)
                    lv = e.laplacian_var(f0)
                ready = [False] * nb
                for k in range(nb):
                    if cut[k]:
                        continue                                    # cut pairs return before detect_hud (:244-254)
                    hud.update(f0[k], mask_out=masks[k], cov_out=covs[k:k + 1])
                    ready[k] = len(hud) >= 5
                if any(ready):                                      # router.hud_mask = the newest mask, as after analyze()
                    r._hud_mask_dev, r.hud_mask = masks[max(k for k in range(nb) if ready[k])].clone(), None
                if st is not None:
                    host = torch.cat([st.reshape(-1), lv, covs[:nb]]).tolist()      # the batch's one device->host read
                    st_h, lv_h, cov_h = host[:3 * nb], host[3 * nb:4 * nb], host[4 * nb:]
                batch_an = []
                for k in range(nb):
                    sc = scores[j0 + k]
                    if cut[k]:
                        a = ContentAnalysis(0, 0, 0, False, True, 0, 'scene_change', 1.0 - sc)
                        masks[k].zero_()
                    else:
                        mean, mx, sd = st_h[3 * k:3 * k + 3]
                        cov = cov_h[k] if ready[k] else 0.0
                        route, conf, hp, _ = _abi.route_decide(sc, mean, mx, sd, lv_h[k], r.scene_change_threshold,
                                                               r.particle_threshold, r.motion_threshold_low,
                                                               r.motion_threshold_high)
                        a = ContentAnalysis(motion_mean=mean, motion_max=mx, motion_std=sd, has_particles=hp,
                                            is_scene_change=False, hud_coverage=cov,
                                            recommended_model=_abi.ROUTE_NAMES[route], confidence=conf)
                        if not (ready[k] and cov > 0.01):
                            masks[k].zero_()                        # no composite (:398)
                    self.stats.add('scene_change' if cut[k] else 'rife')   # VFIMamba disabled: RIFE runs the non-cut pairs
                    batch_an.append(a)
                    local_routes.append(code[a.recommended_model])
                    local_cov.append(a.hud_coverage)
                dst = out[j0 - lo:j0 - lo + nb] if out is not None else None
                if not all(cut):
                    res = e.infer(f0, f1, ts, scale, prec, hud_mask=masks[:nb], out=dst)
                else:
                    res = dst if dst is not None else torch.empty((nb, T, H, W, 3), dtype=torch.uint8, device=dev)
                for k in range(nb):
                    if cut[k]:
                        res[k] = f0[k]                              # n copies of frame0 (:383-385), broadcast over T
                analyses += batch_an
                if on_batch is not None:
                    on_batch(j0, res, batch_an)
        # the recommendation counters (what RoutingStats would hold with VFIMamba enabled is the route histogram)
        vec = torch.tensor([len(local_routes)] + [sum(1 for c in local_routes if c == k) for k in
                                                  (_abi.ROUTE_RIFE, _abi.ROUTE_VFIMAMBA, _abi.ROUTE_SCENE_CHANGE)],
                           dtype=torch.int64, device=dev)
        if collect:
            vec = sharding.reduce_counters(vec, group).tolist()
            routes = sharding.gather_flags(torch.tensor(local_routes, dtype=torch.int8, device=dev), n_pairs, rank, world, group)
            cov = sharding.gather_flags(torch.tensor(local_cov, dtype=torch.float64, device=dev), n_pairs, rank, world, group)
        else:
            vec = vec.tolist()
            routes = torch.tensor(local_routes, dtype=torch.int8, device=dev)
            cov = torch.tensor(local_cov, dtype=torch.float64, device=dev)
        routes = routes.cpu().numpy()
        return {'range': (lo, hi), 'routes': routes, 'cuts': (routes == _abi.ROUTE_SCENE_CHANGE),
                'hud_coverage': cov.cpu().numpy(), 'analyses': analyses,
                'stats': {'total': vec[0], 'recommended_rife': vec[1], 'recommended_vfimamba': vec[2], 'scene_change': vec[3]}}

    def get_stats(self) -> dict:
        return self.stats.to_dict()

    def reset_stats(self) -> None:
        self.stats = RoutingStats()
