"""Frame-pair sharding of a clip across GPUs (one process per GPU).

RIFE interpolation, PSNR and SSIM are pure functions of one pair, so ranks take contiguous pair ranges and
never talk inside the loop.  The adaptive router carries one piece of sequential state: the HUD mask is the
temporal variance of the last five gray thumbnails pushed by *non-cut* pairs (reference
models/novel/adaptive_pipeline.py:205-212; cut pairs return before detect_hud, :244-254).  A rank therefore
replays only the HUD pushes of the pairs just before its range until it holds the same four previous pushes
the serial reference would, scanning back across cuts.  The single collective is an all-reduce (NCCL on the
GPUs, gloo in the CPU tests) of the route counters / metric sums at the end of the clip, plus an optional
all-gather of the per-pair route flags for the bit-exact comparison.
"""
from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

HUD_WINDOW = 5  # np.var over the last five thumbnails (adaptive_pipeline.py:212)


def shard_range(n_pairs: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous ranges: rank r owns [r*ceil(n/world), min(n, (r+1)*ceil(n/world)))."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    per = -(-n_pairs // world)
    lo = min(n_pairs, rank * per)
    return lo, min(n_pairs, lo + per)


def warmup_start(lo: int, is_cut: Callable[[int], bool], pushes_per_pair: int = 1) -> Tuple[int, List[int]]:
    """First pair a rank starting at `lo` must replay, and the non-cut pairs in [start, lo) to push.

    The serial run's window at pair `lo` holds the HUD_WINDOW-1 most recent pushes before it; every non-cut
    pair pushes its frame0 `pushes_per_pair` times (2 under process_pair, which analyses twice:
    adaptive_pipeline.py:443 + :382)."""
    need = HUD_WINDOW - 1
    j, have, replay = lo, 0, []
    while j > 0 and have < need:
        j -= 1
        if not is_cut(j):
            replay.append(j)
            have += pushes_per_pair
    replay.reverse()
    return (replay[0] if replay else lo), replay


def run_sharded(n_pairs: int, rank: int, world: int, is_cut: Callable[[int], bool],
                push_hud: Callable[[int], None], process: Callable[[int], object],
                pushes_per_pair: int = 1) -> Tuple[Tuple[int, int], list]:
    """Replay the HUD warm-up for this rank's range, then process its pairs in order."""
    lo, hi = shard_range(n_pairs, rank, world)
    if lo < hi:
        _, replay = warmup_start(lo, is_cut, pushes_per_pair)
        for j in replay:
            for _ in range(pushes_per_pair):
                push_hud(j)
    return (lo, hi), [process(j) for j in range(lo, hi)]


def _active(group=None) -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


def reduce_counters(vec: torch.Tensor, group=None) -> torch.Tensor:
    """Sum a small int64/float64 vector over ranks (the only data-path collective of the clip)."""
    if _active(group):
        dist.all_reduce(vec, op=dist.ReduceOp.SUM, group=group)
    return vec


def gather_flags(local: torch.Tensor, n_pairs: int, rank: int, world: int, group=None) -> torch.Tensor:
    """All-gather per-pair int8 flags (route / cut) into clip order."""
    per = -(-n_pairs // world)
    buf = torch.full((per,), -1, dtype=local.dtype, device=local.device)
    buf[: local.numel()] = local
    if not _active(group):
        return buf[:n_pairs]
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    return torch.cat(out)[: n_pairs] if per * world == n_pairs else torch.cat(
        [o[: max(0, min(per, n_pairs - r * per))] for r, o in enumerate(out)])
