"""Stream sharding over GPUs: a stream's state never interacts with another's, so streams are dealt
round-robin to ranks, every rank holds a full weight replica, and there is no collective on the data path.  Only the
final per-stream score lists are gathered (outside any timed region)."""
from __future__ import annotations

from typing import Dict, List, Optional


def shard_streams(n_streams: int, world_size: int, rank: int) -> List[int]:
    """Stream s lives on rank s mod world_size."""
    if world_size <= 0 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank {rank} for world size {world_size}")
    if n_streams < 0:
        raise ValueError("n_streams must be non-negative")
    return list(range(rank, n_streams, world_size))


def gather_stream_results(local: Dict[int, list], dist=None, dst: int = 0) -> Optional[Dict[int, list]]:
    """Collects {stream id: per-frame results} from every rank onto `dst` (all ranks call it)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return dict(local)
    world = dist.get_world_size()
    bucket = [None] * world if dist.get_rank() == dst else None
    dist.gather_object(local, bucket, dst=dst)
    if dist.get_rank() != dst:
        return None
    merged: Dict[int, list] = {}
    for part in bucket:
        for k, v in part.items():
            if k in merged:
                raise RuntimeError(f"stream {k} reported by two ranks")
            merged[k] = v
    return merged
