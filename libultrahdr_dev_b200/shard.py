"""Image-batch sharding across GPUs: images are independent, so image i goes to rank
i mod G (or a contiguous block) with NO data-path collective. The only cross-rank traffic is the
control-plane reduction of counters/timings (max over ranks) done by the caller through
torch.distributed; nothing here touches CUDA.

This replaces the reference's only parallelism, the JobQueue of row blocks on <= 4 host threads
(lib/src/ultrahdr.cpp:131-183, 304-354, 500-513), at the granularity that makes sense on an
8 x B200 box: whole images per device, rows per thread block inside a device.
"""
from dataclasses import dataclass
from typing import List, Sequence


@dataclass(frozen=True)
class Shard:
    rank: int
    world: int
    indices: List[int]  # global image indices owned by this rank, ascending


def shard_round_robin(n_images: int, rank: int, world: int) -> Shard:
    """image i -> rank i mod world (balanced to within one image for any n)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    return Shard(rank, world, list(range(rank, n_images, world)))


def shard_contiguous(n_images: int, rank: int, world: int) -> Shard:
    """contiguous blocks, the first (n mod world) ranks get one extra image."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(n_images, world)
    start = rank * base + min(rank, extra)
    return Shard(rank, world, list(range(start, start + base + (1 if rank < extra else 0))))


def owner_round_robin(index: int, world: int) -> int:
    return index % world


def merge_checksums(per_rank: Sequence[Sequence[tuple]]) -> List[int]:
    """per_rank[r] = [(global_index, checksum), ...] -> checksums ordered by global index.
    Raises if an image is missing or owned twice (a sharding bug)."""
    flat = {}
    for part in per_rank:
        for idx, cs in part:
            if idx in flat:
                raise ValueError(f"image {idx} processed by two ranks")
            flat[idx] = cs
    n = len(flat)
    if sorted(flat) != list(range(n)):
        raise ValueError("images missing from the gathered result")
    return [flat[i] for i in range(n)]


def whole_job_throughput(images_per_rank: Sequence[int], pixels_per_image: int, seconds_per_rank: Sequence[float]) -> float:
    """pixels of ALL ranks / the slowest rank's time (pixels per second)."""
    return sum(images_per_rank) * pixels_per_image / max(seconds_per_rank)
