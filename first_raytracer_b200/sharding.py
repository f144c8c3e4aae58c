"""How the path shards over ranks: by SAMPLE RANGE for rendering, by BLOCK for ray
casting; geometry is replicated; the one exchange step is the sum of the partial films.
Used by bench.py (NCCL, one process per GPU) and tested on CPU with gloo (tests/test_sharding.py).
The in-process multi-device context (frt_create(n > 1)) applies the same split in frt_render_tile."""
from __future__ import annotations


def spp_share(spp_begin: int, spp_end: int, rank: int, world: int) -> tuple[int, int]:
    """Rank's [begin, end) of the sample range; the shares tile the range and differ by <= 1.
    Same arithmetic as frt_render_tile's device split (csrc/frt_render.inl)."""
    if world < 1 or not 0 <= rank < world or spp_end < spp_begin:
        raise ValueError(f"bad share request: range [{spp_begin},{spp_end}) rank {rank} of {world}")
    n = spp_end - spp_begin
    return spp_begin + n * rank // world, spp_begin + n * (rank + 1) // world


def ray_share(n_rays: int, rank: int, world: int) -> tuple[int, int]:
    """Rank's block of a ray array (frt_trace_rays's device partition): no exchange needed."""
    if world < 1 or not 0 <= rank < world or n_rays < 0:
        raise ValueError(f"bad share request: {n_rays} rays rank {rank} of {world}")
    return n_rays * rank // world, n_rays * (rank + 1) // world


def reduce_film(film, dst: int = 0):
    """Sum the ranks' partial radiance films onto `dst` in place: the single collective of the path
    (ncclReduce over NVLink under NCCL; the film tensor aliases libfrt's device film)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return film
