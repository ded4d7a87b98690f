"""Multi-GPU plumbing of the BDPT render: shard by SAMPLE INDEX, one reduce of the frame buffers.

The reference has no multi-device code; samples are independent and the frame buffer is additive
(contributions are normalised by the TOTAL spp: bdpt.h:330-332, renderer.cpp:201), so rank g of G renders sample
indices [begin_g, end_g) of every pixel with the scene replicated, RNG streams keyed by sample index (hence the
result does not depend on G), and ONE collective ends the render: reduce(sum) of the W*H*3 float frame buffer to
rank 0 (NCCL over NVLink on GPUs; the same code runs on gloo for the CPU tests).  No exchange during rendering.
"""
from __future__ import annotations

from typing import Callable

import torch
import torch.distributed as dist


def sample_range(spp_total: int, rank: int, world: int) -> tuple[int, int]:
    """[begin, end) of rank's sample indices: contiguous, disjoint, covering [0, spp_total), sizes differ by <= 1."""
    if spp_total < 0 or world < 1 or not (0 <= rank < world):
        raise ValueError("bad partition arguments")
    return spp_total * rank // world, spp_total * (rank + 1) // world


def reduce_framebuffer(fb: torch.Tensor, dst: int = 0) -> torch.Tensor:
    """Sum of the partial frame buffers on rank `dst` (in place); a no-op for a single process."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(fb, dst=dst, op=dist.ReduceOp.SUM)
    return fb


def render_sharded(render_range: Callable[[int, int, torch.Tensor], None], fb: torch.Tensor, spp_total: int,
                   rank: int | None = None, world: int | None = None, dst: int = 0) -> torch.Tensor:
    """render_range(begin, end, fb) must write this rank's partial frame buffer (normalised by spp_total) into fb;
    afterwards rank `dst` holds the full image.  With the product: render_range = lambda b, e, fb:
    ctx.render_bdpt_device(render_desc(spp=spp_total, spp_begin=b, spp_end=e, ...), fb.data_ptr())."""
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
    b, e = sample_range(spp_total, rank, world)
    if e > b:
        render_range(b, e, fb)
    else:
        fb.zero_()
    return reduce_framebuffer(fb, dst)
