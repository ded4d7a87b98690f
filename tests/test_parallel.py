"""Host logic of the multi-GPU path on CPU: world_size-2 gloo processes (the N>1 path of bench.py uses the same
functions with NCCL).  The stand-in renderer is additive over sample indices like the real one, so the sharded result
must equal the single-process result."""
import importlib
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _parallel():
    sys.path.insert(0, str(ROOT))
    return importlib.import_module("bidirectional-path-tracing_b200.parallel")


def _fake_render(spp_total, h=6, w=8):
    def render_range(b, e, fb):
        fb.zero_()
        for s in range(b, e):  # a per-sample image that depends only on the sample index, normalised by the total spp
            g = torch.Generator().manual_seed(1000 + s)
            fb += torch.rand(h, w, 3, generator=g) / spp_total
    return render_range


def _worker(rank, world, port, spp, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    par = _parallel()
    fb = torch.zeros(6, 8, 3)
    par.render_sharded(_fake_render(spp), fb, spp)
    if rank == 0:
        np.save(os.path.join(out_dir, "fb.npy"), fb.numpy())
    dist.destroy_process_group()


def test_sample_range_partition():
    par = _parallel()
    for spp in (0, 1, 7, 16, 4096):
        for world in (1, 2, 3, 8):
            r = [par.sample_range(spp, g, world) for g in range(world)]
            assert r[0][0] == 0 and r[-1][1] == spp
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        par.sample_range(4, 2, 2)


@pytest.mark.parametrize("spp", [5, 16])
def test_two_rank_gloo_reduce_equals_single(tmp_path, spp):
    par = _parallel()
    single = torch.zeros(6, 8, 3)
    par.render_sharded(_fake_render(spp), single, spp, rank=0, world=1)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_worker, args=(2, port, spp, str(tmp_path)), nprocs=2, join=True)
    got = np.load(tmp_path / "fb.npy")
    assert np.allclose(got, single.numpy(), rtol=1e-6, atol=1e-7)
