"""trb_trace_any_device_noalloc (include/trb.h): the occlusion query with the traversal's scratch in a fixed device array inside the
library - an entry point that calls no cudaMalloc (no counterpart in the reference).
Checked through the C ABI against the existing entry point trb_trace_any_device and the golden flags of the unmodified reference:
occlusion flags are exact (the tolerance every level-1 test here uses), on calls in a row that share the scratch array.
This file is named to be collected LAST."""
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = Path(__file__).parent / "golden"


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()


def test_trace_any_noalloc_matches_existing_entry_point(trb, gpu_ctx):
    import torch
    ctx = gpu_ctx("caustic")
    g = np.load(GOLD / "trace_caustic.npz")
    rays = np.zeros(len(g["rays"]), trb.RAY_DTYPE)
    rays["o"], rays["d"], rays["min_t"], rays["max_t"] = g["rays"][:, 0:3], g["rays"][:, 3:6], g["rays"][:, 6], g["rays"][:, 7]
    n = len(rays)
    degenerate = (g["rays"][:, 3:6] == 0).any(1)
    assert degenerate.sum() > 500          # the set holds rays that go through the scratch queue (handed to the binary tree)
    rng = np.random.default_rng(3)
    perm = rng.permutation(n)
    sets = [rays, rays[perm[: n // 3]], rays[perm[::-1]]]      # a big call, a smaller one after it (stale queue entries beyond its length), a big one again
    gold = [g["occluded"], g["occluded"][perm[: n // 3]], g["occluded"][perm[::-1]]]
    d_rays = [_dev(torch, s) for s in sets]
    d_ref = [torch.zeros(len(s), dtype=torch.uint8, device="cuda") for s in sets]
    d_out = [torch.full((len(s),), 7, dtype=torch.uint8, device="cuda") for s in sets]
    for r, o, s in zip(d_rays, d_ref, sets):
        ctx.trace_any_device(r.data_ptr(), len(s), o.data_ptr())               # the existing entry point (scratch from cudaMalloc)
    ctx.trace_any_device_noalloc(d_rays[0].data_ptr(), 1, d_out[0].data_ptr())  # set-up: the scratch address is asked for once per device
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for r, o, s in zip(d_rays, d_out, sets):                                    # calls in a row on the shared scratch
        ctx.trace_any_device_noalloc(r.data_ptr(), len(s), o.data_ptr())
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    for k in range(len(sets)):
        got, ref = d_out[k].cpu().numpy(), d_ref[k].cpu().numpy()
        assert np.array_equal(got, ref), f"call {k}: {int((got != ref).sum())} of {len(ref)} flags differ from trb_trace_any_device"
        assert np.array_equal(got, gold[k]), f"call {k}: flags differ from the reference's"
    assert free_after >= free_before, "device memory was allocated inside trb_trace_any_device_noalloc"
    # the stated maximum, exactly: TRB_NOALLOC_MAX_RAYS rays in one call
    m = trb.TRB_NOALLOC_MAX_RAYS
    big = np.tile(rays, m // n + 1)[:m]
    d_big = _dev(torch, big)
    a = torch.zeros(m, dtype=torch.uint8, device="cuda"); b = torch.full((m,), 7, dtype=torch.uint8, device="cuda")
    ctx.trace_any_device(d_big.data_ptr(), m, a.data_ptr())
    ctx.trace_any_device_noalloc(d_big.data_ptr(), m, b.data_ptr())
    assert torch.equal(a, b)
    # beyond it: refused with a clear error, nothing launched, the output untouched
    b.fill_(9)
    with pytest.raises(trb.TrbError) as e:
        ctx.trace_any_device_noalloc(d_big.data_ptr(), m + 1, b.data_ptr())
    assert e.value.code == trb.TRB_ERR_INVALID and "TRB_NOALLOC_MAX_RAYS" in str(e.value)
    assert bool((b == 9).all())
    # empty input is a no-op; and the shared scratch still serves an ordinary call afterwards
    ctx.trace_any_device_noalloc(0, 0, 0)
    ctx.trace_any_device_noalloc(d_rays[1].data_ptr(), len(sets[1]), d_out[1].data_ptr())
    assert np.array_equal(d_out[1].cpu().numpy(), gold[1])
