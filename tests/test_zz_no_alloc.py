"""dppo_gae_scan_noalloc: the GAE scan for callers that cannot allocate scratch memory
(a host that owns no device allocator, or one that must not call cudaMalloc mid-run).
Called through the C ABI and compared with dppo_gae_scan on two consecutive calls - the
second call reuses the library's built-in array with the first call's leftovers in it.

This file sorts last on purpose: with `pytest -x` every other test has run before it."""
import numpy as np
import pytest
import torch

from util import assert_close_scaled, random_segments

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _inputs(T, seed):
    rng = np.random.default_rng(seed)
    r = torch.from_numpy(np.clip(rng.normal(size=T), -1, 1).astype(np.float32)).to(DEV)
    v = torch.from_numpy(rng.normal(size=T).astype(np.float32)).to(DEV)
    se, boot = random_segments(T, rng, mean_len=700)
    return r, v, torch.from_numpy(boot).to(DEV), torch.from_numpy(se).to(DEV)


def test_noalloc_matches_scratch_variant_twice():
    import b200_native as nv
    # two calls in a row, the second on a different, shorter buffer: stale flags and
    # carries of the first call are still in the shared array and must not leak through
    for T, seed in ((300000, 1), (70001, 2)):
        r, v, boot, se = _inputs(T, seed)
        a0, q0 = torch.empty(T, device=DEV), torch.empty(T, device=DEV)
        a1, q1 = torch.empty(T, device=DEV), torch.empty(T, device=DEV)
        scratch = torch.empty(nv.gae_scratch_bytes(T), dtype=torch.uint8, device=DEV)
        nv.gae_scan(r, v, boot, se, a0, q0, 0.99, 0.95, scratch)
        before = torch.cuda.memory_allocated()
        nv.gae_scan_noalloc(r, v, boot, se, a1, q1, 0.99, 0.95)
        assert torch.cuda.memory_allocated() == before           # nothing allocated through torch either
        torch.cuda.synchronize()
        assert_close_scaled(a1.cpu(), a0.cpu(), 1e-5, "advantages (T=%d)" % T)   # the suite's fp32 tolerance
        assert_close_scaled(q1.cpu(), q0.cpu(), 1e-5, "returns (T=%d)" % T)
        assert torch.equal(a1, a0) and torch.equal(q1, q0)       # same kernel, same arithmetic


def test_noalloc_rejects_oversize():
    import b200_native as nv
    T = nv.GAE_NOALLOC_MAX_T + 1
    x = torch.zeros(T, device=DEV)
    se = torch.zeros(T, dtype=torch.uint8, device=DEV)
    with pytest.raises(nv.DppoError, match="built-in scratch"):
        nv.gae_scan_noalloc(x, x, x, se, torch.empty_like(x), torch.empty_like(x), 0.99, 0.95)
    # the limit itself is accepted
    T = nv.GAE_NOALLOC_MAX_T
    out = torch.empty(T, device=DEV)
    nv.gae_scan_noalloc(x[:T], x[:T], x[:T], se[:T], out, torch.empty_like(out), 0.99, 0.95)
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0
