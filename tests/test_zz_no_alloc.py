"""dppo_gae_scan_noalloc: the GAE scan for callers that cannot allocate scratch memory
(a host that owns no device allocator, or one that must not call cudaMalloc mid-run).
Called through the C ABI and compared with dppo_gae_scan and with the oracle on calls in a
row - every later call finds the earlier call's flags and carries in the library's
built-in array.

The array is only touched when the scan chains its tiles (decoupled look-back), which
dppo_gae_scan does when gamma*lam is too close to 1 for independent tiles (>= ~0.987):
the chained cases below use gamma*lam = 0.99 and 0.998.  With the reference's 0.99*0.95
the tiles are independent and the array is not used at all; that mode is covered too.

This file sorts last on purpose: with `pytest -x` every other test has run before it."""
import numpy as np
import pytest
import torch

from oracle import ppo_oracle as orc
from util import assert_close_scaled, random_segments

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _inputs(T, seed, mean_len=700):
    rng = np.random.default_rng(seed)
    r = np.clip(rng.normal(size=T), -1, 1).astype(np.float32)
    v = rng.normal(size=T).astype(np.float32)
    se, boot = random_segments(T, rng, mean_len=mean_len)
    return r, v, boot, se


def _both(nv, r, v, boot, se, gamma, lam):
    """(adv, ret) of dppo_gae_scan with caller scratch and of dppo_gae_scan_noalloc."""
    T = len(r)
    R, V, Bt, SE = (torch.from_numpy(a).to(DEV) for a in (r, v, boot, se))
    a0, q0 = torch.empty(T, device=DEV), torch.empty(T, device=DEV)
    a1, q1 = torch.full((T,), float("nan"), device=DEV), torch.full((T,), float("nan"), device=DEV)
    scratch = torch.empty(nv.gae_scratch_bytes(T), dtype=torch.uint8, device=DEV)
    nv.gae_scan(R, V, Bt, SE, a0, q0, gamma, lam, scratch)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    nv.gae_scan_noalloc(R, V, Bt, SE, a1, q1, gamma, lam)
    assert torch.cuda.memory_allocated() == before               # nothing allocated through torch either
    torch.cuda.synchronize()
    return a0, q0, a1, q1


@pytest.mark.parametrize("gamma,lam", [(0.99, 1.0), (0.999, 0.999)])
def test_noalloc_chained_scan_uses_the_builtin_array_on_consecutive_calls(gamma, lam):
    """Look-back mode: ticket, tile flags and carries live in the built-in array.  A long
    buffer, then a shorter one, then a long one again: the second and third calls find the
    flags of the previous call set to "resolved" and its carries in place - if the per-call
    clear did not happen, or the symbol's device address were wrong, the tiles would pick
    up stale carries (or hang) and the results would differ."""
    import b200_native as nv
    for T, seed, mean_len in ((300000, 1, 9000), (70001, 2, 20000), (200003, 3, 100000)):
        r, v, boot, se = _inputs(T, seed, mean_len)              # segments span several 4096-step tiles
        a0, q0, a1, q1 = _both(nv, r, v, boot, se, gamma, lam)
        assert torch.equal(a1, a0) and torch.equal(q1, q0), "T=%d" % T      # same kernel, same arithmetic
        a_ref, q_ref = orc.gae(r, v, boot, se, gamma, lam)
        assert_close_scaled(a1.cpu(), a_ref, 1e-5, "advantages vs oracle (T=%d)" % T)   # the suite's fp32 tolerance
        assert_close_scaled(q1.cpu(), q_ref, 1e-5, "returns vs oracle (T=%d)" % T)


def test_noalloc_chained_scan_at_the_size_limit():
    """T == DPPO_GAE_NOALLOC_MAX_T with chained tiles uses every slot of the array."""
    import b200_native as nv
    T = nv.GAE_NOALLOC_MAX_T
    r, v, boot, se = _inputs(T, 4, mean_len=50000)
    a0, q0, a1, q1 = _both(nv, r, v, boot, se, 0.99, 1.0)
    assert torch.equal(a1, a0) and torch.equal(q1, q0)
    # and a short call right after the full-size one
    r, v, boot, se = _inputs(5000, 5, mean_len=5000)
    a0, q0, a1, q1 = _both(nv, r, v, boot, se, 0.99, 1.0)
    assert torch.equal(a1, a0) and torch.equal(q1, q0)
    a_ref, q_ref = orc.gae(r, v, boot, se, 0.99, 1.0)
    assert_close_scaled(a1.cpu(), a_ref, 1e-5, "advantages vs oracle")
    assert_close_scaled(q1.cpu(), q_ref, 1e-5, "returns vs oracle")


def test_noalloc_independent_tiles_twice():
    """The reference's gamma*lam = 0.9405: tiles are independent, the array is not needed."""
    import b200_native as nv
    for T, seed in ((300000, 1), (70001, 2)):
        r, v, boot, se = _inputs(T, seed)
        a0, q0, a1, q1 = _both(nv, r, v, boot, se, 0.99, 0.95)
        assert_close_scaled(a1.cpu(), a0.cpu(), 1e-5, "advantages (T=%d)" % T)
        assert_close_scaled(q1.cpu(), q0.cpu(), 1e-5, "returns (T=%d)" % T)
        assert torch.equal(a1, a0) and torch.equal(q1, q0)


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
