"""Shared helpers for the tests (not collected)."""
from collections import OrderedDict

import numpy as np
import torch

from oracle import ppo_oracle as orc


def to_flat(pdict, offsets, sizes, p_pad, device="cpu"):
    flat = torch.zeros(p_pad, dtype=torch.float32)
    for i, k in enumerate(orc.PARAM_ORDER):
        flat[offsets[i]:offsets[i] + sizes[i]] = pdict[k].reshape(-1)
    return flat.to(device)


def from_flat(flat, offsets, sizes, shapes):
    flat = flat.detach().cpu()
    out = OrderedDict()
    for i, k in enumerate(orc.PARAM_ORDER):
        out[k] = flat[offsets[i]:offsets[i] + sizes[i]].reshape(shapes[k]).clone()
    return out


def unpadded(flat, offsets, sizes):
    flat = flat.detach().cpu()
    return torch.cat([flat[offsets[i]:offsets[i] + sizes[i]] for i in range(len(sizes))])


# Two criteria, both asserted by every comparison:
#   per tensor     max|got-ref| <= rel * max|ref|                                   (north_star's 1e-5 relative)
#   element-wise   |got_i-ref_i| <= ELEM_RTOL*|ref_i| + elem_atol * max|ref|        with elem_atol = rel/2 by default
# The survey proposed elem_atol = 1e-8.  That is below what fp32 can deliver for the results of long reductions
# (gradients, means, normalised rows): an entry that is 1e-3 of the tensor's largest carries the absolute rounding error
# of the whole reduction (~1e-7 * scale), i.e. 1e-4 relative to itself - for the reference's own fp32 run just as for
# ours: tests/test_tolerance_cpu.py shows the reference missing 1e-8 against its own fp64 evaluation and against
# itself with another thread count.  Measured on the B200 over the whole GPU suite (closeness_report.json, written by
# conftest): the largest elem_atol any tensor needs is 2.4e-6 (p_fc2.bias gradient), the largest per-tensor error
# 2.4e-6; the asserted 5e-6 / 1e-5 leave a factor 2 / 4.  Every comparison records what it needed (CLOSENESS).
ELEM_RTOL = 1e-5
ELEM_ATOL_SURVEY = 1e-8
ELEM_ATOL_REDUCTION = 5e-6
CLOSENESS = []


def closeness(got, ref):
    got = torch.as_tensor(got, dtype=torch.float64).reshape(-1)
    ref = torch.as_tensor(ref, dtype=torch.float64).reshape(-1)
    if ref.numel() == 0:
        return dict(scale=0.0, max_err=0.0, rel=0.0, elem_atol_needed=0.0, n=0, n_beyond_survey=0)
    scale = float(ref.abs().max())
    d = (got - ref).abs()
    excess = (d - ELEM_RTOL * ref.abs()).clamp_(min=0.0)
    s = max(scale, 1e-30)
    return dict(scale=scale, max_err=float(d.max()), rel=float(d.max()) / s, elem_atol_needed=float(excess.max()) / s,
                n=int(ref.numel()), n_beyond_survey=int((excess > ELEM_ATOL_SURVEY * s).sum()))


def assert_close_scaled(got, ref, rel=1e-5, name="", elem_atol="default"):
    """Per-tensor bound and element-wise bound (elem_atol = rel/2 unless given; None switches the latter off)."""
    if elem_atol == "default":
        elem_atol = 0.5 * rel
    got_t = torch.as_tensor(got, dtype=torch.float64).reshape(-1)
    ref_t = torch.as_tensor(ref, dtype=torch.float64).reshape(-1)
    assert got_t.shape == ref_t.shape, (name, got_t.shape, ref_t.shape)
    assert torch.isfinite(got_t).all(), name
    c = closeness(got_t, ref_t)
    CLOSENESS.append(dict(name=name, **c))
    assert c["max_err"] <= rel * max(c["scale"], 1e-30), \
        "%s: max err %.3e > %.1e * scale %.3e" % (name, c["max_err"], rel, c["scale"])
    if elem_atol is not None:
        assert c["elem_atol_needed"] <= elem_atol, \
            "%s: element-wise |d| <= %.0e|ref| + atol*scale needs atol %.3e > %.1e (scale %.3e, %d of %d beyond 1e-8)" % (
                name, ELEM_RTOL, c["elem_atol_needed"], elem_atol, c["scale"], c["n_beyond_survey"], c["n"])


def synth_batch(O, A, T, seed=0):
    g = torch.Generator().manual_seed(seed)
    obs = torch.randn(T, O, generator=g).clamp_(-5, 5)
    act = torch.randn(T, A, generator=g)
    adv = torch.randn(T, 1, generator=g)
    ret = torch.randn(T, 1, generator=g)
    return obs, act, adv, ret


def random_segments(T, rng, mean_len=300):
    seg_end = np.zeros(T, np.uint8)
    boot = np.zeros(T, np.float32)
    i = 0
    while i < T:
        L = int(rng.integers(1, 2 * mean_len))
        j = min(T, i + L) - 1
        seg_end[j] = 1
        boot[j] = 0.0 if rng.random() < 0.5 else np.float32(rng.normal())
        i = j + 1
    seg_end[T - 1] = 1
    return seg_end, boot
