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


def assert_close_scaled(got, ref, rel=1e-5, name=""):
    """max |got-ref| <= rel * max |ref|  (per-tensor tolerance)."""
    got = torch.as_tensor(got, dtype=torch.float64).reshape(-1)
    ref = torch.as_tensor(ref, dtype=torch.float64).reshape(-1)
    assert got.shape == ref.shape, (name, got.shape, ref.shape)
    assert torch.isfinite(got).all(), name
    scale = float(ref.abs().max()) if ref.numel() else 0.0
    err = float((got - ref).abs().max()) if ref.numel() else 0.0
    assert err <= rel * max(scale, 1e-30), "%s: max err %.3e > %.1e * scale %.3e" % (name, err, rel, scale)


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
