"""Parity of each CUDA kernel, called through the C ABI, against the oracle and
the golden fixtures.  Needs a B200 (run with -m gpu)."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import ppo_oracle as orc
from util import assert_close_scaled, from_flat, random_segments, synth_batch, to_flat, unpadded

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def nv():
    import b200_native
    return b200_native


def _u8(n):
    return torch.empty(max(int(n), 16), dtype=torch.uint8, device=DEV)


def run_gae(nv, r, v, boot, se, gamma=0.99, lam=0.95, offset=0):
    T = len(r)
    def dev(a, dt):
        t = torch.zeros(T + offset, dtype=dt, device=DEV)
        t[offset:] = torch.from_numpy(np.asarray(a)).to(DEV)
        return t[offset:]
    R, V, Bt, SE = dev(r, torch.float32), dev(v, torch.float32), dev(boot, torch.float32), dev(se, torch.uint8)
    adv = torch.empty(T + offset, device=DEV)[offset:]
    ret = torch.empty(T + offset, device=DEV)[offset:]
    nv.gae_scan(R, V, Bt, SE, adv, ret, gamma, lam, _u8(nv.gae_scratch_bytes(T)))
    torch.cuda.synchronize()
    return adv.cpu().numpy(), ret.cpu().numpy()


@pytest.mark.parametrize("T", [1, 2, 7, 8, 9, 255, 2048, 2049, 5000, 65536 + 3])
@pytest.mark.parametrize("offset", [0, 1])
def test_gae_vs_oracle(nv, T, offset):
    rng = np.random.default_rng(T)
    r = np.clip(rng.normal(size=T), -1, 1).astype(np.float32)
    v = rng.normal(size=T).astype(np.float32)
    se, boot = random_segments(T, rng, mean_len=max(2, min(300, T // 3 + 1)))
    a_ref, r_ref = orc.gae(r, v, boot, se)
    a, rr = run_gae(nv, r, v, boot, se, offset=offset)
    assert_close_scaled(a, a_ref, 1e-5, "adv")      # fp32 tolerance of north_star: 1e-5 relative
    assert_close_scaled(rr, r_ref, 1e-5, "ret")


def test_gae_single_long_segment(nv):
    T = 40000                                   # no segment end except the last step: carries cross 20 tiles
    rng = np.random.default_rng(5)
    r = np.clip(rng.normal(size=T), -1, 1).astype(np.float32)
    v = rng.normal(size=T).astype(np.float32)
    se = np.zeros(T, np.uint8); se[-1] = 1
    boot = np.zeros(T, np.float32); boot[-1] = 0.7
    for lam in (0.95, 1.0):                     # gamma*lam = 0.99: slow decay, long look-back
        a_ref, r_ref = orc.gae(r, v, boot, se, 0.99, lam)
        a, rr = run_gae(nv, r, v, boot, se, 0.99, lam)
        assert_close_scaled(a, a_ref, 1e-5, "adv")
        assert_close_scaled(rr, r_ref, 1e-5, "ret")


def test_gae_every_step_a_segment(nv):
    T = 4097
    rng = np.random.default_rng(6)
    r = rng.normal(size=T).astype(np.float32); v = rng.normal(size=T).astype(np.float32)
    se = np.ones(T, np.uint8); boot = rng.normal(size=T).astype(np.float32)
    a_ref, r_ref = orc.gae(r, v, boot, se)
    a, rr = run_gae(nv, r, v, boot, se)
    assert np.array_equal(a, a_ref) and np.array_equal(rr, r_ref)      # no carries at all: bit-exact


@pytest.mark.parametrize("name,prefixes", [("dppo_w2_o17a6", ["w0_", "w1_"]), ("ppo_o4a2", [""])])
def test_gae_golden(nv, golden_dir, name, prefixes):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    for pre in prefixes:
        a, rr = run_gae(nv, g[pre + "gae_rewards"], g[pre + "gae_values"], g[pre + "gae_boot"], g[pre + "gae_seg_end"],
                        float(g["gamma"]), float(g["gae_param"]))
        assert_close_scaled(a, g[pre + "gae_advantages"], 1e-5, "adv")
        assert_close_scaled(rr, g[pre + "gae_returns"], 1e-5, "ret")


def test_gae_property_large(nv):
    """BASELINE size sweep (T = 2^22): returns - advantages == values exactly, and the
    recurrence A_i = td_i + gl*A_{i+1} holds step by step in fp32."""
    T = 1 << 22
    rng = np.random.default_rng(9)
    r = np.clip(rng.normal(size=T), -1, 1).astype(np.float32)
    v = rng.normal(size=T).astype(np.float32)
    se, boot = random_segments(T, rng, mean_len=1000)
    a, rr = run_gae(nv, r, v, boot, se)
    assert np.array_equal((a + v).astype(np.float32), rr)
    g, gl = np.float32(0.99), np.float32(0.99 * 0.95)
    vn = np.where(se.astype(bool), boot, np.append(v[1:], np.float32(0)))
    td = ((r + g * vn).astype(np.float32) - v).astype(np.float32)
    an = np.where(se.astype(bool), np.float32(0), np.append(a[1:], np.float32(0)))
    rec = (td + (gl * an).astype(np.float32)).astype(np.float32)
    assert_close_scaled(a, rec, 1e-6, "recurrence")


# ---------------------------------------------------------------- K4 Welford
def run_welford(nv, obs, st=None):
    T, O = obs.shape
    X = torch.as_tensor(obs, dtype=torch.float32).to(DEV).contiguous()
    out = torch.empty_like(X)
    if st is None:
        st = [torch.zeros(O, device=DEV) for _ in range(4)]
    nv.welford(X, out, st[0], st[1], st[2], st[3], _u8(nv.welford_scratch_bytes(T, O)))
    torch.cuda.synchronize()
    return out.cpu(), st


@pytest.mark.parametrize("T,O", [(1, 3), (5, 17), (481, 17), (482, 17), (3000, 17), (2048, 4), (700, 376), (1000, 1), (300, 300)])
def test_welford_vs_oracle(nv, T, O):
    g = torch.Generator().manual_seed(T * 1000 + O)
    loc = torch.rand(O, generator=g) * 4 - 2
    scale = torch.rand(O, generator=g) * 3 + 0.05
    obs = loc + scale * torch.randn(T, O, generator=g)
    st = orc.welford_new(O)
    ref = orc.welford_observe_normalize(st, obs)
    out, dst = run_welford(nv, obs)
    assert_close_scaled(out, ref, 1e-5, "normalised")           # outputs live in [-5,5]
    assert torch.equal(dst[0].cpu(), st["n"])
    assert_close_scaled(dst[1].cpu(), st["mean"], 1e-5, "mean")
    assert_close_scaled(dst[2].cpu(), st["mean_diff"], 1e-5, "mean_diff")
    assert_close_scaled(dst[3].cpu(), st["var"], 1e-5, "var")


def test_welford_continues_state(nv):
    O = 17
    g = torch.Generator().manual_seed(3)
    a = 1.5 + 0.3 * torch.randn(900, O, generator=g)
    b = 1.4 + 0.4 * torch.randn(1300, O, generator=g)
    st = orc.welford_new(O)
    orc.welford_observe_normalize(st, a)
    ref_b = orc.welford_observe_normalize(st, b)
    _, dst = run_welford(nv, a)
    out_b, dst = run_welford(nv, b, dst)
    assert_close_scaled(out_b, ref_b, 1e-5, "second rollout")
    assert_close_scaled(dst[2].cpu(), st["mean_diff"], 1e-5, "mean_diff")


def test_welford_golden(nv, golden_dir):
    for name, pre in (("ppo_o4a2", ""), ("dppo_w1_o3a2", "w0_")):
        g = np.load(os.path.join(golden_dir, name + ".npz"))
        raw = g[pre + "raw_obs"]
        out, dst = run_welford(nv, raw)
        Tn = g[pre + "states"].shape[0]
        assert_close_scaled(out[-Tn:], g[pre + "states"], 1e-5, "states")
        for i, k in enumerate(("n", "mean", "mean_diff", "var")):
            assert_close_scaled(dst[i].cpu(), g[pre + "stats_" + k], 1e-5, k)


def test_welford_in_place_and_large(nv):
    T, O = 1 << 18, 17
    g = torch.Generator().manual_seed(1)
    obs = (0.5 + 2.0 * torch.randn(T, O, generator=g))
    X = obs.to(DEV)
    st = [torch.zeros(O, device=DEV) for _ in range(4)]
    nv.welford(X, X, st[0], st[1], st[2], st[3], _u8(nv.welford_scratch_bytes(T, O)))     # out aliases obs
    torch.cuda.synchronize()
    # size-independent property: fp64 prefix statistics
    x64 = obs.double()
    n = torch.arange(1, T + 1, dtype=torch.float64)[:, None]
    mean = x64.cumsum(0) / n
    m2 = (x64 ** 2).cumsum(0) - n * mean ** 2
    ref = ((x64 - mean) / (m2 / n).clamp_min(1e-2).sqrt()).clamp(-5, 5)
    assert_close_scaled(X.cpu(), ref, 2e-5, "prefix-normalised")
    assert float(st[0][0]) == T


@pytest.mark.parametrize("T,O,inplace", [((1 << 19) + 13, 17, False), (1 << 19, 17, True), ((1 << 21) + 460, 4, False),
                                         (300001, 31, True)])
def test_welford_long_rollouts(nv, T, O, inplace):
    """Rollouts of thousands of tiles (three and four levels of the look-back hierarchy, tiles moved by bulk async
    copies, the ragged last tile by the threads): checked against fp64 prefix statistics, against a second run bit
    for bit, and - on the first rows - against the sequential oracle."""
    g = torch.Generator().manual_seed(T + O)
    obs = (0.5 + 2.0 * torch.randn(T, O, generator=g))
    outs = []
    for rep in range(2):
        X = obs.to(DEV)
        Y = X if inplace else torch.empty_like(X)
        st = [torch.zeros(O, device=DEV) for _ in range(4)]
        nv.welford(X, Y, st[0], st[1], st[2], st[3], _u8(nv.welford_scratch_bytes(T, O)))
        torch.cuda.synchronize()
        outs.append((Y.clone(), torch.stack(st)))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    x64 = obs.double()
    n = torch.arange(1, T + 1, dtype=torch.float64)[:, None]
    mean = x64.cumsum(0) / n
    m2 = (x64 ** 2).cumsum(0) - n * mean ** 2
    ref = ((x64 - mean) / (m2 / n).clamp_min(1e-2).sqrt()).clamp(-5, 5)
    assert_close_scaled(outs[0][0].cpu(), ref, 2e-5, "prefix-normalised")
    assert float(outs[0][1][0][0]) == T
    assert_close_scaled(outs[0][1][1].cpu(), mean[-1], 1e-5, "final mean")
    ost = orc.welford_new(O)
    want = orc.welford_observe_normalize(ost, obs[:3000])
    assert_close_scaled(outs[0][0][:3000].cpu(), want, 1e-5, "first rows vs the sequential oracle")


# ---------------------------------------------------------------- K3 reduce + clip + Adam
@pytest.mark.parametrize("nparts,clipnorm", [(1, 0.0), (5, 0.0), (147, 0.0), (3, 0.5), (3, 1e9)])
def test_reduce_adam_vs_oracle(nv, nparts, clipnorm):
    P = 71720
    g = torch.Generator().manual_seed(nparts)
    theta = torch.randn(P, generator=g)
    m = torch.zeros(P); v = torch.zeros(P)
    th_d, m_d, v_d = theta.to(DEV), m.to(DEV), v.to(DEV)
    gsh = torch.ones(P, device=DEV)                                   # model.py:51
    gsh_ref = torch.ones(P)
    prev = torch.empty(P, device=DEV); gout = torch.empty(P, device=DEV)
    scratch = _u8(nv.reduce_adam_scratch_bytes(P))
    for step in range(1, 5):
        parts = torch.randn(nparts, P, generator=g) * (0.01 * step)
        G = gsh_ref.clone()
        for p in range(nparts):
            G += parts[p]
        if clipnorm > 0:
            cl, _ = orc.clip_grad_norm({"x": G}, clipnorm)
            G = cl["x"]
        before = theta.clone()
        orc.adam_step(theta, G, m, v, step, 3e-4)
        gsh_ref.zero_()
        nv.reduce_adam(parts.to(DEV), nparts, gsh, th_d, prev, m_d, v_d, step, 3e-4, max_grad_norm=clipnorm,
                       grad_out=gout, scratch=scratch)
        torch.cuda.synchronize()
        assert torch.equal(prev.cpu(), before)
        assert float(gsh.abs().max()) == 0.0                           # chief.py:19
        # the partials are summed in a fixed but different order than the oracle's loop (8 interleaved
        # groups, then the groups): an fp32 re-association, bounded by the suite's 1e-5
        tol = 1e-6 if nparts <= 2 else 1e-5
        assert_close_scaled(gout.cpu(), G, tol, "G step %d" % step)
        assert_close_scaled(th_d.cpu(), theta, tol, "theta step %d" % step)
        assert_close_scaled(m_d.cpu(), m, tol, "m")
        assert_close_scaled(v_d.cpu(), v, tol, "v")
        # keep the two trajectories from drifting apart through the (non-IEEE) torch sqrt
        theta.copy_(th_d.cpu()); m.copy_(m_d.cpu()); v.copy_(v_d.cpu())


def test_adam_drift_100_steps(nv):
    P = 4096
    g = torch.Generator().manual_seed(0)
    theta = torch.randn(P, generator=g); m = torch.zeros(P); v = torch.zeros(P)
    th_d, m_d, v_d = theta.to(DEV), m.to(DEV), v.to(DEV)
    for step in range(1, 101):
        G = torch.randn(P, generator=g) * 0.05
        orc.adam_step(theta, G, m, v, step, 7e-4)
        nv.reduce_adam(G.to(DEV).view(1, P), 1, None, th_d, None, m_d, v_d, step, 7e-4)
    torch.cuda.synchronize()
    assert_close_scaled(th_d.cpu(), theta, 1e-5, "theta after 100 updates")


# ---------------------------------------------------------------- K2 fused update
def _setup_update(nv, O, A, widths, T, B, seed, flavour):
    h1, h2, g1, g2 = widths
    dims = nv.Dims(O, A, h1, h2, g1, g2)
    off, size, p_pad = nv.param_layout(dims)
    torch.manual_seed(seed)
    p = orc.init_params(O, A, h1, h2, g1, g2)
    p["log_std"] = torch.randn(1, A) * 0.3
    for k in p:                                         # biases are zero at init: make them matter
        if k.endswith("bias"):
            p[k] = torch.randn_like(p[k]) * 0.1
    p_old = {k: t + 0.02 * torch.randn_like(t) for k, t in p.items()}
    obs, act, adv, ret = synth_batch(O, A, T, seed)
    with torch.no_grad():
        mu, sig, v = orc.forward(p_old, obs)
        if flavour == nv.FLAVOUR_DPPO:
            act = mu + sig.sqrt() * torch.randn(T, A)                   # train.py:89
        else:
            act = mu + sig * torch.randn(T, A)                          # ppo.py:93
        ret = v + 0.5 * ret
        logp = (-0.5 * ((act - mu) / sig).pow(2) - 0.5 * np.log(2 * np.pi) - p_old["log_std"]).sum(-1, keepdim=True)
    rng = random.Random(seed)
    idx = orc.sample_indices(T, B, rng)
    return dims, off, size, p_pad, p, p_old, obs, act, adv * 2.0, ret, logp, idx


def _run_update(nv, flavour, dims, off, size, p_pad, p, p_old, obs, act, adv, ret, logp, idx, clip=0.2, ent=0.0):
    B = len(idx)
    nparts = nv.update_nparts(dims, B)
    partials = torch.full((nparts, p_pad), float("nan"), device=DEV)
    lparts = torch.full((nparts, 4), float("nan"), device=DEV)
    th = to_flat(p, off, size, p_pad, DEV)
    tho = to_flat(p_old, off, size, p_pad, DEV)
    nv.update_fused(flavour, dims, obs.to(DEV).contiguous(), act.to(DEV).contiguous(),
                    adv.reshape(-1).to(DEV).contiguous(), ret.reshape(-1).to(DEV).contiguous(),
                    logp.reshape(-1).to(DEV).contiguous() if flavour == nv.FLAVOUR_PPO else None,
                    torch.tensor(idx, dtype=torch.int32, device=DEV), th,
                    tho if flavour == nv.FLAVOUR_DPPO else None, clip, ent, partials, lparts)
    torch.cuda.synchronize()
    return partials.double().sum(0).float().cpu(), lparts.double().sum(0).cpu()


CASES = [
    # O, A, (h1,h2,g1,g2), T, B
    (4, 1, (100, 100, 500, 100), 512, 64),             # C1 shape, reference arch
    (17, 6, (100, 100, 500, 100), 2048, 2048),         # C2: B == T (a permutation)
    (15, 3, (100, 100, 500, 100), 1500, 1000),         # C3 shape, main.py batch 1000
    (17, 6, (64, 64, 64, 64), 4096, 4096),             # 64x64 nets
    (3, 2, (8, 12, 20, 4), 100, 37),                   # ragged: B not a multiple of the tile
    (17, 6, (100, 100, 500, 100), 8192, 6000),         # more tiles than SMs: CTAs loop and accumulate
    (33, 5, (256, 256, 256, 256), 600, 300),           # wide
    (376, 17, (256, 256, 256, 256), 700, 512),         # C5: Humanoid shape, 256x256 nets
    (15, 3, (100, 100, 500, 100), 2048, 2048),         # C3: HopperBullet shape
    (17, 6, (128, 128, 128, 128), 1024, 1024),         # 128-wide: every hidden layer on the tcgen05 path
    (11, 3, (96, 96, 160, 96), 700, 450),              # odd numbers of 32-wide k-blocks (3 and 5) in the tensor-core pipeline
    # the exact shapes bench.py times (BASELINE configs): tile size and CTA count as in the benchmark
    (17, 6, (100, 100, 500, 100), 65536, 4096),        # C4: TM = 28, 147 CTAs
    (376, 17, (256, 256, 256, 256), 32768, 4096),      # C5 at its full minibatch
    (4, 1, (64, 64, 64, 64), 2048, 64),                # C1 with BASELINE's 64x64 nets
    (17, 6, (64, 64, 64, 64), 2048, 2048),             # C2 with 64x64 nets
    # tensor-core layers with several 128-row tiles, first layers on the tensor cores (observation width >= 64), trunk-serial tiles
    (64, 3, (128, 128, 192, 128), 700, 300),           # first layers eligible (K = 64), 192 outputs = two tiles, trunks side by side
    (40, 4, (300, 200, 260, 132), 900, 400),           # odd widths over several tiles: 300 (3 tiles), 200, 260, 132
    (376, 17, (256, 256, 256, 256), 4096, 4096 - 17),  # Humanoid, more tiles than SMs is false here but the last tile is ragged
    (128, 8, (512, 256, 512, 256), 800, 333),          # 512-wide first layers: four tiles, trunk-serial
]


@pytest.mark.parametrize("case", CASES)
def test_update_dppo_vs_oracle(nv, case):
    O, A, widths, T, B = case
    s = _setup_update(nv, O, A, widths, T, B, 7, nv.FLAVOUR_DPPO)
    dims, off, size, p_pad, p, p_old, obs, act, adv, ret, logp, idx = s
    ii = torch.tensor(idx)
    for ent in (0.0, 0.01):
        terms, grads = orc.dppo_loss_grads(p, p_old, obs[ii], act[ii], ret[ii], adv[ii], 0.2, ent)
        G, L = _run_update(nv, nv.FLAVOUR_DPPO, *s, ent=ent)
        assert float(L[0]) == pytest.approx(terms["loss_clip"], rel=1e-5, abs=1e-7)
        assert float(L[1]) == pytest.approx(terms["loss_value"], rel=1e-5)
        assert float(L[2]) == pytest.approx(terms["loss_ent"], rel=1e-5, abs=1e-9)
        got = from_flat(G, off, size, dims.shapes())
        for k in orc.PARAM_ORDER:
            assert_close_scaled(got[k], grads[k], 1e-5, k)


@pytest.mark.parametrize("case,flavour", [((17, 6, (100, 100, 500, 100), 65536, 4096), "dppo"),
                                          ((17, 6, (100, 100, 500, 100), 8192, 6000), "dppo"),     # CTAs loop over tiles
                                          ((15, 3, (100, 100, 500, 100), 2048, 2048), "ppo"),
                                          ((17, 6, (64, 64, 64, 64), 4096, 1024 + 5), "dppo")])      # ragged last tile
def test_update_two_launch_form_matches_single_launch(nv, case, flavour):
    """Large minibatches take two launches (the update kernel leaves layer inputs and dz in the library's buffer, a
    second pass forms every weight gradient as a GEMM over a few ranges of samples).  Same gradients and loss sums as
    the single-launch form (per-CTA dW, one partial per CTA) to fp32 reduction-order accuracy, at most 8 partial rows
    instead of one per CTA, and bit-repeatable."""
    O, A, widths, T, B = case
    fl = nv.FLAVOUR_DPPO if flavour == "dppo" else nv.FLAVOUR_PPO
    s = _setup_update(nv, O, A, widths, T, B, 11, fl)
    dims = s[0]
    prev = nv.update_set_batch_dw(True)
    try:
        assert nv.update_launches(dims, B) == 2 and 1 <= nv.update_nparts(dims, B) <= 8
        G2, L2 = _run_update(nv, fl, *s, ent=0.01)
        G2b, L2b = _run_update(nv, fl, *s, ent=0.01)
        nv.update_set_batch_dw(False)
        assert nv.update_launches(dims, B) == 1 and nv.update_nparts(dims, B) > 8
        G1, L1 = _run_update(nv, fl, *s, ent=0.01)
    finally:
        nv.update_set_batch_dw(bool(prev))
    assert torch.equal(G2, G2b) and torch.equal(L2, L2b)
    off, size = s[1], s[2]
    a, b = from_flat(G2, off, size, dims.shapes()), from_flat(G1, off, size, dims.shapes())
    for k in orc.PARAM_ORDER:
        assert_close_scaled(a[k], b[k], 2e-6, k)
    assert torch.allclose(L2, L1, rtol=1e-5, atol=1e-9)         # fp32 sums of the per-CTA loss terms in different groupings
    # small minibatches and the FP32-pipe path keep the single-launch form
    assert nv.update_launches(dims, 512) == 1
    p0 = nv.update_set_path(nv.K2_PATH_FP32)
    try:
        assert nv.update_launches(dims, B) == 1
    finally:
        nv.update_set_path(p0)


@pytest.mark.parametrize("case", CASES)
def test_update_ppo_vs_oracle(nv, case):
    O, A, widths, T, B = case
    s = _setup_update(nv, O, A, widths, T, B, 8, nv.FLAVOUR_PPO)
    dims, off, size, p_pad, p, p_old, obs, act, adv, ret, logp, idx = s
    ii = torch.tensor(idx)
    terms, grads = orc.ppo_loss_grads(p, obs[ii], act[ii], ret[ii], adv[ii], logp[ii], 0.2, 0.01)
    G, L = _run_update(nv, nv.FLAVOUR_PPO, *s, ent=0.01)
    assert float(L[0]) == pytest.approx(terms["loss_clip"], rel=1e-5, abs=1e-7)
    assert float(L[1]) == pytest.approx(terms["loss_value"], rel=1e-5)
    assert float(L[2]) == pytest.approx(terms["loss_ent"], rel=1e-5)
    got = from_flat(G, off, size, dims.shapes())
    for k in orc.PARAM_ORDER:
        assert_close_scaled(got[k], grads[k], 1e-5, k)


def test_update_epoch0_ratio_is_one(nv):
    """theta_old == theta (epoch 0, train.py:128-129): every ratio is p/(1e-10+p)."""
    s = list(_setup_update(nv, 17, 6, (100, 100, 500, 100), 1024, 512, 3, nv.FLAVOUR_DPPO))
    s[5] = {k: t.clone() for k, t in s[4].items()}
    dims, off, size, p_pad, p, p_old, obs, act, adv, ret, logp, idx = s
    ii = torch.tensor(idx)
    terms, grads = orc.dppo_loss_grads(p, p_old, obs[ii], act[ii], ret[ii], adv[ii])
    G, L = _run_update(nv, nv.FLAVOUR_DPPO, *s)
    got = from_flat(G, off, size, dims.shapes())
    for k in orc.PARAM_ORDER:
        assert_close_scaled(got[k], grads[k], 1e-5, k)


@pytest.mark.parametrize("flavour_name", ["dppo", "ppo"])
def test_update_fp32_pipe_path_vs_oracle_and_tensor_path(nv, flavour_name):
    """The reference-sized layers have two implementations: tcgen05 through tensor memory (default) and the
    register-tiled FP32-pipe layers (every other shape; forced here with dppo_update_set_path).  Both must meet
    the oracle at 1e-5 and each other."""
    fl = nv.FLAVOUR_DPPO if flavour_name == "dppo" else nv.FLAVOUR_PPO
    s = _setup_update(nv, 17, 6, (100, 100, 500, 100), 4096, 3000, 11, fl)
    dims, off, size, p_pad, p, p_old, obs, act, adv, ret, logp, idx = s
    ii = torch.tensor(idx)
    if fl == nv.FLAVOUR_DPPO:
        terms, grads = orc.dppo_loss_grads(p, p_old, obs[ii], act[ii], ret[ii], adv[ii], 0.2, 0.01)
    else:
        terms, grads = orc.ppo_loss_grads(p, obs[ii], act[ii], ret[ii], adv[ii], logp[ii], 0.2, 0.01)
    G_ts, L_ts = _run_update(nv, fl, *s, ent=0.01)
    prev = nv.update_set_path(nv.K2_PATH_FP32)
    try:
        G_fp, L_fp = _run_update(nv, fl, *s, ent=0.01)
    finally:
        nv.update_set_path(prev)
    assert not torch.equal(G_ts, G_fp)                   # different arithmetic: the switch did switch
    got_ts, got_fp = from_flat(G_ts, off, size, dims.shapes()), from_flat(G_fp, off, size, dims.shapes())
    for k in orc.PARAM_ORDER:
        assert_close_scaled(got_fp[k], grads[k], 1e-5, k + " (fp32 pipe)")
        assert_close_scaled(got_ts[k], grads[k], 1e-5, k + " (tensor cores)")
        assert_close_scaled(got_ts[k], got_fp[k], 1e-5, k + " (paths)")
    for j in range(3):
        assert float(L_ts[j]) == pytest.approx(float(L_fp[j]), rel=1e-5, abs=1e-7)


def test_update_is_deterministic(nv):
    s = _setup_update(nv, 17, 6, (100, 100, 500, 100), 4096, 4096, 5, nv.FLAVOUR_DPPO)
    G1, L1 = _run_update(nv, nv.FLAVOUR_DPPO, *s)
    G2, L2 = _run_update(nv, nv.FLAVOUR_DPPO, *s)
    assert torch.equal(G1, G2) and torch.equal(L1, L2)


def test_update_golden_first_epoch(nv, golden_dir):
    """Gradients of epoch 0 of the reference run (train.py:140-169), worker by worker."""
    g = np.load(os.path.join(golden_dir, "dppo_w2_o17a6.npz"))
    dims = nv.Dims(17, 6, 100, 100, 500, 100)
    off, size, p_pad = nv.param_layout(dims)
    sh = orc.param_shapes(17, 6)
    p = orc.unflatten(torch.from_numpy(g["theta0"]), sh)
    for w in range(2):
        pre = "w%d_" % w
        idx = g[pre + "idx"][0].tolist()
        G, L = _run_update(nv, nv.FLAVOUR_DPPO, dims, off, size, p_pad, p, p, torch.from_numpy(g[pre + "states"]),
                           torch.from_numpy(g[pre + "actions"]), torch.from_numpy(g[pre + "advantages"]),
                           torch.from_numpy(g[pre + "returns"]), None, idx, clip=float(g["clip"]), ent=float(g["ent_coeff"]))
        assert_close_scaled(unpadded(G, off, size), g[pre + "grad_0"], 1e-5, "grad_0")
        t = g[pre + "terms"][0]
        assert float(L[0]) == pytest.approx(t[0], rel=1e-5, abs=1e-7)
        assert float(L[1]) == pytest.approx(t[1], rel=1e-5)
