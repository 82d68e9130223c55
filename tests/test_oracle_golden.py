"""The oracle (oracle/ppo_oracle.py) against the fixtures recorded from the
unmodified reference (tests/golden/make_golden.py).  CPU only."""
import os
import random
from collections import OrderedDict

import numpy as np
import pytest
import torch

from oracle import ppo_oracle as orc

torch.set_num_threads(1)


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


def shapes_of(g):
    return orc.param_shapes(int(g["O"]), int(g["A"]))


def T(a):
    return torch.from_numpy(np.asarray(a))


def test_param_order_and_init(golden_dir):
    g = load(golden_dir, "dppo_w2_o17a6")
    assert tuple(g["names"]) == orc.PARAM_ORDER
    torch.manual_seed(1)                       # main.py:32,44
    p = orc.init_params(17, 6)
    assert torch.equal(orc.flatten(p), T(g["theta0"]))
    assert orc.flatten(p).numel() == 71713


@pytest.mark.parametrize("name,prefixes", [("dppo_w2_o17a6", ["w0_", "w1_"]), ("dppo_w1_o3a2", ["w0_"]), ("ppo_o4a2", [""])])
def test_gae_bit_exact(golden_dir, name, prefixes):
    g = load(golden_dir, name)
    for pre in prefixes:
        adv, ret = orc.gae(g[pre + "gae_rewards"], g[pre + "gae_values"], g[pre + "gae_boot"], g[pre + "gae_seg_end"],
                           float(g["gamma"]), float(g["gae_param"]))
        assert np.array_equal(adv, g[pre + "gae_advantages"])
        assert np.array_equal(ret, g[pre + "gae_returns"])
        # the replay buffer keeps the LAST T steps (train.py:22-26)
        Tn = g[pre + "returns"].shape[0]
        assert np.array_equal(ret[-Tn:], g[pre + "returns"].reshape(-1))
        assert np.array_equal(adv[-Tn:], g[pre + "advantages"].reshape(-1))


def test_welford_bit_exact(golden_dir):
    # single worker runs: the statistics are only touched by one process
    for name, pre in (("ppo_o4a2", ""), ("dppo_w1_o3a2", "w0_")):
        g = load(golden_dir, name)
        raw = g[pre + "raw_obs"]
        st = orc.welford_new(raw.shape[1])
        out = orc.welford_observe_normalize(st, raw)
        Tn = g[pre + "states"].shape[0]
        assert torch.equal(out[-Tn:], T(g[pre + "states"]))
        for k in ("n", "mean", "mean_diff", "var"):
            assert torch.equal(st[k], T(g[pre + "stats_" + k])), k


def test_welford_two_workers_sequential(golden_dir):
    # W=2: the harness rolls the workers out one after the other on ONE shared
    # Shared_obs_stats (main.py:54), so worker 1 continues worker 0's statistics
    g = load(golden_dir, "dppo_w2_o17a6")
    st = orc.welford_new(17)
    for pre in ("w0_", "w1_"):
        out = orc.welford_observe_normalize(st, g[pre + "raw_obs"])
        Tn = g[pre + "states"].shape[0]
        assert torch.equal(out[-Tn:], T(g[pre + "states"]))


def test_sample_indices_bit_exact(golden_dir):
    g = load(golden_dir, "dppo_w2_o17a6")
    for r in range(2):
        rng = random.Random(7 + r)
        for k in range(int(g["E"])):
            idx = orc.sample_indices(int(g["T"]), int(g["B"]), rng)
            assert np.array_equal(np.asarray(idx, np.int32), g["w%d_idx" % r][k])
    g = load(golden_dir, "ppo_o4a2")
    rng = random.Random(11)
    for k in range(int(g["E"])):
        assert np.array_equal(np.asarray(orc.sample_indices(int(g["T"]), int(g["B"]), rng), np.int32), g["idx"][k])


def _bufs(g, pre, cols):
    return {c: T(g[pre + c]) for c in cols}


@pytest.mark.parametrize("name,W", [("dppo_w2_o17a6", 2), ("dppo_w1_o3a2", 1)])
def test_dppo_cycle(golden_dir, name, W):
    g = load(golden_dir, name)
    sh = shapes_of(g)
    E = int(g["E"])
    p = orc.unflatten(T(g["theta0"]), sh)
    m = orc.zeros_like_params(p)
    v = orc.zeros_like_params(p)
    gsh = orc.zeros_like_params(p, 1.0)          # model.py:51 ones
    workers = [_bufs(g, "w%d_" % r, ("states", "actions", "returns", "advantages")) for r in range(W)]
    idx = [[g["w%d_idx" % r][k] for k in range(E)] for r in range(W)]
    seen = {}

    def record(k, w, terms, grads):
        seen[(k, w)] = (terms, orc.flatten(grads))

    # step the cycle one epoch at a time so that G and theta can be compared per epoch
    p_old = OrderedDict((k, t.clone()) for k, t in p.items())
    step = 0
    for k in range(E):
        snap = OrderedDict((kk, t.clone()) for kk, t in p.items())
        for w in range(W):
            ii = torch.as_tensor(idx[w][k], dtype=torch.long)
            b = workers[w]
            terms, grads = orc.dppo_loss_grads(snap, p_old, b["states"][ii], b["actions"][ii], b["returns"][ii],
                                               b["advantages"][ii], float(g["clip"]), float(g["ent_coeff"]))
            gt = g["w%d_terms" % w][k]
            assert terms["loss_clip"] == pytest.approx(gt[0], rel=1e-6, abs=1e-9)
            assert terms["loss_value"] == pytest.approx(gt[1], rel=1e-6)
            assert terms["total"] == pytest.approx(gt[3], rel=1e-6)
            gg = T(g["w%d_grad_%d" % (w, k)])
            torch.testing.assert_close(orc.flatten(grads), gg, rtol=1e-6, atol=1e-7 * float(gg.abs().max()))
            for kk in orc.PARAM_ORDER:
                gsh[kk] += grads[kk]
        if "G_%d" % k in g.files:
            torch.testing.assert_close(orc.flatten(gsh), T(g["G_%d" % k]), rtol=1e-6, atol=1e-7)
        p_old = snap
        step += 1
        orc.adam_step(p, gsh, m, v, step, float(g["lr"]))
        for kk in orc.PARAM_ORDER:
            gsh[kk].fill_(0)
        torch.testing.assert_close(orc.flatten(p), T(g["theta_%d" % k]), rtol=1e-6, atol=1e-8)
    torch.testing.assert_close(orc.flatten(m), T(g["m_final"]), rtol=1e-6, atol=1e-9)
    torch.testing.assert_close(orc.flatten(v), T(g["v_final"]), rtol=1e-6, atol=1e-12)
    assert step == int(g["step_final"])


def test_dppo_cycle_helper_matches(golden_dir):
    g = load(golden_dir, "dppo_w2_o17a6")
    sh = shapes_of(g)
    E = int(g["E"])
    p = orc.unflatten(T(g["theta0"]), sh)
    m, v, gsh = orc.zeros_like_params(p), orc.zeros_like_params(p), orc.zeros_like_params(p, 1.0)
    workers = [_bufs(g, "w%d_" % r, ("states", "actions", "returns", "advantages")) for r in range(2)]
    idx = [[g["w%d_idx" % r][k] for k in range(E)] for r in range(2)]
    step = orc.dppo_update_cycle(p, m, v, 0, workers, idx, gsh, lr=float(g["lr"]), num_epoch=E)
    assert step == E
    torch.testing.assert_close(orc.flatten(p), T(g["theta_%d" % (E - 1)]), rtol=1e-6, atol=1e-8)


def test_ppo_cycle(golden_dir):
    g = load(golden_dir, "ppo_o4a2")
    sh = shapes_of(g)
    E = int(g["E"])
    p = orc.unflatten(T(g["theta0"]), sh)
    m, v = orc.zeros_like_params(p), orc.zeros_like_params(p)
    buf = _bufs(g, "", ("states", "actions", "returns", "advantages", "logprobs"))
    idx = [g["idx"][k] for k in range(E)]
    checked = []

    def record(k, terms, grads, clipped, norm):
        gt = g["terms"][k]
        assert terms["loss_clip"] == pytest.approx(gt[0], rel=2e-6, abs=1e-8)
        assert terms["loss_value"] == pytest.approx(gt[1], rel=2e-6)
        assert terms["loss_ent"] == pytest.approx(gt[2], rel=2e-6)
        assert norm == pytest.approx(float(g["norm"][k]), rel=2e-6)
        if "g_raw_%d" % k in g.files:
            gr = T(g["g_raw_%d" % k])
            torch.testing.assert_close(orc.flatten(grads), gr, rtol=1e-5, atol=1e-6 * float(gr.abs().max()))
            gc = T(g["g_clip_%d" % k])
            torch.testing.assert_close(orc.flatten(clipped), gc, rtol=1e-5, atol=1e-6 * float(gc.abs().max()))
            checked.append(k)

    orc.ppo_update_cycle(p, m, v, 0, buf, idx, lr=float(g["lr"]), clip=float(g["clip"]), ent_coeff=float(g["ent_coeff"]),
                         max_grad_norm=float(g["max_grad_norm"]), num_epoch=E, record=record)
    assert checked == [0, 1, 9]
    torch.testing.assert_close(orc.flatten(p), T(g["theta_9"]), rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(orc.flatten(m), T(g["m_final"]), rtol=1e-5, atol=1e-8)
    torch.testing.assert_close(orc.flatten(v), T(g["v_final"]), rtol=1e-5, atol=1e-11)
