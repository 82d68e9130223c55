"""Whole update cycles on the GPU (engine + CUDA graph + drop-in surface) against
the fixtures recorded from the reference and against the oracle."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import ppo_oracle as orc
from util import assert_close_scaled, to_flat, unpadded

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _load_model(g, O, A):
    from model import Model
    m = Model(O, A, device=DEV)
    sh = orc.param_shapes(O, A)
    p = orc.unflatten(torch.from_numpy(g["theta0"]), sh)
    m.load_state_dict({k: v.to(DEV) for k, v in p.items()})
    return m


def _buffer(g, pre, O, A, with_logp=False):
    from b200_engine import RolloutBuffer
    T = g[pre + "states"].shape[0]
    b = RolloutBuffer(T, O, A, DEV)
    b.obs.copy_(torch.from_numpy(g[pre + "states"]))
    b.act.copy_(torch.from_numpy(g[pre + "actions"]))
    b.adv.copy_(torch.from_numpy(g[pre + "advantages"]).reshape(-1))
    b.ret.copy_(torch.from_numpy(g[pre + "returns"]).reshape(-1))
    if with_logp:
        b.logp.copy_(torch.from_numpy(g[pre + "logprobs"]).reshape(-1))
    return b


@pytest.mark.parametrize("use_graph", [False, True])
def test_dppo_w1_cycle_matches_reference(golden_dir, use_graph):
    """train.py:128-175 + chief.py:13-20 for one worker: parameters after every epoch,
    Adam moments, the ones-initialised shared buffer (model.py:51)."""
    import b200_native as nv
    from b200_engine import UpdateEngine
    from model import Shared_grad_buffers
    g = np.load(os.path.join(golden_dir, "dppo_w1_o3a2.npz"))
    O, A, E, B = int(g["O"]), int(g["A"]), int(g["E"]), int(g["B"])
    m = _load_model(g, O, A)
    sgb = Shared_grad_buffers(m)
    buf = _buffer(g, "w0_", O, A)
    eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, 1, float(g["lr"]), clip=float(g["clip"]), ent_coeff=float(g["ent_coeff"]))
    # E = 1 per call so that every epoch's parameters can be compared; theta_old must
    # then lag across calls exactly as inside one call
    for k in range(E):
        eng.idx_host[0] = torch.from_numpy(g["w0_idx"][k])
        prev = m.flat.clone()
        if k > 0:
            eng.theta_old.copy_(prev_prev)
            eng.flavour_keep_old = True
        eng_launch(eng, buf, sgb.flat, use_graph, first=(k == 0))
        prev_prev = prev
        torch.cuda.synchronize()
        assert_close_scaled(unpadded(eng.grad, m.offsets, m.sizes), g["G_%d" % k], 1e-5, "G_%d" % k)
        assert_close_scaled(unpadded(m.flat, m.offsets, m.sizes), g["theta_%d" % k], 1e-5, "theta_%d" % k)
        t = g["w0_terms"][k]
        L = eng.loss_out[0].cpu()
        assert float(L[0]) == pytest.approx(t[0], rel=1e-5, abs=1e-7)
        assert float(L[1]) == pytest.approx(t[1], rel=1e-5)
    assert_close_scaled(unpadded(eng.m, m.offsets, m.sizes), g["m_final"], 1e-5, "m")
    assert_close_scaled(unpadded(eng.v, m.offsets, m.sizes), g["v_final"], 1e-5, "v")
    assert float(sgb.flat.abs().max()) == 0.0


def eng_launch(eng, buf, gshared, use_graph, first):
    """One epoch through the engine without resetting theta_old (the engine copies
    theta -> theta_old at the start of a cycle; for k > 0 we restore the lagged copy)."""
    import b200_native as nv
    eng.idx_dev.copy_(eng.idx_host)
    eng._stage_scalars()
    if first:
        eng.theta_old.copy_(eng.theta)
    nv.update_fused(eng.flavour, eng.dims, buf.obs, buf.act, buf.adv, buf.ret, None, eng.idx_dev[0], eng.theta,
                    eng.theta_old, eng.clip, eng.ent_coeff, eng.partials, eng.loss_partials)
    nv.reduce_adam(eng.partials, eng.nparts, gshared, eng.theta, None, eng.m, eng.v, 0, eng.lr,
                   grad_out=eng.grad, loss_partials=eng.loss_partials, loss_out=eng.loss_out[0],
                   scalars_dev=eng.scalars_dev[0])
    eng.step += 1


def test_dppo_w1_full_cycle_graph_vs_reference(golden_dir):
    """The engine's own E-epoch cycle (graph replay, theta_old written by K3) must land on
    the reference's final parameters."""
    import b200_native as nv
    from b200_engine import UpdateEngine
    from model import Shared_grad_buffers
    g = np.load(os.path.join(golden_dir, "dppo_w1_o3a2.npz"))
    O, A, E, B = int(g["O"]), int(g["A"]), int(g["E"]), int(g["B"])
    for use_graph in (False, True):
        m = _load_model(g, O, A)
        sgb = Shared_grad_buffers(m)
        buf = _buffer(g, "w0_", O, A)
        eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, E, float(g["lr"]), clip=float(g["clip"]), ent_coeff=float(g["ent_coeff"]))
        eng.idx_host.copy_(torch.from_numpy(g["w0_idx"]))
        losses = eng.run_epochs(buf, gshared=sgb.flat, use_graph=use_graph)
        torch.cuda.synchronize()
        assert_close_scaled(unpadded(m.flat, m.offsets, m.sizes), g["theta_%d" % (E - 1)], 1e-5, "theta final")
        assert_close_scaled(unpadded(eng.m, m.offsets, m.sizes), g["m_final"], 1e-5, "m")
        assert_close_scaled(unpadded(eng.v, m.offsets, m.sizes), g["v_final"], 1e-5, "v")
        assert eng.step == int(g["step_final"])
        for k in range(E):
            assert float(losses[k, 1]) == pytest.approx(g["w0_terms"][k][1], rel=1e-5)


def test_dppo_w2_cycle_matches_reference(golden_dir):
    """Two workers' gradients summed into the shared buffer in rank order, then Adam
    (model.py:53-55, chief.py:13-20).  Both workers run on this one GPU; K3 sums the
    concatenated partials (the multi-GPU exchange is tested in test_gpu_comm.py)."""
    import b200_native as nv
    from model import Shared_grad_buffers
    g = np.load(os.path.join(golden_dir, "dppo_w2_o17a6.npz"))
    O, A, E, B = int(g["O"]), int(g["A"]), int(g["E"]), int(g["B"])
    m = _load_model(g, O, A)
    sgb = Shared_grad_buffers(m)
    bufs = [_buffer(g, "w%d_" % w, O, A) for w in range(2)]
    npart = nv.update_nparts(m.dims, B)
    partials = torch.empty(2 * npart, m.P_pad, device=DEV)
    lparts = torch.empty(2 * npart, 4, device=DEV)
    theta_old = m.flat.clone()
    mm, vv = torch.zeros_like(m.flat), torch.zeros_like(m.flat)
    G = torch.empty_like(m.flat)
    for k in range(E):
        for w in range(2):
            idx = torch.from_numpy(g["w%d_idx" % w][k]).to(DEV)
            nv.update_fused(nv.FLAVOUR_DPPO, m.dims, bufs[w].obs, bufs[w].act, bufs[w].adv, bufs[w].ret, None, idx,
                            m.flat, theta_old, float(g["clip"]), float(g["ent_coeff"]),
                            partials[w * npart:(w + 1) * npart], lparts[w * npart:(w + 1) * npart])
            torch.cuda.synchronize()
            gw = partials[w * npart:(w + 1) * npart].double().sum(0).float()
            assert_close_scaled(unpadded(gw, m.offsets, m.sizes), g["w%d_grad_%d" % (w, k)], 1e-5, "grad w%d e%d" % (w, k))
            L = lparts[w * npart:(w + 1) * npart].double().sum(0).cpu()
            assert float(L[0]) == pytest.approx(g["w%d_terms" % w][k][0], rel=1e-5, abs=1e-7)
            assert float(L[1]) == pytest.approx(g["w%d_terms" % w][k][1], rel=1e-5)
        nv.reduce_adam(partials, 2 * npart, sgb.flat, m.flat, theta_old, mm, vv, k + 1, float(g["lr"]), grad_out=G)
        torch.cuda.synchronize()
        if "G_%d" % k in g.files:
            assert_close_scaled(unpadded(G, m.offsets, m.sizes), g["G_%d" % k], 1e-5, "G_%d" % k)
        assert_close_scaled(unpadded(m.flat, m.offsets, m.sizes), g["theta_%d" % k], 1e-5, "theta_%d" % k)
    assert_close_scaled(unpadded(mm, m.offsets, m.sizes), g["m_final"], 1e-5, "m")
    assert_close_scaled(unpadded(vv, m.offsets, m.sizes), g["v_final"], 1e-5, "v")


@pytest.mark.parametrize("use_graph", [False, True])
def test_ppo_cycle_matches_reference(golden_dir, use_graph):
    """ppo.py:139-174, ten epochs: losses, clipped gradient norm path, final parameters."""
    import ppo as ppo_mod
    g = np.load(os.path.join(golden_dir, "ppo_o4a2.npz"))
    O, A, E, B = int(g["O"]), int(g["A"]), int(g["E"]), int(g["B"])
    m = _load_model(g, O, A)
    buf = _buffer(g, "", O, A, with_logp=True)
    p = ppo_mod.Params()
    p.batch_size, p.num_epoch, p.num_steps = B, E, int(g["T"])
    eng = ppo_mod.make_engine(p, m)
    eng.idx_host.copy_(torch.from_numpy(g["idx"]))
    losses = eng.run_epochs(buf, use_graph=use_graph)
    torch.cuda.synchronize()
    for k in range(E):
        t = g["terms"][k]
        assert float(losses[k, 0]) == pytest.approx(t[0], rel=2e-5, abs=1e-7)
        assert float(losses[k, 1]) == pytest.approx(t[1], rel=2e-5)
        assert float(losses[k, 2]) == pytest.approx(t[2], rel=2e-5)
    assert_close_scaled(unpadded(eng.grad, m.offsets, m.sizes), g["g_clip_9"], 2e-5, "clipped grad, epoch 9")
    assert_close_scaled(unpadded(m.flat, m.offsets, m.sizes), g["theta_9"], 1e-5, "theta after 10 epochs")
    assert_close_scaled(unpadded(eng.m, m.offsets, m.sizes), g["m_final"], 2e-5, "m")
    assert_close_scaled(unpadded(eng.v, m.offsets, m.sizes), g["v_final"], 2e-5, "v")


def test_engine_vs_oracle_c2_shape_ten_epochs():
    """C2 (HalfCheetah shape, reference arch, T = B = 2048, 10 epochs) end to end through
    K4 -> K1 -> 10 x (K2 -> K3) from a host rollout, against the oracle on the same data
    and the same random.sample stream."""
    import b200_native as nv
    from b200_engine import UpdateEngine
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from synthetic import make_rollout
    torch.manual_seed(1)
    m = Model(17, 6, device=DEV)
    stats = Shared_obs_stats(17)
    T = B = 2048
    E = 10
    host, dev = make_rollout(m, stats, T, seed=3, flavour=nv.FLAVOUR_DPPO, seg_len=700)
    p0 = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    sgb = Shared_grad_buffers(m)
    eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, E, 3e-4)
    rng = random.Random(5)
    loss, h2d, d2h = eng.update_from_host(host, dev, obs_stats=stats, rng=rng)
    assert h2d > T * 17 * 4 and d2h == E * 4 * 4
    # oracle on the same inputs
    st = orc.welford_new(17)
    states = orc.welford_observe_normalize(st, host.raw_obs.clone())
    assert_close_scaled(dev.obs.cpu(), states, 1e-5, "states")
    adv, ret = orc.gae(host.rew.numpy(), host.val.numpy(), host.boot.numpy(), host.seg_end.numpy())
    assert_close_scaled(dev.adv.cpu(), adv, 1e-5, "adv")
    rng = random.Random(5)
    idx = [orc.sample_indices(T, B, rng) for _ in range(E)]
    assert np.array_equal(np.asarray(idx, np.int32), eng.idx_host.numpy())        # indices bit-exact
    p = {k: v.clone() for k, v in p0.items()}
    mo, vo = orc.zeros_like_params(p), orc.zeros_like_params(p)
    gsh = orc.zeros_like_params(p, 0.0)
    worker = dict(states=dev.obs.cpu(), actions=host.act.clone(), returns=dev.ret.cpu().reshape(-1, 1),
                  advantages=dev.adv.cpu().reshape(-1, 1))
    terms = []
    orc.dppo_update_cycle(p, mo, vo, 0, [worker], [idx], gsh, lr=3e-4, num_epoch=E,
                          record=lambda k, w, t, gr: terms.append(t))
    for k in range(E):
        assert float(loss[k, 0]) == pytest.approx(terms[k]["loss_clip"], rel=1e-4, abs=1e-6)
        assert float(loss[k, 1]) == pytest.approx(terms[k]["loss_value"], rel=1e-4)
    assert_close_scaled(unpadded(m.flat, m.offsets, m.sizes), orc.flatten(p), 1e-5, "theta after 10 epochs")
    assert eng.step == E


def test_prefetched_indices_are_the_same_stream():
    """update_from_host(prefetch_next=True) draws the next call's positions early; the stream of
    random.sample results must be exactly the one Python would have produced."""
    import b200_native as nv
    from b200_engine import UpdateEngine
    from model import Model, Shared_obs_stats
    from synthetic import make_rollout
    torch.manual_seed(1)
    m = Model(4, 2, device=DEV)
    stats = Shared_obs_stats(4)
    T, B, E = 512, 64, 3
    host, dev = make_rollout(m, stats, T, seed=1, flavour=nv.FLAVOUR_DPPO, seg_len=200)
    eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, E, 3e-4)
    rng = random.Random(9)
    ref = random.Random(9)
    for call in range(3):
        eng.update_from_host(host, dev, obs_stats=None, rng=rng, prefetch_next=True)
        want = [ref.sample(range(T), B) for _ in range(E)]
        assert eng.idx_host.tolist() == want, call


def test_update_stream_matches_step_by_step():
    """The pipelined host->device stream (copy stream + two device buffers + early index sampling) must produce
    exactly what step-by-step update_from_host calls produce."""
    import b200_native as nv
    from b200_engine import RolloutBuffer, UpdateEngine
    from model import Model, Shared_obs_stats
    from synthetic import make_rollout
    T, B, E, n = 2048, 256, 3, 5
    res = []
    for mode in ("stream", "steps"):
        torch.manual_seed(1)
        m = Model(17, 6, device=DEV)
        stats = Shared_obs_stats(17)
        hosts = [make_rollout(m, stats, T, seed=40 + i, flavour=nv.FLAVOUR_DPPO, seg_len=500)[0] for i in range(3)]
        devs = [RolloutBuffer(T, 17, 6, DEV), RolloutBuffer(T, 17, 6, DEV)]
        eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, E, 3e-4)
        rng = random.Random(4)
        if mode == "stream":
            losses, h2d, d2h = eng.update_stream(hosts, devs, n, obs_stats=stats, rng=rng)
            losses = [l.clone() for l in losses[-2:]]
        else:
            losses = []
            for i in range(n):
                l, _, _ = eng.update_from_host(hosts[i % 3], devs[0], obs_stats=stats, rng=rng)
                losses.append(l.clone())
            losses = losses[-2:]
        torch.cuda.synchronize()
        res.append((m.flat.clone(), eng.m.clone(), stats.mean.clone(), losses))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1]) and torch.equal(res[0][2], res[1][2])
    for a, b in zip(res[0][3], res[1][3]):
        assert torch.equal(a, b)


@pytest.mark.gpu
def test_engine_is_bit_repeatable():
    """Two identical runs of the engine (CUDA-graph replay, K2 and K3 back to back over several rollouts) must give
    bit-identical parameters and Adam state.  A kernel-level repeat of K2 alone (one launch, then a synchronise) did
    NOT expose a race in the tensor-core pipeline that this did: keep it at the engine level."""
    import b200_native as nv
    from b200_engine import RolloutBuffer, UpdateEngine
    from model import Model, Shared_obs_stats
    from synthetic import make_rollout
    T, B, E, n = 2048, 256, 3, 5
    for rep in range(3):
        res = []
        for _ in range(2):
            torch.manual_seed(1)
            m = Model(17, 6, device=DEV)
            stats = Shared_obs_stats(17)
            hosts = [make_rollout(m, stats, T, seed=40 + i, flavour=nv.FLAVOUR_DPPO, seg_len=500)[0] for i in range(3)]
            dev = RolloutBuffer(T, 17, 6, DEV)
            eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, E, 3e-4)
            rng = random.Random(4)
            for i in range(n):
                eng.update_from_host(hosts[i % 3], dev, obs_stats=stats, rng=rng)
            torch.cuda.synchronize()
            res.append((m.flat.clone(), eng.m.clone()))
        assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1]), "repeat %d differs" % rep


def _replay_run(widths, O, A, T, B, E, replays, seed=1):
    """`replays` consecutive graph replays of the whole step (K4, K1, E x (K2, K3)) on one rollout with fresh
    minibatch positions every time; returns the final parameters and Adam state."""
    import b200_native as nv
    from b200_engine import UpdateEngine
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from synthetic import make_rollout
    torch.manual_seed(seed)
    m = Model(O, A, *widths, device=DEV)
    stats = Shared_obs_stats(O)
    host, dev = make_rollout(m, stats, T, seed=7, flavour=nv.FLAVOUR_DPPO, seg_len=1000)
    sgb = Shared_grad_buffers(m)
    eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, E, 3e-4)
    rng = random.Random(3)
    stats0 = [t.clone() for t in (stats.n, stats.mean, stats.mean_diff, stats.var)]
    for i in range(replays):
        for t, s in zip((stats.n, stats.mean, stats.mean_diff, stats.var), stats0):
            t.copy_(s)
        eng.draw_indices(T, rng)
        eng.run_epochs(dev, gshared=sgb.flat, use_graph=True, obs_stats=stats, with_gae=True)
    torch.cuda.synchronize()
    assert torch.isfinite(m.flat).all()
    return m.flat.clone(), eng.m.clone(), eng.v.clone()


def test_engine_is_bit_repeatable_at_the_bench_shape():
    """C4 exactly as bench.py runs it (reference nets, 65536-step rollout, minibatch 4096 -> 28 samples per CTA on
    147 CTAs, 10 epochs), 24 back-to-back graph replays, twice: bit-identical parameters and Adam moments.  This is
    the guard on the ordering of the tensor-core pipeline's barriers (csrc/dppo_update.cu, ts_layer)."""
    a = _replay_run((100, 100, 500, 100), 17, 6, 65536, 4096, 10, 24)
    b = _replay_run((100, 100, 500, 100), 17, 6, 65536, 4096, 10, 24)
    for x, y in zip(a, b):
        assert torch.equal(x, y)


@pytest.mark.parametrize("widths,O,A", [((100, 100, 500, 100), 17, 6), ((64, 64, 64, 64), 17, 6),
                                        ((128, 128, 128, 128), 17, 6), ((96, 96, 160, 96), 11, 3)])
def test_engine_replay_stress_is_bit_repeatable(widths, O, A):
    """120 graph replays x 4 epochs per shape (every tcgen05 layer variant: 3..16 k-blocks, one and several tiles),
    two runs, compared bit for bit."""
    a = _replay_run(widths, O, A, 8192, 1024, 4, 120)
    b = _replay_run(widths, O, A, 8192, 1024, 4, 120)
    for x, y in zip(a, b):
        assert torch.equal(x, y)


def test_engine_vs_oracle_c4_shape_full_cycle():
    """The benchmark's own configuration end to end (K4 -> K1 -> 10 x (K2 -> K3) on a 65536-step rollout, minibatch
    4096) against the oracle on the same data and the same random.sample stream."""
    import b200_native as nv
    from b200_engine import UpdateEngine
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from synthetic import make_rollout
    torch.manual_seed(1)
    O, A, T, B, E = 17, 6, 65536, 4096, 10
    m = Model(O, A, device=DEV)
    stats = Shared_obs_stats(O)
    host, dev = make_rollout(m, stats, T, seed=2, flavour=nv.FLAVOUR_DPPO)
    p = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    sgb = Shared_grad_buffers(m)
    eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, E, 3e-4)
    eng.draw_indices(T, random.Random(11))
    loss = eng.run_epochs(dev, gshared=sgb.flat, use_graph=True, obs_stats=stats, with_gae=True).cpu()
    torch.cuda.synchronize()
    rng = random.Random(11)
    idx = [orc.sample_indices(T, B, rng) for _ in range(E)]
    assert np.array_equal(np.asarray(idx, np.int32), eng.idx_host.numpy())
    mo, vo = orc.zeros_like_params(p), orc.zeros_like_params(p)
    gsh = orc.zeros_like_params(p, 1.0)                               # model.py:51
    worker = dict(states=dev.obs.cpu(), actions=host.act.clone(), returns=dev.ret.cpu().reshape(-1, 1),
                  advantages=dev.adv.cpu().reshape(-1, 1))
    terms = []
    orc.dppo_update_cycle(p, mo, vo, 0, [worker], [idx], gsh, lr=3e-4, num_epoch=E,
                          record=lambda k, w, t, g: terms.append(t))
    for k in range(E):
        assert float(loss[k, 0]) == pytest.approx(terms[k]["loss_clip"], rel=1e-5, abs=1e-7)
        assert float(loss[k, 1]) == pytest.approx(terms[k]["loss_value"], rel=1e-5)
    got = unpadded(m.flat, m.offsets, m.sizes)
    assert_close_scaled(got, orc.flatten(p), 1e-5, "theta after 10 epochs (C4)")
    o = 0
    for k in orc.PARAM_ORDER:                                         # and tensor by tensor
        n = p[k].numel()
        assert_close_scaled(got[o:o + n], p[k].reshape(-1), 1e-5, "theta[%s] (C4)" % k)
        o += n
    assert_close_scaled(unpadded(eng.m, m.offsets, m.sizes), orc.flatten(mo), 1e-5, "Adam m (C4)")


def test_update_stream_with_device_sampler_equals_host_sampler():
    """Minibatch positions drawn on the GPU (DeviceSampler, one step ahead on a side stream) give bit-identical
    training to positions drawn by the host sampler from the same random.Random stream."""
    import b200_native as nv
    from b200_engine import RolloutBuffer, UpdateEngine
    from model import Model, Shared_obs_stats
    from synthetic import make_rollout
    T, B, E, n = 4096, 512, 3, 6
    res = []
    for mode in ("host", "device"):
        torch.manual_seed(1)
        m = Model(17, 6, device=DEV)
        stats = Shared_obs_stats(17)
        hosts = [make_rollout(m, stats, T, seed=40 + i, flavour=nv.FLAVOUR_DPPO, seg_len=500)[0] for i in range(3)]
        devs = [RolloutBuffer(T, 17, 6, DEV), RolloutBuffer(T, 17, 6, DEV)]
        eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, E, 3e-4)
        rng = random.Random(4)
        if mode == "host":
            losses, h2d_host, _ = eng.update_stream(hosts, devs, n, obs_stats=stats, rng=rng)
        else:
            losses, h2d_dev, _ = eng.update_stream(hosts, devs, n, obs_stats=stats, sampler=nv.DeviceSampler(rng))
        torch.cuda.synchronize()
        res.append((m.flat.clone(), eng.m.clone(), losses[-1].clone()))
    assert h2d_dev == h2d_host - E * B * 4                      # the index block no longer crosses PCIe
    for a, b in zip(res[0], res[1]):
        assert torch.equal(a, b)
