"""The reference's entry points, used the way main.py / ppo.py use them, on the GPU with a synthetic
environment (gym is not installed): train.train, chief.chief, ppo.train, Shared_obs_stats.observes/normalize."""
import threading

import numpy as np
import pytest
import torch

from oracle import ppo_oracle as orc
from util import assert_close_scaled

pytestmark = pytest.mark.gpu
DEV = "cuda"


class _Space:
    def __init__(self, n):
        self.shape = (n,)


class FakeEnv:
    def __init__(self, O, A, seed, ep_len=37):
        self.observation_space, self.action_space = _Space(O), _Space(A)
        self.rng = np.random.RandomState(seed)
        self.O, self.ep_len, self.t = O, ep_len, 0

    def reset(self):
        self.t = 0
        return self.rng.randn(self.O).astype(np.float32) * 2 + 1

    def step(self, action):
        assert np.all(np.isfinite(action))
        self.t += 1
        return self.rng.randn(self.O).astype(np.float32) * 2 + 1, float(self.rng.randn()), self.t >= self.ep_len, {}


def _params(O, A):
    import main as main_mod
    p = main_mod.Params()
    p.batch_size, p.num_epoch, p.num_steps, p.exploration_size, p.num_processes = 64, 3, 50, 128, 1
    p.update_treshold = 0
    p.env_fn = lambda: FakeEnv(O, A, 5)
    p.max_outer_iters = 2
    return p


def test_train_runs_with_fused_chief():
    """train.train(rank, params, traffic_light, counter, shared_model, shared_grad_buffers, shared_obs_stats,
    test_n) - the reference's worker signature (train.py:46) - two outer iterations."""
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from train import train
    from utils import Counter, TrafficLight
    O, A = 5, 2
    torch.manual_seed(1)
    model = Model(O, A, device=DEV)
    before = model.flat.clone()
    sgb = Shared_grad_buffers(model)
    stats = Shared_obs_stats(O)
    light, counter = TrafficLight(), Counter()
    test_n = torch.zeros(1)
    p = _params(O, A)
    train(0, p, light, counter, model, sgb, stats, test_n)
    torch.cuda.synchronize()
    assert float(test_n) == 2.0                                   # train.py:177
    assert float(stats.n[0]) >= 2 * 129                           # every env step was observed (train.py:84)
    assert not torch.equal(model.flat, before) and torch.isfinite(model.flat).all()
    assert float(sgb.flat.abs().max()) == 0.0                     # chief.py:19 after every update
    assert counter.get() == 0 and light.get() is False            # 6 switches
    # Adam moved every parameter by about lr per update (6 updates of 3e-4)
    assert float((model.flat - before).abs().max()) < 10 * 6 * 3e-4


def test_train_with_external_chief_thread():
    """The reference's hand-shake: worker adds gradients and spins on the traffic light (train.py:169-175), the
    chief polls the counter and steps (chief.py:9-20) - here as two threads of one process."""
    import torch.optim as optim
    from chief import chief
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from train import train
    from utils import Counter, TrafficLight
    O, A = 4, 1
    torch.manual_seed(1)
    model = Model(O, A, device=DEV)
    before = model.flat.clone()
    sgb = Shared_grad_buffers(model)
    stats = Shared_obs_stats(O)
    light, counter = TrafficLight(), Counter()
    p = _params(O, A)
    p.fused_chief = False
    p.max_outer_iters = 1
    p.max_chief_steps = p.num_epoch
    optimizer = optim.Adam(model.parameters(), lr=p.lr)           # main.py:57
    th = threading.Thread(target=chief, args=(1, p, light, counter, model, sgb, optimizer))
    th.start()
    train(0, p, light, counter, model, sgb, stats, torch.zeros(1))
    th.join(60)
    assert not th.is_alive()
    assert not torch.equal(model.flat, before)
    st = optimizer.state[next(iter(model.parameters()))]
    assert float(st["step"]) == p.num_epoch and st["exp_avg"].shape == (1, A)
    assert float(sgb.flat.abs().max()) == 0.0


def test_ppo_train_runs():
    """ppo.train(env, model, optimizer, shared_obs_stats) with a module-level params (ppo.py:62-63,187)."""
    import torch.optim as optim
    import ppo as ppo_mod
    from model import Model, Shared_obs_stats
    O, A = 6, 3
    torch.manual_seed(1)
    model = Model(O, A, device=DEV)
    before = model.flat.clone()
    p = ppo_mod.Params()
    p.batch_size, p.num_epoch, p.num_steps, p.time_horizon = 32, 4, 96, 2
    ppo_mod.params = p
    try:
        eng = ppo_mod.train(FakeEnv(O, A, 9), model, optim.Adam(model.parameters(), lr=p.lr), Shared_obs_stats(O))
    finally:
        ppo_mod.params = None
    torch.cuda.synchronize()
    assert eng.step == 8 and torch.isfinite(model.flat).all() and not torch.equal(model.flat, before)


def _train_dppo_n(outer, use_graph):
    import random
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from train import train
    from utils import Counter, TrafficLight
    O, A = 5, 2
    torch.manual_seed(1)
    random.seed(11)                                              # the reference never seeds random.sample; we do, to compare
    model = Model(O, A, device=DEV)
    sgb, stats = Shared_grad_buffers(model), Shared_obs_stats(O)
    p = _params(O, A)
    p.max_outer_iters, p.use_graph = outer, use_graph
    p.engine = None
    train(0, p, TrafficLight(), Counter(), model, sgb, stats, torch.zeros(1))
    torch.cuda.synchronize()
    return model.flat.clone(), stats.mean.clone()


def test_train_six_outer_iterations_graph_replay_equals_plain_launches():
    """The loop of train.py:60-178 run for six outer iterations.  Every iteration refills the engine's one rollout
    buffer, so the cycle captured in the first iteration is replayed on the right memory in all later ones; the
    result must be bit-identical to the same run with plain kernel launches."""
    a = _train_dppo_n(6, True)
    b = _train_dppo_n(6, False)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_ppo_train_six_horizons_graph_replay_equals_plain_launches():
    import random
    import torch.optim as optim
    import ppo as ppo_mod
    from model import Model, Shared_obs_stats
    O, A = 6, 3
    res = []
    for use_graph in (True, False):
        torch.manual_seed(1)
        random.seed(12)
        model = Model(O, A, device=DEV)
        p = ppo_mod.Params()
        p.batch_size, p.num_epoch, p.num_steps, p.time_horizon, p.use_graph = 32, 4, 96, 6, use_graph
        ppo_mod.params = p
        try:
            eng = ppo_mod.train(FakeEnv(O, A, 9), model, optim.Adam(model.parameters(), lr=p.lr), Shared_obs_stats(O))
        finally:
            ppo_mod.params = None
        torch.cuda.synchronize()
        assert eng.step == 24
        res.append((model.flat.clone(), eng.m.clone()))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])


def test_graph_cache_is_keyed_by_addresses_not_object_ids():
    """A NEW RolloutBuffer per outer iteration (what a caller that ignores rollout_buffer() does): Python reuses
    the ids of freed objects, the allocator reuses or moves their memory.  Every cycle must still run on the
    buffer that was passed in - compared with plain launches - and the cache must stay bounded."""
    import random
    import b200_native as nv
    from b200_engine import RolloutBuffer, UpdateEngine
    from model import Model
    T, B, E, O, A = 1024, 128, 2, 17, 6
    res = []
    for use_graph in (True, False):
        torch.manual_seed(1)
        m = Model(O, A, device=DEV)
        eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, E, 3e-4)
        rng = random.Random(2)
        g = torch.Generator().manual_seed(5)
        keep = []
        for it in range(8):
            buf = RolloutBuffer(T, O, A, DEV)
            buf.obs.copy_(torch.randn(T, O, generator=g)); buf.act.copy_(torch.randn(T, A, generator=g))
            buf.adv.copy_(torch.randn(T, generator=g)); buf.ret.copy_(torch.randn(T, generator=g))
            if it % 3 == 0:
                keep.append(torch.empty(T * (it + 1), device=DEV))    # perturb the allocator's free lists
            eng.draw_indices(T, rng)
            eng.run_epochs(buf, use_graph=use_graph)
            del buf
        torch.cuda.synchronize()
        assert len(eng._graphs) <= eng.MAX_GRAPHS
        res.append(m.flat.clone())
    assert torch.equal(res[0], res[1])


def test_obs_stats_single_step_api_matches_reference_semantics():
    from model import Shared_obs_stats
    s = Shared_obs_stats(3)
    ref = orc.welford_new(3)
    g = torch.Generator().manual_seed(0)
    for _ in range(20):
        x = torch.randn(1, 3, generator=g) * 3 + 1
        s.observes(x)                                             # model.py:68-75, one row
        out = s.normalize(x.to(DEV))                              # model.py:77-80
        want = orc.welford_observe_normalize(ref, x)
        assert_close_scaled(out.cpu(), want, 1e-5, "normalised row")
    assert_close_scaled(s.mean.cpu(), ref["mean"], 1e-5, "mean")


def test_checkpoint_roundtrip(tmp_path):
    import b200_native as nv
    from b200_engine import UpdateEngine, load_checkpoint, save_checkpoint
    from model import Model, Shared_obs_stats
    from synthetic import make_rollout
    import random
    torch.manual_seed(1)
    m = Model(4, 2, device=DEV)
    stats = Shared_obs_stats(4)
    host, dev = make_rollout(m, stats, 512, seed=1, flavour=nv.FLAVOUR_DPPO, seg_len=200)
    eng = UpdateEngine(m, nv.FLAVOUR_DPPO, 64, 2, 3e-4)
    rng = random.Random(3)
    eng.update_from_host(host, dev, obs_stats=stats, rng=rng)
    path = str(tmp_path / "ckpt.pt")
    save_checkpoint(path, eng, obs_stats=stats, rng=rng)
    eng.update_from_host(host, dev, obs_stats=stats, rng=rng)
    want = (m.flat.clone(), eng.m.clone(), eng.v.clone(), eng.step, stats.mean.clone())
    # fresh objects, restored, same second step
    torch.manual_seed(2)
    m2 = Model(4, 2, device=DEV)
    stats2 = Shared_obs_stats(4)
    eng2 = UpdateEngine(m2, nv.FLAVOUR_DPPO, 64, 2, 3e-4)
    rng2 = random.Random(0)
    load_checkpoint(path, eng2, obs_stats=stats2, rng=rng2)
    host2, dev2 = host, dev
    eng2.update_from_host(host2, dev2, obs_stats=stats2, rng=rng2)
    assert torch.equal(m2.flat, want[0]) and torch.equal(eng2.m, want[1]) and torch.equal(eng2.v, want[2])
    assert eng2.step == want[3] and torch.equal(stats2.mean, want[4])


def test_evaluator_role_runs_on_the_shared_model(capsys):
    """test.test(rank, params, shared_model, shared_obs_stats, test_n) - the reference's evaluator (test.py:24-65):
    three episodes on the shared model, the reference's print format, statistics fed per step."""
    import importlib.util
    import os
    from model import Model, Shared_obs_stats
    spec = importlib.util.spec_from_file_location("dppo_test_role", os.path.join(os.path.dirname(os.path.dirname(
        os.path.abspath(__file__))), "pytorch-dppo_b200", "test.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    O, A = 5, 2
    torch.manual_seed(1)
    model = Model(O, A, device=DEV)
    stats = Shared_obs_stats(O)
    p = _params(O, A)
    p.max_eval_episodes, p.eval_sleep_s = 3, 0.0
    res = mod.test(1, p, model, stats, torch.zeros(1))
    out = capsys.readouterr().out
    assert len(res) == 3 and all(r[1] == 37 for r in res)              # FakeEnv episodes last 37 steps
    assert out.count("episode reward") == 3 and "episode length 37" in out
    assert float(stats.n[0]) == 3 * 37                                 # test.py:45: every step observed
