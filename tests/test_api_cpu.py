"""Host-side surface and the C-ABI boundary, no GPU needed."""
import ctypes
import os
import random
import re

import numpy as np
import pytest
import torch

from oracle import ppo_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import b200_native as nv
    header = open(os.path.join(ROOT, "include", "dppo_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(dppo_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 12
    lib = ctypes.CDLL(nv.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), "symbol %s declared in include/dppo_b200.h is not exported" % name
    assert declared == set(nv.SIGNATURES), declared ^ set(nv.SIGNATURES)
    assert nv.lib.dppo_version() >= 100
    assert b"scratch" in nv.lib.dppo_error_string(-3)


def test_param_layout_matches_reference_shapes():
    import b200_native as nv
    d = nv.Dims(17, 6, 100, 100, 500, 100)
    off, size, p_pad = nv.param_layout(d)
    sh = orc.param_shapes(17, 6)
    assert [int(np.prod(sh[k])) for k in orc.PARAM_ORDER] == size
    assert sum(size) == 71713                          # HalfCheetah, reference arch
    assert all(o % 4 == 0 for o in off) and p_pad % 4 == 0
    assert nv.param_layout(nv.Dims(4, 1, 64, 64, 64, 64))[1] and sum(nv.param_layout(nv.Dims(4, 1, 64, 64, 64, 64))[1]) == 9091
    with pytest.raises(nv.DppoError):
        nv.param_layout(nv.Dims(4, 1, 4096, 64, 64, 64))


def test_update_planning_abi_without_a_gpu():
    """dppo_update_nparts / _launches / _partials_bytes are host-side planning (tiling, shared-memory plan, choice of
    the two-launch form) and must agree with each other without a device."""
    import b200_native as nv
    ref = nv.Dims(17, 6, 100, 100, 500, 100)
    off, size, p_pad = nv.param_layout(ref)
    prev = nv.update_set_batch_dw(True)
    try:
        # large minibatch, every hidden layer on the tensor-core routines: update kernel + batch-level dW pass
        assert nv.update_launches(ref, 4096) == 2
        n = nv.update_nparts(ref, 4096)
        assert 1 <= n <= 8
        assert nv.lib.dppo_update_partials_bytes(ctypes.byref(ref), 4096) == 4 * n * p_pad
        # small minibatches keep one launch and one partial per CTA
        assert nv.update_launches(ref, 64) == 1 and nv.update_nparts(ref, 64) == 16
        # a 8-wide net has no tensor-core layers: single launch whatever the minibatch
        tiny = nv.Dims(3, 2, 8, 12, 20, 4)
        assert nv.update_launches(tiny, 4096) == 1 and nv.update_nparts(tiny, 4096) > 8
        # Humanoid 256x256 (trunk-serial tiles) takes the two-launch form too
        hum = nv.Dims(376, 17, 256, 256, 256, 256)
        assert nv.update_launches(hum, 4096) == 2 and 1 <= nv.update_nparts(hum, 4096) <= 8
        nv.update_set_batch_dw(False)
        assert nv.update_launches(ref, 4096) == 1 and nv.update_nparts(ref, 4096) == 128
    finally:
        nv.update_set_batch_dw(bool(prev))
    with pytest.raises(nv.DppoError):
        nv.update_nparts(ref, 0)


def test_no_cpu_path():
    import b200_native as nv
    x = torch.zeros(8)
    with pytest.raises(nv.DppoError, match="no CPU path"):
        nv.gae_scan(x, x, x, torch.zeros(8, dtype=torch.uint8), x, x, 0.99, 0.95, torch.zeros(64, dtype=torch.uint8))


def test_model_surface_and_init_parity():
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    torch.manual_seed(1)
    m = Model(17, 6, device="cpu")
    torch.manual_seed(1)
    ref = orc.init_params(17, 6)
    names = [n for n, _ in m.named_parameters()]
    assert tuple(names) == orc.PARAM_ORDER
    assert list(m.state_dict().keys()) == list(orc.PARAM_ORDER)
    for n, p in m.named_parameters():
        assert tuple(p.shape) == tuple(ref[n].shape)
        assert torch.equal(p.detach(), ref[n]), n
    # parameters are views into the flat bucket
    m.flat.zero_()
    assert all(float(p.detach().abs().sum()) == 0 for p in m.parameters())
    m.load_state_dict({k: v for k, v in ref.items()})
    assert torch.equal(m.p_fc2.weight.detach(), ref["p_fc2.weight"])
    x = torch.randn(5, 17)
    mu, sig, v = m(x)
    rmu, rsig, rv = orc.forward(ref, x)
    assert mu.shape == (5, 6) and sig.shape == (1, 6) and v.shape == (5, 1)
    torch.testing.assert_close(mu, rmu); torch.testing.assert_close(v, rv)
    m.share_memory(); m.zero_grad(); m.train()
    # widths are arguments (BASELINE configs), defaults are the reference's
    assert Model(4, 1, 64, 64, 64, 64, device="cpu").num_params == 9091

    g = Shared_grad_buffers(m)
    assert set(g.grads) == {n + "_grad" for n in orc.PARAM_ORDER}
    assert all(float(t.min()) == 1.0 and float(t.max()) == 1.0 for t in g.grads.values())     # model.py:51
    (mu.sum() + v.sum() + sig.sum()).backward()
    g.add_gradient(m)
    assert torch.equal(g.grads["mu.bias_grad"], 1.0 + m.mu.bias.grad)
    g.reset()
    assert float(g.flat.abs().sum()) == 0.0

    s = Shared_obs_stats(17, device="cpu")
    assert s.n.shape == (17,) and float(s.var.sum()) == 0.0


def test_traffic_light_and_counter():
    from utils import Counter, TrafficLight
    t = TrafficLight(True)
    assert t.get() is False                            # utils.py:7: the argument is ignored
    t.switch(); assert t.get() is True
    c = Counter(5)
    assert c.get() == 0
    c.increment(); c.increment(); assert c.get() == 2
    c.reset(); assert c.get() == 0


def test_replay_memory_semantics():
    from train import ReplayMemory, normal
    mem = ReplayMemory(5)
    cols = lambda a, b: [[torch.full((1, 2), float(i)) for i in range(a, b)], [torch.full((1, 1), float(i)) for i in range(a, b)]]
    mem.push(cols(0, 4))
    mem.push(cols(4, 8))                               # FIFO: keeps the last 5 (train.py:25-26)
    assert len(mem) == 5 and len(mem.memory) == 5
    assert mem.columns[1].reshape(-1).tolist() == [3, 4, 5, 6, 7]
    random.seed(3)
    want = random.sample(range(5), 3)
    random.seed(3)
    s, r = mem.sample(3)
    assert r.reshape(-1).tolist() == [3.0 + i for i in want]
    assert s.shape == (3, 2)
    mem.clear(); assert len(mem) == 0
    x, mu, var = torch.randn(4, 3), torch.randn(4, 3), torch.rand(1, 3) + 0.5
    torch.testing.assert_close(normal(x, mu, var), orc.normal_pdf(x, mu, var))


def test_params_surface():
    import main as main_mod
    import ppo as ppo_mod
    p = main_mod.Params()
    assert (p.batch_size, p.lr, p.num_epoch, p.num_processes, p.update_treshold) == (1000, 3e-4, 10, 4, 3)
    q = ppo_mod.Params()
    assert (q.batch_size, q.lr, q.ent_coeff, q.max_grad_norm, q.num_steps) == (64, 7e-4, 0.01, 0.5, 2048)
