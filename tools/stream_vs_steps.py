"""Localise a stream-vs-steps mismatch: same set-up as tests/test_gpu_engine.py::test_update_stream_matches_step_by_step,
but n is swept so the first diverging step shows, and theta / Adam m / obs mean are reported separately."""
import random, sys, torch
sys.path.insert(0, '/root/repo/pytorch-dppo_b200'); sys.path.insert(0, '/root/repo')
import b200_native as nv
from b200_engine import RolloutBuffer, UpdateEngine
from model import Model, Shared_obs_stats
from synthetic import make_rollout
DEV = 'cuda:0'
T, B, E = 2048, 256, 3
def run(mode, n):
    torch.manual_seed(1)
    m = Model(17, 6, device=DEV); stats = Shared_obs_stats(17)
    hosts = [make_rollout(m, stats, T, seed=40 + i, flavour=nv.FLAVOUR_DPPO, seg_len=500)[0] for i in range(3)]
    devs = [RolloutBuffer(T, 17, 6, DEV), RolloutBuffer(T, 17, 6, DEV)]
    eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, E, 3e-4); rng = random.Random(4)
    if mode == 'stream': eng.update_stream(hosts, devs, n, obs_stats=stats, rng=rng)
    else:
        for i in range(n): eng.update_from_host(hosts[i % 3], devs[0], obs_stats=stats, rng=rng)
    torch.cuda.synchronize()
    return m.flat.clone(), eng.m.clone(), stats.mean.clone()
for rep in range(3):
    out = []
    for n in (1, 2, 3, 5):
        a, b = run('stream', n), run('steps', n)
        out.append('n=%d theta:%s m:%s mean:%s' % (n, torch.equal(a[0], b[0]), torch.equal(a[1], b[1]), torch.equal(a[2], b[2])))
    s1, s2 = run('steps', 5), run('steps', 5)
    out.append('steps-vs-steps theta:%s' % torch.equal(s1[0], s2[0]))
    print('rep %d | ' % rep + ' | '.join(out))
