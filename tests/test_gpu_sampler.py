"""dppo_sample_indices (device) against CPython's random.sample itself (train.py:32 / ppo.py:59): positions and the
generator state left behind must be bit-exact, for CPython's pool branch and its set branch."""
import random

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,k", [(1, 1), (5, 5), (10, 3), (21, 5), (22, 6), (100, 6), (2048, 64), (2048, 2048),
                                 (1000, 1000), (65536, 4096), (32768, 4096), (16405, 4096), (16406, 4096),
                                 (100000, 7), (1 << 20, 4096), (65537, 1), (3, 0), (60000, 20000)])
def test_device_sampler_matches_random_sample(n, k):
    import b200_native as nv
    for seed in (0, 1, 12345):
        a = random.Random(seed)
        b = random.Random(seed)
        if seed == 1:                                    # start in the middle of a generated batch
            [a.random() for _ in range(100)]; [b.random() for _ in range(100)]
        want = [a.sample(range(n), k) for _ in range(3)]
        ds = nv.DeviceSampler(b)
        got = ds.sample(n, k, 3)
        torch.cuda.synchronize()
        assert got.cpu().tolist() == want
        ds.sync_to(b)                                    # the generator is left exactly where Python would have left it
        assert a.getstate() == b.getstate()
        assert a.random() == b.random()


def test_device_sampler_consecutive_calls_continue_the_stream():
    import b200_native as nv
    a, b = random.Random(7), random.Random(7)
    ds = nv.DeviceSampler(b)
    for n, k, r in ((65536, 4096, 10), (2048, 2048, 10), (65536, 4096, 2), (300, 17, 5)):
        want = [a.sample(range(n), k) for _ in range(r)]
        got = ds.sample(n, k, r)
        assert got.cpu().tolist() == want
    ds.sync_to(b)
    assert a.getstate() == b.getstate()


def test_device_sampler_same_as_host_sampler():
    import b200_native as nv
    a, b = random.Random(3), random.Random(3)
    host = nv.sample_indices(65536, 4096, 10, rng=a)
    dev = nv.DeviceSampler(b).sample(65536, 4096, 10)
    assert torch.equal(host, dev.cpu())
