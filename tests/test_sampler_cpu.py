"""dppo_sample_indices_host against CPython's random.sample itself (the call the
reference makes at train.py:32 / ppo.py:59).  Indices must be bit-exact."""
import random

import pytest
import torch

import b200_native as nv


@pytest.mark.parametrize("n,k", [(1, 1), (5, 5), (10, 3), (21, 5), (22, 6), (100, 6), (2048, 64), (2048, 2048),
                                 (1000, 1000), (65536, 4096), (32768, 4096), (16405, 4096), (16406, 4096),
                                 (100000, 7), (1 << 20, 4096), (65537, 1), (3, 0)])
def test_matches_random_sample(n, k):
    for seed in (0, 1, 12345):
        a = random.Random(seed)
        b = random.Random(seed)
        want = [a.sample(range(n), k) for _ in range(3)]
        got = nv.sample_indices(n, k, 3, rng=b)
        assert got.tolist() == want
        # the generator is left exactly where Python would have left it
        assert a.getstate() == b.getstate()
        assert a.random() == b.random()


def test_module_level_generator_and_list_population():
    random.seed(7)
    pop = list("abcdefghijklmnopqrstuvwxyz")
    want = random.sample(pop, 10)                       # the reference samples the list itself (train.py:32)
    random.seed(7)
    got = nv.sample_indices(len(pop), 10)[0].tolist()
    assert [pop[i] for i in got] == want


def test_rejects_bad_arguments():
    with pytest.raises(nv.DppoError):
        nv.sample_indices(5, 6)
    with pytest.raises(nv.DppoError):
        nv.sample_indices(10, 2, out=torch.empty(1, 2, dtype=torch.int64))
