"""torchrun target: the launcher of pytorch-dppo_b200/main.py (one process per GPU, main.py:41-73 equivalent) with a
synthetic environment; every rank trains one outer iteration, replicas must end bit-identical."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-dppo_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


class _Space:
    def __init__(self, n):
        self.shape = (n,)


class FakeEnv:
    def __init__(self, O, A, seed):
        self.observation_space, self.action_space = _Space(O), _Space(A)
        self.rng, self.O, self.t = np.random.RandomState(seed), O, 0

    def reset(self):
        self.t = 0
        return self.rng.randn(self.O).astype(np.float32)

    def step(self, action):
        self.t += 1
        return self.rng.randn(self.O).astype(np.float32), float(self.rng.randn()), self.t >= 41, {}


def main():
    import main as main_mod
    import model as model_mod
    rank = int(os.environ["RANK"])
    p = main_mod.Params()
    p.batch_size, p.num_epoch, p.num_steps, p.exploration_size = 64, 3, 50, 128
    p.env_fn = lambda: FakeEnv(5, 2, 100 + rank)
    p.max_outer_iters = 1
    captured = {}
    orig = model_mod.Model

    class Capturing(orig):
        def __init__(self, *a, **k):
            super().__init__(*a, **k)
            captured["model"] = self

    model_mod.Model = Capturing
    main_mod.run(p, 5, 2)
    flat = captured["model"].flat.clone()
    world = dist.get_world_size()
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    assert all(torch.equal(gathered[0], g) for g in gathered), "replicas diverged"
    assert torch.isfinite(flat).all()
    if rank == 0:
        print("MAIN_LAUNCHER_OK world=%d" % world, flush=True)
    dist.barrier()
    p.comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
