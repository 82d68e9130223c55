"""Launcher (reference main.py) - one process per GPU instead of forked CPU
workers plus a chief.  ``Params`` keeps the reference's attributes
(main.py:18-39, including the ``update_treshold`` spelling).

    torchrun --nproc-per-node N pytorch-dppo_b200/main.py     # N = number of GPUs

Every rank builds the same model (same seed, main.py:32,44), owns one worker's
rollout and runs train.train; the gradient sum over workers and the Adam step
(chief.py) happen inside K3 on every rank.  Needs an environment:
``Params.env_fn`` or gym."""
import os

import torch


class Params():
    def __init__(self):
        self.batch_size = 1000
        self.lr = 3e-4
        self.gamma = 0.99
        self.gae_param = 0.95
        self.clip = 0.2
        self.ent_coeff = 0.
        self.num_epoch = 10
        self.num_steps = 1000
        self.exploration_size = 1000
        self.num_processes = 4
        self.update_treshold = self.num_processes - 1
        self.max_episode_length = 10000
        self.seed = 1
        self.env_name = 'InvertedPendulum-v1'


def run(params, num_inputs, num_outputs):
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from train import train
    from utils import Counter, TrafficLight
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        import torch.distributed as dist
        from b200_comm import PeerComm
        dist.init_process_group("nccl")
        params.comm = PeerComm.create(dist.group.WORLD)
    params.num_processes = world
    params.update_treshold = world - 1
    torch.manual_seed(params.seed)
    shared_model = Model(num_inputs, num_outputs)
    shared_obs_stats = Shared_obs_stats(num_inputs)
    test_n = torch.zeros(1)
    evaluator = None
    if rank == 0 and getattr(params, "with_evaluator", False):
        # main.py:62-64: the test process.  A thread of rank 0 here; it normalises with its own copy of the statistics
        # (in the reference its observations enter the shared statistics, unlocked: test.py:45)
        import threading
        from test import test
        evaluator = threading.Thread(target=test, args=(world, params, shared_model, shared_obs_stats.clone(), test_n),
                                     daemon=True)
        evaluator.start()
    train(rank, params, TrafficLight(), Counter(), shared_model, Shared_grad_buffers(shared_model),
          shared_obs_stats, test_n)
    if evaluator is not None:
        evaluator.join(getattr(params, "eval_join_s", 60))


if __name__ == '__main__':
    params = Params()
    env = params.env_fn() if getattr(params, "env_fn", None) else __import__("gym").make(params.env_name)
    run(params, env.observation_space.shape[0], env.action_space.shape[0])
