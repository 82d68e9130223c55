"""Evaluator role of the reference (test.py:24-65): an endless loop that plays episodes with the CURRENT shared
model and prints wall time, episode reward and episode length (test.py:56-59).  Kept: the signature
``test(rank, params, shared_model, shared_obs_stats, test_n)``, the seed (``params.seed + rank``, test.py:25), the
per-step order observes -> normalize -> forward -> ``action = mu + sqrt(sigma_sq) * eps`` (test.py:45-50), the
print format and the pause after an episode (10 s, test.py:63; ``params.eval_sleep_s`` overrides it).

In the reference it is one more forked process reading the shared-memory model; here it runs as a thread of a GPU
process (main.run starts it on rank 0) on its own CUDA stream and reads the replica that K3 keeps up to date, so
there is no ``load_state_dict`` copy per step.  It needs an environment (``params.env_fn`` or gym) and is outside
the accelerated update path.  ``params.max_eval_episodes`` bounds the loop (tests)."""
import time

import numpy as np
import torch

from train import _make_env


def test(rank, params, shared_model, shared_obs_stats, test_n):
    torch.manual_seed(params.seed + rank)                                  # test.py:25
    env = _make_env(params)
    model = shared_model
    dev = model.device
    stream = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None
    state = torch.as_tensor(np.asarray(env.reset()), dtype=torch.float32).unsqueeze(0)
    reward_sum, episode_length, episodes = 0.0, 0, 0
    start_time = time.time()
    results = []
    limit = getattr(params, "max_eval_episodes", None)
    while limit is None or episodes < limit:
        episode_length += 1
        with torch.cuda.stream(stream) if stream is not None else _null():
            shared_obs_stats.observes(state)                               # test.py:45
            s = shared_obs_stats.normalize(state.to(dev))                  # test.py:47
            with torch.no_grad():
                mu, sigma_sq, _ = model(s)                                 # test.py:48
                eps = torch.randn(mu.size()).to(dev)
                action = mu + sigma_sq.sqrt() * eps                        # test.py:50
            env_action = action.reshape(-1).cpu().numpy()
        state, reward, done, _ = env.step(env_action)
        reward_sum += float(reward)
        if done:
            print("Time {}, episode reward {}, episode length {}".format(
                time.strftime("%Hh %Mm %Ss", time.gmtime(time.time() - start_time)), reward_sum, episode_length))
            results.append((reward_sum, episode_length, float(test_n.reshape(-1)[0])))
            reward_sum, episode_length = 0.0, 0
            episodes += 1
            state = env.reset()
            time.sleep(getattr(params, "eval_sleep_s", 10))                # test.py:63
        state = torch.as_tensor(np.asarray(state), dtype=torch.float32).unsqueeze(0)
    return results


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
