"""Parameter-server step of the reference (chief.py:7-21) on K3.

``chief(rank, params, traffic_light, counter, shared_model, shared_grad_buffers,
optimizer)`` keeps the reference's loop: poll the counter, and once every worker
has contributed (``counter.get() > params.update_treshold`` - the reference's
spelling) run ONE Adam step on the summed gradient, reset counter and buffers,
flip the traffic light.  The step itself is ``dppo_reduce_adam`` on the flat
buckets (no per-tensor loop, no autograd); hyper-parameters are read from the
torch optimizer the caller built (main.py:57) and the moments live in flat
buffers exposed through ``optimizer.state`` so ``state_dict()`` keeps working.
The reference sleeps one second per poll (chief.py:10), which caps DPPO at one
update per second; ``params.chief_poll_s`` (default 0.001) replaces it."""
import time

import torch

import b200_native as nv


class _FlatAdamState:
    def __init__(self, model, optimizer):
        self.m = torch.zeros_like(model.flat)
        self.v = torch.zeros_like(model.flat)
        self.step = 0
        self.step_t = torch.zeros((), dtype=torch.float32)
        for i, (name, p) in enumerate(model.named_parameters()):
            sl = slice(model.offsets[i], model.offsets[i] + model.sizes[i])
            optimizer.state[p] = {"step": self.step_t, "exp_avg": self.m[sl].view(p.shape),
                                  "exp_avg_sq": self.v[sl].view(p.shape)}


def chief_step(shared_model, shared_grad_buffers, optimizer):
    """chief.py:15-19 without the polling: Adam on the shared gradient sum, then reset."""
    st = getattr(optimizer, "_b200_state", None)
    if st is None:
        st = optimizer._b200_state = _FlatAdamState(shared_model, optimizer)
    grp = optimizer.param_groups[0]
    st.step += 1
    st.step_t.fill_(st.step)
    nv.reduce_adam(None, 0, shared_grad_buffers.flat, shared_model.flat, None, st.m, st.v, st.step, grp["lr"],
                   grp["betas"][0], grp["betas"][1], grp["eps"])


def chief(rank, params, traffic_light, counter, shared_model, shared_grad_buffers, optimizer):
    poll = getattr(params, "chief_poll_s", 0.001)
    steps = 0
    while getattr(params, "max_chief_steps", None) is None or steps < params.max_chief_steps:
        time.sleep(poll)
        if counter.get() > params.update_treshold:
            chief_step(shared_model, shared_grad_buffers, optimizer)
            torch.cuda.synchronize()
            counter.reset()
            traffic_light.switch()
            steps += 1
