"""CPU oracle for the data-parallel PPO update path of kibeomKim/Pytorch-DPPO.

THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it.  The shipped path (``pytorch-dppo_b200/``)
never does: it calls the sm_100a kernels through the C-ABI and fails loudly
when the library is missing.

What it is: a restatement, in torch CPU fp32 (the arithmetic library the
reference itself runs on), of each function on the hot path, one plain
function per reference item, each citing the reference file:line it follows.
Autograd is used exactly where the reference uses it (``backward`` at
train.py:166 / ppo.py:172), so the gradients are the reference's gradients.

Pinning: the reference has no tests or golden vectors of its own
(SURVEY.md section 4).  The oracle is pinned instead against outputs of the
UNMODIFIED reference code executed in the build container
(``tests/golden/make_golden.py`` imports /root/reference under stub gym
modules and records rollouts, index streams, losses, gradients, post-Adam
parameters and moments); the fixtures live in ``tests/golden/*.npz`` and
``tests/test_oracle_golden.py`` checks every function here against them.
Third-party arithmetic (torch.optim.Adam, clip_grad_norm, autograd) is pinned
to torch 2.11.0 CPU, single-tensor Adam.
"""
from __future__ import annotations

import math
import random
from collections import OrderedDict

import numpy as np
import torch
import torch.nn.functional as F

# named_parameters() order of model.Model (model.py:14-22; log_std is the
# module's own Parameter so it is yielded first, then submodules in
# registration order).
PARAM_ORDER = (
    "log_std",
    "p_fc1.weight", "p_fc1.bias",
    "p_fc2.weight", "p_fc2.bias",
    "v_fc1.weight", "v_fc1.bias",
    "v_fc2.weight", "v_fc2.bias",
    "mu.weight", "mu.bias",
    "v.weight", "v.bias",
)


def param_shapes(O, A, h1=100, h2=100, g1=None, g2=None):
    """Shapes of the 13 parameters (model.py:11-22).  Reference widths are
    h1=h2=100, value trunk 5*h1 -> h2."""
    g1 = 5 * h1 if g1 is None else g1
    g2 = h2 if g2 is None else g2
    return OrderedDict([
        ("log_std", (1, A)),
        ("p_fc1.weight", (h1, O)), ("p_fc1.bias", (h1,)),
        ("p_fc2.weight", (h2, h1)), ("p_fc2.bias", (h2,)),
        ("v_fc1.weight", (g1, O)), ("v_fc1.bias", (g1,)),
        ("v_fc2.weight", (g2, g1)), ("v_fc2.bias", (g2,)),
        ("mu.weight", (A, h2)), ("mu.bias", (A,)),
        ("v.weight", (1, g2)), ("v.bias", (1,)),
    ])


def init_params(O, A, h1=100, h2=100, g1=None, g2=None):
    """model.py:9-33: nn.Linear default init consumed from the global torch
    RNG in CONSTRUCTION order (p_fc1, p_fc2, v_fc1, v_fc2, mu, v), biases then
    zero-filled (model.py:24-27), log_std zeros (model.py:21)."""
    g1 = 5 * h1 if g1 is None else g1
    g2 = h2 if g2 is None else g2
    lin = OrderedDict()
    lin["p_fc1"] = torch.nn.Linear(O, h1)
    lin["p_fc2"] = torch.nn.Linear(h1, h2)
    lin["v_fc1"] = torch.nn.Linear(O, g1)
    lin["v_fc2"] = torch.nn.Linear(g1, g2)
    lin["mu"] = torch.nn.Linear(h2, A)
    lin["v"] = torch.nn.Linear(g2, 1)
    out = OrderedDict()
    out["log_std"] = torch.zeros(1, A)
    for name in ("p_fc1", "p_fc2", "v_fc1", "v_fc2", "mu", "v"):
        out[name + ".weight"] = lin[name].weight.detach().clone()
        out[name + ".bias"] = torch.zeros_like(lin[name].bias)
    return OrderedDict((k, out[k]) for k in PARAM_ORDER)


def forward(p, x):
    """model.py:35-45.  Returns (mu, sigma_sq = exp(log_std), v)."""
    h = torch.tanh(F.linear(x, p["p_fc1.weight"], p["p_fc1.bias"]))
    h = torch.tanh(F.linear(h, p["p_fc2.weight"], p["p_fc2.bias"]))
    mu = F.linear(h, p["mu.weight"], p["mu.bias"])
    sigma_sq = torch.exp(p["log_std"])
    g = torch.tanh(F.linear(x, p["v_fc1.weight"], p["v_fc1.bias"]))
    g = torch.tanh(F.linear(g, p["v_fc2.weight"], p["v_fc2.bias"]))
    v = F.linear(g, p["v.weight"], p["v.bias"])
    return mu, sigma_sq, v


# --------------------------------------------------------------------------
# a4  minibatch indices
# --------------------------------------------------------------------------
def sample_indices(T, B, rng=random):
    """train.py:31-33 / ppo.py:58-60: ``random.sample(self.memory, B)``.
    Sampling ``range(T)`` consumes the Mersenne Twister identically to
    sampling a length-T list, so these are the positions the reference
    gathers (bit-exact)."""
    return rng.sample(range(T), B)


# --------------------------------------------------------------------------
# a5  GAE / returns
# --------------------------------------------------------------------------
def gae(rewards, values, boot, seg_end, gamma=0.99, lam=0.95):
    """train.py:113-122 / ppo.py:124-133, all segments of a rollout buffer in
    one call.  ``seg_end[i]`` marks the last step of a segment and ``boot[i]``
    is then V_{i+1} (0 if the segment ended by ``done`` else the bootstrap
    value, train.py:109-112); inside a segment V_{i+1} = values[i+1].

    fp32 with the reference's op order and roundings: gamma is applied to a
    0-dim fp32 tensor (-> fp32(gamma)), gamma*gae_param is a Python double
    product applied to an fp32 tensor (-> fp32(gamma*gae_param)); no FMA."""
    r = np.asarray(rewards, dtype=np.float32)
    v = np.asarray(values, dtype=np.float32)
    b = np.asarray(boot, dtype=np.float32)
    e = np.asarray(seg_end).astype(bool)
    T = r.shape[0]
    adv = np.empty(T, np.float32)
    ret = np.empty(T, np.float32)
    g = np.float32(gamma)
    gl = np.float32(gamma * lam)
    A = np.float32(0.0)
    for i in range(T - 1, -1, -1):
        if e[i]:
            A = np.float32(0.0)
            vnext = b[i]
        else:
            vnext = v[i + 1]
        td = np.float32(np.float32(r[i] + np.float32(g * vnext)) - v[i])
        A = np.float32(td + np.float32(gl * A))
        adv[i] = A
        ret[i] = np.float32(A + v[i])
    return adv, ret


# --------------------------------------------------------------------------
# a13  running observation statistics
# --------------------------------------------------------------------------
def welford_new(O):
    """model.py:62-66."""
    return dict(n=torch.zeros(O), mean=torch.zeros(O), mean_diff=torch.zeros(O), var=torch.zeros(O))


def welford_observe_normalize(st, obs):
    """model.py:68-80 applied to every row of ``obs`` (T,O) in order, as the
    rollout loop does (train.py:84-85): observe, then normalise THAT row with
    the statistics that already include it.  Returns the normalised rows;
    ``st`` is updated in place."""
    obs = torch.as_tensor(obs, dtype=torch.float32)
    out = torch.empty_like(obs)
    n, mean, md = st["n"], st["mean"], st["mean_diff"]
    var = st["var"]
    for t in range(obs.shape[0]):
        x = obs[t]
        n += 1.0                                    # model.py:71
        last_mean = mean.clone()                    # model.py:72
        mean += (x - mean) / n                      # model.py:73
        md += (x - last_mean) * (x - mean)          # model.py:74
        var = torch.clamp(md / n, min=1e-2)         # model.py:75
        out[t] = torch.clamp((x - mean) / torch.sqrt(var), -5.0, 5.0)   # model.py:77-80
    st["var"] = var
    return out


def welford_normalize(st, x):
    """model.py:77-80 without an update."""
    return torch.clamp((x - st["mean"]) / torch.sqrt(st["var"]), -5.0, 5.0)


# --------------------------------------------------------------------------
# a7  DPPO (train.py) loss, a8 backward
# --------------------------------------------------------------------------
def normal_pdf(x, mu, std):
    """train.py:41-44 (``std`` is exp(log_std) used as a VARIANCE)."""
    a = (-1 * (x - mu).pow(2) / (2 * std)).exp()
    b = 1 / (2 * std * np.pi).sqrt()
    return a * b


def _leaf(p):
    return OrderedDict((k, t.detach().clone().requires_grad_(True)) for k, t in p.items())


def dppo_loss_grads(p, p_old, states, actions, returns, advantages, clip=0.2, ent_coeff=0.0):
    """train.py:140-166 on detached minibatch tensors.  Returns
    (dict of loss terms, OrderedDict of 13 gradients)."""
    q = _leaf(p)
    with torch.no_grad():
        mu_old, sig_old, v_old = forward(p_old, states)           # train.py:142
        probs_old = normal_pdf(actions, mu_old, sig_old)           # train.py:143
    A = actions.shape[1]
    mu, sig, v_pred = forward(q, states)                           # train.py:145
    probs = normal_pdf(actions, mu, sig)                           # train.py:146
    ratio = probs / (1e-10 + probs_old)                            # train.py:148
    adv = torch.cat([advantages] * A, 1)                           # train.py:150
    surr1 = ratio * adv
    surr2 = ratio.clamp(1 - clip, 1 + clip) * adv                  # train.py:151
    loss_clip = -torch.mean(torch.min(surr1, surr2))               # train.py:152
    vfloss1 = (v_pred - returns) ** 2                              # train.py:154
    v_clipped = v_old + (v_pred - v_old).clamp(-clip, clip)        # train.py:155
    vfloss2 = (v_clipped - returns) ** 2                           # train.py:156
    loss_value = 0.5 * torch.mean(torch.max(vfloss1, vfloss2))     # train.py:157
    loss_ent = -ent_coeff * torch.mean(probs * torch.log(probs + 1e-5))   # train.py:159
    total = loss_clip + loss_value + loss_ent                      # train.py:161
    total.backward()                                               # train.py:166
    grads = OrderedDict((k, q[k].grad.detach().clone()) for k in PARAM_ORDER)
    terms = dict(loss_clip=float(loss_clip.detach()), loss_value=float(loss_value.detach()),
                 loss_ent=float(loss_ent.detach()), total=float(total.detach()))
    return terms, grads


# --------------------------------------------------------------------------
# a6  ppo.py loss, a8 backward, a9 grad clip
# --------------------------------------------------------------------------
def ppo_loss_grads(p, states, actions, returns, advantages, logp_old, clip=0.2, ent_coeff=0.01):
    """ppo.py:147-172.  sigma_sq = exp(log_std) is used as the STD here."""
    q = _leaf(p)
    mu, sigma, v_pred = forward(q, states)                          # ppo.py:148
    log_std = q["log_std"]                                          # ppo.py:149
    logp = -0.5 * ((actions - mu) / sigma).pow(2) - 0.5 * math.log(2 * math.pi) - log_std   # :150
    logp = logp.sum(-1, keepdim=True)                               # ppo.py:151
    ent = (0.5 + 0.5 * math.log(2 * math.pi) + log_std).sum(-1).mean()   # ppo.py:152-153
    ratio = torch.exp(logp - logp_old)                              # ppo.py:156
    surr1 = ratio * advantages.expand_as(ratio)                     # ppo.py:159
    surr2 = ratio.clamp(1 - clip, 1 + clip) * advantages.expand_as(ratio)   # ppo.py:160
    loss_clip = -torch.mean(torch.min(surr1, surr2))                # ppo.py:161
    loss_value = (v_pred - returns).pow(2).mean()                   # ppo.py:164
    loss_ent = -ent_coeff * ent                                     # ppo.py:167
    total = loss_clip + loss_value + loss_ent                       # ppo.py:170
    total.backward()                                                # ppo.py:172
    grads = OrderedDict((k, q[k].grad.detach().clone()) for k in PARAM_ORDER)
    terms = dict(loss_clip=float(loss_clip.detach()), loss_value=float(loss_value.detach()),
                 loss_ent=float(loss_ent.detach()), total=float(total.detach()))
    return terms, grads


def clip_grad_norm(grads, max_norm):
    """ppo.py:173 -> torch.nn.utils.clip_grad_norm_ (torch 2.11): total L2
    norm = norm of the per-tensor norms; coef = max_norm/(total+1e-6) clamped
    to 1; every gradient multiplied by it.  Returns (clipped grads, norm)."""
    norms = torch.stack([torch.linalg.vector_norm(g, 2.0) for g in grads.values()])
    total = torch.linalg.vector_norm(norms, 2.0)
    coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
    out = OrderedDict((k, g * coef) for k, g in grads.items())
    return out, float(total)


# --------------------------------------------------------------------------
# a10  Adam  (torch 2.11 torch/optim/adam.py _single_tensor_adam, CPU path)
# --------------------------------------------------------------------------
def adam_step(p, g, m, v, step, lr, b1=0.9, b2=0.999, eps=1e-8):
    """One update, in place on p/m/v (dicts or flat tensors of equal layout).
    ``step`` is the 1-based update count.  Op-for-op the single-tensor path:
    lerp_, mul_/addcmul_, sqrt/div/add_, addcdiv_ with Python-double scalars."""
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    step_size = lr / bc1
    bc2_sqrt = bc2 ** 0.5
    items = p.keys() if isinstance(p, dict) else [None]
    for k in items:
        pk, gk, mk, vk = (p[k], g[k], m[k], v[k]) if k is not None else (p, g, m, v)
        mk.lerp_(gk, 1 - b1)
        vk.mul_(b2).addcmul_(gk, gk, value=1 - b2)
        denom = (vk.sqrt() / bc2_sqrt).add_(eps)
        pk.addcdiv_(mk, denom, value=-step_size)


# --------------------------------------------------------------------------
# a11/a12  whole update cycle  (worker epochs + shared sum + chief Adam)
# --------------------------------------------------------------------------
def zeros_like_params(p, fill=0.0):
    return OrderedDict((k, torch.full_like(t, fill)) for k, t in p.items())


def dppo_update_cycle(p, m, v, step0, workers, idx_streams, gshared, lr=3e-4, clip=0.2,
                      ent_coeff=0.0, num_epoch=10, record=None):
    """One outer iteration of DPPO (train.py:128-175 + chief.py:13-20) for W
    workers in lock-step.

    workers[w] = dict(states, actions, returns, advantages) SoA tensors;
    idx_streams[w][k] = minibatch positions of worker w in epoch k;
    gshared = the Shared_grad_buffers content (ones before the first update,
    model.py:51; zeros after every chief step, model.py:57-59).  Gradients are
    SUMMED into it in worker order (model.py:53-55) and fed to Adam unscaled
    (chief.py:15-17).  theta_old of epoch k is theta of epoch k-1
    (train.py:128-129,164).  p/m/v/gshared are updated in place; returns the
    step count after the cycle."""
    p_old = OrderedDict((k, t.clone()) for k, t in p.items())           # train.py:128-129
    step = step0
    for k in range(num_epoch):
        snap = OrderedDict((kk, t.clone()) for kk, t in p.items())      # model at train.py:135
        for w, buf in enumerate(workers):
            idx = torch.as_tensor(idx_streams[w][k], dtype=torch.long)
            terms, grads = dppo_loss_grads(
                snap, p_old, buf["states"][idx], buf["actions"][idx],
                buf["returns"][idx], buf["advantages"][idx], clip, ent_coeff)
            for kk in PARAM_ORDER:
                gshared[kk] += grads[kk]                                 # model.py:53-55
            if record is not None:
                record(k, w, terms, grads)
        p_old = snap                                                     # train.py:164
        step += 1
        adam_step(p, gshared, m, v, step, lr)                            # chief.py:15-17
        for kk in PARAM_ORDER:
            gshared[kk].fill_(0)                                         # chief.py:19
    return step


def ppo_update_cycle(p, m, v, step0, buf, idx_stream, lr=7e-4, clip=0.2, ent_coeff=0.01,
                     max_grad_norm=0.5, num_epoch=10, record=None):
    """ppo.py:139-174 for one horizon."""
    step = step0
    for k in range(num_epoch):
        idx = torch.as_tensor(idx_stream[k], dtype=torch.long)
        terms, grads = ppo_loss_grads(p, buf["states"][idx], buf["actions"][idx], buf["returns"][idx],
                                      buf["advantages"][idx], buf["logprobs"][idx], clip, ent_coeff)
        clipped, norm = clip_grad_norm(grads, max_grad_norm)
        if record is not None:
            record(k, terms, grads, clipped, norm)
        step += 1
        adam_step(p, clipped, m, v, step, lr)
    return step


# --------------------------------------------------------------------------
# flat <-> dict helpers (test convenience)
# --------------------------------------------------------------------------
def flatten(d):
    return torch.cat([d[k].reshape(-1) for k in PARAM_ORDER])


def unflatten(flat, shapes):
    out = OrderedDict()
    o = 0
    for k, s in shapes.items():
        n = int(np.prod(s))
        out[k] = flat[o:o + n].reshape(s).clone()
        o += n
    return out
