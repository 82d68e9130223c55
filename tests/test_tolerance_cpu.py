"""Where the element-wise tolerance of SURVEY 7.3-2 (|d_i| <= 1e-5 |ref_i| + 1e-8 * max|ref|) stops being
meaningful: the REFERENCE'S OWN fp32 gradients do not meet it against themselves when only the number of
MKL threads changes (a different summation order), nor against their fp64 evaluation.  Small entries of a
gradient tensor carry the absolute rounding error of a 4096-term reduction.  The GPU tests therefore assert
the per-tensor bound (1e-5 * max|ref|) and an element-wise bound with atol = 5e-6 * max|ref| (util.py); the floor
measured here is what that constant is set against."""
import torch

from oracle import ppo_oracle as orc
from util import ELEM_ATOL_REDUCTION, ELEM_ATOL_SURVEY, closeness, synth_batch


def _case(B=4096, O=17, A=6):
    torch.manual_seed(1)
    p = orc.init_params(O, A)
    po = {k: v + 0.01 * torch.randn_like(v) for k, v in p.items()}
    obs, act, adv, ret = synth_batch(O, A, B, 3)
    with torch.no_grad():
        mu, s, _ = orc.forward(p, obs)
        act = mu + s.sqrt() * torch.randn(B, A)
    return p, po, obs, act, ret, adv


def test_elementwise_floor_of_the_reference():
    p, po, obs, act, ret, adv = _case()
    threads = torch.get_num_threads()
    try:
        torch.set_num_threads(max(2, threads))
        _, g_many = orc.dppo_loss_grads(p, po, obs, act, ret, adv)
        torch.set_num_threads(1)
        _, g_one = orc.dppo_loss_grads(p, po, obs, act, ret, adv)
    finally:
        torch.set_num_threads(threads)
    d = lambda x: {k: v.double() for k, v in x.items()}
    _, g64 = orc.dppo_loss_grads(d(p), d(po), obs.double(), act.double(), ret.double(), adv.double())
    worst_self, worst_64 = 0.0, 0.0
    for k in orc.PARAM_ORDER:
        c64 = closeness(g_one[k], g64[k])
        assert c64["rel"] < 1e-5                                  # the per-tensor bound holds with a wide margin
        assert c64["elem_atol_needed"] < ELEM_ATOL_REDUCTION
        worst_64 = max(worst_64, c64["elem_atol_needed"])
        worst_self = max(worst_self, closeness(g_many[k], g_one[k])["elem_atol_needed"])
    # fp32 evaluation vs exact: the weight gradients need more than the survey's 1e-8
    assert worst_64 > ELEM_ATOL_SURVEY
    # (thread-count dependence is machine dependent: with a single core both runs are identical)
    assert worst_self < ELEM_ATOL_REDUCTION
