"""N2: the rollout's batched forward (dppo_policy_forward through Model.act) - microseconds per environment step for N
environments stepped together, against the reference's way (one torch-eager model(state) call per environment on the
same GPU).    python tools/rollout_probe.py"""
import os, sys, time, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-dppo_b200")); sys.path.insert(0, ROOT)
from model import Model
import torch.nn as nn
torch.manual_seed(0)
m = Model(17, 6)
class Ref(nn.Module):                               # the reference's layers (model.py:11-45) in plain torch, for the timing only
    def __init__(s):
        super().__init__()
        s.p1, s.p2, s.v1, s.v2, s.mu, s.v = nn.Linear(17, 100), nn.Linear(100, 100), nn.Linear(17, 500), nn.Linear(500, 100), nn.Linear(100, 6), nn.Linear(100, 1)
        s.ls = nn.Parameter(torch.zeros(6))
    def forward(s, x):
        h = torch.tanh(s.p2(torch.tanh(s.p1(x)))); g = torch.tanh(s.v2(torch.tanh(s.v1(x))))
        return s.mu(h), torch.exp(s.ls).expand(x.shape[0], 6), s.v(g)
ref = Ref().cuda()
for N in (1, 64, 1024, 8192):
    st = torch.randn(N, 17, device="cuda"); eps = torch.randn(N, 6, device="cuda")
    mu = torch.empty(N, 6, device="cuda"); act = torch.empty(N, 6, device="cuda"); v = torch.empty(N, device="cuda"); lp = torch.empty(N, device="cuda")
    for _ in range(5): m.act(st, eps, True, mu, v, act, lp)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20): m.act(st, eps, True, mu, v, act, lp)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20
    with torch.no_grad():
        for _ in range(3): ref(st)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(20):
            mu_r, sg, v_r = ref(st); a_r = mu_r + sg.sqrt() * eps
        torch.cuda.synchronize(); us_ref = (time.perf_counter() - t0) * 1e6 / 20
    print("N = %5d environments: dppo_policy_forward %.1f us per step (%.1f M states/s)   torch eager on the same GPU %.0f us" % (N, us, N / us, us_ref))
