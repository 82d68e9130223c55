"""Which kernel is not bit-repeatable?  Each piece of the step is run many times back to back on identical inputs and
every result is compared bit for bit with the first:  make_rollout (torch eager, twice), K4 alone, K1 alone, K2 alone
(both layer paths, launches back to back inside one CUDA graph), K3 alone.    python tools/repeat_probe.py"""
import os, random, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-dppo_b200")); sys.path.insert(0, ROOT)
import b200_native as nv
from b200_engine import UpdateEngine
from model import Model, Shared_obs_stats
from synthetic import make_rollout
DEV = "cuda"
R = int(os.environ.get("PROBE_REPS", "200"))

def probe(widths, O, A, T, B):
    tag = "%s O=%d T=%d B=%d" % (widths, O, T, B)
    torch.manual_seed(1)
    m = Model(O, A, *widths, device=DEV)
    stats = Shared_obs_stats(O)
    h1, d1 = make_rollout(m, stats, T, seed=7, flavour=nv.FLAVOUR_DPPO)
    h2, d2 = make_rollout(m, stats, T, seed=7, flavour=nv.FLAVOUR_DPPO)
    same = all(torch.equal(getattr(h1, f), getattr(h2, f)) for f in ("raw_obs", "obs", "act", "val", "rew", "boot", "logp"))
    print("%s | make_rollout twice identical: %s" % (tag, same))
    # K4 alone
    outs = []
    for r in range(20):
        st = [torch.zeros(O, device=DEV) for _ in range(4)]
        out = torch.empty_like(d1.raw_obs)
        nv.welford(d1.raw_obs, out, st[0], st[1], st[2], st[3], torch.empty(nv.welford_scratch_bytes(T, O), dtype=torch.uint8, device=DEV))
        outs.append((out, torch.stack(st)))
    torch.cuda.synchronize()
    print("%s | K4 20x identical: %s" % (tag, all(torch.equal(o[0], outs[0][0]) and torch.equal(o[1], outs[0][1]) for o in outs)))
    eng = UpdateEngine(m, nv.FLAVOUR_DPPO, B, 1, 3e-4)
    eng.gae(d1)
    a0 = d1.adv.clone()
    ok = True
    for r in range(20):
        eng.gae(d1); ok = ok and torch.equal(d1.adv, a0)
    print("%s | K1 20x identical: %s" % (tag, ok))
    eng.draw_indices(T, random.Random(3))
    eng.idx_dev.copy_(eng.idx_host)
    old = (m.flat + 0.01 * torch.randn_like(m.flat)).contiguous()
    for i, (o, s) in enumerate(zip(m.offsets, m.sizes)):                     # keep the padding floats zero
        end = m.offsets[i + 1] if i + 1 < len(m.offsets) else m.P_pad
        old[o + s:end] = 0
    for path, name in ((nv.K2_PATH_AUTO, "tensor"), (nv.K2_PATH_FP32, "fp32")):
        nv.update_set_path(path)
        log = torch.zeros(R, eng.nparts, eng.P_pad, device=DEV)
        def k2():
            nv.update_fused(nv.FLAVOUR_DPPO, eng.dims, d1.obs, d1.act, d1.adv, d1.ret, None, eng.idx_dev[0], m.flat, old,
                            0.2, 0.0, eng.partials, eng.loss_partials)
        k2(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for r in range(R):
                k2()
                log[r].copy_(eng.partials)
        bad_total = 0
        for rep in range(3):
            g.replay(); torch.cuda.synchronize()
            bad = [r for r in range(R) if not torch.equal(log[r], log[0])]
            bad_total += len(bad)
            if bad:
                r = bad[0]
                diff = (log[r] != log[0])
                cta = diff.any(1).nonzero().reshape(-1).tolist()[:8]
                pos = diff.any(0).nonzero().reshape(-1)
                which = sorted(set(int(sum(1 for o in m.offsets if o <= int(p)) - 1) for p in pos[:2000].tolist()))
                print("%s | K2 %s replay %d: %d of %d launches differ; first %d: CTAs %s tensors %s maxdiff %.3e" % (
                    tag, name, rep, len(bad), R, r, cta, [nv.PARAM_NAMES[w] for w in which], float((log[r] - log[0]).abs().max())))
        print("%s | K2 %s %dx3 identical: %s" % (tag, name, R, bad_total == 0))
        del log
    nv.update_set_path(nv.K2_PATH_AUTO)
    # K3 alone on fixed partials
    res = []
    for r in range(20):
        th, mm, vv = m.flat.clone(), torch.zeros_like(m.flat), torch.zeros_like(m.flat)
        nv.reduce_adam(eng.partials, eng.nparts, None, th, None, mm, vv, 1, 3e-4)
        res.append(th)
    torch.cuda.synchronize()
    print("%s | K3 20x identical: %s" % (tag, all(torch.equal(t, res[0]) for t in res)))

for cfg in (((100, 100, 500, 100), 17, 6, 65536, 4096), ((64, 64, 64, 64), 17, 6, 8192, 1024),
            ((128, 128, 128, 128), 17, 6, 8192, 1024), ((96, 96, 160, 96), 11, 3, 8192, 1024)):
    probe(*cfg)
