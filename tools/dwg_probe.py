"""The two-launch update (batch-level dW) against the single-launch form: run once with DPPO_K2_DWG=0 and once without,
compare the gradients (sum of the partial rows) and time both.    python tools/dwg_probe.py [B]"""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 2 and sys.argv[2] == "child":
    import torch
    sys.path.insert(0, os.path.join(ROOT, "pytorch-dppo_b200")); sys.path.insert(0, ROOT)
    import b200_native as nv
    from model import Model
    torch.manual_seed(1)
    B = int(sys.argv[1]); T = 65536
    m = Model(17, 6)
    obs = torch.randn(T, 17, device="cuda"); act = torch.randn(T, 6, device="cuda")
    adv = torch.randn(T, device="cuda"); ret = torch.randn(T, device="cuda")
    idx = torch.randperm(T, device="cuda")[:B].int()
    old = m.flat.clone() + 0.01 * torch.randn_like(m.flat)
    nv.update_prepare(m.dims, B)
    n = nv.update_nparts(m.dims, B)
    parts = torch.full((n, m.P_pad), float("nan"), device="cuda"); lp = torch.full((n, 4), float("nan"), device="cuda")
    for _ in range(3):
        nv.update_fused(0, m.dims, obs, act, adv, ret, None, idx, m.flat, old, 0.2, 0.0, parts, lp)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20):
            nv.update_fused(0, m.dims, obs, act, adv, ret, None, idx, m.flat, old, 0.2, 0.0, parts, lp)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g.replay(); torch.cuda.synchronize()
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    G = parts.double().sum(0)
    torch.save({"G": G.cpu(), "L": lp.double().sum(0).cpu()}, sys.argv[3])
    print(json.dumps({"nparts": n, "us": e0.elapsed_time(e1) * 1e3 / 20, "nan": bool(torch.isnan(parts).any())}))
    sys.exit(0)
import torch
B = sys.argv[1] if len(sys.argv) > 1 else "4096"
res = {}
for tag, env in (("two-launch", {}), ("first-only", {"DPPO_K2_DWG": "2"}), ("single", {"DPPO_K2_DWG": "0"})):
    out = subprocess.run([sys.executable, __file__, B, "child", "/tmp/dwg_%s.pt" % tag], env=dict(os.environ, **env),
                         capture_output=True, text=True)
    print(tag, out.stdout.strip().splitlines()[-1] if out.stdout.strip() else out.stderr[-2000:])
    res[tag] = torch.load("/tmp/dwg_%s.pt" % tag)
a, b = res["two-launch"], res["single"]
print("grad max|d| / max|ref| = %.3e   loss sums %s vs %s" % (float((a["G"] - b["G"]).abs().max() / b["G"].abs().max()),
                                                           a["L"].tolist(), b["L"].tolist()))
