"""One small K2 launch (reference nets, tensor-core path) for compute-sanitizer:
    compute-sanitizer --tool racecheck|synccheck|memcheck python tools/k2_sanitize.py [B]
Checks the result against the FP32-pipe path so that a tool-induced timing change that breaks the pipeline shows."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-dppo_b200")); sys.path.insert(0, ROOT)
import b200_native as nv
B = int(sys.argv[1]) if len(sys.argv) > 1 else 96
O, A, T = 17, 6, 512
dims = nv.Dims(O, A, 100, 100, 500, 100)
off, size, P = nv.param_layout(dims)
g = torch.Generator().manual_seed(0)
theta = (torch.randn(P, generator=g) * 0.1).cuda(); old = (theta.cpu() + 0.01 * torch.randn(P, generator=g)).cuda()
obs = torch.randn(T, O, generator=g).cuda(); act = torch.randn(T, A, generator=g).cuda()
adv = torch.randn(T, generator=g).cuda(); ret = torch.randn(T, generator=g).cuda()
idx = torch.randperm(T, generator=g)[:B].to(torch.int32).cuda()
out = []
for path in (nv.K2_PATH_AUTO, nv.K2_PATH_FP32):
    nv.update_set_path(path)
    n = nv.update_nparts(dims, B)
    parts = torch.zeros(n, P, device="cuda"); lp = torch.zeros(n, 4, device="cuda")
    nv.update_fused(nv.FLAVOUR_DPPO, dims, obs, act, adv, ret, None, idx, theta, old, 0.2, 0.0, parts, lp)
    torch.cuda.synchronize()
    out.append(parts.double().sum(0))
err = float((out[0] - out[1]).abs().max() / out[1].abs().max())
print("k2_sanitize: B=%d nparts=%d tensor-vs-fp32 rel %.2e" % (B, n, err))
assert err < 1e-5
