"""K1 / K4 / K3 at sizes far beyond L2 (HBM study).  Usage: python tools/hbm_study.py [reps]"""
import sys, torch
sys.path.insert(0, '/root/repo/pytorch-dppo_b200'); sys.path.insert(0, '/root/repo')
import b200_native as nv
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
dev = 'cuda'
def timed(fn, reps):
    fn(); torch.cuda.synchronize(); tot = 0.0
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize(); tot += e0.elapsed_time(e1)
    return tot / reps * 1e3
T1 = 1 << 26
r = torch.rand(T1, device=dev); v = torch.rand(T1, device=dev); b = torch.zeros(T1, device=dev)
se = (torch.arange(T1, device=dev) % 1000 == 999).to(torch.uint8)
a = torch.empty(T1, device=dev); q = torch.empty(T1, device=dev)
scr = torch.empty(nv.gae_scratch_bytes(T1), dtype=torch.uint8, device=dev)
us = timed(lambda: nv.gae_scan(r, v, b, se, a, q, 0.99, 0.95, scr), reps)
print('K1 gae T=2^26: %.1f us  %.0f GB/s' % (us, 21 * T1 / us / 1e3))
del r, v, b, se, a, q
T4, O = 1 << 22, 17
x = torch.randn(T4, O, device=dev); y = torch.empty_like(x)
s4 = [torch.zeros(O, device=dev) for _ in range(4)]
scr = torch.empty(nv.welford_scratch_bytes(T4, O), dtype=torch.uint8, device=dev)
us = timed(lambda: nv.welford(x, y, s4[0], s4[1], s4[2], s4[3], scr), reps)
print('K4 welford T=2^22 O=17: %.1f us  %.0f GB/s' % (us, 8 * T4 * O / us / 1e3))
