"""K4 alone: time at the HBM-study size (2^22 x 17) and at C4 (65536 x 17), for ncu captures.
    python tools/k4_probe.py [T] [O]"""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-dppo_b200")); sys.path.insert(0, ROOT)
import b200_native as nv
T = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
O = int(sys.argv[2]) if len(sys.argv) > 2 else 17
x = torch.randn(T, O, device="cuda"); y = torch.empty_like(x)
st = [torch.zeros(O, device="cuda") for _ in range(4)]
scr = torch.empty(nv.welford_scratch_bytes(T, O), dtype=torch.uint8, device="cuda")
for rep in range(4):
    for t in st: t.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); nv.welford(x, y, st[0], st[1], st[2], st[3], scr); e1.record(); e1.synchronize()
    us = e0.elapsed_time(e1) * 1e3
    print("K4 T=%d O=%d: %.1f us, %.0f GB/s" % (T, O, us, 8.0 * T * O / us / 1e3))

if os.environ.get("DPPO_LIB", "").endswith("k4timing.so"):
    import ctypes
    lib = ctypes.CDLL(os.environ["DPPO_LIB"])
    buf = (ctypes.c_ulonglong * 8)()
    lib.dppo_welford_phase_ns(buf, 1)
    for t in st: t.zero_()
    nv.welford(x, y, st[0], st[1], st[2], st[3], scr); torch.cuda.synchronize()
    lib.dppo_welford_phase_ns(buf, 0)
    ntiles = -(-T // 492) if O == 17 else None
    names = ["ticket", "load", "pass A", "publish", "look-back", "pass B", "store"]
    tot = sum(buf[:7])
    print("phase ns per tile (thread 0), %s tiles:" % ntiles, {k: round(buf[i] / (ntiles or 1)) for i, k in enumerate(names)}, "sum", round(tot / (ntiles or 1)))
