"""Build an A/B variant of libdppo_b200.so: dppo_update.cu recompiled with extra nvcc flags, every other object taken
from the normal build.    python tools/build_variant.py early1 -DDPPO_K2_EARLY_RELEASE=1
Another source file:      python tools/build_variant.py --file dppo_welford.cu k4timing -DDPPO_K4_TIMING
The variant lands in tools/variants/libdppo_b200_<name>.so (git-ignored); run anything with DPPO_LIB=<that path>."""
import glob, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pytorch-dppo_b200")
sys.path.insert(0, PKG)
import build as b  # noqa: E402
args = sys.argv[1:]
src = "dppo_update.cu"
if args[0] == "--file":
    src, args = args[1], args[2:]
stem = src[:-3]
name, flags = args[0], args[1:]
b.build()
out = os.path.join(ROOT, "tools", "variants")
os.makedirs(out, exist_ok=True)
obj = os.path.join(out, "%s_%s.o" % (stem, name))
subprocess.check_call([b.NVCC] + [f for f in b.FLAGS if f not in ("-Xptxas", "-v")] + flags +
                      ["-c", os.path.join(b.CSRC, src), "-o", obj])
objs = [o for o in glob.glob(os.path.join(PKG, "build", "*.o")) if not o.endswith(stem + ".o")] + [obj]
lib = os.path.join(out, "libdppo_b200_%s.so" % name)
subprocess.check_call([b.NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib] + objs)
print(lib)
