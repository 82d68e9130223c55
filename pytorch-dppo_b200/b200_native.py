"""ctypes binding of libdppo_b200.so (include/dppo_b200.h).

The product path goes through here and nowhere else: there is no CPU or
PyTorch fallback.  If the library is missing it is built once with nvcc
(build.py); if that is impossible the import fails loudly."""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# DPPO_LIB: load another build of the same library (A/B runs of kernel variants); default is the in-tree build
LIB_PATH = os.environ.get("DPPO_LIB") or os.path.join(_HERE, "libdppo_b200.so")

FLAVOUR_DPPO = 0
FLAVOUR_PPO = 1
N_PARAMS = 13

PARAM_NAMES = (
    "log_std",
    "p_fc1.weight", "p_fc1.bias", "p_fc2.weight", "p_fc2.bias",
    "v_fc1.weight", "v_fc1.bias", "v_fc2.weight", "v_fc2.bias",
    "mu.weight", "mu.bias", "v.weight", "v.bias",
)


class Dims(ctypes.Structure):
    """dppo_dims: policy O->H1->H2->A, value O->G1->G2->1 (model.py:11-22)."""
    _fields_ = [("O", ctypes.c_int), ("A", ctypes.c_int), ("H1", ctypes.c_int), ("H2", ctypes.c_int),
                ("G1", ctypes.c_int), ("G2", ctypes.c_int)]

    def shapes(self):
        return {
            "log_std": (1, self.A),
            "p_fc1.weight": (self.H1, self.O), "p_fc1.bias": (self.H1,),
            "p_fc2.weight": (self.H2, self.H1), "p_fc2.bias": (self.H2,),
            "v_fc1.weight": (self.G1, self.O), "v_fc1.bias": (self.G1,),
            "v_fc2.weight": (self.G2, self.G1), "v_fc2.bias": (self.G2,),
            "mu.weight": (self.A, self.H2), "mu.bias": (self.A,),
            "v.weight": (1, self.G2), "v.bias": (1,),
        }


def _load():
    if not os.path.exists(LIB_PATH):
        import importlib.util
        spec = importlib.util.spec_from_file_location("_dppo_build", os.path.join(_HERE, "build.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.build()
    lib = ctypes.CDLL(LIB_PATH)
    c = ctypes
    vp, i32, i64, dbl, flt, sz = c.c_void_p, c.c_int, c.c_int64, c.c_double, c.c_float, c.c_size_t
    dp = c.POINTER(Dims)
    sig = {
        "dppo_version": (i32, []),
        "dppo_error_string": (c.c_char_p, [i32]),
        "dppo_param_layout": (i32, [dp, c.POINTER(i32), c.POINTER(i32)]),
        "dppo_gae_scratch_bytes": (sz, [i64]),
        "dppo_gae_scan": (i32, [vp, vp, vp, vp, vp, vp, i64, dbl, dbl, vp, sz, vp]),
        "dppo_gae_scan_noalloc": (i32, [vp, vp, vp, vp, vp, vp, i64, dbl, dbl, vp]),
        "dppo_welford_scratch_bytes": (sz, [i64, i32]),
        "dppo_welford": (i32, [vp, vp, i64, i32, vp, vp, vp, vp, vp, sz, vp]),
        "dppo_welford_totals_scratch_bytes": (sz, [i64, i32]),
        "dppo_welford_totals": (i32, [vp, i64, i32, vp, vp, vp, sz, vp]),
        "dppo_welford_shared": (i32, [vp, vp, i64, i32, vp, vp, vp, vp, vp, i32, i32, vp, sz, vp]),
        "dppo_comm_allgather": (i32, [vp, vp, vp, i32, vp]),
        "dppo_comm_allgather_loopback": (i32, [c.POINTER(vp), i32, c.POINTER(vp), c.POINTER(vp), i32, vp]),
        "dppo_update_prepare": (i32, [dp, i32]),
        "dppo_update_launches": (i32, [dp, i32]),
        "dppo_update_set_batch_dw": (i32, [i32]),
        "dppo_update_nparts": (i32, [dp, i32]),
        "dppo_update_partials_bytes": (sz, [dp, i32]),
        "dppo_update_fused": (i32, [i32, dp, vp, vp, vp, vp, vp, vp, i32, vp, vp, flt, flt, vp, vp, vp]),
        "dppo_policy_forward": (i32, [dp, vp, i32, vp, vp, i32, vp, vp, vp, vp, vp, i64, i64, vp]),
        "dppo_reduce_adam_scratch_bytes": (sz, [i32]),
        "dppo_comm_create": (i32, [i32, i32, i32, c.POINTER(vp)]),
        "dppo_comm_handle_bytes": (i32, []),
        "dppo_comm_local_handle": (i32, [vp, vp]),
        "dppo_comm_open_peers": (i32, [vp, vp]),
        "dppo_comm_create_loopback": (i32, [i32, i32, c.POINTER(vp)]),
        "dppo_comm_destroy": (i32, [vp]),
        "dppo_comm_check": (i32, [vp]),
        "dppo_comm_set_algorithm": (i32, [vp, i32]),
        "dppo_comm_get_algorithm": (i32, [vp]),
        "dppo_update_set_path": (i32, [i32]),
        "dppo_allreduce_adam": (i32, [vp, vp, i32, vp, vp, vp, vp, vp, i32, i32, dbl, dbl, dbl, dbl, vp, vp, vp, vp, vp]),
        "dppo_allreduce_adam_loopback": (i32, [c.POINTER(vp), i32, c.POINTER(vp), c.POINTER(i32), c.POINTER(vp),
                                               c.POINTER(vp), c.POINTER(vp), c.POINTER(vp), c.POINTER(vp), i32, i32,
                                               dbl, dbl, dbl, dbl, c.POINTER(vp), vp]),
        "dppo_reduce_partials": (i32, [vp, i32, vp, i32, vp]),
        "dppo_sample_indices_host": (i32, [c.POINTER(c.c_uint32), c.POINTER(i32), i64, i32, i32, c.POINTER(c.c_int32)]),
        "dppo_sample_scratch_bytes": (sz, [i64, i32]),
        "dppo_sample_indices": (i32, [vp, vp, i64, i32, i32, vp, vp, sz, vp]),
        "dppo_fp32_probe": (i32, [i32, vp, c.POINTER(dbl), vp]),
        "dppo_adam_scalars_host": (i32, [i32, dbl, dbl, dbl, dbl, c.POINTER(flt)]),
        "dppo_reduce_adam": (i32, [vp, i32, vp, vp, vp, vp, vp, i32, i32, dbl, dbl, dbl, dbl, vp, dbl, vp, vp, vp, vp, sz, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)            # AttributeError = the library is stale: rebuild
        fn.restype = res
        fn.argtypes = args
    return lib, sig


lib, SIGNATURES = _load()


class DppoError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise DppoError("libdppo_b200: %s (code %d)" % (lib.dppo_error_string(rc).decode(), rc))


def _ptr(t, dtype=torch.float32, allow_none=False):
    if t is None:
        if allow_none:
            return None
        raise DppoError("libdppo_b200: missing tensor")
    if not t.is_cuda:
        raise DppoError("libdppo_b200: tensor is not on a CUDA device - there is no CPU path")
    if t.dtype != dtype or not t.is_contiguous():
        raise DppoError("libdppo_b200: expected a contiguous %s tensor, got %s" % (dtype, t.dtype))
    return t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def param_layout(dims):
    off = (ctypes.c_int * N_PARAMS)()
    size = (ctypes.c_int * N_PARAMS)()
    p_pad = lib.dppo_param_layout(ctypes.byref(dims), off, size)
    if p_pad < 0:
        check(p_pad)
    return list(off), list(size), p_pad


def gae_scratch_bytes(T):
    return lib.dppo_gae_scratch_bytes(T)


def gae_scan(rewards, values, boot, seg_end, adv, ret, gamma, lam, scratch):
    T = rewards.numel()
    check(lib.dppo_gae_scan(_ptr(rewards), _ptr(values), _ptr(boot), _ptr(seg_end, torch.uint8), _ptr(adv), _ptr(ret),
                            T, gamma, lam, _ptr(scratch, torch.uint8), scratch.numel(), _stream()))


GAE_NOALLOC_MAX_T = 4194304


def gae_scan_noalloc(rewards, values, boot, seg_end, adv, ret, gamma, lam):
    """K1 without caller scratch, for users who cannot allocate device memory at call time
    (the look-back chain sits in an array built into the library).  At most GAE_NOALLOC_MAX_T
    steps; calls on one device must not overlap; first call per device outside stream capture."""
    check(lib.dppo_gae_scan_noalloc(_ptr(rewards), _ptr(values), _ptr(boot), _ptr(seg_end, torch.uint8), _ptr(adv),
                                    _ptr(ret), rewards.numel(), gamma, lam, _stream()))


def welford_scratch_bytes(T, O):
    return lib.dppo_welford_scratch_bytes(T, O)


def welford(obs, out, n, mean, mean_diff, var, scratch):
    T, O = obs.shape
    check(lib.dppo_welford(_ptr(obs), _ptr(out), T, O, _ptr(n), _ptr(mean), _ptr(mean_diff), _ptr(var),
                           _ptr(scratch, torch.uint8), scratch.numel(), _stream()))


def welford_totals_scratch_bytes(T, O):
    return lib.dppo_welford_totals_scratch_bytes(T, O)


def welford_totals(obs, mean, totals, scratch):
    """This worker's (rows, S1[O], S2[O]) relative to ``mean`` into ``totals`` (1+2*O doubles)."""
    T, O = obs.shape
    check(lib.dppo_welford_totals(_ptr(obs), T, O, _ptr(mean), _ptr(totals, torch.float64),
                                  _ptr(scratch, torch.uint8), scratch.numel(), _stream()))


def welford_shared(obs, out, n, mean, mean_diff, var, totals_all, rank, scratch):
    """K4 with statistics shared by the workers in rank order; ``totals_all`` is (W, 1+2*O) float64."""
    T, O = obs.shape
    check(lib.dppo_welford_shared(_ptr(obs), _ptr(out), T, O, _ptr(n), _ptr(mean), _ptr(mean_diff), _ptr(var),
                                  _ptr(totals_all, torch.float64), rank, totals_all.shape[0],
                                  _ptr(scratch, torch.uint8), scratch.numel(), _stream()))


def comm_allgather(comm, src, dst):
    check(lib.dppo_comm_allgather(comm, _ptr(src, torch.float64), _ptr(dst, torch.float64), src.numel(), _stream()))


def comm_allgather_loopback(comms, srcs, dsts):
    W = len(comms)
    vp = ctypes.c_void_p
    cc = (vp * W)(*[c.value for c in comms])
    s = (vp * W)(*[_ptr(t, torch.float64) for t in srcs])
    d = (vp * W)(*[_ptr(t, torch.float64) for t in dsts])
    check(lib.dppo_comm_allgather_loopback(cc, W, s, d, srcs[0].numel(), _stream()))


def update_prepare(dims, B):
    """Allocate the library's per-device buffer of the two-launch update (large minibatches); call before a graph capture."""
    check(lib.dppo_update_prepare(ctypes.byref(dims), B))


def update_set_batch_dw(on):
    """Two-launch update with the batch-level dW pass (default for large minibatches) on/off; returns the previous
    setting.  Changes update_nparts: set it before partial buffers are sized."""
    return lib.dppo_update_set_batch_dw(1 if on else 0)


def update_launches(dims, B):
    """Kernels one update_fused call launches (2 for the large-minibatch form with the batch-level dW pass)."""
    n = lib.dppo_update_launches(ctypes.byref(dims), B)
    if n < 0:
        check(n)
    return n


def update_nparts(dims, B):
    n = lib.dppo_update_nparts(ctypes.byref(dims), B)
    if n < 0:
        check(n)
    return n


def update_fused(flavour, dims, obs, act, adv, ret, logp_old, idx, theta, theta_old, clip, ent_coeff, partials,
                 loss_partials):
    check(lib.dppo_update_fused(flavour, ctypes.byref(dims), _ptr(obs), _ptr(act), _ptr(adv), _ptr(ret),
                                _ptr(logp_old, allow_none=True), _ptr(idx, torch.int32), idx.numel(), _ptr(theta),
                                _ptr(theta_old, allow_none=True), clip, ent_coeff, _ptr(partials), _ptr(loss_partials),
                                _stream()))


def policy_forward(dims, theta, states, eps=None, std_is_variance=True, mu=None, v=None, act=None, logp=None,
                   states_out=None, row0=0, row_stride=1):
    """Batched Model.forward + action sampling of N states (train.py:84-90 / ppo.py:90-98); batch row r goes to row
    ``row0 + r*row_stride`` of every output given."""
    N = states.shape[0]
    check(lib.dppo_policy_forward(ctypes.byref(dims), _ptr(states), N, _ptr(theta), _ptr(eps, allow_none=True),
                                  1 if std_is_variance else 0, _ptr(mu, allow_none=True), _ptr(v, allow_none=True),
                                  _ptr(act, allow_none=True), _ptr(logp, allow_none=True),
                                  _ptr(states_out, allow_none=True), row0, row_stride, _stream()))


K2_PATH_AUTO, K2_PATH_FP32 = 0, 1


def update_set_path(path):
    """Layer routines of K2: K2_PATH_AUTO (tensor cores where eligible) or K2_PATH_FP32.  Returns the previous setting."""
    prev = lib.dppo_update_set_path(path)
    if prev < 0:
        check(prev)
    return prev


def reduce_adam_scratch_bytes(P_pad):
    return lib.dppo_reduce_adam_scratch_bytes(P_pad)


def adam_scalars(step, lr, beta1=0.9, beta2=0.999, eps=1e-8):
    """The six fp32 scalars of one Adam update (host side of scalars_dev)."""
    out = (ctypes.c_float * 6)()
    check(lib.dppo_adam_scalars_host(step, lr, beta1, beta2, eps, out))
    return list(out)


def reduce_adam(partials, nparts, gshared, theta, theta_prev, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-8,
                max_grad_norm=0.0, grad_out=None, loss_partials=None, loss_out=None, scratch=None, scalars_dev=None):
    check(lib.dppo_reduce_adam(_ptr(partials, allow_none=True), nparts, _ptr(gshared, allow_none=True), _ptr(theta),
                               _ptr(theta_prev, allow_none=True), _ptr(m), _ptr(v), theta.numel(), step, lr, beta1,
                               beta2, eps, _ptr(scalars_dev, allow_none=True), max_grad_norm,
                               _ptr(grad_out, allow_none=True),
                               _ptr(loss_partials, allow_none=True), _ptr(loss_out, allow_none=True),
                               _ptr(scratch, torch.uint8, allow_none=True), 0 if scratch is None else scratch.numel(),
                               _stream()))


def sample_indices(n, k, rounds=1, rng=None, out=None):
    """``rounds`` x ``random.sample(range(n), k)`` as an int32 (rounds,k) CPU tensor,
    bit-identical to Python's and leaving ``rng`` (default: the ``random`` module's
    generator) in the state Python's own calls would have left it in."""
    import random as _random
    rng = _random if rng is None else rng
    ver, internal, gauss = rng.getstate()
    state = (ctypes.c_uint32 * 624)(*internal[:624])
    index = ctypes.c_int(internal[624])
    if out is None:
        out = torch.empty(rounds, k, dtype=torch.int32)
    if out.dtype != torch.int32 or not out.is_contiguous() or out.numel() < rounds * k or out.is_cuda:
        raise DppoError("sample_indices: out must be a contiguous int32 CPU tensor of rounds*k elements")
    if rounds * k == 0:
        if k > n or n < 1:
            raise DppoError("sample_indices: need 1 <= n and k <= n")
        return out                                  # random.sample(pop, 0) draws nothing
    check(lib.dppo_sample_indices_host(state, ctypes.byref(index), n, k, rounds,
                                       ctypes.cast(out.data_ptr(), ctypes.POINTER(ctypes.c_int32))))
    rng.setstate((ver, tuple(state) + (index.value,), gauss))
    return out


def sample_uses_pool(n, k):
    """True when CPython's random.sample(population of n, k) takes its pool branch (n <= setsize, Lib/random.py)."""
    import math
    setsize = 21
    if k > 5:
        setsize += 4 ** math.ceil(math.log(k * 3, 4))
    return n <= setsize


class DeviceSampler:
    """``random.sample`` on the GPU: the generator state of a Python ``random.Random`` (or of the ``random`` module)
    is uploaded once; every ``sample`` call writes ``rounds`` x k positions into a device tensor and advances the
    state on the device exactly as Python would have.  ``sync_to(rng)`` stores the state back into a Python generator."""

    def __init__(self, rng=None, device=None):
        import random as _random
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.state = torch.zeros(624, dtype=torch.int32, device=self.device)     # uint32 words
        self.index = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._scratch = None
        self.seed_from(_random if rng is None else rng)

    def seed_from(self, rng):
        ver, internal, gauss = rng.getstate()
        words = torch.tensor([w - (1 << 32) if w >= (1 << 31) else w for w in internal[:624]], dtype=torch.int32)
        self.state.copy_(words)
        self.index.fill_(internal[624])

    def sync_to(self, rng):
        ver, internal, gauss = rng.getstate()
        words = [int(w) & 0xffffffff for w in self.state.cpu().tolist()]
        rng.setstate((ver, tuple(words) + (int(self.index.item()),), gauss))

    def sample(self, n, k, rounds=1, out=None):
        if out is None:
            out = torch.empty(rounds, k, dtype=torch.int32, device=self.device)
        if out.dtype != torch.int32 or not out.is_contiguous() or out.numel() < rounds * k or not out.is_cuda:
            raise DppoError("DeviceSampler.sample: out must be a contiguous int32 CUDA tensor of rounds*k elements")
        if k > n or n < 1:
            raise DppoError("DeviceSampler.sample: need 1 <= n and k <= n")
        need = lib.dppo_sample_scratch_bytes(n, k)
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
        check(lib.dppo_sample_indices(self.state.data_ptr(), self.index.data_ptr(), n, k, rounds, out.data_ptr(),
                                      self._scratch.data_ptr(), self._scratch.numel(), _stream()))
        return out


def fp32_probe(iters=4096):
    """Measured FP32 FMA throughput of this GPU in TFLOP/s (best of 5)."""
    out = torch.zeros(4, device="cuda")
    flops = ctypes.c_double(0)
    best = 0.0
    for _ in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(lib.dppo_fp32_probe(iters, _ptr(out), ctypes.byref(flops), _stream()))
        e1.record()
        e1.synchronize()
        best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best


# ---- multi-GPU exchange (csrc/dppo_comm.cu) -----------------------------------------------------
def comm_create(rank, world, P_pad):
    h = ctypes.c_void_p()
    check(lib.dppo_comm_create(rank, world, P_pad, ctypes.byref(h)))
    return h


def comm_local_handle(comm):
    buf = ctypes.create_string_buffer(lib.dppo_comm_handle_bytes())
    check(lib.dppo_comm_local_handle(comm, buf))
    return bytes(buf.raw)


def comm_open_peers(comm, handles):
    blob = b"".join(handles)
    check(lib.dppo_comm_open_peers(comm, ctypes.c_char_p(blob)))


def comm_destroy(comm):
    lib.dppo_comm_destroy(comm)


COMM_ALGO_PUSH_ALL, COMM_ALGO_OWNER = 1, 2


def comm_set_algorithm(comm, algo):
    check(lib.dppo_comm_set_algorithm(comm, algo))


def comm_get_algorithm(comm):
    return lib.dppo_comm_get_algorithm(comm)


def comm_check(comm):
    """Raises DppoError once a poll of this communicator has timed out (a peer never pushed)."""
    check(lib.dppo_comm_check(comm))


def allreduce_adam(comm, partials, nparts, gshared, theta, theta_prev, m, v, step, lr, beta1=0.9, beta2=0.999,
                   eps=1e-8, grad_out=None, loss_partials=None, loss_out=None, scalars_dev=None):
    check(lib.dppo_allreduce_adam(comm, _ptr(partials, allow_none=True), nparts, _ptr(gshared, allow_none=True),
                                  _ptr(theta), _ptr(theta_prev, allow_none=True), _ptr(m), _ptr(v), theta.numel(), step,
                                  lr, beta1, beta2, eps, _ptr(scalars_dev, allow_none=True),
                                  _ptr(grad_out, allow_none=True), _ptr(loss_partials, allow_none=True),
                                  _ptr(loss_out, allow_none=True), _stream()))


def comm_create_loopback(world, P_pad):
    arr = (ctypes.c_void_p * world)()
    check(lib.dppo_comm_create_loopback(world, P_pad, arr))
    return [ctypes.c_void_p(arr[r]) for r in range(world)]


def allreduce_adam_loopback(comms, partials, gshared, theta, theta_prev, m, v, step, lr, grad_out=None,
                            beta1=0.9, beta2=0.999, eps=1e-8):
    """W virtual ranks on one GPU (test helper): every argument is a list indexed by rank."""
    W = len(comms)
    vp = ctypes.c_void_p

    def arr(ts):
        if ts is None:
            return None
        return (vp * W)(*[_ptr(t, allow_none=True) for t in ts])

    cc = (vp * W)(*[c.value for c in comms])
    np_ = (ctypes.c_int * W)(*[p.shape[0] for p in partials])
    check(lib.dppo_allreduce_adam_loopback(cc, W, arr(partials), np_, arr(gshared), arr(theta), arr(theta_prev), arr(m),
                                           arr(v), theta[0].numel(), step, lr, beta1, beta2, eps, arr(grad_out),
                                           _stream()))


def reduce_partials(partials, nparts, out):
    check(lib.dppo_reduce_partials(_ptr(partials), nparts, _ptr(out), out.numel(), _stream()))
