#!/usr/bin/env python
"""Benchmark of the data-parallel PPO update path (BASELINE.json metric:
"PPO update samples/sec at 1/2/4/8 B200; % of HBM roofline vs host-CPU reference").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (config.workload): BASELINE configs[3], the large-batch sweep the metric is
quoted on - one worker per GPU, HalfCheetah shape (obs 17, act 6), reference network
(model.py:11-22: 100,100 / 500,100), 65536-step synthetic rollout per worker, minibatch
4096, 10 epochs (weak scaling: every GPU owns one worker's rollout).
One "step" = one outer iteration of the reference worker on rollouts already in HBM:
K4 obs normalisation -> K1 GAE -> 10 x (K2 gather+fwd/bwd+losses -> K3 sum [+exchange] +
Adam).  samples/s = W * E * B / t_step.

Prints ONE JSON line (rank 0).  Timing: CUDA events on the launching stream, per-step
event pairs with an L2 flush (256 MiB write) between steps outside the events, barrier +
synchronize on both sides, max over ranks.
"""
import argparse
import json
import os
import random
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "pytorch-dppo_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

WORKLOAD = dict(name="C4 large-batch sweep (BASELINE configs[3])", O=17, A=6, widths=(100, 100, 500, 100),
                T=65536, B=4096, E=10, lr=3e-4, flavour="dppo")
METRIC = "ppo_update_samples_per_sec"
UNIT = "samples/s"
# algorithmic work per unit, reference arch, DPPO flavour
K2_FLOP_PER_SAMPLE = 425400          # 6 x fwd MACs + 2 x MACs for the theta_old forward (O=17, A=6)
K2_BYTES_PER_SAMPLE = 4 * (17 + 6 + 2) + 4
K1_BYTES_PER_STEP = 21               # r, V, boot fp32 + seg_end u8 in; A, R out
K4_BYTES_PER_ELEM = 8                # obs in, normalised out
K3_BYTES_PER_PARAM = 28              # g, theta, m, v in; theta, m, v out


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def tensor_peak_tf32():
    """Dense TF32 tensor-core ceiling in TFLOP/s: half of the measured dense bf16 figure (same MMA unit, twice the
    operand width); the burst figure, K2 is timed alone."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return 0.5 * float(json.load(open(p))["bf16_tflops"]), "0.5 x measured dense bf16 (MEASURED_PEAKS.json)"
    return 0.5 * 2250.0, "0.5 x nominal dense bf16 (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons during the timed region (NVML, 5 ms period)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz, self.ok = index, [], set(), False, None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:                                     # noqa: BLE001
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:                                    # noqa: BLE001
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:                                     # noqa: BLE001
                pass
            time.sleep(0.005)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


# --------------------------------------------------------------------------------------
# the reference arm: the oracle port (torch CPU, autograd - what train.py/chief.py do)
# --------------------------------------------------------------------------------------
def cpu_update_cycle_time(W, steps, warmup, threads):
    """Times `steps` update cycles (train.py:133-175 + chief.py:15-19) of W workers on the
    host cores.  Returns (seconds per cycle, description)."""
    from oracle import ppo_oracle as orc
    wl = WORKLOAD
    torch.set_num_threads(threads)
    torch.manual_seed(1)
    p = orc.init_params(wl["O"], wl["A"])
    g = torch.Generator().manual_seed(0)
    buf = dict(states=torch.randn(wl["T"], wl["O"], generator=g).clamp_(-5, 5), actions=torch.randn(wl["T"], wl["A"], generator=g),
               returns=torch.randn(wl["T"], 1, generator=g), advantages=torch.randn(wl["T"], 1, generator=g))
    m, v, gsh = orc.zeros_like_params(p), orc.zeros_like_params(p), orc.zeros_like_params(p)
    rng = random.Random(1)
    times = []
    step = 0
    for it in range(warmup + steps):
        idx = [[orc.sample_indices(wl["T"], wl["B"], rng) for _ in range(wl["E"])] for _ in range(W)]
        t0 = time.perf_counter()
        step = orc.dppo_update_cycle(p, m, v, step, [buf] * W, idx, gsh, lr=wl["lr"], num_epoch=wl["E"])
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    return sum(times) / len(times), ("oracle port (torch %s CPU fp32 autograd, %d threads): %d worker(s) x %d epochs x "
                                     "minibatch %d gathered from a %d-step rollout, gather + 2 forwards + backward + "
                                     "gradient sum + Adam (train.py:133-175, chief.py:15-19); GAE/obs-stats loops excluded"
                                     % (torch.__version__, threads, W, wl["E"], wl["B"], wl["T"]))


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    W = max(1, args.gpus)
    steps = max(1, min(args.steps, 10))               # bounded: one cycle is ~0.3-1 s per worker
    warmup = min(args.warmup, 2)
    t, desc = cpu_update_cycle_time(W, steps, warmup, threads)
    wl = WORKLOAD
    value = W * wl["E"] * wl["B"] / t
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
           "warmup": warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic", "config": config_dict(args.gpus),
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": desc},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


def config_dict(n):
    wl = WORKLOAD
    return {"workload": "%s: %d worker(s) (one per GPU) x %d-step HalfCheetah-shape rollout (obs %d, act %d), reference "
                        "net %s, minibatch %d, %d epochs; step = K4 obs-norm + K1 GAE + %d x (K2 + K3)"
                        % (wl["name"], n, wl["T"], wl["O"], wl["A"], "100,100/500,100", wl["B"], wl["E"], wl["E"]),
            "workers": n, "steps_per_worker": wl["T"], "minibatch": wl["B"], "epochs": wl["E"],
            "l2": "flushed between timed steps (256 MiB write outside the event pairs)",
            "parallelism": "dp%d" % n}


# --------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)

    rank, local_rank, world = dist_env()
    if world != args.gpus and world > 1:
        args.gpus = world
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the update path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    import b200_native as nv
    from b200_engine import UpdateEngine
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from synthetic import make_rollout

    comm = None
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        from b200_comm import PeerComm
        comm = PeerComm.create(dist.group.WORLD)

    wl = WORKLOAD
    torch.manual_seed(1)                                           # main.py:32,44 - same model on every rank
    h1, h2, g1, g2 = wl["widths"]
    model = Model(wl["O"], wl["A"], h1, h2, g1, g2, device=dev)
    stats = Shared_obs_stats(wl["O"], device=dev)
    sgb = Shared_grad_buffers(model)
    engine = UpdateEngine(model, nv.FLAVOUR_DPPO, wl["B"], wl["E"], wl["lr"], comm=comm)
    host, devbuf = make_rollout(model, stats, wl["T"], seed=1 + rank, flavour=nv.FLAVOUR_DPPO)
    rng = random.Random(1 + rank)
    use_graph = not args.no_graph and not (comm is not None and comm.mode == "nccl")   # NCCL baseline: plain launches
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stats0 = [t.clone() for t in (stats.n, stats.mean, stats.mean_diff, stats.var)]

    def reset_stats():
        for t, s in zip((stats.n, stats.mean, stats.mean_diff, stats.var), stats0):
            t.copy_(s)

    def step_resident():
        """inputs already in HBM: K4 -> K1 -> E x (K2 -> K3).  (minibatch positions are host
        data by nature - random.sample - and are staged before the timed events.)"""
        engine.run_epochs(devbuf, gshared=sgb.flat, use_graph=use_graph, indices_staged=True, obs_stats=stats,
                          with_gae=True)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up -------------------------------------------------------------------------
    for _ in range(args.warmup):
        reset_stats()
        engine.draw_indices(wl["T"], rng)
        engine.idx_dev.copy_(engine.idx_host, non_blocking=True)
        step_resident()
    launches_per_step = engine.launches_per_cycle - (1 if engine.flavour == nv.FLAVOUR_DPPO else 0)
    barrier()

    # ---- timed region: K steps, device-resident inputs -------------------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    evs = []
    barrier()
    for _ in range(args.steps):
        engine.draw_indices(wl["T"], rng)
        engine.idx_dev.copy_(engine.idx_host, non_blocking=True)
        flush.fill_(1)                                             # evict L2 (126 MB) between steps
        reset_stats()                                              # keep n below 2^24 (fp32 counter, model.py:71)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step_resident()
        e1.record()
        evs.append((e0, e1))
    barrier()
    t_ms = sum(a.elapsed_time(b) for a, b in evs)
    sampler.stop_flag = True
    sampler.join()

    # ---- end to end: host rollouts through the public call, copies inside the timed region --
    # engine.update_stream is the call a training loop makes: every step uploads its rollout from pinned
    # host memory and downloads its losses; the upload of step i+1 and the host-side index sampling overlap
    # the kernels of step i (two device buffers, copy stream).
    from b200_engine import RolloutBuffer
    devbuf2 = RolloutBuffer(wl["T"], wl["O"], wl["A"], dev)
    devs2 = [devbuf, devbuf2]
    reset_stats()
    engine.update_stream([host], devs2, 3, obs_stats=stats, rng=rng, gshared=sgb.flat, on_step=None)
    barrier()
    n_e2e = max(3, min(args.steps, 100))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reset_stats()
    e0.record()
    _, h2d, d2h = engine.update_stream([host], devs2, n_e2e, obs_stats=stats, rng=rng, gshared=sgb.flat,
                                       on_step=lambda i: reset_stats() if i % 64 == 63 else None)
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)

    # ---- per-kernel durations (instrumented pass, plain launches, same data) ----------------
    kt = kernel_times(nv, engine, stats, devbuf, sgb, wl, flush)

    tt = torch.tensor([t_ms, e2e_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_ms, e2e_ms = float(tt[0]), float(tt[1])
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    samples_per_step = world * wl["E"] * wl["B"]
    value = samples_per_step * args.steps / (t_ms * 1e-3)
    e2e_value = samples_per_step * n_e2e / (e2e_ms * 1e-3)
    hbm_peak, peak_src = peaks()
    fp32_peak = nv.fp32_probe()
    k2_s = kt["k2_us"] * 1e-6
    P = model.num_params
    k2_bytes = wl["B"] * K2_BYTES_PER_SAMPLE + 4 * P
    k2_flops = wl["B"] * K2_FLOP_PER_SAMPLE
    traffic = None
    tp = os.path.join(ROOT, "profiles", "k2_dram_traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get("dram_bytes_per_launch")
    roofline = {
        "kernel": "update_fused_kernel (K2)", "bound": "hbm", "achieved": k2_bytes / k2_s / 1e9, "peak": hbm_peak,
        "unit": "GB/s", "frac": k2_bytes / k2_s / 1e9 / hbm_peak, "traffic": traffic, "peak_source": peak_src,
        "note": "K2 does %.0f FLOP per algorithmic byte, so HBM is not its bound; its layers run on the tensor cores as "
                "3xTF32 (three MMA flops per algorithmic flop) with 28 samples = 32 MMA columns per SM, and the kernel is "
                "latency bound (eight warps per SM, profiles/r01f_k2_summary.md): see 'compute' and 'tensor'"
                % (k2_flops / k2_bytes),
        "compute": {"bound": "fp32_fma", "achieved": k2_flops / k2_s / 1e12, "peak": fp32_peak, "unit": "TFLOP/s",
                    "frac": k2_flops / k2_s / 1e12 / fp32_peak, "peak_source": "dppo_fp32_probe, measured in this run",
                    "note": "algorithmic fp32 flops against the FP32 FMA pipe, the ceiling of a CUDA-core implementation"},
        "tensor": {"bound": "tensor", "achieved": 3.0 * k2_flops / k2_s / 1e12, "peak": tensor_peak_tf32()[0],
                   "unit": "TFLOP/s", "frac": 3.0 * k2_flops / k2_s / 1e12 / tensor_peak_tf32()[0],
                   "peak_source": tensor_peak_tf32()[1],
                   "note": "3xTF32 MMA flops; an MMA with N = 32..64 columns occupies the unit as long as its 128 rows "
                           "need, so 25-50 % of this peak is the most this tiling can reach"},
        "share_of_step": wl["E"] * kt["k2_us"] / (t_ms / args.steps * 1e3),
        "kernels": {
            "k2_update_fused": {"us": kt["k2_us"], "launches_per_step": wl["E"]},
            "k3_reduce_adam": {"us": kt["k3_us"], "launches_per_step": wl["E"],
                               "hbm_gbs": (K3_BYTES_PER_PARAM * P + 4.0 * P * engine.nparts) / (kt["k3_us"] * 1e-6) / 1e9},
            "k1_gae_scan": {"us": kt["k1_us"], "launches_per_step": 1,
                            "hbm_gbs": K1_BYTES_PER_STEP * wl["T"] / (kt["k1_us"] * 1e-6) / 1e9},
            "k4_welford": {"us": kt["k4_us"], "launches_per_step": 3,
                           "hbm_gbs": K4_BYTES_PER_ELEM * wl["T"] * wl["O"] / (kt["k4_us"] * 1e-6) / 1e9},
            "hbm_study": kt["hbm_study"],
        },
    }
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": t_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic", "config": config_dict(world), "clocks": sampler.summary(),
           "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                   "ms_per_step": e2e_ms / n_e2e, "steps": n_e2e},
           "gpu_launches": launches_per_step * args.steps, "roofline": roofline,
           "graph": use_graph, "comm": None if comm is None else comm.describe()}
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        t, desc = cpu_update_cycle_time(1, 8, 1, threads)
        out["cpu_baseline"] = {"value": wl["E"] * wl["B"] / t, "unit": UNIT, "cores": threads, "kind": "port",
                               "sample": desc + "; 8 cycles after 1 warm-up"}
    print(json.dumps(out), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def kernel_times(nv, engine, stats, buf, sgb, wl, flush):
    """Average device time of each kernel with CUDA events around plain launches (the timed
    region replays them from a CUDA graph, where events cannot be interleaved), plus the
    HBM study of the streaming kernels at sizes beyond L2."""
    def timed(fn, reps):
        tot = 0.0
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / reps * 1e3                                   # us

    E = engine.E
    saved = [t.clone() for t in (engine.theta, engine.theta_old, engine.m, engine.v)]
    old = engine.theta_old if engine.flavour == nv.FLAVOUR_DPPO else None

    def k2():
        nv.update_fused(engine.flavour, engine.dims, buf.obs, buf.act, buf.adv, buf.ret, None, engine.idx_dev[0],
                        engine.theta, old, engine.clip, engine.ent_coeff, engine.partials, engine.loss_partials)

    def k3():
        if engine.comm is not None:                        # every rank runs this the same number of times
            engine.comm.allreduce_adam(engine, 0, None)
            return
        nv.reduce_adam(engine.partials, engine.nparts, None, engine.theta, old, engine.m, engine.v, 0, engine.lr,
                       grad_out=engine.grad, loss_partials=engine.loss_partials, loss_out=engine.loss_out[0],
                       scalars_dev=engine.scalars_dev[0])

    for f in (k2, k3):
        f()
    torch.cuda.synchronize()
    res = {"k2_us": timed(k2, 3 * E), "k3_us": timed(k3, 3 * E)}
    for t, s in zip((engine.theta, engine.theta_old, engine.m, engine.v), saved):
        t.copy_(s)
    res["k1_us"] = timed(lambda: engine.gae(buf), 20)
    st = [t.clone() for t in (stats.n, stats.mean, stats.mean_diff, stats.var)]
    res["k4_us"] = timed(lambda: stats.observes_batch(buf.raw_obs, out=buf.obs), 20)
    for t, s in zip((stats.n, stats.mean, stats.mean_diff, stats.var), st):
        t.copy_(s)
    # HBM study: K1 at 2^26 steps (1.4 GB of traffic), K4 at 2^22 x 17 (570 MB) - far beyond the 126 MB L2
    study = {}
    try:
        T1 = 1 << 26
        dev = buf.obs.device
        r = torch.rand(T1, device=dev); v = torch.rand(T1, device=dev); b = torch.zeros(T1, device=dev)
        se = (torch.arange(T1, device=dev) % 1000 == 999).to(torch.uint8)
        a = torch.empty(T1, device=dev); q = torch.empty(T1, device=dev)
        scr = torch.empty(nv.gae_scratch_bytes(T1), dtype=torch.uint8, device=dev)
        fn = lambda: nv.gae_scan(r, v, b, se, a, q, 0.99, 0.95, scr)
        fn(); torch.cuda.synchronize()
        us = timed(fn, 5)
        study["k1_gae_T2^26"] = {"us": us, "hbm_gbs": K1_BYTES_PER_STEP * T1 / (us * 1e-6) / 1e9}
        del r, v, b, se, a, q
        T4, O = 1 << 22, wl["O"]
        x = torch.randn(T4, O, device=dev); y = torch.empty_like(x)
        s4 = [torch.zeros(O, device=dev) for _ in range(4)]
        scr = torch.empty(nv.welford_scratch_bytes(T4, O), dtype=torch.uint8, device=dev)
        fn = lambda: nv.welford(x, y, s4[0], s4[1], s4[2], s4[3], scr)
        fn(); torch.cuda.synchronize()
        us = timed(fn, 5)
        study["k4_welford_T2^22xO17"] = {"us": us, "hbm_gbs": K4_BYTES_PER_ELEM * T4 * O / (us * 1e-6) / 1e9}
        del x, y
        P = 1 << 26
        th = torch.zeros(P, device=dev); mm = torch.zeros(P, device=dev); vv = torch.zeros(P, device=dev)
        g = torch.randn(1, P, device=dev)
        fn = lambda: nv.reduce_adam(g, 1, None, th, None, mm, vv, 1, 3e-4)
        fn(); torch.cuda.synchronize()
        us = timed(fn, 5)
        study["k3_adam_P2^26"] = {"us": us, "hbm_gbs": K3_BYTES_PER_PARAM * P / (us * 1e-6) / 1e9}
    except Exception as e:                                        # noqa: BLE001
        study["error"] = repr(e)
    res["hbm_study"] = study
    return res


if __name__ == "__main__":
    main()
