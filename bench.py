#!/usr/bin/env python
"""Benchmark of the data-parallel PPO update path (BASELINE.json metric:
"PPO update samples/sec at 1/2/4/8 B200; % of HBM roofline vs host-CPU reference").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c1|c2|c3|c4|c5] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workloads (--config, default c4 = BASELINE configs[3], the large-batch sweep the metric is quoted on):
  c1  ppo.py update, InvertedPendulum shape (obs 4, act 1), 64x64 nets, 2048-step rollout, 10 epochs x minibatch 64
  c2  DPPO, HalfCheetah shape (17/6), reference nets (model.py:11-22: 100,100 / 500,100), 2048 steps, minibatch 2048
  c3  DPPO, HopperBullet shape (15/3), reference nets, 2048 steps, minibatch 2048
  c4  DPPO, HalfCheetah shape, reference nets, 65536-step rollout per worker, minibatch 4096, 10 epochs
  c5  DPPO, Humanoid shape (376/17), 256x256 nets, 32768 steps, minibatch 4096
One worker per GPU (weak scaling: every GPU owns one worker's rollout).
One "step" = one outer iteration of the reference worker on rollouts already in HBM:
K4 obs normalisation -> K1 GAE -> E x (K2 gather+fwd/bwd+losses -> K3 sum [+exchange] + Adam).
samples/s = W * E * B / t_step.

Prints ONE JSON line (rank 0).  Timing: CUDA events on the launching stream, per-step event pairs with an L2
flush (256 MiB write) between steps outside the events, barrier + synchronize on both sides, max over ranks.
After the timed region ONE more step is run from a known state and compared with the CPU oracle on the same
rollouts and the same random.sample stream ("parity"); at W > 1 the replicas are also compared bit for bit.
"""
import argparse
import json
import os
import random
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))

import torch  # noqa: E402

CONFIGS = {
    "c1": dict(name="C1 ppo.py single-agent update (BASELINE configs[0])", env="InvertedPendulum", O=4, A=1,
               widths=(64, 64, 64, 64), T=2048, B=64, E=10, lr=7e-4, flavour="ppo", ent=0.01, max_grad_norm=0.5),
    "c2": dict(name="C2 DPPO 4 workers (BASELINE configs[1])", env="HalfCheetah", O=17, A=6,
               widths=(100, 100, 500, 100), T=2048, B=2048, E=10, lr=3e-4, flavour="dppo", ent=0.0, max_grad_norm=0.0),
    "c3": dict(name="C3 DPPO 8 workers (BASELINE configs[2])", env="HopperBullet", O=15, A=3,
               widths=(100, 100, 500, 100), T=2048, B=2048, E=10, lr=3e-4, flavour="dppo", ent=0.0, max_grad_norm=0.0),
    "c4": dict(name="C4 large-batch sweep (BASELINE configs[3])", env="HalfCheetah", O=17, A=6,
               widths=(100, 100, 500, 100), T=65536, B=4096, E=10, lr=3e-4, flavour="dppo", ent=0.0, max_grad_norm=0.0),
    "c5": dict(name="C5 Humanoid 256x256 (BASELINE configs[4])", env="Humanoid", O=376, A=17,
               widths=(256, 256, 256, 256), T=32768, B=4096, E=10, lr=3e-4, flavour="dppo", ent=0.0, max_grad_norm=0.0),
}
METRIC = "ppo_update_samples_per_sec"
UNIT = "samples/s"
K1_BYTES_PER_STEP = 21               # r, V, boot fp32 + seg_end u8 in; A, R out
K4_BYTES_PER_ELEM = 8                # obs in, normalised out
K3_BYTES_PER_PARAM = 28              # g, theta, m, v in; theta, m, v out


def algorithmic(wl):
    """K2 FLOPs and bytes per sample from the layer widths."""
    h1, h2, g1, g2 = wl["widths"]
    macs = wl["O"] * h1 + h1 * h2 + h2 * wl["A"] + wl["O"] * g1 + g1 * g2 + g2
    flop = 6 * macs + (2 * macs if wl["flavour"] == "dppo" else 0)       # fwd 2 + bwd 4 per MAC [+ theta_old forward]
    byts = 4 * (wl["O"] + wl["A"] + (2 if wl["flavour"] == "dppo" else 3)) + 4
    return macs, flop, byts


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


def config_dict(wl, n):
    h = wl["widths"]
    return {"workload": "%s: %d worker(s) (one per GPU) x %d-step %s-shape rollout (obs %d, act %d), nets %d,%d/%d,%d, "
                        "%s loss, minibatch %d, %d epochs; step = K4 obs-norm + K1 GAE + %d x (K2 + K3)"
                        % (wl["name"], n, wl["T"], wl["env"], wl["O"], wl["A"], h[0], h[1], h[2], h[3],
                           "train.py" if wl["flavour"] == "dppo" else "ppo.py", wl["B"], wl["E"], wl["E"]),
            "workers": n, "steps_per_worker": wl["T"], "minibatch": wl["B"], "epochs": wl["E"],
            "l2": "flushed between timed steps (256 MiB write outside the event pairs)",
            "parallelism": "dp%d" % n}


# --------------------------------------------------------------------------------------
# CPU arms: the reference itself (oracle/_ref through oracle/ref_runner.py) and the oracle port
# --------------------------------------------------------------------------------------
def ref_available(wl):
    # the reference hard-codes its widths (model.py:11-18) and ref_runner drives its DPPO programs (train.py + chief.py)
    return (wl["flavour"] == "dppo" and tuple(wl["widths"]) == (100, 100, 500, 100) and
            os.path.exists(os.path.join(ROOT, "oracle", "_ref", "train.py")))


def run_ref(wl, W, epochs, as_is=False, timeout=900):
    """The unmodified reference: W forked train.train workers + chief.chief, one thread each.  Returns ref_runner's
    JSON.  The rollout phase (untimed) is shortened to B steps per worker; the update phase is the config's."""
    cmd = [sys.executable, os.path.join(ROOT, "oracle", "ref_runner.py"), "--O", str(wl["O"]), "--A", str(wl["A"]),
           "--W", str(W), "--B", str(wl["B"]), "--E", str(wl["E"]), "--T", str(wl["B"]), "--epochs", str(epochs)]
    if as_is:
        cmd.append("--as-is")
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT")}
    env["OMP_NUM_THREADS"] = "1"                                   # main.py:42
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
    if out.returncode != 0:
        raise RuntimeError("oracle/ref_runner.py failed: " + out.stderr[-2000:])
    return json.loads(out.stdout.strip().splitlines()[-1])


def port_cycle_time(wl, W, cycles, warmup, threads):
    """The oracle port (torch CPU autograd, all host threads on one stream of W workers): seconds per cycle."""
    sys.path.insert(0, ROOT)
    from oracle import ppo_oracle as orc
    torch.set_num_threads(threads)
    torch.manual_seed(1)
    h1, h2, g1, g2 = wl["widths"]
    p = orc.init_params(wl["O"], wl["A"], h1, h2, g1, g2)
    g = torch.Generator().manual_seed(0)
    buf = dict(states=torch.randn(wl["T"], wl["O"], generator=g).clamp_(-5, 5), actions=torch.randn(wl["T"], wl["A"], generator=g),
               returns=torch.randn(wl["T"], 1, generator=g), advantages=torch.randn(wl["T"], 1, generator=g),
               logprobs=torch.randn(wl["T"], 1, generator=g) - 3)
    m, v, gsh = orc.zeros_like_params(p), orc.zeros_like_params(p), orc.zeros_like_params(p)
    rng = random.Random(1)
    times, step = [], 0
    for it in range(warmup + cycles):
        idx = [[orc.sample_indices(wl["T"], wl["B"], rng) for _ in range(wl["E"])] for _ in range(W)]
        t0 = time.perf_counter()
        if wl["flavour"] == "dppo":
            step = orc.dppo_update_cycle(p, m, v, step, [buf] * W, idx, gsh, lr=wl["lr"], num_epoch=wl["E"])
        else:
            step = orc.ppo_update_cycle(p, m, v, step, buf, idx[0], lr=wl["lr"], num_epoch=wl["E"])
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    return sum(times) / len(times)


def cpu_arms(wl, W, steps, warmup, full):
    """Returns (value, ms_per_step, steps_done, warmup_done, cpu_baseline dict).  A "step" of the CPU arm is ONE
    reference epoch (train.py:133-175 once + chief.py:15-19 once) - a bounded sample of the update cycle, so that
    --steps/--warmup are honoured and the run still ends within minutes."""
    cores = os.cpu_count() or 1
    samples_per_epoch = W * wl["B"]
    if ref_available(wl):
        r = run_ref(wl, W, warmup + steps)
        ep = r["epoch_s"][warmup:warmup + steps] if len(r["epoch_s"]) > warmup else r["epoch_s"]
        t_ep = sum(ep) / len(ep)
        cb = {"value": samples_per_epoch / t_ep, "unit": UNIT, "cores": r["processes"], "kind": "reference",
              "sample": "the unmodified reference (oracle/_ref: train.train x %d forked workers + chief.chief, 1 thread per "
                        "process as main.py:42 sets, torch %s CPU) on the synthetic rollouts of the GPU arm (same seeds); "
                        "%d epochs timed after %d warm-up, each = random.sample + 2 forwards + backward + add_gradient of "
                        "minibatch %d per worker + the chief's Adam step, chief.py:10's sleep(1) shortened to 2 ms "
                        "(compute-only); the untimed rollout phase is shortened to %d steps per worker"
                        % (W, r["torch"], len(ep), warmup, wl["B"], r["rollout_steps_per_worker"]),
              "host_cores": cores}
        if full:
            try:                                                   # the as-is figure, one cycle
                a = run_ref(wl, W, wl["E"], as_is=True)
                cb["as_is_value"] = W * wl["E"] * wl["B"] / a["cycle_wall_s"][0]
                cb["as_is_note"] = "one update cycle with chief.py:10's sleep(1) left in: %.1f s for %d epochs" % (
                    a["cycle_wall_s"][0], wl["E"])
                nw = max(1, cores - 1)                             # the reference's way to use a whole host: more workers
                if nw != W:
                    b = run_ref(wl, nw, wl["E"])
                    tb = sum(b["epoch_s"]) / len(b["epoch_s"])
                    cb["all_cores_value"] = nw * wl["B"] / tb
                    cb["all_cores_note"] = "%d workers + chief on %d host cores, compute-only" % (nw, cores)
                tp = port_cycle_time(wl, W, 4, 1, cores)
                cb["port_value"] = W * wl["E"] * wl["B"] / tp
                cb["port_note"] = "oracle port, torch autograd with %d threads on one stream of %d worker(s)" % (cores, W)
            except Exception as e:                                 # noqa: BLE001
                cb["extras_error"] = repr(e)[:300]
        return cb["value"], t_ep * 1e3, len(ep), warmup, cb
    cyc = max(1, min(-(-steps // wl["E"]), 10))
    t = port_cycle_time(wl, W, cyc, 1, cores)
    t_ep = t / wl["E"]
    cb = {"value": samples_per_epoch / t_ep, "unit": UNIT, "cores": cores, "kind": "port",
          "sample": "oracle port (torch %s CPU fp32 autograd, %d threads): %d worker(s), %d update cycles of %d epochs x "
                    "minibatch %d after 1 warm-up cycle (oracle/_ref %s)"
                    % (torch.__version__, cores, W, cyc, wl["E"], wl["B"],
                       "absent" if wl["flavour"] == "dppo" and tuple(wl["widths"]) == (100, 100, 500, 100)
                       else "runs train.py/chief.py with the reference's hard-coded widths only")}
    return cb["value"], t_ep * 1e3, cyc * wl["E"], wl["E"], cb


def run_reference(args, wl):
    rank, _, world = dist_env()
    if rank != 0:
        return
    W = max(1, args.gpus)
    value, ms, steps, warmup, cb = cpu_arms(wl, W, args.steps, args.warmup, full=False)
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
           "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic", "config": config_dict(wl, args.gpus), "cpu_baseline": cb,
           "step_unit": "one reference epoch (minibatch of %d per worker)" % wl["B"],
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


# --------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--config", default="c4", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--host-sampler", action="store_true",
                    help="e2e: draw the minibatch positions with the host sampler (worker thread + upload) instead of on the GPU")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    wl = CONFIGS[args.config]
    if args.impl == "reference":
        return run_reference(args, wl)

    for _p in (ROOT, os.path.join(ROOT, "pytorch-dppo_b200")):
        if _p not in sys.path:
            sys.path.insert(0, _p)
    rank, local_rank, world = dist_env()
    if world != args.gpus and world > 1:
        args.gpus = world
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the update path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    import b200_native as nv
    from b200_engine import RolloutBuffer, UpdateEngine
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from synthetic import make_rollout

    comm = None
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        from b200_comm import PeerComm
        comm = PeerComm.create(dist.group.WORLD)

    flavour = nv.FLAVOUR_DPPO if wl["flavour"] == "dppo" else nv.FLAVOUR_PPO
    torch.manual_seed(1)                                           # main.py:32,44 - same model on every rank
    h1, h2, g1, g2 = wl["widths"]
    model = Model(wl["O"], wl["A"], h1, h2, g1, g2, device=dev)
    theta0 = model.flat.clone()
    stats = Shared_obs_stats(wl["O"], device=dev)
    sgb = Shared_grad_buffers(model)
    gsh = sgb.flat if flavour == nv.FLAVOUR_DPPO else None
    engine = UpdateEngine(model, flavour, wl["B"], wl["E"], wl["lr"], ent_coeff=wl["ent"], max_grad_norm=wl["max_grad_norm"],
                          comm=comm)
    host, devbuf = make_rollout(model, stats, wl["T"], seed=1 + rank, flavour=flavour)
    rng = random.Random(1 + rank)
    use_graph = not args.no_graph and not (comm is not None and comm.mode == "nccl")   # NCCL baseline: plain launches
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stat_t = (stats.n, stats.mean, stats.mean_diff, stats.var)
    stats0 = [t.clone() for t in stat_t]

    def reset_stats():
        for t, s in zip(stat_t, stats0):
            t.copy_(s)

    def step_resident():
        """inputs already in HBM: K4 -> K1 -> E x (K2 -> K3).  (minibatch positions are host
        data by nature - random.sample - and are staged before the timed events.)"""
        engine.run_epochs(devbuf, gshared=gsh, use_graph=use_graph, indices_staged=True, obs_stats=stats, with_gae=True)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up -------------------------------------------------------------------------
    for _ in range(args.warmup):
        reset_stats()
        engine.draw_indices(wl["T"], rng)
        engine._upload_indices()
        step_resident()
    launches_per_step = engine.launches_per_cycle - (1 if flavour == nv.FLAVOUR_DPPO else 0)   # minus the theta_old copy
    barrier()

    # ---- timed region: K steps, device-resident inputs -------------------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    evs = []
    barrier()
    for _ in range(args.steps):
        engine.draw_indices(wl["T"], rng)
        engine._upload_indices()                                   # (draw_indices waits for the previous upload of the block)
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
    devbuf2 = RolloutBuffer(wl["T"], wl["O"], wl["A"], dev)
    devs2 = [devbuf, devbuf2]
    reset_stats()
    # random.sample on the GPU (same stream of positions) where CPython's algorithm parallelises: its set branch
    # (population larger than ~4x the sample, C4/C5).  Its pool branch (C1-C3: a permutation of the whole buffer) is a
    # sequential swap chain - one GPU thread is slower at that than a host core - so those configs keep the host
    # sampler (a worker thread one step ahead, 4*E*B bytes uploaded per step).
    use_dev_sampler = not args.host_sampler and not nv.sample_uses_pool(wl["T"], wl["B"])
    idx_sampler = nv.DeviceSampler(rng) if use_dev_sampler else None
    engine.update_stream([host], devs2, 3, obs_stats=stats, rng=rng, gshared=gsh, on_step=None, sampler=idx_sampler)
    barrier()
    n_e2e = max(3, min(args.steps, 100))
    keep_n = max(1, (1 << 23) // wl["T"] // max(world, 1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reset_stats()
    e0.record()
    _, h2d, d2h = engine.update_stream([host], devs2, n_e2e, obs_stats=stats, rng=rng, gshared=gsh, sampler=idx_sampler,
                                       on_step=lambda i: reset_stats() if i % keep_n == keep_n - 1 else None)
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)

    # ---- per-kernel durations: each kernel replayed back to back from its own CUDA graph ----------------
    kt = kernel_times(nv, engine, stats, devbuf, gsh, wl, use_graph, barrier)

    tt = torch.tensor([t_ms, e2e_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_ms, e2e_ms = float(tt[0]), float(tt[1])

    # ---- parity: one more step from a known state against the CPU oracle (every rank takes part) ----------
    parity = None
    if not args.no_parity:
        parity = parity_check(nv, engine, model, theta0, stats, sgb, host, devbuf, wl, flavour, use_graph, rank, world,
                              dist, barrier)
    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    samples_per_step = world * wl["E"] * wl["B"]
    value = samples_per_step * args.steps / (t_ms * 1e-3)
    e2e_value = samples_per_step * n_e2e / (e2e_ms * 1e-3)
    step_us = t_ms / args.steps * 1e3
    hbm_peak, peak_src = peaks()
    fp32_peak = nv.fp32_probe()
    macs, flop_per_sample, bytes_per_sample = algorithmic(wl)
    k2_s = kt["k2_us"] * 1e-6
    P = model.num_params
    k2_bytes = wl["B"] * bytes_per_sample + 4 * P
    k2_flops = wl["B"] * flop_per_sample
    tpeak, tsrc = tensor_peak_tf32()
    traffic = None
    tp = os.path.join(ROOT, "profiles", "k2_dram_traffic.json")
    if os.path.exists(tp) and args.config == "c4":
        traffic = json.load(open(tp)).get("dram_bytes_per_launch")
    k3_launches = 2 if wl["max_grad_norm"] > 0 else 1
    sum_us = wl["E"] * (kt["k2_us"] + kt["k3_us"]) + kt["k1_us"] + kt["k4_us"]
    roofline = {
        "kernel": "update_fused_kernel (K2)", "bound": "tensor", "achieved": k2_flops / k2_s / 1e12, "peak": tpeak / 3.0,
        "unit": "TFLOP/s", "frac": k2_flops / k2_s / 1e12 / (tpeak / 3.0), "traffic": traffic,
        "peak_source": tsrc + " / 3: the layers run as 3xTF32, three tensor-core flops per algorithmic fp32 flop "
                              "(SURVEY.md 8(d): tensor peak / split factor)",
        "flop_per_launch": k2_flops, "bytes_per_launch": k2_bytes,
        "note": "K2 does %.0f FLOP per algorithmic byte, so HBM is not its bound (hbm object beside); an MMA with N = 32..64 "
                "columns occupies the unit as long as its 128 rows need, so 25-50 %% of the TF32 peak is the most this tiling "
                "can reach" % (k2_flops / k2_bytes),
        "hbm": {"bound": "hbm", "achieved": k2_bytes / k2_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": k2_bytes / k2_s / 1e9 / hbm_peak, "peak_source": peak_src},
        "fp32": {"bound": "fp32_fma", "achieved": k2_flops / k2_s / 1e12, "peak": fp32_peak, "unit": "TFLOP/s",
                 "frac": k2_flops / k2_s / 1e12 / fp32_peak, "peak_source": "dppo_fp32_probe, measured in this run",
                 "note": "algorithmic fp32 flops against the FP32 FMA pipe, the ceiling of a CUDA-core implementation"},
        "share_of_step": wl["E"] * kt["k2_us"] / step_us,
        "sum_of_kernels_us": sum_us, "step_us": step_us, "sum_over_step": sum_us / step_us,
        "timing": "each kernel replayed %d x back to back from its own CUDA graph (events around the replay); the step is "
                  "one graph of all of them" % kt["reps"],
        "kernels": {
            "k2_update_fused": {"us": kt["k2_us"], "launches_per_step": wl["E"] * engine.k2_launches,
                                "what": "one dppo_update_fused call" + (" = the update kernel + the batch-level dW pass (same kernel, second launch)" if engine.k2_launches == 2 else "")},
            "k3_reduce_adam": {"us": kt["k3_us"], "launches_per_step": wl["E"] * k3_launches,
                               "hbm_gbs": (K3_BYTES_PER_PARAM * P + 4.0 * P * engine.nparts) / (kt["k3_us"] * 1e-6) / 1e9},
            "k1_gae_scan": {"us": kt["k1_us"], "launches_per_step": 1,
                            "hbm_gbs": K1_BYTES_PER_STEP * wl["T"] / (kt["k1_us"] * 1e-6) / 1e9},
            "k4_welford": {"us": kt["k4_us"], "launches_per_step": 1 if world == 1 else 4,
                           "hbm_gbs": K4_BYTES_PER_ELEM * wl["T"] * wl["O"] / (kt["k4_us"] * 1e-6) / 1e9},
            "hbm_study": kt["hbm_study"],
        },
    }
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": t_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic", "config": config_dict(wl, world), "clocks": sampler.summary(),
           "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                   "ms_per_step": e2e_ms / n_e2e, "steps": n_e2e,
                   "index_sampler": "host (C restatement of random.sample, worker thread, uploaded per step)" if idx_sampler is None
                   else "device (dppo_sample_indices: random.sample on the GPU, no index upload)"},
           "gpu_launches": launches_per_step * args.steps, "roofline": roofline, "parity": parity,
           "update_latency_us": step_us, "graph": use_graph, "comm": None if comm is None else comm.describe()}
    if world == 1 and not args.no_cpu_baseline:
        try:
            out["cpu_baseline"] = cpu_arms(wl, 1, 12, 3, full=True)[4]
        except Exception as e:                                     # noqa: BLE001
            out["cpu_baseline"] = {"error": repr(e)[:300]}
    print(json.dumps(out), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def parity_check(nv, engine, model, theta0, stats, sgb, host, devbuf, wl, flavour, use_graph, rank, world, dist, barrier):
    """One update step from (theta0, zero Adam state, fresh statistics, ones-initialised shared buffer) on every rank's
    rollout, then on rank 0 the CPU oracle over the same W rollouts and random.sample streams (train.py:128-175 +
    chief.py:13-20 / ppo.py:139-174).  Returns the dict that goes into the JSON line."""
    sys.path.insert(0, ROOT)
    from oracle import ppo_oracle as orc
    dev = model.flat.device
    T, B, E, O, A = wl["T"], wl["B"], wl["E"], wl["O"], wl["A"]
    model.flat.copy_(theta0)
    engine.m.zero_(); engine.v.zero_(); engine.step = 0
    for t in (stats.n, stats.mean, stats.mean_diff, stats.var):
        t.zero_()
    if flavour == nv.FLAVOUR_DPPO:
        for g in sgb.grads.values():
            g.fill_(1.0)                                           # model.py:51
    devbuf.copy_from(host, ["raw_obs", "act", "rew", "val", "boot", "seg_end", "logp"])
    prng = random.Random(4242 + rank)
    engine.draw_indices(T, prng)
    barrier()
    losses = engine.run_epochs(devbuf, gshared=sgb.flat if flavour == nv.FLAVOUR_DPPO else None, use_graph=use_graph,
                               obs_stats=stats, with_gae=True).clone()
    barrier()
    res = {"replicas_identical": None}
    pack = torch.cat([devbuf.obs.reshape(-1), devbuf.act.reshape(-1), devbuf.adv, devbuf.ret, devbuf.logp])
    idx = engine.idx_host.to(dev)
    if world > 1:
        thetas = [torch.empty_like(model.flat) for _ in range(world)]
        dist.all_gather(thetas, model.flat)
        res["replicas_identical"] = all(torch.equal(thetas[0], t) for t in thetas)
        packs = [torch.empty_like(pack) for _ in range(world)]
        dist.all_gather(packs, pack)
        idxs = [torch.empty_like(idx) for _ in range(world)]
        dist.all_gather(idxs, idx)
    else:
        packs, idxs = [pack], [idx]
    if rank != 0:
        return None
    t0 = time.perf_counter()
    h1, h2, g1, g2 = wl["widths"]
    shapes = orc.param_shapes(O, A, h1, h2, g1, g2)
    unpad = lambda flat: torch.cat([flat[o:o + s] for o, s in zip(model.offsets, model.sizes)]).cpu()
    p = orc.unflatten(unpad(theta0), shapes)
    rel = lambda got, ref: float((got.double() - ref.double()).abs().max()) / max(float(ref.double().abs().max()), 1e-30)
    # K4 and K1 of this rank against the sequential loops (rank 0 comes first in the rank-ordered statistics)
    st = orc.welford_new(O)
    want_obs = orc.welford_observe_normalize(st, host.raw_obs.clone())
    res["obs_rel"] = rel(devbuf.obs.cpu(), want_obs)
    a_ref, r_ref = orc.gae(host.rew.numpy(), host.val.numpy(), host.boot.numpy(), host.seg_end.numpy())
    res["gae_rel"] = max(rel(devbuf.adv.cpu(), torch.from_numpy(a_ref)), rel(devbuf.ret.cpu(), torch.from_numpy(r_ref)))
    workers, streams = [], []
    for w in range(world):
        x = packs[w].cpu()
        o = 0
        cols = {}
        for name, n, shape in (("states", T * O, (T, O)), ("actions", T * A, (T, A)), ("advantages", T, (T, 1)),
                               ("returns", T, (T, 1)), ("logprobs", T, (T, 1))):
            cols[name] = x[o:o + n].reshape(shape).clone()
            o += n
        workers.append(cols)
        streams.append(idxs[w].cpu().tolist())
        if w == 0:                                                 # the index stream itself: bit-exact with random.sample
            check = random.Random(4242)
            res["indices_bit_exact"] = streams[0] == [check.sample(range(T), B) for _ in range(E)]
    mo, vo = orc.zeros_like_params(p), orc.zeros_like_params(p)
    terms = []
    if flavour == nv.FLAVOUR_DPPO:
        gshared = orc.zeros_like_params(p, 1.0)                    # ones before the first update (model.py:51)
        orc.dppo_update_cycle(p, mo, vo, 0, workers, streams, gshared, lr=wl["lr"], ent_coeff=wl["ent"], num_epoch=E,
                              record=lambda k, w, t, g: terms.append(t) if w == 0 else None)
    else:
        orc.ppo_update_cycle(p, mo, vo, 0, workers[0], streams[0], lr=wl["lr"], ent_coeff=wl["ent"],
                             max_grad_norm=wl["max_grad_norm"], num_epoch=E, record=lambda k, t, g, c, n: terms.append(t))
    got = unpad(model.flat)
    res["theta_rel"] = rel(got, orc.flatten(p))
    per = []
    o = 0
    for k in orc.PARAM_ORDER:
        n = p[k].numel()
        per.append(rel(got[o:o + n], p[k].reshape(-1)))
        o += n
    res["theta_rel_worst_tensor"] = max(per)
    lo = losses.cpu()
    lr_ = 0.0
    for k in range(E):
        for j, name in enumerate(("loss_clip", "loss_value", "loss_ent")):
            ref = terms[k][name]
            if abs(ref) > 1e-6:
                lr_ = max(lr_, abs(float(lo[k, j]) - ref) / abs(ref))
    res["loss_rel"] = lr_
    res["m_rel"] = rel(unpad(engine.m), orc.flatten(mo))
    res["tolerance"] = 1e-5
    res["oracle_seconds"] = round(time.perf_counter() - t0, 2)
    res["what"] = ("one extra step after the timed region, from theta0 / zero Adam state / fresh statistics: K4, K1, %d x (K2, K3) "
                   "of %d worker(s) vs oracle/ppo_oracle.py on the same rollouts and random.sample streams" % (E, world))
    res["ok"] = bool(res["theta_rel"] <= 1e-5 and res["theta_rel_worst_tensor"] <= 1e-5 and res["loss_rel"] <= 1e-5 and
                     res["obs_rel"] <= 1e-5 and res["gae_rel"] <= 1e-5 and res.get("indices_bit_exact", False) and
                     res["replicas_identical"] in (None, True))
    return res


def kernel_times(nv, engine, stats, buf, gsh, wl, use_graph, barrier):
    """Average device time of each kernel, measured the way the step runs them: REPS launches back to back inside one
    CUDA graph, events around the replay (no host launch latency in the figure), plus the HBM study of the streaming
    kernels at sizes beyond L2."""
    REPS = 50
    E = engine.E

    def graph_us(fn, reps=REPS):
        fn()
        torch.cuda.synchronize()
        if not use_graph:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                fn()
            e1.record()
            e1.synchronize()
            return e0.elapsed_time(e1) / reps * 1e3
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(reps):
                fn()
        g.replay()
        torch.cuda.synchronize()
        best = None
        for _ in range(3):
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.replay()
            e1.record()
            e1.synchronize()
            t = e0.elapsed_time(e1) / reps * 1e3
            best = t if best is None else min(best, t)
        return best

    state = [engine.theta, engine.theta_old, engine.m, engine.v] + ([gsh] if gsh is not None else [])
    saved = [t.clone() for t in state]
    old = engine.theta_old if engine.flavour == nv.FLAVOUR_DPPO else None
    logp = buf.logp if engine.flavour == nv.FLAVOUR_PPO else None

    def k2():
        nv.update_fused(engine.flavour, engine.dims, buf.obs, buf.act, buf.adv, buf.ret, logp, engine.idx_dev[0],
                        engine.theta, old, engine.clip, engine.ent_coeff, engine.partials, engine.loss_partials)

    def k3():
        if engine.comm is not None:                        # every rank runs this the same number of times
            engine.comm.allreduce_adam(engine, 0, gsh)
            return
        nv.reduce_adam(engine.partials, engine.nparts, gsh, engine.theta, old, engine.m, engine.v, 0, engine.lr,
                       max_grad_norm=engine.max_grad_norm, grad_out=engine.grad, loss_partials=engine.loss_partials,
                       loss_out=engine.loss_out[0], scratch=engine._adam_scratch, scalars_dev=engine.scalars_dev[0])

    res = {"reps": REPS}
    res["k2_us"] = graph_us(k2)
    res["k3_us"] = graph_us(k3)
    for t, s in zip(state, saved):
        t.copy_(s)
    res["k1_us"] = graph_us(lambda: engine.gae(buf))
    st = [t.clone() for t in (stats.n, stats.mean, stats.mean_diff, stats.var)]
    reps4 = max(1, min(REPS, (1 << 23) // wl["T"] // (engine.comm.world if engine.comm is not None else 1)))
    res["k4_us"] = graph_us(lambda: stats.observes_batch(buf.raw_obs, out=buf.obs, comm=engine.comm), reps4)
    for t, s in zip((stats.n, stats.mean, stats.mean_diff, stats.var), st):
        t.copy_(s)
    # HBM study: K1 at 2^26 steps (1.4 GB of traffic), K4 at 2^22 x 17 (570 MB) - far beyond the 126 MB L2
    study = {}

    def timed(fn, reps):
        tot = 0.0
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / reps * 1e3

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
        T4, O = 1 << 22, 17
        x = torch.randn(T4, O, device=dev); y = torch.empty_like(x)
        s4 = [torch.zeros(O, device=dev) for _ in range(4)]
        scr = torch.empty(nv.welford_scratch_bytes(T4, O), dtype=torch.uint8, device=dev)
        fn = lambda: nv.welford(x, y, s4[0], s4[1], s4[2], s4[3], scr)
        fn(); torch.cuda.synchronize()
        for t in s4:
            t.zero_()
        us = timed(fn, 3)
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
