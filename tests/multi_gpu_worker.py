"""Run under torchrun with W GPUs: every rank owns one worker's synthetic rollout, runs E epochs
through the engine with the peer-memory exchange, then rank 0 checks that (a) all replicas are
bit-identical and (b) the result equals a single-GPU emulation of the same W workers (K2 per
worker, partials concatenated, one K3) within the suite's 1e-5 tolerance."""
import os
import random
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-dppo_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def main():
    import b200_native as nv
    from b200_comm import PeerComm
    from b200_engine import UpdateEngine
    from model import Model, Shared_grad_buffers, Shared_obs_stats
    from synthetic import make_rollout
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    O, A, T, B, E = 17, 6, 4096, 1024, 4

    def build(r):
        torch.manual_seed(1)
        m = Model(O, A, device=dev)
        stats = Shared_obs_stats(O, device=dev)
        host, buf = make_rollout(m, stats, T, seed=10 + r, flavour=nv.FLAVOUR_DPPO, seg_len=500)
        return m, buf

    model, buf = build(rank)
    sgb = Shared_grad_buffers(model)
    comm = PeerComm.create()
    eng = UpdateEngine(model, nv.FLAVOUR_DPPO, B, E, 3e-4, comm=comm)
    eng.gae(buf)
    for cycle in range(2):                                   # second cycle replays the captured graph
        eng.draw_indices(T, random.Random(100 * cycle + rank))
        eng.run_epochs(buf, gshared=sgb.flat, use_graph=True)
    torch.cuda.synchronize()
    flat = model.flat.clone()
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    ok = all(torch.equal(gathered[0], g) for g in gathered)
    if rank == 0:
        assert ok, "replicas diverged"
        # single-GPU emulation of the same W workers
        torch.manual_seed(1)
        ref = Model(O, A, device=dev)
        bufs = []
        for r in range(world):
            _, b = build(r)
            e = UpdateEngine(ref, nv.FLAVOUR_DPPO, B, E, 3e-4)
            e.gae(b)
            bufs.append(b)
        npart = nv.update_nparts(ref.dims, B)
        partials = torch.empty(world * npart, ref.P_pad, device=dev)
        lparts = torch.empty(world * npart, 4, device=dev)
        gsh = torch.ones_like(ref.flat)
        for i, (o, s) in enumerate(zip(ref.offsets, ref.sizes)):
            gsh[o + s:ref.offsets[i + 1] if i + 1 < len(ref.offsets) else ref.P_pad] = 0
        mm, vv = torch.zeros_like(ref.flat), torch.zeros_like(ref.flat)
        step = 0
        for cycle in range(2):
            idx = [nv.sample_indices(T, B, E, rng=random.Random(100 * cycle + r)).to(dev) for r in range(world)]
            old = ref.flat.clone()
            for k in range(E):
                for r in range(world):
                    nv.update_fused(nv.FLAVOUR_DPPO, ref.dims, bufs[r].obs, bufs[r].act, bufs[r].adv, bufs[r].ret, None,
                                    idx[r][k].contiguous(), ref.flat, old, 0.2, 0.0, partials[r * npart:(r + 1) * npart],
                                    lparts[r * npart:(r + 1) * npart])
                step += 1
                nv.reduce_adam(partials, world * npart, gsh, ref.flat, old, mm, vv, step, 3e-4)
        torch.cuda.synchronize()
        err = float((ref.flat - flat).abs().max()) / float(flat.abs().max())
        assert err < 1e-5, err
        print("MULTI_GPU_OK world=%d mode=%s algo=%s rel_err_vs_single_gpu=%.2e" % (world, comm.mode, comm.describe().get("algorithm"), err), flush=True)
    # ---- observation statistics shared by the workers in rank order (main.py:54, model.py:61-80) ----
    stats = Shared_obs_stats(O, device=dev)
    norm = stats.observes_batch(buf.raw_obs, comm=comm)
    torch.cuda.synchronize()
    finals = [torch.empty(4, O, device=dev) for _ in range(world)]
    dist.all_gather(finals, torch.stack([stats.n, stats.mean, stats.mean_diff, stats.var]))
    assert all(torch.equal(finals[0], f) for f in finals), "statistics replicas diverged"
    last = [torch.empty_like(norm) for _ in range(world)]
    dist.all_gather(last, norm)
    if rank == 0:
        seq = Shared_obs_stats(O, device=dev)                      # one statistics object fed rank after rank
        for r in range(world):
            _, b = build(r)
            want = seq.observes_batch(b.raw_obs)
            err = float((want - last[r]).abs().max()) / 5.0
            assert err < 1e-5, (r, err)
        for got, want in zip(finals[0], (seq.n, seq.mean, seq.mean_diff, seq.var)):
            assert float((got - want).abs().max()) <= 1e-5 * float(want.abs().max()), "final statistics"
        print("SHARED_STATS_OK world=%d" % world, flush=True)
    dist.barrier()
    if comm.mode == "p2p":
        import b200_native as nv2
        nv2.comm_check(comm.handle)                                # no poll timed out
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
