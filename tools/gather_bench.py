"""tools/gather_bench.py -- rollouts for the learner over NVLink (SURVEY 8e "Gather"), run under torchrun on N >= 2 GPUs:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      tools/gather_bench.py [--envs 131072] [--workload qbb|qq]

1. parity: the learner's buffer filled by the rollout kernels through peer-mapped memory (sharding.RecordGather) equals
   the NCCL gather of the ranks' local record buffers (sharding.gather_records), bit for bit;
2. timing (CUDA events, max over ranks): rollout with local records / + NCCL gather afterwards / with direct peer writes.
Prints one JSON line on rank 0."""
import argparse
import json
import os
import os.path as osp
import sys

ROOT = osp.dirname(osp.dirname(osp.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=1 << 17)
    ap.add_argument("--workload", default="qbb")
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist

    from bench import RolloutRunner, make_workload
    from simurlacra_b200 import _lib as L
    from simurlacra_b200 import engine as E
    from simurlacra_b200.sampling import sharding

    rank, ws, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    wl = make_workload(a.workload, E, L, torch, dev, 64, 1)
    wl["record"] = True
    n, T = a.envs, wl["H"] + 1
    W = E.record_layout(wl["kind"])["width"]
    run = RolloutRunner(wl, E, L, torch, dev, n, rank=rank)
    gat = sharding.RecordGather(T, n * ws, W, torch.float32, dev, dst=0)

    def local_step(i):
        return run.step(i)

    def peer_step(i):
        st = E.rollout(wl["cfg"], run.pol, n, T, device=dev, distr=wl["distr"], seed=i, env_offset=rank * n, auto_reset=True,
                       record=True, **gat.target())["stats"]
        gat.finish()
        return st

    # ---- parity
    local_step(7)
    ref_t, ref_f = sharding.gather_records(run.extra["traj"], run.extra["row_flag"], n * ws, dst=0)
    peer_step(7)
    torch.cuda.synchronize()
    dist.barrier()
    mine_ok = bool(torch.equal(gat.traj[:, gat.lo:gat.hi].view(torch.int32), run.extra["traj"].view(torch.int32)))
    print(f"rank {rank}: its columns of the learner buffer equal its local records: {mine_ok}", file=sys.stderr, flush=True)
    ok = True
    if rank == 0:  # peer_step already ran the barrier on every rank: rank 0's own buffer is complete
        # bit patterns: a lane that diverged holds NaN records on both routes
        bad_t = (gat.traj.view(torch.int32) != ref_t.view(torch.int32)).any(dim=2).any(dim=0)
        bad_f = (gat.row_flag != ref_f).any(dim=0)
        ok = not bool(bad_t.any()) and not bool(bad_f.any())
        if not ok:
            d = (gat.traj.view(torch.int32) != ref_t.view(torch.int32))
            idx = d.nonzero()[0].tolist()
            print("first mismatch (row, lane, col)", idx, float(gat.traj[tuple(idx)]), float(ref_t[tuple(idx)]), "frac", float(d.float().mean()),
                  "ref peer half all zero:", bool((ref_t[:, n:] == 0).all()), file=sys.stderr, flush=True)
            print("mismatching lanes: traj own", int(bad_t[:n].sum()), "peer", int(bad_t[n:].sum()), "flag own",
                  int(bad_f[:n].sum()), "peer", int(bad_f[n:].sum()), file=sys.stderr, flush=True)
    dist.barrier()  # the other ranks must not start the next launches (which write into rank 0's buffer) before the comparison is done
    del ref_t, ref_f

    def timed(fn, k):
        tot = 0.0
        for i in range(k):
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn(20 + i)
            e1.record()
            torch.cuda.synchronize()
            tt = torch.tensor([e0.elapsed_time(e1) * 1e-3], device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            tot += float(tt)
        return tot / k

    def nccl_step(i):
        local_step(i)
        return sharding.gather_records(run.extra["traj"], run.extra["row_flag"], n * ws, dst=0)

    for i in range(2):
        local_step(i), peer_step(i), nccl_step(i)
    t_local, t_peer, t_nccl = timed(local_step, a.reps), timed(peer_step, a.reps), timed(nccl_step, a.reps)
    steps = n * ws * wl["H"]
    rec_bytes = steps * (W * 4 + 1)
    if rank == 0:
        print(json.dumps(dict(workload=wl["label"], n_gpus=ws, envs_per_gpu=n, parity_peer_equals_nccl_gather=ok,
                              env_steps_per_s=dict(local_records=steps / t_local, nccl_gather_after=steps / t_nccl,
                                                   peer_writes_in_kernel=steps / t_peer),
                              ms=dict(local=1e3 * t_local, nccl_gather_after=1e3 * t_nccl, peer=1e3 * t_peer),
                              learner_ingress_gbs=rec_bytes * (ws - 1) / ws / t_peer / 1e9)), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if not ok:
        sys.exit(1)


if __name__ == "__main__":
    main()
