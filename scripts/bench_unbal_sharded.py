"""Per-iteration time of the row-sharded unbalanced scaling iteration (torchrun, one rank per GPU) against the
single-device solve of the same problem.  Usage:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/bench_unbal_sharded.py [n] [m]

Synthetic data; the time of (200 - 100) iterations, fixed count (stopThr = 0), device events, max over ranks."""
import json
import os
import sys
import warnings

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pot_b200 as ot  # noqa: E402
from pot_b200.sharded import shard_rows, sinkhorn_knopp_unbalanced_sharded  # noqa: E402


def timed(fn):
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
    m = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
    r0, r1 = shard_rows(n, world, rank)
    g = torch.Generator(device=dev).manual_seed(1 + rank)
    Ml = torch.rand(r1 - r0, m, device=dev, generator=g, dtype=torch.float32)
    al = torch.full((r1 - r0,), 1.0 / n, device=dev)
    b = torch.full((m,), 1.2 / m, device=dev)
    out = {"n": n, "m": m, "dtype": "f32", "ranks": world}
    warnings.simplefilter("ignore")
    for exchange in ("peer", "nccl"):
        def run(its):
            sinkhorn_knopp_unbalanced_sharded(al, b, Ml, 0.1, 1.0, numItermax=its, stopThr=0.0, group=dist.group.WORLD,
                                              exchange=exchange, return_plan=False)
        run(20)
        t1 = min(timed(lambda: run(100)) for _ in range(2))
        t2 = min(timed(lambda: run(200)) for _ in range(2))
        out[f"us_per_iter_{exchange}"] = round((t2 - t1) * 10.0, 2)
    if rank == 0:  # the same rows on one device, no exchange
        def run1(its):
            ot.unbalanced.sinkhorn_knopp_unbalanced(al, b, Ml, 0.1, 1.0, numItermax=its, stopThr=0.0)
        run1(20)
    t1 = min(timed(lambda: run1(100) if rank == 0 else None) for _ in range(2))
    t2 = min(timed(lambda: run1(200) if rank == 0 else None) for _ in range(2))
    out["us_per_iter_single_device_local_rows"] = round((t2 - t1) * 10.0, 2)
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
