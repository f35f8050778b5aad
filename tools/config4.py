"""BASELINE config 4's grid across GPUs (under torchrun): 2-D 9-point box fp32 on 16384 x 16384,
y-slabs of 16384/N rows, ghost rows exchanged every fused block of k = 4 steps.

    torchrun --nproc-per-node N tools/config4.py [--exchange p2p|nccl] [--steps T] [--reps R]

Prints one JSON line (rank 0): GStencil/s (device time, max over ranks) and the checksum of the
whole grid — equal across N and across exchange modes (bitwise independence of the GPU count)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from stencil_b200 import capi, slab

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=48)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--time-block", type=int, default=0)
ap.add_argument("--exchange", default="auto")
ap.add_argument("--n", type=int, default=16384)
args = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
capi.check(capi.lib().sb200_init(local))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
d = capi.make_desc((args.n, args.n // world), 1, capi.BOX, capi.F32, args.steps, time_block=args.time_block)
sw = slab.SlabSweep(d, rank, world, exchange=args.exchange)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


best = 1e30
for rep in range(args.reps + 1):
    sw.fill(capi.INIT_RANDOM, seed=1)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    sw.sweep(args.steps)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rep > 0:  # rep 0 is the warm-up
        best = min(best, float(ms.item()))
c = sw.checksum()   # raises if a fused exchange timed out
cs = torch.tensor([c & 0x7FFFFFFF, (c >> 31) & 0x7FFFFFFF, c >> 62], dtype=torch.int64, device="cuda")
if world > 1:
    dist.all_reduce(cs)
total = (int(cs[0]) + (int(cs[1]) << 31) + (int(cs[2]) << 62)) % (1 << 64)
if rank == 0:
    gs = args.n * args.n * args.steps / best / 1e6
    print(json.dumps({"config": 4, "n_gpus": world, "grid": [args.n, args.n], "steps": args.steps,
                      "time_block": sw.time_block, "exchange": sw.exchange, "ms": best, "GStencil/s": gs,
                      "alg_GBps_per_gpu": gs * 8 / world, "checksum": f"{total:#018x}"}), flush=True)
sw.close()
if world > 1:
    dist.destroy_process_group()
