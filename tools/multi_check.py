"""Multi-GPU parity check (run under torchrun on N GPUs): the slab-decomposed sweep must be
bitwise identical to the same grid swept on ONE GPU (checksum of checksums), for the 3-D
TMA kernel and the 2-D warp kernel, including a remainder block."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from stencil_b200 import capi, slab, sweep

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
capi.check(capi.lib().sb200_init(local))
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ok = True
CASES = [  # n_local, radius, shape, dtype, T, k, exchange
    ((256, 200, 64), 1, capi.STAR, capi.F64, 7, 2, "p2p"),
    ((256, 200, 64), 1, capi.STAR, capi.F64, 7, 2, "nccl"),
    ((256, 200, 64), 1, capi.STAR, capi.F32, 6, 3, "p2p"),
    ((130, 61, 12), 1, capi.STAR, capi.F64, 31, 2, "p2p"),
    ((1000, 256), 1, capi.BOX, capi.F32, 19, 4, "nccl"),
    ((1000, 256), 1, capi.BOX, capi.F32, 19, 4, "p2p"),    # 2-D fused exchange: faces are rows
    ((4096, 512), 1, capi.STAR, capi.F64, 26, 6, "p2p"),
    ((300, 128, 40), 1, capi.BOX, capi.F64, 3, 1, "nccl"),
    ((300, 128, 40), 1, capi.BOX, capi.F64, 9, 2, "p2p"),
    ((1000, 256), 2, capi.STAR, capi.F32, 11, 4, "p2p"),   # radius > 1: warp2dr, faces are r*k rows
    ((1000, 256), 2, capi.STAR, capi.F32, 11, 4, "nccl"),
    ((900, 200), 2, capi.BOX, capi.F32, 7, 2, "p2p"),      # 25-point box
    ((640, 256), 3, capi.STAR, capi.F64, 5, 2, "p2p"),
    ((200, 160, 48), 2, capi.STAR, capi.F64, 5, 1, "nccl"),  # 13-point star (tma3dr)
    ((200, 160, 48), 2, capi.STAR, capi.F64, 5, 1, "p2p"),
    ((200, 160, 48), 2, capi.BOX, capi.F32, 4, 1, "p2p"),    # 125-point box (tma3db)
    ((200, 160, 48), 3, capi.STAR, capi.F64, 4, 1, "p2p"),   # 19-point star (fp64: where AUTO takes the TMA kernel)
    ((200, 160, 48), 2, capi.BOX, capi.F32, 3, 1, "nccl"),   # 125-point box: generic kernel, NCCL only
]
for n_local, radius, shape, dtype, T, k, exch in CASES:
    d = capi.make_desc(n_local, radius, shape, dtype, T, time_block=k)
    sw = slab.SlabSweep(d, rank, world, exchange=exch)
    sw.fill(capi.INIT_RANDOM, seed=3)
    sw.sweep(T)
    if exch == "p2p" and sw.eng.link_timed_out():
        print(f"rank {rank}: a face CTA timed out waiting for its neighbour", flush=True)
        ok = False
    cs = torch.tensor([sw.checksum() & 0x7FFFFFFFFFFFFFFF, sw.checksum() >> 63], dtype=torch.int64, device="cuda")
    parts = [torch.zeros_like(cs) for _ in range(world)]
    dist.all_gather(parts, cs)
    total = sum(int(p[0].item()) + (int(p[1].item()) << 63) for p in parts) % (1 << 64)
    if rank == 0:
        n_full = list(n_local)
        n_full[-1] *= world
        df = capi.make_desc(n_full, radius, shape, dtype, T, time_block=k)
        g = sweep.DeviceGrids(df)
        g.fill(capi.INIT_RANDOM, seed=3)
        g.sweep(T)
        want = g.checksum()
        good = want == total
        ok &= good
        print(f"n_local={n_local} r={radius} shape={shape} dtype={dtype} T={T} k={k} world={world} exchange={exch}: "
              f"{'bitwise identical' if good else 'MISMATCH'} ({want:#x} vs {total:#x})", flush=True)
        del g
    dist.barrier()
    sw.close()
okt = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(okt, op=dist.ReduceOp.MIN)
ok = bool(okt.item())
dist.destroy_process_group()
sys.exit(0 if ok else 1)
