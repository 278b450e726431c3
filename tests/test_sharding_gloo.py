"""Multi-GPU host logic on CPU: two gloo ranks shard the frame sequence exactly like bench.py does
(rank r owns frames [r*B, (r+1)*B), no data-path collective) and the max-over-ranks timing
reduction works."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, B, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from bench import shifts_for
    from orb_slam3_vio_fixes_b200 import synthetic as S
    t0 = rank * B
    frames = S.make_frames("C3", B, t0=t0, width=160, height=120)
    sh = shifts_for(t0, B)
    t = torch.tensor([1.0 + rank], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    chk = torch.tensor([int(frames.astype(np.int64).sum()), int(np.abs(sh).sum() * 10)], dtype=torch.int64)
    gathered = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(gathered, chk)
    dist.barrier()
    if rank == 0:
        out.put((float(t.item()), [g.tolist() for g in gathered]))
    dist.destroy_process_group()


def test_two_rank_sharding():
    world, B = 2, 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, B, q)) for r in range(world)]
    for p in procs:
        p.start()
    tmax, sums = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert tmax == 2.0  # MAX over ranks
    from orb_slam3_vio_fixes_b200 import synthetic as S
    whole = S.make_frames("C3", world * B, width=160, height=120)
    for r in range(world):  # each rank generated exactly its own stretch of the same sequence
        assert sums[r][0] == int(whole[r * B:(r + 1) * B].astype(np.int64).sum())
    assert sums[0][0] != sums[1][0]
