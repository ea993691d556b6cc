"""world_size-2 gloo test of the N > 1 host logic: sharding of independent videos across ranks,
barrier, max-over-ranks timing and whole-job throughput -- what bench.py does under torchrun."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from rnb_b200 import dist as rdist
    from rnb_b200.handoff import segment_ranges
    videos = list(range(11))
    lo, hi = rdist.shard_range(len(videos), rank, world)
    mine = videos[lo:hi]
    rdist.barrier()
    # rank 1 is "slower": the job time is the max over ranks, the work is the sum
    ms = 10.0 * (rank + 1)
    clips = sum(15 if v % 11 == 0 else 1 for v in mine)
    value, job_ms = rdist.whole_job_throughput(clips, ms)
    assert segment_ranges(len(videos), world)[rank] == (lo, hi)
    out_q.put((rank, mine, value, job_ms))
    rdist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_replicas():
    world = 2
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, v0, val0, ms0), (r1, v1, val1, ms1) = results
    assert v0 + v1 == list(range(11)) and len(v0) == 6 and len(v1) == 5   # remainder to rank 0
    assert ms0 == ms1 == 20.0                                              # max over ranks
    total_clips = 15 + 10
    assert abs(val0 - total_clips / 20e-3) < 1e-6 and val0 == val1         # whole-job aggregate


def test_single_process_helpers_need_no_process_group():
    from rnb_b200 import dist as rdist
    assert rdist.shard_range(10, 0, 3) == (0, 4) and rdist.shard_range(10, 2, 3) == (7, 10)
    value, ms = rdist.whole_job_throughput(160, 10.0)
    assert ms == 10.0 and abs(value - 16000.0) < 1e-9
    rdist.barrier()
