"""CPU, world_size 2 over gloo: the multi-GPU path's host logic (round-robin stream sharding, result gather, no
collective on the data path)."""
import os
import socket

import torch.multiprocessing as mp


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from aha_b200.sharding import gather_stream_results, shard_streams
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = shard_streams(5, world, rank)
    local = {s: [dict(time=float(t), relevance_score=s + 0.1 * t) for t in range(3)] for s in mine}
    merged = gather_stream_results(local, dist, dst=0)
    dist.barrier()
    if rank == 0:
        q.put((mine, sorted(merged), merged[3][2]["relevance_score"]))
    else:
        assert merged is None
        q.put((mine, None, None))
    dist.destroy_process_group()


def test_two_rank_stream_sharding_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort(key=lambda r: r[0])
    assert res[0][0] == [0, 2, 4] and res[1][0] == [1, 3]
    full = [r for r in res if r[1] is not None][0]
    assert full[1] == [0, 1, 2, 3, 4] and abs(full[2] - 3.2) < 1e-12
