"""N > 1 host path on CPU: two gloo ranks shard one request stream, each serves its shard with its
own replica (fake engine), rank 0 gathers; the result equals the single-replica run and the
timing reduction is the max over ranks."""
import os
import socket
import sys

import torch.multiprocessing as mp

from conftest import ROOT, pkg

sharding = pkg("sharding")


def test_shard_bounds_are_contiguous_balanced_and_cover_everything():
    for n in (0, 1, 7, 64, 65, 257):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, queue):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist

    from conftest import pkg as _pkg
    from fakes import FakeEngine

    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh, sched, eng_mod = _pkg("sharding"), _pkg("scheduler"), _pkg("engine")
    stream = [[i, 3 * i + 1, 7] for i in range(37)]
    mine = sh.shard_requests(stream, rank, world)
    router = sched.Router([sched.EngineWorker(FakeEngine(max_seqs=4))])
    sp = eng_mod.SamplingParams(max_new_tokens=5)
    futs = [router.submit(p, sp) for p in mine]
    tokens = [f.result(timeout=20).tokens for f in futs]
    router.shutdown()
    dist.barrier()
    times = sh.max_over_ranks([1.0 + rank, 5.0 - rank])
    gathered = sh.gather_to_rank0(tokens)
    if rank == 0:
        queue.put(([t for part in gathered for t in part], times))
    dist.barrier()
    dist.destroy_process_group()


def test_two_gloo_ranks_reproduce_the_single_replica_result():
    from fakes import FakeEngine

    sched, eng_mod = pkg("scheduler"), pkg("engine")
    stream = [[i, 3 * i + 1, 7] for i in range(37)]
    router = sched.Router([sched.EngineWorker(FakeEngine(max_seqs=4))])
    sp = eng_mod.SamplingParams(max_new_tokens=5)
    expected = [router.generate(p, sp, timeout=20).tokens for p in stream]
    router.shutdown()

    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, queue)) for r in range(2)]
    [p.start() for p in procs]
    got, times = queue.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert got == expected
    assert times == [2.0, 5.0]
