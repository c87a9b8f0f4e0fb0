"""pgpu_comm rendezvous (host logic, no GPU): ranks are separate processes that meet in a POSIX shared-memory segment;
device < 0 joins the rendezvous only. The device side is covered by tests/test_gpu_exchange.py."""
import multiprocessing as mp
import os
import time

import pytest

from presto_b200 import exchange as X
from presto_b200.operators import TrinoException


def _worker(rank, world, name, q):
    try:
        if rank == 1:
            time.sleep(0.3)                       # the others have to wait for the late rank inside pgpu_comm_init
        before = time.time()
        c = X.Comm(world, rank, -1, name)
        after = time.time()
        for _ in range(50):
            c.barrier()
        c.close()
        q.put((rank, before, after, None))
    except BaseException as e:                    # noqa: BLE001
        q.put((rank, 0.0, 0.0, repr(e)))


@pytest.mark.parametrize("world", [2, 3])
def test_ranks_rendezvous_and_barrier(world):
    name = "pgpu_test_comm_%d_%d" % (os.getpid(), time.time_ns() % 1_000_000_000)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, name, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert all(err is None for _, _, _, err in got), got
    assert min(after for _, _, after, _ in got) >= max(before for _, before, _, _ in got) - 1e-3   # nobody left init before the last rank entered it
    assert not os.path.exists("/dev/shm/" + name)            # rank 0 unlinked the name once everybody had it mapped


def test_single_rank_needs_nobody():
    c = X.Comm(1, 0, -1, "pgpu_test_comm_solo_%d" % os.getpid())
    c.barrier()
    c.close()


def test_invalid_arguments():
    with pytest.raises(TrinoException):
        X.Comm(2, 2, -1, "pgpu_test_comm_bad")
    with pytest.raises(TrinoException):
        X.Comm(0, 0, -1, "pgpu_test_comm_bad")
