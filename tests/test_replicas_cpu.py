"""N>1 host logic on CPU: replica dispatch policy and the rank aggregation bench.py uses (gloo, world_size 2)."""
import os
import socket
import threading

import pytest

from localai_b200.replicas import Ranks, ReplicaPool


class FakeEngine:
    """Stands in for capi.Engine in the dispatch tests (the policy never looks at the device)."""

    def __init__(self, dev):
        self.dev, self.next_id, self.released, self.closed = dev, 0, [], False

    def submit(self, **req):
        if req.get("boom"):
            raise RuntimeError("submit failed")
        self.next_id += 1
        return self.next_id

    def events(self, rid, timeout_ms=0):
        yield {"text": b"x", "final": False}
        yield {"text": b"", "final": True}

    def release(self, rid):
        self.released.append(rid)

    def cancel(self, rid):
        pass

    def close(self):
        self.closed = True


def make_pool(n):
    return ReplicaPool("unused.gguf", devices=range(n), engine_factory=FakeEngine)


def test_least_loaded_then_lru():
    pool = make_pool(4)
    hs = [pool.submit(prompt="a") for _ in range(4)]
    assert sorted(h[0] for h in hs) == [0, 1, 2, 3]  # one each before anyone gets a second
    assert pool.load() == [1, 1, 1, 1]
    list(pool.events(hs[2]))  # replica 2 drains -> it is the least loaded now
    assert pool.load() == [1, 1, 0, 1]
    assert pool.submit(prompt="b")[0] == 2
    # all equal again: the tie goes to the replica used longest ago (0), not to the lowest index by accident
    for h in (hs[0], hs[1], hs[3]):
        list(pool.events(h))
    assert pool.load() == [0, 0, 1, 0]
    assert pool.submit(prompt="c")[0] == 0
    assert pool.submit(prompt="d")[0] == 1
    assert pool.submit(prompt="e")[0] == 3


def test_release_and_failed_submit_do_not_leak_load():
    pool = make_pool(2)
    with pytest.raises(RuntimeError):
        pool.submit(boom=True)
    assert pool.load() == [0, 0]
    h = pool.submit(prompt="a")
    evs = list(pool.events(h))
    assert evs[-1]["final"] and pool.load() == [0, 0]
    assert pool.engines[h[0]].released == [h[1]]
    pool.close()


def test_concurrent_dispatch_balances():
    pool = make_pool(8)
    picks = []

    def worker():
        for _ in range(50):
            picks.append(pool.pick())

    ts = [threading.Thread(target=worker) for _ in range(8)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert pool.assigned == [50] * 8  # nothing drains, so least-loaded = round robin, exactly


def test_no_device_is_an_error():
    with pytest.raises(RuntimeError):
        ReplicaPool("unused.gguf", devices=[], engine_factory=FakeEngine)


def test_single_rank_is_identity():
    r = Ranks()
    r.barrier()
    assert r.max(3.5) == 3.5 and r.sum(2.0) == 2.0 and r.throughput(10, 2.0) == 5.0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    r = Ranks(rank, world, rank, backend="gloo")
    r.barrier()
    # rank 0 is the slow one: 100 tokens in 2 s; rank 1: 300 tokens in 1 s -> 400 tokens / 2 s
    units, secs = (100.0, 2.0) if rank == 0 else (300.0, 1.0)
    out.put((rank, r.max(secs), r.sum(units), r.throughput(units, secs)))
    r.barrier()
    r.close()


def test_gloo_world2_aggregation():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_rank_main, args=(i, 2, port, q)) for i in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=120) for _ in ps)
    [p.join(60) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    for rank, mx, sm, thr in res:
        assert mx == 2.0 and sm == 400.0 and thr == 200.0
