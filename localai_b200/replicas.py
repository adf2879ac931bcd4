"""Replica serving and rank aggregation: the whole multi-GPU story of this path ("replicas only").

Every model of configs 1-4 fits one B200, requests are independent, and LocalAI keeps ONE backend process per model
id (pkg/model/loader.go:117-141).  So more GPUs = more engines inside that process, one per device, each with its own
weights, KV pool and stream; a new request goes to the least-loaded engine -- the policy the reference uses one level
down for slots (`get_slot`, grpc-server.cpp:609-628: a free slot, the least recently used one on a tie) and one level
up for federated workers (`SelectLeastUsedServer`, core/p2p/federated.go:77-99).  There is no data-path collective.

`ReplicaPool` is that dispatcher (the Go shim in backend/go/llm/b200 holds the same few lines); the functions below it
are the torch.distributed plumbing bench.py uses when the driver launches one rank per GPU: a barrier and a
max-over-ranks time, nothing else.
"""
from __future__ import annotations

import itertools
import threading
from typing import Callable, Iterator, Sequence


class ReplicaPool:
    """One engine per device; least-loaded dispatch; request handles are (replica, request id)."""

    def __init__(self, model_path: str, devices: Sequence[int] | None = None, engine_factory: Callable | None = None, **opts):
        if engine_factory is None:
            from . import capi

            if devices is None:
                devices = range(capi.device_count())
            engine_factory = lambda dev: capi.Engine(model_path, device=dev, **opts)  # noqa: E731
        devices = list(devices if devices is not None else [0])
        if not devices:
            raise RuntimeError("ReplicaPool: no sm_100 device visible (there is no CPU fallback)")
        self.devices = devices
        self.engines = [engine_factory(d) for d in devices]
        self._lock = threading.Lock()
        self._inflight = [0] * len(devices)
        self._last_used = [0] * len(devices)
        self._tick = itertools.count(1)
        self.assigned = [0] * len(devices)

    # -- dispatch -----------------------------------------------------------------------------------------------
    def pick(self) -> int:
        """Least in-flight requests; ties go to the replica used longest ago (LRU, as get_slot does for slots)."""
        with self._lock:
            i = min(range(len(self.engines)), key=lambda j: (self._inflight[j], self._last_used[j]))
            self._inflight[i] += 1
            self._last_used[i] = next(self._tick)
            self.assigned[i] += 1
            return i

    def _done(self, i: int):
        with self._lock:
            self._inflight[i] -= 1

    def submit(self, **req) -> tuple[int, int]:
        i = self.pick()
        try:
            return i, self.engines[i].submit(**req)
        except Exception:
            self._done(i)
            raise

    def events(self, handle: tuple[int, int], timeout_ms: int = 60000) -> Iterator[dict]:
        i, rid = handle
        try:
            yield from self.engines[i].events(rid, timeout_ms)
        finally:
            self.engines[i].release(rid)
            self._done(i)

    def cancel(self, handle: tuple[int, int]):
        self.engines[handle[0]].cancel(handle[1])

    def tokenize(self, text, **kw):
        return self.engines[0].tokenize(text, **kw)  # every replica holds the same vocabulary

    def load(self) -> list[int]:
        with self._lock:
            return list(self._inflight)

    def close(self):
        for e in self.engines:
            e.close()
        self.engines = []


# ---- rank plumbing for `torchrun bench.py --gpus N` (one replica per rank) -------------------------------------------
class Ranks:
    """Barrier + reductions over the launched ranks.  backend "nccl" on the GPU box, "gloo" in the CPU tests."""

    def __init__(self, rank: int = 0, world: int = 1, local_rank: int = 0, backend: str | None = None):
        self.rank, self.world, self.local_rank = rank, world, local_rank
        self.dist = None
        self.device = "cpu"
        if world > 1:
            import torch
            import torch.distributed as dist

            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            if not dist.is_initialized():
                if backend == "nccl":
                    torch.cuda.set_device(local_rank)
                    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
                else:
                    dist.init_process_group(backend)
            self.dist = dist
            self.device = f"cuda:{local_rank}" if backend == "nccl" else "cpu"

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def host_barrier(self):
        """A barrier that keeps the GPUs free while ranks wait (a gloo group): for the legs where ONE rank drives several
        devices -- the waiting ranks' NCCL barrier kernel would otherwise spin on the very GPUs being measured."""
        if self.dist is None:
            return
        if getattr(self, "_gloo", None) is None:
            self._gloo = self.dist.new_group(backend="gloo")
        self.dist.barrier(group=self._gloo)

    def _reduce(self, x: float, op: str) -> float:
        if self.dist is None:
            return float(x)
        import torch

        t = torch.tensor([x], dtype=torch.float64, device=self.device)
        self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op))
        return float(t.item())

    def max(self, x: float) -> float:
        return self._reduce(x, "MAX")

    def sum(self, x: float) -> float:
        return self._reduce(x, "SUM")

    def throughput(self, units_local: float, seconds_local: float) -> float:
        """Whole-job rate: the units ALL ranks processed / the slowest rank's time (never a mean of rates)."""
        return self.sum(units_local) / self.max(seconds_local)

    def close(self):
        if self.dist is not None and self.dist.is_initialized():
            self.dist.destroy_process_group()
            self.dist = None
