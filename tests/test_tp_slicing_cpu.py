"""The slicing arithmetic of the tensor-parallel loader (csrc/model.cu: Model::Model with tp_size),
checked on the CPU with two gloo ranks and the oracle's block dots:

  * column-parallel matrices (wq | wk | wv | ffn_gate | ffn_up | output): rank r keeps rows [r R/tp, (r+1) R/tp); the
    ranks' outputs concatenated are the full output, bit for bit;
  * row-parallel matrices (attn_output | ffn_down): rank r keeps the K-slice [r K/tp, (r+1) K/tp) of every row -- whole
    256-weight blocks, so the slice of the quantised activations it multiplies is exactly the full row's slice -- and the
    all-reduced sum of the partial rows is the full product up to the fp32 summation order.
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import torch
    import torch.distributed as dist

    import make_gguf as mg
    from oracle import oracle as O

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        res = {}
        for wtype, name in ((12, "q4_k"), (14, "q6_k"), (8, "q8_0")):
            bs, bb = mg.BLOCK[wtype]
            rows, K = 64, 2048
            rng = np.random.default_rng(wtype)  # the same matrix and activations on every rank
            raw = mg.random_blocks(rng, wtype, rows * K // bs, 1 / np.sqrt(K)).reshape(rows, K // bs, bb)
            x = rng.standard_normal((1, K)).astype(np.float32)
            full = O.mul_mat(wtype, raw.reshape(-1), rows, K, x)
            # column-parallel: my rows
            r0, r1 = rank * rows // world, (rank + 1) * rows // world
            mine = O.mul_mat(wtype, raw[r0:r1].reshape(-1), r1 - r0, K, x)
            parts = [torch.zeros(1, rows // world) for _ in range(world)]
            dist.all_gather(parts, torch.from_numpy(mine.copy()))
            res[name + "_col"] = bool(np.array_equal(torch.cat(parts, 1).numpy(), full))
            # row-parallel: my K-slice of every row (blocks [k0/bs, k1/bs)) times my slice of the activations
            k0, k1 = rank * K // world, (rank + 1) * K // world
            part = O.mul_mat(wtype, np.ascontiguousarray(raw[:, k0 // bs:k1 // bs]).reshape(-1), rows, k1 - k0, np.ascontiguousarray(x[:, k0:k1]))
            t = torch.from_numpy(part.copy())
            dist.all_reduce(t)
            res[name + "_row_err"] = float(np.abs(t.numpy() - full).max() / np.abs(full).max())
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def test_sharded_products_recombine_world_size_2():
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    out = dict(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank in (0, 1):
        r = out[rank]
        for name in ("q4_k", "q6_k", "q8_0"):
            assert r[name + "_col"], (rank, name)
            assert r[name + "_row_err"] <= 2e-6, (rank, name, r[name + "_row_err"])  # integer block dots are exact; fp32 sums differ in order only
