"""N>1 host logic on CPU: two gloo ranks exchange key-hash-partitioned records with the same
exchange functions the GPU path uses (data-accelerator_b200/parallel.py); the partitioning and the
aggregation on each rank are done by the oracle, so no GPU is needed.  The union of the per-rank
aggregates must equal the single-process aggregate (keys are owned by exactly one rank)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _make_records(rank, n=20_000):
    rng = np.random.default_rng(100 + rank)
    dev = rng.integers(1, 500, n)
    home = dev % 37
    ts = rng.integers(0, 10**6, n)
    val = rng.integers(0, 2, n)
    valid = np.full(n, 0b1111, dtype=np.int64)
    null_dev = rng.random(n) < 0.01
    valid[null_dev] &= ~1
    dev = np.where(null_dev, 0, dev)
    return np.stack([dev, home, ts, val, valid], axis=1).astype(np.int64)      # 4 columns + validity word


def _agg(recs):
    out = {}
    for dev, home, ts, val, valid in recs:
        k = (int(dev) if valid & 1 else None, int(home))
        c = out.setdefault(k, [0, 0, None])
        c[0] += 1
        c[1] += int(val)
        c[2] = int(ts) if c[2] is None else max(c[2], int(ts))
    return out


def _worker(rank, world, port, q):
    for p in (ROOT, os.path.join(ROOT, "oracle")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import dxgpu
    import dxo_flow
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    recs = _make_records(rank)
    owner = dxo_flow.owner_of([(recs[:, 0], (recs[:, 4] & 1).astype(bool)), (recs[:, 1], np.ones(len(recs), bool))], world)
    order = np.argsort(owner, kind="stable")
    send = torch.from_numpy(np.ascontiguousarray(recs[order]).reshape(-1))
    counts = np.bincount(owner, minlength=world)
    recv_counts = dxgpu.parallel.exchange_counts(counts)
    recv = dxgpu.parallel.exchange_records(send, counts, recv_counts, 5).numpy().reshape(-1, 5)
    assert len(recv) == recv_counts.sum()
    # every received record is owned by this rank
    own2 = dxo_flow.owner_of([(recv[:, 0], (recv[:, 4] & 1).astype(bool)), (recv[:, 1], np.ones(len(recv), bool))], world)
    assert (own2 == rank).all()
    q.put((rank, _agg(recv), int(counts.sum()), int(recv_counts.sum())))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_key_hash_exchange_over_gloo():
    world, port = 2, 29533
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    merged = {}
    sent = recv = 0
    for rank, agg, s, r in results:
        assert not (set(agg) & set(merged)), "a key landed on two ranks"
        merged.update(agg)
        sent += s
        recv += r
    assert sent == recv == 40_000
    expected = _agg(np.concatenate([_make_records(r) for r in range(world)]))
    assert merged == expected


def test_block_offsets_tile_every_receive_buffer():
    """Fused put bookkeeping: the blocks the senders are assigned in one receiver's buffer are disjoint, in sender order,
    and cover exactly the records that receiver is due."""
    import numpy as np
    import dxgpu as dx
    rng = np.random.default_rng(3)
    for world in (1, 2, 8):
        M = rng.integers(0, 1000, size=(world, world))
        firsts = [dx.parallel.block_offsets(M, r) for r in range(world)]
        for d in range(world):
            at = 0
            for s in range(world):
                assert firsts[s][0][d] == at
                at += M[s][d]
            assert at == firsts[0][1][d] == M[:, d].sum()
