"""K5 (key-hash pack/unpack) on one GPU, with the ranks of an 8-GPU box emulated as partitions:
ownership must agree bit-for-bit with the oracle's hash, and running the windowed flow per partition
must give the same groups as running it on everything (each key lives on exactly one rank)."""
import numpy as np
import pytest

import dxo
import dxo_flow
from flowcmp import gpu_rows

pytestmark = pytest.mark.gpu

PROJ = "CAST(Raw.deviceDetails.eventTime / 1000 AS timestamp) AS eventTimeStamp\nRaw.*\n"
T0, I = 1_700_000_040_000, 60_000
TR = """--DataXQuery--
Tumbling = SELECT deviceId, homeId, COUNT(*) AS Count, SUM(status) AS S, MAX(eventTimeStamp) AS Last FROM DataXProcessedInput GROUP BY deviceId, homeId
--DataXQuery--
Sliding = SELECT deviceId, homeId, COUNT(*) AS Count, MIN(status) AS Mn FROM DataXProcessedInput_5minutes GROUP BY deviceId, homeId
"""
COLS = ["deviceDetails.deviceId", "deviceDetails.homeId", "eventTimeStamp", "deviceDetails.status"]
RECV_NAMES = ["deviceId", "homeId", "eventTimeStamp", "status"]
RECV_TYPES = [1, 1, 5, 1]


def test_pack_ownership_matches_oracle_and_sharded_windows_equal_unsharded(dx, ctx):
    world = 4
    conf = dx.flowgen.conf_text("shard", timestamp_column="eventTimeStamp", watermark="60 second",
                                windows=[("DataXProcessedInput_5minutes", "5 minutes")])
    schema = dxo.iot_schema_json(1)
    decode_flow = dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("dec"), schema, [PROJ], "")
    whole = dx.CommonProcessorFactory.createProcessor(ctx, conf, schema, ["eventTimeStamp\ndeviceId\nhomeId\nstatus\n"], TR)   # never decodes: route only
    shards = [dx.CommonProcessorFactory.createProcessor(ctx, conf, schema, [PROJ], TR) for _ in range(world)]
    for b in range(5):
        T = T0 + b * I
        p = dxo.GenParams(77, 3000, 300, T, I, 60_000, 1, 20, 5, 0)
        data = np.frombuffer(dxo.gen_events(p, b * 20_000, 20_000), dtype=np.uint8)
        src = ctx.alloc(data.size + 64)
        ctx.h2d(src, data)
        projected = dx.parallel.project_device(decode_flow, src, data.size, 20_000, T)
        n = projected.num_rows
        buf = ctx.alloc(n * 8 * (len(COLS) + 1))
        counts, rec_bytes = dx.parallel.pack_by_key(projected, COLS, COLS[:2], world, buf)
        assert rec_bytes == 40 and counts.sum() == n
        recs = ctx.d2h(buf, n * rec_bytes).view(np.int64).reshape(n, 5)
        # ownership: every record sits in the partition the oracle's hash assigns
        own = dxo_flow.owner_of([(recs[:, 0], (recs[:, 4] & 1).astype(bool)), (recs[:, 1], ((recs[:, 4] >> 1) & 1).astype(bool))], world)
        bounds = np.concatenate([[0], np.cumsum(counts)])
        for r in range(world):
            assert (own[bounds[r]:bounds[r + 1]] == r).all()
        # the multiset of records equals the projected rows
        cols = [projected.column(c)[0] for c in COLS]
        exp = np.stack(cols, axis=1)
        assert sorted(map(tuple, recs[:, :4].tolist())) == sorted(map(tuple, exp.tolist()))
        # per-partition flows vs the whole
        all_recv = dx.parallel.table_from_packed(ctx, RECV_NAMES, RECV_TYPES, buf, n)
        rw = dx.parallel.process_table(whole, all_recv, T, I)
        merged = {"Tumbling": [], "Sliding": []}
        for r in range(world):
            part = dx.parallel.table_from_packed(ctx, RECV_NAMES, RECV_TYPES, buf + int(bounds[r]) * rec_bytes, int(counts[r]))
            rr = dx.parallel.process_table(shards[r], part, T, I)
            for name in merged:
                merged[name] += gpu_rows(rr.table(name), dx)
            rr.release()
        for name in merged:
            assert sorted(merged[name]) == sorted(gpu_rows(rw.table(name), dx)), (b, name)
        rw.release()
        ctx.free(buf)
        ctx.free(src)


def test_shuffle_refuses_rank_local_columns(dx, ctx):
    schema = dxo.iot_schema_json(1)
    f = dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("dec"), schema, ["Raw.*\n"], "")
    data = np.frombuffer(dxo.gen_events(dxo.GenParams(1, 10, 5, T0, I, 60_000, 1, 0, 0, 0), 0, 100), dtype=np.uint8)
    src = ctx.alloc(data.size + 64)
    ctx.h2d(src, data)
    t = dx.parallel.project_device(f, src, data.size, 100, T0)
    buf = ctx.alloc(100 * 64)
    with pytest.raises(dx.DxError):
        dx.parallel.pack_by_key(t, ["deviceDetails.deviceId", "deviceDetails.deviceType"], ["deviceDetails.deviceId"], 2, buf)
    ctx.free(buf)
    ctx.free(src)


def test_fused_put_lands_every_record_in_its_owners_buffer(dx, ctx):
    """The fused pack+exchange kernel with the four owners' receive buffers emulated on one GPU, and two "sending ranks"
    writing their blocks one after the other at the offsets the count exchange assigns: every buffer ends up with
    exactly the records its owner is due (the same multiset the separate pack produces), nothing overlaps."""
    world = 4
    schema = dxo.iot_schema_json(1)
    f = dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("dec"), schema, [PROJ], "")
    tables, srcs = [], []
    for s in range(2):      # two senders with different events
        p = dxo.GenParams(500 + s, 3000, 300, T0, I, 60_000, 1, 20, 5, 0)
        data = np.frombuffer(dxo.gen_events(p, 0, 15_000 + 1000 * s), dtype=np.uint8)
        src = ctx.alloc(data.size + 64)
        ctx.h2d(src, data)
        srcs.append(src)
        tables.append(dx.parallel.project_device(f, src, data.size, 15_000 + 1000 * s, T0))
    M = np.stack([dx.parallel.count_by_key(t, COLS, COLS[:2], world) for t in tables])      # M[s][d]
    assert (M.sum(axis=1) == [t.num_rows for t in tables]).all()
    recv_total = M.sum(axis=0)
    bufs = [ctx.alloc(int(recv_total[d]) * 40 + 64) for d in range(world)]
    for s, t in enumerate(tables):
        first = M[:s].sum(axis=0) if s else np.zeros(world, dtype=np.int64)
        dx.parallel.put_by_key(t, COLS, COLS[:2], world, bufs, first)
    for d in range(world):
        n = int(recv_total[d])
        recs = ctx.d2h(bufs[d], n * 40).view(np.int64).reshape(n, 5)
        own = dxo_flow.owner_of([(recs[:, 0], (recs[:, 4] & 1).astype(bool)), (recs[:, 1], ((recs[:, 4] >> 1) & 1).astype(bool))], world)
        assert (own == d).all()
        # the separate pack gives the same multiset for this owner
        want = []
        for t in tables:
            buf = ctx.alloc(t.num_rows * 40 + 64)
            counts, _ = dx.parallel.pack_by_key(t, COLS, COLS[:2], world, buf)
            b = np.concatenate([[0], np.cumsum(counts)])
            allr = ctx.d2h(buf, t.num_rows * 40).view(np.int64).reshape(t.num_rows, 5)
            want += list(map(tuple, allr[b[d]:b[d + 1]].tolist()))
            ctx.free(buf)
        assert sorted(map(tuple, recs.tolist())) == sorted(want)
    for b in bufs + srcs:
        ctx.free(b)


def test_put_into_sender_blocks_and_unpack_from_blocks(dx, ctx):
    """The exchange without a count round: each of two senders owns a fixed block in each of four owners' buffers, puts
    without knowing the counts, reports what it sent, and the owner unpacks the blocks in sender order.  The unpacked table
    must hold exactly the records the separate pack assigns to that owner; a block that is too small fails the call and
    writes nothing outside it."""
    world, senders, block = 4, 2, 6000
    schema = dxo.iot_schema_json(1)
    f = dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("dec"), schema, [PROJ], "")
    tables, srcs = [], []
    for s in range(senders):
        p = dxo.GenParams(900 + s, 3000, 300, T0, I, 60_000, 1, 20, 5, 0)
        n = 14_000 + 3000 * s
        data = np.frombuffer(dxo.gen_events(p, 0, n), dtype=np.uint8)
        src = ctx.alloc(data.size + 64)
        ctx.h2d(src, data)
        srcs.append(src)
        tables.append(dx.parallel.project_device(f, src, data.size, n, T0))
    bufs = [ctx.alloc(senders * block * 40 + 64) for _ in range(world)]
    sent = []
    for s, t in enumerate(tables):
        sent.append(dx.parallel.put_blocks_by_key(t, COLS, COLS[:2], world, bufs, [s * block] * world, [block] * world))
        assert (sent[-1] == dx.parallel.count_by_key(t, COLS, COLS[:2], world)).all()
    M = np.stack(sent)
    for d in range(world):
        part = dx.parallel.table_from_packed_blocks(ctx, RECV_NAMES, RECV_TYPES, bufs[d], [s * block for s in range(senders)], M[:, d])
        assert part.num_rows == int(M[:, d].sum())
        got = sorted(gpu_rows(part, dx))
        want = []
        for t in tables:
            buf = ctx.alloc(t.num_rows * 40 + 64)
            counts, _ = dx.parallel.pack_by_key(t, COLS, COLS[:2], world, buf)
            b = np.concatenate([[0], np.cumsum(counts)])
            one = dx.parallel.table_from_packed(ctx, RECV_NAMES, RECV_TYPES, buf + int(b[d]) * 40, int(counts[d]))
            want += gpu_rows(one, dx)
            one.release(); ctx.free(buf)
        assert got == sorted(want)
        part.release()
    # a block one record too small for the fullest owner: the call fails, the record after the block stays untouched
    tight = int(sent[0].max()) - 1
    guard = ctx.alloc((tight + 1) * 40 + 64)
    ctx.h2d(guard, np.full((tight + 1) * 5, -7, dtype=np.int64))
    with pytest.raises(dx.DxError) as e:
        dx.parallel.put_blocks_by_key(tables[0], COLS, COLS[:2], world, [guard] * world, [0] * world, [tight] * world)
    assert "exceed this rank's block" in str(e.value)
    assert (ctx.d2h(guard + tight * 40, 40).view(np.int64) == -7).all()
    for b in bufs + srcs + [guard]:
        ctx.free(b)
    f.close()
