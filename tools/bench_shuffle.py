#!/usr/bin/env python3
"""C5-style multi-GPU pipeline under torchrun (one process per GPU, NCCL over NVLink):
every rank decodes + projects its own events, packs (deviceId, homeId, eventTimeStamp, status) records
by owner = hash64(deviceId, homeId) % G (K5), exchanges them with one NCCL all-to-all, unpacks, and runs
tumbling + 5-minute sliding group-bys on the keys it owns.  Prints one JSON line on rank 0:
events/s, the all-to-all's achieved GB/s per GPU per direction against NVLink, K5 kernel times.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
      tools/bench_shuffle.py --batches 10 --events-per-batch 20000000
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

T0, I = 1_700_000_040_000, 60_000
PROJ = "CAST(Raw.deviceDetails.eventTime / 1000 AS timestamp) AS eventTimeStamp\nRaw.*\n"
COLS = ["deviceDetails.deviceId", "deviceDetails.homeId", "eventTimeStamp", "deviceDetails.status"]
NAMES, TYPES = ["deviceId", "homeId", "eventTimeStamp", "status"], [1, 1, 5, 1]
TR = """--DataXQuery--
Tumbling = SELECT deviceId, homeId, COUNT(*) AS Count, SUM(status) AS S, MIN(status) AS Mn, MAX(status) AS Mx FROM DataXProcessedInput GROUP BY deviceId, homeId
--DataXQuery--
Sliding5m = SELECT deviceId, homeId, COUNT(*) AS Count, SUM(status) AS S FROM DataXProcessedInput_5minutes GROUP BY deviceId, homeId
"""


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", type=int, default=10)
    ap.add_argument("--events-per-batch", type=int, default=20_000_000)
    ap.add_argument("--keys", type=int, default=1_000_000)
    ap.add_argument("--put", action="store_true", help="pack fused with the exchange: records written into the owners' buffers over NVLink peer memory (no NCCL all-to-all)")
    ap.add_argument("--rules", type=int, default=0, help="C5: evaluate this many generated rules on the local events before the shuffle (S2 events)")
    args = ap.parse_args()
    import numpy as np
    import torch
    import torch.distributed as dist
    import dxgpu as dx
    import dxo
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = dx.Context(local)
    shape = 2 if args.rules else 1
    schema = dxo.iot_schema_json(shape)
    rules_tr = dx.flowgen.process_rules_sql(dx.flowgen.make_rules(args.rules, seed=20260105, n_keys=args.keys, n_homes=max(1, args.keys // 10))) if args.rules else ""
    dec = dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("dec"), schema, [PROJ], rules_tr)
    conf = dx.flowgen.conf_text("c5", timestamp_column="eventTimeStamp", watermark="60 second", windows=[("DataXProcessedInput_5minutes", "5 minutes")])
    agg = dx.CommonProcessorFactory.createProcessor(ctx, conf, schema, [PROJ], TR)
    n = args.events_per_batch
    rec_words = len(COLS) + 1
    send = torch.empty(n * rec_words, dtype=torch.int64, device="cuda") if not args.put else None
    px = dx.parallel.PeerExchange(ctx, int(n * 1.25) + 1024, rec_words) if args.put else None
    a2a_ms, pack_ms, total_ms, wire = 0.0, 0.0, 0.0, 0
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    groups = 0
    for b in range(args.batches):
        T = T0 + b * I
        gp = dx.GenParams(9000 + rank, args.keys, max(1, args.keys // 10), T, I, 60_000, shape, 10, 1, 0)
        ptr, nb, _ = ctx.gen_events(gp, b * n, n)
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        ev[0].record()
        if args.rules:      # decode + projection + rules -> (columns, flag bitmap); the flags stay on the rank that decoded the event
            r0 = dec.process_device(ptr, nb, T, I, n_events=n)
            proj = r0.table("Rules")
        else:
            proj = dx.parallel.project_device(dec, ptr, nb, n, T)
        if args.put:
            ev[1].record()
            recv_ptr, n_recv, counts = px.exchange(proj, COLS, COLS[:2])
            rec_bytes = 8 * rec_words
            ev[2].record()
            part = dx.parallel.table_from_packed(ctx, NAMES, TYPES, recv_ptr if n_recv else 0, n_recv)
        else:
            counts, rec_bytes = dx.parallel.pack_by_key(proj, COLS, COLS[:2], world, send.data_ptr())
            ev[1].record()
            recv_counts = dx.parallel.exchange_counts(counts)
            recv = dx.parallel.exchange_records(send, counts, recv_counts, rec_words)
            ev[2].record()
            torch.cuda.synchronize()      # the collective runs on torch's stream, the unpack on the library's: order them
            n_recv = int(recv_counts.sum())
            part = dx.parallel.table_from_packed(ctx, NAMES, TYPES, recv.data_ptr() if n_recv else 0, n_recv)
        r = dx.parallel.process_table(agg, part, T, I)
        ev[3].record()
        torch.cuda.synchronize()
        if b >= 1:       # first batch warms allocators, NCCL channels and the hash-table size hints
            t = torch.tensor([ev[0].elapsed_time(ev[3]), ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total_ms += float(t[0]); pack_ms += float(t[1]); a2a_ms += float(t[2])
            wire += int((counts.sum() - counts[rank]) * rec_bytes)
        groups = r.table("Tumbling").num_rows
        r.release(); ctx.free(ptr)
        if args.rules:
            r0.release()
    timed = args.batches - 1
    host = {k: round(ctx.take_counter("host_us:" + k) / 1e3 / args.batches, 3) for k in ("route", "stmt:Tumbling", "stmt:Sliding5m")}
    g = torch.tensor([groups], device="cuda", dtype=torch.int64)
    dist.all_reduce(g)
    if rank == 0:
        print(json.dumps({"metric": "events/sec per micro-batch (decode + key-hash all-to-all + windowed group-by)", "n_gpus": world,
                          "value": round(world * n * timed / (total_ms / 1e3), 1), "unit": "events/s", "batches_timed": timed,
                          "events_per_batch_per_gpu": n, "ms_per_batch": round(total_ms / timed, 3),
                          "decode_project_pack_ms": round(pack_ms / timed, 3), "all_to_all_ms": round(a2a_ms / timed, 3),
                          "wire_bytes_out_per_gpu_per_batch": wire // max(timed, 1), "rec_bytes": 8 * rec_words,
                          "all_to_all_gbs_per_gpu_per_direction": round(wire / max(timed, 1) / max(a2a_ms / timed, 1e-9) / 1e6, 2),
                          "nvlink_peak_gbs": 900, "frac_of_nvlink": round(wire / max(timed, 1) / max(a2a_ms / timed, 1e-9) / 1e6 / 900, 4),
                          "tumbling_groups_all_ranks": int(g.item()), "rules": args.rules, "exchange": "fused put over peer memory" if args.put else "K5 pack + NCCL all-to-all", "rank0_host_ms_per_batch": host, "scaling": "weak"}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
