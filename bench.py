#!/usr/bin/env python3
"""bench.py — events/sec of the per-micro-batch hot path (decode + rules -> alert flags).

Workload (BASELINE.json configs[1]): 100 M synthetic S2 IoT events
(compact NDJSON, ~151 B/event) as 4 micro-batches of 25 M, 200 generated threshold/AND-OR rules
in the shape CodegenRules emits (ProcessRules -> filterNull(Array(IF(cond, MAP(...), NULL)...))),
output = the 200-bit alert-flag bitmap per event.  A "step" is one pass over all 100 M events.

  python bench.py --gpus N --steps K --warmup W            our arm (one process per GPU; torchrun for N>1)
  python bench.py --impl reference --steps K --warmup W    the CPU restatement of the reference path on the
                                                           host cores (the reference itself needs JVM+Spark)

Prints ONE JSON line (rank 0).  See DESIGN.md §6 for every field's definition.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

N_EVENTS = 100_000_000
N_CHUNKS = 4
N_RULES = 200
N_KEYS, N_HOMES = 1_000_000, 100_000
BATCH_MS, INTERVAL_MS = 1_700_000_040_000, 60_000
SEED = 20260102
METRIC = "events/sec per micro-batch (decode+rules -> alert flags), 100M events, 200 rules"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_of(kernel, events_per_launch):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (profiles/ncu_traffic.json holds
    bytes per event of the captured launch), scaled to the events one launch of this run processes; None if absent."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        d = json.load(open(p))
        return round(d[kernel]["dram_bytes_per_event"] * events_per_launch)
    except (OSError, KeyError, ValueError):
        return None


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region."""

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def artefacts(dx):
    import dxo
    rules = dx.flowgen.make_rules(N_RULES, seed=SEED, n_keys=N_KEYS, n_homes=N_HOMES)
    return (dx.flowgen.conf_text("c2rules", outputs=["Rules"]), dxo.iot_schema_json(2), ["Raw.*\n"],
            dx.flowgen.process_rules_sql(rules))


def gen_params(cls, rank):
    return cls(SEED + rank, N_KEYS, N_HOMES, BATCH_MS, INTERVAL_MS, 60_000, 2, 10, 1, 0)


# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import numpy as np
    import torch
    import dxgpu as dx

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = dx.Context(local)
    conf, schema, proj, transform = artefacts(dx)
    flow = dx.CommonProcessorFactory.createProcessor(ctx, conf, schema, proj, transform)
    n_events = args.events
    per_chunk = n_events // N_CHUNKS
    # ---- inputs: generated on the device (each rank its own events), also staged to pinned host memory for e2e
    chunks = []
    gp = gen_params(dx.GenParams, rank)
    for c in range(N_CHUNKS):
        ptr, nb, _ = ctx.gen_events(gp, c * per_chunk, per_chunk)
        chunks.append((ptr, nb))
    total_bytes = sum(nb for _, nb in chunks)
    L = ctx.L
    L.dx_host_alloc.restype = C.c_int
    L.dx_host_alloc.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]
    L.dx_host_free.argtypes = [C.c_void_p, C.c_void_p]
    L.dx_copy_d2h_async.restype = C.c_int
    L.dx_copy_d2h_async.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
    host_in, host_out = [], []
    words = (N_RULES + 31) // 32
    for ptr, nb in chunks:
        hp = C.c_void_p()
        ctx.check(L.dx_host_alloc(ctx.h, nb, C.byref(hp)))
        ctx.check(L.dx_copy_d2h(ctx.h, hp, C.c_void_p(ptr), nb))
        host_in.append(hp)
        ho = C.c_void_p()
        ctx.check(L.dx_host_alloc(ctx.h, per_chunk * words * 4, C.byref(ho)))
        host_out.append(ho)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", local))

    def step_device():
        for ptr, nb in chunks:
            r = flow.process_device(ptr, nb, BATCH_MS, INTERVAL_MS, n_events=per_chunk)
            r.release()

    def step_e2e():
        for (ptr, nb), hp, ho in zip(chunks, host_in, host_out):
            buf = (C.c_uint8 * nb).from_address(hp.value)
            r = flow.process_json(np.frombuffer(buf, dtype=np.uint8), BATCH_MS, INTERVAL_MS, n_events=per_chunk)
            t = r.table("Rules")
            d, _, _, _ = t.device_view("Rules")
            ctx.check(L.dx_copy_d2h_async(ctx.h, ho, C.c_void_p(d), per_chunk * words * 4))
            ctx.synchronize()
            t.release()
            r.release()

    # e2e runs the batch's partitions through TWO processors on two streams from two host threads (the way the
    # reference runs a micro-batch's partitions on several executor cores): one partition's decode, rules and
    # result read-back overlap the other's host->device copy on the full-duplex link.
    lanes = [(ctx, flow)]
    for _ in range(max(1, args.e2e_partitions) - 1):
        cx = dx.Context(local)
        lanes.append((cx, dx.CommonProcessorFactory.createProcessor(cx, conf, schema, proj, transform)))

    def e2e_worker(k, errs):
        try:
            cx, fl = lanes[k]
            for c in range(k, N_CHUNKS, len(lanes)):
                (ptr, nb), hp, ho = chunks[c], host_in[c], host_out[c]
                buf = (C.c_uint8 * nb).from_address(hp.value)
                r = fl.process_json(np.frombuffer(buf, dtype=np.uint8), BATCH_MS, INTERVAL_MS, n_events=per_chunk)
                t = r.table("Rules")
                d, _, _, _ = t.device_view("Rules")
                cx.check(L.dx_copy_d2h_async(cx.h, ho, C.c_void_p(d), per_chunk * words * 4))
                cx.synchronize()
                t.release()
                r.release()
        except Exception as ex:      # surfaced by the caller: a failed partition must fail the bench
            errs.append(ex)

    def step_e2e_overlapped():
        errs = []
        ths = [threading.Thread(target=e2e_worker, args=(k, errs)) for k in range(len(lanes))]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        if errs:
            raise errs[0]

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        e1.synchronize()
        barrier()
        ms = e0.elapsed_time(e1)
        if dist:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(args.warmup):
        step_device()
    ctx.take_launches()
    sampler = ClockSampler(local)
    sampler.start()
    ms = timed(step_device, args.steps)
    clocks = sampler.stop()
    launches = ctx.take_launches()
    value = world * n_events * args.steps / (ms / 1e3)
    # ---- per-kernel times (separate pass: per-kernel events add host syncs, so never inside the headline region)
    ctx.set_timing(True)
    step_device()
    ctx.set_timing(False)
    k1_ms, k1_n = ctx.take_stage_ms("k1_decode")
    k2_ms, k2_n = ctx.take_stage_ms("k2_rules_bucket")
    k2_name = "k2_rules_bucket"
    if k2_n == 0:
        k2_ms, k2_n = ctx.take_stage_ms("k2_eval_flags")
        k2_name = "k2_eval_flags"
    peak, peak_src = peaks()
    k1_bytes = total_bytes + n_events * (6 * 8 + 4 + 1)                 # JSON bytes + 6x8 B + 4 B code + 1 B validity per event
    k2_bytes = n_events * (5 * 8 + 4 + 1 + (N_RULES + 7) // 8)          # 6 referenced columns (44 B) + validity + ceil(R/8) flag bytes = 70 B
    dominant = (k2_name, k2_bytes, k2_ms) if k2_ms >= k1_ms else ("k1_decode", k1_bytes, k1_ms)
    achieved = dominant[1] / (dominant[2] / 1e3) / 1e9 if dominant[2] > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": dominant[0], "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": traffic_of(dominant[0], per_chunk), "alg_bytes_per_launch": int(dominant[1] // N_CHUNKS),
                "launch_ms": round(dominant[2] / N_CHUNKS, 4), "peak_source": peak_src,
                "kernels": {"k1_decode": {"ms_per_step": round(k1_ms, 3), "launches": k1_n, "alg_bytes": k1_bytes,
                                          "gbs": round(k1_bytes / max(k1_ms, 1e-9) / 1e6, 2)},
                            k2_name: {"ms_per_step": round(k2_ms, 3), "launches": k2_n, "alg_bytes": k2_bytes,
                                              "gbs": round(k2_bytes / max(k2_ms, 1e-9) / 1e6, 2)}}}
    # ---- end to end through the C ABI with host buffers
    step_e2e()
    e2e_steps = max(1, min(args.steps, 3))
    ms_serial = timed(step_e2e, e2e_steps)
    step_e2e_overlapped()
    ms_e2e = timed(step_e2e_overlapped, e2e_steps)
    e2e = {"value": round(world * n_events * e2e_steps / (ms_e2e / 1e3), 1), "unit": "events/s",
           "h2d_bytes_per_step": int(total_bytes), "d2h_bytes_per_step": int(n_events * words * 4), "ms_per_step": round(ms_e2e / e2e_steps, 3),
           "partitions_in_flight": len(lanes), "one_processor": {"value": round(world * n_events * e2e_steps / (ms_serial / 1e3), 1),
                                                                  "ms_per_step": round(ms_serial / e2e_steps, 3)}}
    out = {"metric": METRIC, "value": round(value, 1), "unit": "events/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64/f64",
           "data": "synthetic", "config": {"workload": "C2 rules-only: %d S2 JSON events/GPU in %d micro-batches, %d rules -> flag bitmap" % (n_events, N_CHUNKS, N_RULES),
                                            "events_per_gpu": n_events, "json_bytes_per_gpu": int(total_bytes), "rules": N_RULES,
                                            "l2": "inputs (%.1f GB) far larger than L2; no flush needed" % (total_bytes / 1e9),
                                            "sharding": "independent event ranges per GPU, no data-path collective"},
           "roofline": roofline, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
    if rank == 0 and not args.no_cpu_baseline and world >= 1:
        out["cpu_baseline"] = cpu_baseline(sample_events=args.cpu_sample)
    for hp in host_in + host_out:
        L.dx_host_free(ctx.h, hp)
    for cx, fl in lanes[1:]:
        fl.close()
        cx.close()
    if rank == 0:
        print(json.dumps(out))
    if dist:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
def cpu_baseline(sample_events=2_000_000, steps=1):
    """The CPU restatement of the same path on every host core: C from_json (oracle/dxo_core.c) + C rule evaluation
    (oracle/dxo_rules.c), both one thread per core; the reference itself needs JVM + Spark 2.4.5."""
    import numpy as np
    import dxgpu as dx
    import dxo
    import dxo_flow
    conf, schema, proj, transform = artefacts(dx)
    o = dxo_flow.Flow(conf, schema, proj, transform)
    threads = os.cpu_count() or 1
    p = gen_params(dxo.GenParams, 0)
    data = dxo.gen_events(p, 0, sample_events)
    cores, _ = dxo_flow.P(o.commands[0][2]).statement()
    rs = dxo_flow._rule_shape(cores[0]["items"][1][0])
    types = [t if isinstance(t, bytes) else t.encode() for t in dxo.TYPES]
    cols = [(name, {dxo.LONG: "long", dxo.DOUBLE: "double", dxo.BOOL: "bool", dxo.STRING: "string"}[ty])
            for name, ty in zip(o.schema.leaf_names, o.schema.leaf_types)]
    prog = dxo.compile_rules([c for c, _ in rs], cols, lambda ci, lit: types.index(lit.encode()) if lit.encode() in types else -2)
    truth = dxo.gen_fields(p, 0, sample_events)        # dictionary codes of the string leaf (the threaded decoder reports token lengths)
    best = None
    for _ in range(steps):
        t0 = time.perf_counter()
        d = dxo.decode(o.schema, data, threads=threads)
        t1 = time.perf_counter()
        arrs, valids = [], []
        for name, ty in zip(o.schema.leaf_names, o.schema.leaf_types):
            v, ok = d[name]
            arrs.append(truth["deviceType"].astype(np.int64) if ty == dxo.STRING else v)
            valids.append(ok)
        t2 = time.perf_counter()
        flags = dxo.eval_rules_mt(prog, arrs, valids, sample_events, len(rs), threads=threads)
        t3 = time.perf_counter()
        assert flags.shape == (sample_events, (len(rs) + 31) // 32)
        sec = (t1 - t0) + (t3 - t2)
        best = sec if best is None else min(best, sec)
    return {"value": round(sample_events / best, 1), "unit": "events/s", "cores": threads, "kind": "port",
            "sample": "%d events of the same workload: C decode (%.2f s) + C rule evaluation (%.2f s), %d threads each" % (
                sample_events, t1 - t0, t3 - t2, threads)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t_all = time.perf_counter()
    vals = []
    info = None
    for i in range(args.warmup + args.steps):
        info = cpu_baseline(sample_events=args.cpu_sample)
        if i >= args.warmup:
            vals.append(info["value"])
        if time.perf_counter() - t_all > 240:
            break
    v = sum(vals) / max(1, len(vals)) if vals else info["value"]
    info["value"] = round(v, 1)
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": round(v, 1), "unit": "events/s", "n_gpus": args.gpus,
                      "steps": len(vals) or 1, "warmup": args.warmup, "ms_per_step": round(args.cpu_sample / v * 1e3, 3),
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64/f64", "data": "synthetic",
                      "config": {"workload": "C2 rules-only: bounded sample of %d S2 JSON events per step, %d rules -> flags; "
                                             "CPU restatement of the reference path (the reference needs JVM + Spark 2.4.5: unavailable on this box)" % (args.cpu_sample, N_RULES)},
                      "cpu_baseline": info,
                      "e2e": {"value": round(v, 1), "unit": "events/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--events", type=int, default=N_EVENTS)
    ap.add_argument("--cpu-sample", type=int, default=10_000_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-partitions", type=int, default=2, help="micro-batch partitions in flight in the end-to-end measurement (one processor + stream each)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
