#!/usr/bin/env python3
"""bench.py — events/sec of the per-micro-batch hot path: decode + rules + window aggregation.

Headline workload (BASELINE.json metric "decode+rules+window agg"; recipe of SURVEY.md §8(d) C2 with C5's windowed
group-by behind it): per GPU and per step 100 M synthetic S2 IoT events (compact NDJSON, ~151 B/event) as 4 micro-batches
of 25 M with consecutive batch times.  Every micro-batch goes through

  decode (K1) -> projection -> 200 generated threshold/AND-OR rules -> flag bitmap + per-rule counts (K2)
  -> key-hash exchange of (deviceId, homeId, eventTimeStamp, status) records, owner = hash64(deviceId, homeId) % G (K5;
     at G > 1 the records cross NVLink: fused put over peer memory, or K5 pack + NCCL all-to-all with --exchange nccl)
  -> on the owner: tumbling (1 interval) and sliding (5 minutes) COUNT/SUM/MIN/MAX group-bys per (deviceId, homeId) (K3P)

with the same code path at every G (at G = 1 every key is local, the exchange kernels still run).  `stages` carries the
other configurations on the same line: c2 (decode + rules only, the round-1 headline), c3 (windows, 1 M keys) and c4
(state-table merge), each with its own clock record.

  python bench.py --gpus N --steps K --warmup W            our arm (one process per GPU; torchrun for N>1)
  python bench.py --impl reference --steps K --warmup W    the CPU restatement of the same path on the host cores
                                                           (the reference itself needs JVM + Spark 2.4.5)

Prints ONE JSON line (rank 0).  DESIGN.md §6 defines every field.
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
sys.path.insert(0, os.path.join(ROOT, "tools"))

N_EVENTS = 100_000_000
N_CHUNKS = 4
N_RULES = 200
N_KEYS, N_HOMES = 1_000_000, 100_000
T0, INTERVAL_MS = 1_700_000_040_000, 60_000
SEED = 20260102
METRIC = "events/sec per micro-batch (decode+rules+window agg)"
PROJ = "CAST(Raw.deviceDetails.eventTime / 1000 AS timestamp) AS eventTimeStamp\nRaw.*\n"
COLS = ["deviceDetails.deviceId", "deviceDetails.homeId", "eventTimeStamp", "deviceDetails.status"]
NAMES, TYPES = ["deviceId", "homeId", "eventTimeStamp", "status"], [1, 1, 5, 1]
AGG_TR = """--DataXQuery--
Tumbling = SELECT deviceId, homeId, COUNT(*) AS Count, SUM(status) AS S, MIN(status) AS Mn, MAX(status) AS Mx FROM DataXProcessedInput GROUP BY deviceId, homeId
--DataXQuery--
Sliding5m = SELECT deviceId, homeId, COUNT(*) AS Count, SUM(status) AS S FROM DataXProcessedInput_5minutes GROUP BY deviceId, homeId
"""
STAGES = ["k1_decode", "k2_rules_bucket", "k2_eval_flags", "k2_eval_column", "k2_rule_counts", "k5_count", "k5_scatter_put", "k5_scatter",
          "k5_unpack", "k3p_scatter", "k3p_aggregate", "k3_hash_agg", "k_gather"]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_of(kernel, events_per_launch):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (profiles/ncu_traffic.json holds bytes
    per event of the captured launch), scaled to the events one launch of this run processes; None if absent."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        return round(d[kernel]["dram_bytes_per_event"] * events_per_launch)
    except (OSError, KeyError, ValueError):
        return None


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING a timed region.  ONE poller per run (rank 0), over the GPUs the run uses:
    a poller per rank at 50 Hz means eight processes taking the driver's lock while every rank is issuing launches and small
    synchronising copies, and the step time of the whole job then follows the unluckiest rank."""

    def __init__(self, gpu_index, n_gpus=1, enabled=True):
        self.rows, self.proc, self.idx, self.n, self.enabled = [], None, gpu_index, n_gpus, enabled

    def start(self):
        if not self.enabled:
            return self
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        ids = ",".join(str(i) for i in range(self.n)) if self.n > 1 else str(self.idx)
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", ids, "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "20" if self.n == 1 else "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None
        return self

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
                "reasons": sorted(reasons), "samples": len(sm), "gpus_sampled": self.n if self.enabled else 0}


def artefacts(dx):
    """(decode+rules flow, owner-side window flow) artefacts; everything from the product package, nothing from oracle/."""
    rules = dx.flowgen.make_rules(N_RULES, seed=SEED, n_keys=N_KEYS, n_homes=N_HOMES)
    schema = dx.flowgen.iot_schema_json(2)
    dec = (dx.flowgen.conf_text("c2rules", outputs=["Rules"]), schema, [PROJ], dx.flowgen.process_rules_sql(rules))
    agg = (dx.flowgen.conf_text("c5windows", timestamp_column="eventTimeStamp", watermark="60 second",
                                windows=[("DataXProcessedInput_5minutes", "5 minutes")]), schema, [PROJ], AGG_TR)
    return dec, agg


def bind_to_gpu_node(ctx):
    """Keeps this rank's host threads on the CPUs of the NUMA node its GPU hangs off (the library already takes the pinned pages
    from that node): with one rank per GPU, host->device copies that cross the inter-socket link share it with every other rank.
    -> {"node": n, "cpus": k} or {"node": -1} when the platform reports no node."""
    fn = ctx.L.dx_ctx_numa_node
    fn.restype, fn.argtypes = C.c_int, [C.c_void_p]
    node = int(fn(ctx.h))
    info = {"node": node}
    try:
        if node >= 0:
            cpus = set()
            for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
            allowed = cpus & os.sched_getaffinity(0)
            if allowed:
                info["all_cpus"] = sorted(os.sched_getaffinity(0))
                os.sched_setaffinity(0, allowed)
                info["cpus"] = len(allowed)
    except (OSError, ValueError):
        pass
    return info


def gen_params(cls, rank, batch):
    return cls(SEED + rank, N_KEYS, N_HOMES, T0 + batch * INTERVAL_MS, INTERVAL_MS, 60_000, 2, 10, 1, 0)


def group_checksum(cols):
    """Order-independent checksum of a group-by result: (groups, sum of the low 31 bits, sum of the next 31 bits) of a
    64-bit mix over each row's values — sums of 31-bit halves cannot overflow an int64 all-reduce."""
    import numpy as np
    n = len(cols[0]) if cols else 0
    h = np.full(n, 0x9E3779B97F4A7C15, dtype=np.uint64)
    with np.errstate(over="ignore"):
        for c in cols:
            h = (h ^ c.astype(np.int64).view(np.uint64)) * np.uint64(0xBF58476D1CE4E5B9)
            h ^= h >> np.uint64(29)
    return np.array([n, int((h & np.uint64(0x7FFFFFFF)).sum()), int(((h >> np.uint64(31)) & np.uint64(0x7FFFFFFF)).sum())], dtype=np.int64)


# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import dxgpu as dx

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(29400 + os.getpid() % 500))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    ctx = dx.Context(local)
    numa = bind_to_gpu_node(ctx)
    L = ctx.L
    dec_art, agg_art = artefacts(dx)
    mk = dx.CommonProcessorFactory.createProcessor
    n_events = args.events
    per_chunk = n_events // N_CHUNKS
    words = (N_RULES + 31) // 32
    rec_words = len(COLS) + 1
    stream = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", local))
    px = dx.parallel.PeerExchange(ctx, int(per_chunk * 1.25) + 4096, rec_words)
    send = torch.empty(per_chunk * rec_words, dtype=torch.int64, device="cuda") if args.exchange == "nccl" or world > 1 else None

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        t = torch.tensor([v], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    class Pipeline:
        """One rank's flows + the per-micro-batch sequence; `mode` picks how the keyed records travel."""

        def __init__(self):
            self.dec, self.agg = mk(ctx, *dec_art), mk(ctx, *agg_art)
            self.wire = 0
            self.a2a_ms = 0.0
            self.ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            self.hits = np.zeros(N_RULES, dtype=np.int64)

        def exchange(self, table, mode):
            if mode == "put":
                recv_ptr, firsts, cnts, counts = px.exchange_blocks(table, COLS, COLS[:2])
                self.keep = None
                self.blocks = (firsts, cnts)
                n_recv = int(cnts.sum())
            else:
                counts, rec_bytes = dx.parallel.pack_by_key(table, COLS, COLS[:2], world, send.data_ptr())
                recv_counts = dx.parallel.exchange_counts(counts)
                ctx.synchronize()          # the pack runs on the library's stream, the collective on torch's
                recv = dx.parallel.exchange_records(send[: int(counts.sum()) * rec_words], counts, recv_counts, rec_words)
                torch.cuda.synchronize()
                n_recv, recv_ptr, self.keep = int(recv_counts.sum()), recv.data_ptr(), recv
                self.blocks = None
            self.wire += int((counts.sum() - counts[rank]) * 8 * rec_words)
            return recv_ptr, n_recv

        def batch(self, b, table_or_result, mode, results=None, time_exchange=False, trace=None):
            """Everything after decode+rules for micro-batch b: per-rule counts, exchange, owner-side windows."""
            T = T0 + b * INTERVAL_MS

            def mark(name):
                if trace is not None:
                    ctx.synchronize(); torch.cuda.synchronize()
                    now = time.perf_counter()
                    trace[name] = trace.get(name, 0.0) + (now - trace["_t"]) * 1e3
                    trace["_t"] = now
            mark("decode+rules")
            t = table_or_result.table("Rules")
            self.hits += dx.rule_hits(t)
            mark("rule_hits")
            if time_exchange:
                self.ev[0].record(stream)
            recv_ptr, n_recv = self.exchange(t, mode)
            if time_exchange:
                self.ev[1].record(stream)
            mark("exchange")
            if self.blocks is not None:
                part = dx.parallel.table_from_packed_blocks(ctx, NAMES, TYPES, recv_ptr, *self.blocks)
            else:
                part = dx.parallel.table_from_packed(ctx, NAMES, TYPES, recv_ptr if n_recv else 0, n_recv)
            mark("unpack")
            r = dx.parallel.process_table(self.agg, part, T, INTERVAL_MS)
            mark("windows")
            if time_exchange:
                self.ev[1].synchronize()
                self.a2a_ms += self.ev[0].elapsed_time(self.ev[1])
            if results is not None:
                results(t, r)
            r.release()
            part.release()
            t.release()

        def close(self):
            self.dec.close()
            self.agg.close()

    # ---- parity of the sharded pipeline (not timed): per-owner checksums of the window results, all-reduced, against the same
    # events run through ONE unsharded flow on rank 0 (every rank's events regenerated there from its seed)
    def parity_check():
        n_par, batches = args.parity_events, 3
        sums = {m: np.zeros(6, dtype=np.int64) for m in (["put", "nccl"] if world > 1 else ["put"])}
        for mode in sums:
            p = Pipeline()
            for b in range(batches):
                ptr, nb, _ = ctx.gen_events(gen_params(dx.GenParams, rank, b), b * n_par, n_par)
                r0 = p.dec.process_device(ptr, nb, T0 + b * INTERVAL_MS, INTERVAL_MS, n_events=n_par)

                def grab(t, r, acc=sums[mode]):
                    tt, ss = r.table("Tumbling"), r.table("Sliding5m")
                    acc[:3] += group_checksum([tt.column(c)[0] for c in ("deviceId", "homeId", "Count", "S", "Mn", "Mx")])
                    acc[3:] += group_checksum([ss.column(c)[0] for c in ("deviceId", "homeId", "Count", "S")])
                    tt.release(); ss.release()
                p.batch(b, r0, mode, results=grab)
                r0.release(); ctx.free(ptr)
            p.close()
        out = {}
        for mode, acc in sums.items():
            t = torch.as_tensor(acc, device="cuda")
            dist.all_reduce(t)
            out[mode] = t.cpu().numpy()
        ok, ref = None, None
        if rank == 0:
            whole = mk(ctx, agg_art[0], agg_art[1], ["eventTimeStamp\ndeviceId\nhomeId\nstatus\n"], AGG_TR)    # route() only
            decs = mk(ctx, dx.flowgen.conf_text("dec"), dec_art[1], [PROJ], "")
            ref = np.zeros(6, dtype=np.int64)
            buf = ctx.alloc(world * n_par * rec_words * 8 + 64)
            for b in range(batches):
                at = 0
                for rk in range(world):
                    ptr, nb, _ = ctx.gen_events(gen_params(dx.GenParams, rk, b), b * n_par, n_par)
                    tab = dx.parallel.project_device(decs, ptr, nb, n_par, T0 + b * INTERVAL_MS)
                    counts, rec_bytes = dx.parallel.pack_by_key(tab, COLS, COLS[:2], 1, buf + at * rec_bytes_of(rec_words))
                    at += int(counts.sum())
                    tab.release(); ctx.free(ptr)
                allrows = dx.parallel.table_from_packed(ctx, NAMES, TYPES, buf, at)
                r = dx.parallel.process_table(whole, allrows, T0 + b * INTERVAL_MS, INTERVAL_MS)
                tt, ss = r.table("Tumbling"), r.table("Sliding5m")
                ref[:3] += group_checksum([tt.column(c)[0] for c in ("deviceId", "homeId", "Count", "S", "Mn", "Mx")])
                ref[3:] += group_checksum([ss.column(c)[0] for c in ("deviceId", "homeId", "Count", "S")])
                tt.release(); ss.release(); r.release(); allrows.release()
            ctx.free(buf)
            whole.close(); decs.close()
            ok = all((v == ref).all() for v in out.values())
        flag = torch.tensor([1 if (ok or rank != 0) else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) != 1:
            raise SystemExit("bench.py: sharded window results differ from the unsharded run: %s vs %s" % (out, ref))
        return {"ok": True, "events_per_rank_per_batch": n_par, "batches": batches, "modes": sorted(out),
                "checksum": "per-owner (groups, two 31-bit sums of mix64(deviceId,homeId,Count,S[,Mn,Mx])) of Tumbling and Sliding5m, summed over ranks, "
                            "equal to one unsharded flow over all ranks' events on rank 0",
                "groups_tumbling": int(out["put"][0]), "groups_sliding": int(out["put"][3])}

    def rec_bytes_of(w):
        return 8 * w

    parity = parity_check()

    # ---- inputs: every step has its own batch times, so its four chunks are generated on the device between the timed steps
    chunks = [None] * N_CHUNKS

    def generate(step):
        for c in range(N_CHUNKS):
            if chunks[c]:
                ctx.free(chunks[c][0])
            ptr, nb, _ = ctx.gen_events(gen_params(dx.GenParams, rank, step * N_CHUNKS + c), (step * N_CHUNKS + c) * per_chunk, per_chunk)
            chunks[c] = (ptr, nb)
        ctx.synchronize()

    pipe = Pipeline()

    def step_device(step, mode=args.exchange, time_exchange=False, trace=None):
        for c, (ptr, nb) in enumerate(chunks):
            b = step * N_CHUNKS + c
            if trace is not None:
                ctx.synchronize(); trace["_t"] = time.perf_counter()
            r0 = pipe.dec.process_device(ptr, nb, T0 + b * INTERVAL_MS, INTERVAL_MS, n_events=per_chunk)
            pipe.batch(b, r0, mode, time_exchange=time_exchange, trace=trace)
            r0.release()

    def timed_step(fn, step):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn(step)
        e1.record(stream)
        e1.synchronize()
        barrier()
        return e0.elapsed_time(e1)

    step_no = 0
    for _ in range(args.warmup):
        generate(step_no)
        step_device(step_no)
        step_no += 1
    ctx.take_launches()
    pipe.wire = 0
    host_keys = ["project", "route", "stmt:Rules", "stmt:Tumbling", "stmt:Sliding5m"]
    for k in host_keys:
        ctx.take_counter("host_us:" + k)
    t_wall0 = time.perf_counter()
    sampler = ClockSampler(local, world, enabled=(rank == 0)).start()
    ms = 0.0
    for _ in range(args.steps):
        generate(step_no)
        ms += timed_step(step_device, step_no)
        step_no += 1
    clocks = sampler.stop()
    launches = ctx.take_launches()
    host_ms = {k: round(ctx.take_counter("host_us:" + k) / 1e3 / args.steps, 3) for k in host_keys}     # host wall time inside the library per step
    host_ms["wall_incl_generation"] = round((time.perf_counter() - t_wall0) * 1e3 / args.steps, 3)
    ms = max_over_ranks(ms)
    total_bytes = sum(nb for _, nb in chunks)
    value = world * n_events * args.steps / (ms / 1e3)
    wire_per_step = pipe.wire // max(args.steps, 1)

    # ---- per-kernel times (separate pass: per-kernel events add host syncs, so never inside the headline region)
    generate(step_no)
    ctx.set_timing(True)
    for s in STAGES:
        ctx.take_stage_ms(s)
    step_device(step_no)
    step_no += 1
    ctx.set_timing(False)
    stage_ms = {}
    for s in STAGES:
        t, n = ctx.take_stage_ms(s)
        if n:
            stage_ms[s] = (t, n)
    if args.trace:       # host wall time per phase of one more step, every phase closed by a synchronize (stderr; not a bench value)
        generate(step_no)
        tr = {}
        step_device(step_no, trace=tr)
        step_no += 1
        tr.pop("_t", None)
        print("trace ms/step:", {k: round(v, 2) for k, v in tr.items()}, "sum", round(sum(tr.values()), 2), file=sys.stderr)
    peak, peak_src = peaks()
    rows = n_events
    alg = {  # algorithmic bytes per step, SURVEY.md §8(d) / DESIGN.md §5
        "k1_decode": total_bytes + rows * (6 * 8 + 4 + 1),
        "k2_rules_bucket": rows * (5 * 8 + 4 + 1 + (N_RULES + 7) // 8),
        "k2_eval_flags": rows * (5 * 8 + 4 + 1 + (N_RULES + 7) // 8),
        "k2_eval_column": rows * (8 + 8 + 1),                         # eventTime in, eventTimeStamp out
        "k2_rule_counts": rows * words * 4,
        "k5_count": rows * (2 * 8 + 1),
        "k5_scatter_put": rows * (4 * 8 + 1 + 8 * rec_words),
        "k5_scatter": rows * (4 * 8 + 1 + 8 * rec_words),
        "k5_unpack": rows * (8 * rec_words + 4 * 8 + 1),
        "k3p_scatter": rows * (2 * (3 * 8 + 1) + 2 * 8 * 6),          # count pass + write pass: keys+ts+value in, [hash|keys|pane|nulls|value] out
    }
    kernels = {}
    for s, (t, n) in stage_ms.items():
        kernels[s] = {"ms_per_step": round(t, 3), "launches": n}
        if s in alg:
            kernels[s].update({"alg_bytes": int(alg[s]), "gbs": round(alg[s] / max(t, 1e-9) / 1e6, 2), "frac": round(alg[s] / max(t, 1e-9) / 1e6 / peak, 4)})
    dom = max((s for s in stage_ms if s in alg), key=lambda s: stage_ms[s][0])
    d_ms, d_n = stage_ms[dom]
    achieved = alg[dom] / (d_ms / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                "traffic": traffic_of(dom, per_chunk), "alg_bytes_per_launch": int(alg[dom] // d_n), "launch_ms": round(d_ms / d_n, 4),
                "peak_source": peak_src, "kernels": kernels,
                "kernel_ms_sum": round(sum(t for t, _ in stage_ms.values()), 3)}

    # ---- the exchange by itself, both ways, on real ranks (NVLink roofline): one untimed + three timed micro-batches per mode
    shuffle = None
    if world > 1:
        shuffle = {"nvlink_peak_gbs_per_direction": 900, "measured_peer_copy_gbs": 770, "rec_bytes": 8 * rec_words,
                   "records_per_gpu_per_batch": per_chunk}
        for mode in ("put", "nccl"):
            generate(step_no)
            step_device(step_no, mode=mode)
            step_no += 1
            generate(step_no)
            pipe.wire, pipe.a2a_ms = 0, 0.0
            barrier()
            step_device(step_no, mode=mode, time_exchange=True)
            step_no += 1
            a2a = max_over_ranks(pipe.a2a_ms) / N_CHUNKS
            wire = pipe.wire / N_CHUNKS
            shuffle[mode] = {"exchange_ms_per_batch": round(a2a, 3), "wire_bytes_out_per_gpu_per_batch": int(wire),
                             "gbs_per_gpu_per_direction": round(wire / max(a2a, 1e-9) / 1e6, 2),
                             "frac_of_nvlink_900": round(wire / max(a2a, 1e-9) / 1e6 / 900, 4),
                             "what": "fused put into per-sender blocks over peer memory + one all-gather of the counts (no count round before the put)" if mode == "put"
                                     else "count exchange + NCCL all_to_all_single (the K5 pack before it is not in this time)"}

    # ---- stage c2: decode + rules only (the round-1 headline; no projection of the timestamp, no exchange, no windows)
    stages = {}
    c2 = mk(ctx, dx.flowgen.conf_text("c2rules", outputs=["Rules"]), dec_art[1], ["Raw.*\n"], dec_art[3])

    def step_c2(_):
        for ptr, nb in chunks:
            r = c2.process_device(ptr, nb, T0, INTERVAL_MS, n_events=per_chunk)
            r.release()
    for _ in range(3):
        step_c2(0)
    smp = ClockSampler(local, world, enabled=(rank == 0)).start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step_c2(0)
    e1.record(stream)
    e1.synchronize()
    barrier()
    c2_ms = max_over_ranks(e0.elapsed_time(e1))
    stages["c2"] = {"what": "decode + %d rules -> flag bitmap, %d events/GPU in %d micro-batches, device-resident" % (N_RULES, n_events, N_CHUNKS),
                    "value": round(world * n_events * args.steps / (c2_ms / 1e3), 1), "unit": "events/s", "ms_per_step": round(c2_ms / args.steps, 3),
                    "clocks": smp.stop()}
    c2.close()

    # ---- end to end through the C ABI with HOST buffers: H2D of every batch's JSON and D2H of the flag words, the per-rule
    # counts and both window tables inside the timed region
    L.dx_host_alloc.restype = C.c_int
    L.dx_host_alloc.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]
    L.dx_host_free.argtypes = [C.c_void_p, C.c_void_p]
    L.dx_copy_d2h_async.restype = C.c_int
    L.dx_copy_d2h_async.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
    host_in, host_out = [], []
    cap_in = int(max(nb for _, nb in chunks) * 1.02) + 4096
    out_bytes = per_chunk * words * 4
    agg_cap = 2 * (N_KEYS + 4096) * 8 * 6        # room for both window tables of one batch
    for _ in range(N_CHUNKS):
        hp, ho = C.c_void_p(), C.c_void_p()
        ctx.check(L.dx_host_alloc(ctx.h, cap_in, C.byref(hp)))
        ctx.check(L.dx_host_alloc(ctx.h, out_bytes + agg_cap, C.byref(ho)))
        host_in.append(hp)
        host_out.append(ho)
    d2h = [0]

    def stage_to_host():
        for (ptr, nb), hp in zip(chunks, host_in):
            assert nb <= cap_in
            ctx.check(L.dx_copy_d2h(ctx.h, hp, C.c_void_p(ptr), nb))

    def results_to_host(ho):
        def grab(t, r):
            d, _, _, _ = t.device_view("Rules")
            ctx.check(L.dx_copy_d2h_async(ctx.h, ho, C.c_void_p(d), out_bytes))
            off = out_bytes
            for name, cols in (("Tumbling", ("deviceId", "homeId", "Count", "S", "Mn", "Mx")), ("Sliding5m", ("deviceId", "homeId", "Count", "S"))):
                tab = r.table(name)
                nr = tab.num_rows
                for cname in cols:
                    dv, _, _, _ = tab.device_view(cname)
                    if nr and off + nr * 8 <= out_bytes + agg_cap:
                        ctx.check(L.dx_copy_d2h_async(ctx.h, C.c_void_p(ho.value + off), C.c_void_p(dv), nr * 8))
                        off += nr * 8
                tab.release()
            d2h[0] += off
            ctx.synchronize()
        return grab

    def step_e2e(step):
        # the next batch's host->device copy is started before the current batch is processed (dx_flow_stage_batch): ingest
        # runs ahead of the processor, as Spark's receivers do, and the PCIe copy overlaps the kernels and the read-back
        depth = 2                                          # batches whose copy is in flight or queued ahead of the one being processed
        staged = [pipe.dec.stage_json(host_in[c].value, chunks[c][1]) for c in range(min(depth, N_CHUNKS))]
        for c, ((_, nb), hp, ho) in enumerate(zip(chunks, host_in, host_out)):
            b = step * N_CHUNKS + c
            if c + depth < N_CHUNKS:
                staged.append(pipe.dec.stage_json(host_in[c + depth].value, chunks[c + depth][1]))
            r0 = pipe.dec.process_staged(staged.pop(0), T0 + b * INTERVAL_MS, INTERVAL_MS, n_events=per_chunk)
            pipe.batch(b, r0, args.exchange, results=results_to_host(ho))
            r0.release()

    e2e_steps = max(1, min(args.steps, 3))
    generate(step_no); stage_to_host(); step_e2e(step_no); step_no += 1
    d2h[0] = 0
    ms_e2e = 0.0
    for _ in range(e2e_steps):
        generate(step_no)
        stage_to_host()
        ms_e2e += timed_step(step_e2e, step_no)
        step_no += 1
    ms_e2e = max_over_ranks(ms_e2e)
    e2e = {"value": round(world * n_events * e2e_steps / (ms_e2e / 1e3), 1), "unit": "events/s",
           "h2d_bytes_per_step": int(total_bytes), "d2h_bytes_per_step": int(d2h[0] // e2e_steps), "ms_per_step": round(ms_e2e / e2e_steps, 3),
           "what": "pinned host NDJSON -> dx_flow_stage_batch (H2D of batches i+1 and i+2 queued while batch i is processed) -> dx_flow_process_staged, per-rule counts, exchange, windows, D2H of flag words + both window tables",
           "numa": {k: v for k, v in numa.items() if k != "all_cpus"}}
    for hp in host_in + host_out:
        L.dx_host_free(ctx.h, hp)
    for c in range(N_CHUNKS):
        ctx.free(chunks[c][0])
        chunks[c] = None
    pipe.close()
    px.close()

    # ---- stages c3 / c4 (BASELINE.json configs[2], [3]) on one GPU, each with its own clock record
    if world == 1 and not args.no_stages:
        import bench_windows as bw
        for name, kw in (("c3", dict(workload="c3", batches=100, events_per_batch=5_000_000, keys=1_000_000)),
                         ("c4", dict(workload="c4", batches=100, events_per_batch=5_000_000, keys=1_000_000)),
                         ("c4_64m_keys", dict(workload="c4", batches=30, events_per_batch=5_000_000, keys=64_000_000))):
            smp = ClockSampler(local).start()
            stages[name] = bw.run(dx, ctx, peak=peak, **kw)
            stages[name]["clocks"] = smp.stop()
        # the sink hand-off: to_json of an output table's rows on the device
        import bench_tojson as bj
        for shape in ("flat", "alerts"):
            smp = ClockSampler(local).start()
            r = bj.run(dx, ctx, 5_000_000, shape)
            r["frac_of_hbm_peak"] = round(r["alg_gbs"] / peak, 4)
            r["clocks"] = smp.stop()
            stages["to_json_" + shape] = r
        # the blob input's gzip files inflated on the device: 2368 files of 20 000 S2 events
        import bench_gunzip as bg
        smp = ClockSampler(local).start()
        r = bg.run(dx, ctx, 2368, 20_000, 6)
        r["clocks"] = smp.stop()
        stages["gunzip"] = r

    out = {"metric": METRIC, "value": round(value, 1), "unit": "events/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64/f64",
           "data": "synthetic",
           "config": {"workload": "C2 recipe + windowed group-by: %d S2 JSON events/GPU/step in %d micro-batches, %d rules -> flags + per-rule counts, "
                                  "key-hash exchange, tumbling + 5-minute sliding COUNT/SUM/MIN/MAX per (deviceId, homeId), %d keys" % (n_events, N_CHUNKS, N_RULES, N_KEYS),
                      "events_per_gpu": n_events, "json_bytes_per_gpu": int(total_bytes), "rules": N_RULES,
                      "l2": "inputs (%.1f GB per step) far larger than L2; no flush needed" % (total_bytes / 1e9),
                      "timing": "each step bracketed by barrier + synchronize and CUDA events on the library's stream; the step's inputs are generated on the device between the brackets",
                      "sharding": "events by range per GPU for decode/rules; keyed rows to owner = hash64(deviceId, homeId) %% %d through the %s" % (
                          world, "fused put over NVLink peer memory (all-to-all)" if args.exchange == "put" else "K5 pack + NCCL all-to-all"),
                      "exchange": args.exchange, "wire_bytes_out_per_gpu_per_step": int(wire_per_step)},
           "roofline": roofline, "host_ms_per_step": host_ms, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "stages": stages, "parity": parity,
           "rule_hits_sample": [int(x) for x in pipe.hits[:4]]}
    if shuffle:
        out["shuffle"] = shuffle
    if rank == 0 and not args.no_cpu_baseline:
        if numa.get("all_cpus"):
            os.sched_setaffinity(0, numa["all_cpus"])      # the CPU leg uses every host core
        out["cpu_baseline"] = cpu_baseline(sample_events=args.cpu_sample)
    if rank == 0:
        print(json.dumps(out))
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
def cpu_baseline(sample_events=8_000_000):
    """The CPU restatement of the same pipeline on every host core: C from_json (oracle/dxo_core.c), C rule
    evaluation on the decoded columns (oracle/dxo_rules.c), and the windowed group-bys the way the reference runs them — the
    cached batches of the window are rescanned every interval (CPF:198-222) — through the C hash aggregate, hash-partitioned
    over the threads like Spark's partial/final aggregation.  The reference itself needs JVM + Spark 2.4.5."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import dxgpu as dx
    import dxo
    import dxo_flow
    (conf, schema, proj, transform), _ = artefacts(dx)
    o = dxo_flow.Flow(conf, schema, proj, transform)
    threads = os.cpu_count() or 1
    per = sample_events // N_CHUNKS
    cores, _ = dxo_flow.P(o.commands[0][2]).statement()
    rs = dxo_flow._rule_shape(cores[0]["items"][1][0])
    cols = [(name, {dxo.LONG: "long", dxo.DOUBLE: "double", dxo.BOOL: "bool", dxo.STRING: "string"}[ty])
            for name, ty in zip(o.schema.leaf_names, o.schema.leaf_types)]
    prog = dxo.compile_rules([c for c, _ in rs], cols, lambda ci, lit: dxo.str_hash(lit.encode()))   # string columns hold FNV-1a identities
    L = dxo.lib()
    vp = C.c_void_p
    cache = []
    t_dec = t_rules = t_agg = 0.0
    for b in range(N_CHUNKS):
        T = T0 + b * INTERVAL_MS
        data = dxo.gen_events(gen_params(dxo.GenParams, 0, b), b * per, per)
        t0 = time.perf_counter()
        d = dxo.decode(o.schema, data, threads=threads)
        t1 = time.perf_counter()
        arrs = [d[name][0] for name in o.schema.leaf_names]
        valids = [d[name][1] for name in o.schema.leaf_names]
        flags = dxo.eval_rules_mt(prog, arrs, valids, per, len(rs), threads=threads)
        assert flags.shape == (per, (len(rs) + 31) // 32)
        t2 = time.perf_counter()
        # windows: late-drop, cache, evict, tumbling + sliding over the cached union (rescan, as the reference does)
        E = (T - 60_000) * 1000
        ts = (d["deviceDetails.eventTime"][0].astype(np.float64) / 1000.0 * 1e6).astype(np.int64)   # CAST(eventTime / 1000 AS timestamp), microseconds
        keep = d["deviceDetails.eventTime"][1] & (ts >= E)
        dev, home, st = d["deviceDetails.deviceId"][0][keep], d["deviceDetails.homeId"][0][keep], d["deviceDetails.status"][0][keep]
        tsk = ts[keep]
        cache = [c for c in cache if c[0] > T - (60_000 + 300_000)]
        use = cache if len(cache) > 1 else []              # the size > 1 quirk (CPF:198-201)
        parts = [(dev, home, st, tsk)] + [c[1] for c in use]
        cache.append((T, (dev, home, st, tsk)))
        adev, ahome, ast_, ats = (np.concatenate([p[i] for p in parts]) for i in range(4))
        part_of = ((adev.view(np.uint64) * np.uint64(0x9E3779B97F4A7C15)) >> np.uint64(40)).astype(np.int64) % threads
        groups = 0
        for lo_us in (E - INTERVAL_MS * 1000, E - 300_000 * 1000):
            inwin = (ats >= lo_us) & (ats < E)
            res = [0] * threads

            def work(k):
                sel = np.ascontiguousarray((inwin & (part_of == k)).astype(np.uint8))
                cap = int(sel.sum()) + 1
                ok_ = np.zeros((2, cap), dtype=np.int64)
                okv = np.zeros((2, cap), dtype=np.uint8)
                agg = (dxo.Agg * cap)()
                keys = (vp * 2)(adev.ctypes.data, ahome.ctypes.data)
                res[k] = L.dxo_group_agg(C.c_int64(len(adev)), C.c_int32(2), keys, None, C.c_int32(1), ast_.ctypes.data_as(vp), None,
                                         sel.ctypes.data_as(vp), ok_.ctypes.data_as(vp), okv.ctypes.data_as(vp), agg, C.c_int64(cap))
            ths = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
            for t in ths:
                t.start()
            for t in ths:
                t.join()
            groups += sum(res)
        t3 = time.perf_counter()
        t_dec += t1 - t0; t_rules += t2 - t1; t_agg += t3 - t2
    sec = t_dec + t_rules + t_agg
    return {"value": round(sample_events / sec, 1), "unit": "events/s", "cores": threads, "kind": "port",
            "sample": "%d events of the same workload in %d micro-batches: C decode (%.2f s) + C rule evaluation on the decoded columns (%.2f s) + "
                      "tumbling and 5-minute group-bys by rescan of the cached batches, C hash aggregate per hash partition (%.2f s), %d threads each; "
                      "built -O3 -march=native" % (sample_events, N_CHUNKS, t_dec, t_rules, t_agg, threads)}


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
        if time.perf_counter() - t_all > 200:
            break
    v = sum(vals) / max(1, len(vals)) if vals else info["value"]
    info["value"] = round(v, 1)
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": round(v, 1), "unit": "events/s", "n_gpus": args.gpus,
                      "steps": len(vals) or 1, "warmup": args.warmup, "ms_per_step": round(args.cpu_sample / v * 1e3, 3),
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64/f64", "data": "synthetic",
                      "config": {"workload": "C2 recipe + windowed group-by: bounded sample of %d S2 JSON events per step in %d micro-batches, %d rules -> flags, "
                                             "tumbling + 5-minute sliding group-bys; CPU restatement of the reference path (the reference needs JVM + "
                                             "Spark 2.4.5: unavailable on this box)" % (args.cpu_sample, N_CHUNKS, N_RULES)},
                      "cpu_baseline": info,
                      "e2e": {"value": round(v, 1), "unit": "events/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--events", type=int, default=N_EVENTS)
    ap.add_argument("--cpu-sample", type=int, default=8_000_000)
    ap.add_argument("--parity-events", type=int, default=1_000_000)
    ap.add_argument("--exchange", default="put", choices=["put", "nccl"], help="how keyed records reach their owner: fused put over peer memory, or K5 pack + NCCL all-to-all")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stages", action="store_true", help="skip the c3 / c4 stage runs")
    ap.add_argument("--trace", action="store_true", help="print the host wall time per pipeline phase of one extra step to stderr")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
