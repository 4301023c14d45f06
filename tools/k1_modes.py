#!/usr/bin/env python3
"""K1 timing by framing mode: NDJSON (structural scan inside the kernel) vs caller-provided offsets (no scan).
  python tools/k1_modes.py [events]"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import bench, dxgpu as dx, dxo
n = int(sys.argv[1]) if len(sys.argv) > 1 else 25_000_000
ctx = dx.Context(0)
dec = dx.Decoder(ctx, dxo.iot_schema_json(2))
gp = bench.gen_params(dx.GenParams, 0, 0)
ptr, nb, off = ctx.gen_events(gp, 0, n, want_offsets=True)
ctx.set_timing(True)
for mode, o in (("ndjson", None), ("offsets", off)):
    for i in range(3):
        t = dec.decode_device(ptr, nb, offsets_ptr=o, n_events=n); t.release()
    ctx.take_stage_ms("k1_decode")
    for i in range(5):
        t = dec.decode_device(ptr, nb, offsets_ptr=o, n_events=n); t.release()
    ms, k = ctx.take_stage_ms("k1_decode")
    print(mode, "k1 ms per launch", round(ms / k, 3), "events/s", round(n / (ms / k) * 1e3 / 1e9, 2), "G", dec.stats())
