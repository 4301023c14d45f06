#!/bin/bash
# tools/variant_sweep.sh NAME... — runs the bench on each prebuilt variant library (tools/build_variant.sh) and on the in-tree build
cd "$(dirname "$0")/.."
for v in base "$@"; do
  if [ "$v" = base ]; then unset DXGPU_LIB; else export DXGPU_LIB=$PWD/data-accelerator_b200/variants/libdxgpu_$v.so; fi
  python bench.py --events 40000000 --steps 3 --warmup 3 --no-cpu-baseline --e2e-partitions 1 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
k=d['roofline']['kernels']
print('$v', 'ms/step', d['ms_per_step'], {n:(x['ms_per_step'], x['gbs']) for n,x in k.items()})"
done
