#!/bin/bash
# usage: tools/k1_sweep.sh "<define variants...>"  — rebuilds dx_decode.o per variant on the GPU box and prints K1/K2 ms
cd /root/repo/data-accelerator_b200/csrc
for v in "$@"; do
  touch dx_decode.cu dx_rules.cu
  make -s EXTRA="$v" 2>&1 | grep -E "error" 
  cd /root/repo
  python bench.py --events 40000000 --steps 3 --warmup 2 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
k=d['roofline']['kernels']
print('$v', 'ms/step', d['ms_per_step'], {n:(x['ms_per_step'], x['gbs']) for n,x in k.items()}, 'e2e', d['e2e']['ms_per_step'])"
  cd /root/repo/data-accelerator_b200/csrc
done
