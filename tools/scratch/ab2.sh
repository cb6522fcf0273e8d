#!/bin/bash
run() { python bench.py --steps 40 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('  ', round(d['value'],1), 'tok/s', round(d['e2e']['value'],1), 'e2e')"; }
for rep in 1 2 3; do
  echo base; DLLM_B200_LIB=$PWD/distributed-llms_b200/lib/variants/libdllm_base.so run
  echo new; run
done
