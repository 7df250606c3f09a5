#!/bin/bash
# tools/stages.sh TAG...: per-stage device times of bench.py's default workload for tuning variants (variants/libghr_TAG.so)
for t in "$@"; do
  GHR_LIB_PATH=$PWD/variants/libghr_$t.so python bench.py --no-cpu-baseline --steps 60 --warmup 5 ${BENCH_ARGS:-} 2>/dev/null | python -c "
import json,sys
b=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('$t', b['value'], b['ms_per_step'], {k:v['us'] for k,v in b['stages'].items() if k.startswith('blend') or k=='sort'})"
done
