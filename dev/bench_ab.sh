#!/bin/bash
# usage: dev/bench_ab.sh VAR "v1 v2 ..." [bench args]   -- runs bench.py once per value of the environment variable
var=$1; vals=$2; shift 2
for v in $vals; do
  env $var=$v timeout 300 python bench.py --steps 200 --warmup 10 --no-e2e "$@" 2>/dev/null | grep '^{' | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$var=$v', round(d['ms_per_step'],4), round(d['value']/1e6,2), d['final_loss'])"
done
