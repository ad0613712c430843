#!/bin/bash
# usage: scripts/bench_batch.sh <batch> <dtype> [steps]   -> "value ms_per_step kernel_ms frac" (one line), env passes through
b=$1; dt=$2; st=${3:-30}
timeout 300 python bench.py --batch $b --dtype $dt --steps $st --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith(chr(123))][-1])
r=d['roofline']
print('batch',d['config']['batch'],'$dt','qps',round(d['value'],1),'ms',round(d['ms_per_step'],3),'kernel_ms',round(r['avg_launch_ms'],3),'frac',round(r['frac'],3))"
