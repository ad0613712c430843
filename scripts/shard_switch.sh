#!/bin/bash
# batch-1 step time on shard-sized stores: the FFMA scan vs the tensor-core scan (b2_scan_tc.cu: the 8 GB switch)
mkdir -p gpurun_out
OUT=gpurun_out/shard_switch.txt
: > $OUT
for rows in 1250000 2500000 5000000; do
  for mode in default tc ffma; do
    case $mode in
      default) envs="";;
      tc) envs="B2_SCAN_TC_MIN_NQ=1";;
      ffma) envs="B2_NO_SCAN_TC=1";;
    esac
    for batch in 1 16; do
      line=$(env $envs python bench.py --rows $rows --batch $batch --steps 200 --warmup 20 --no-cpu-baseline --no-check 2>/dev/null)
      echo "$rows $mode batch=$batch $(echo "$line" | python -c 'import json,sys; j=json.loads(sys.stdin.read()); print(round(j["ms_per_step"]*1e3,1),"us/step kernel", j["roofline"]["kernel"], round(j["roofline"]["avg_launch_ms"]*1e3,1), "us frac", round(j["roofline"]["frac"],3), "e2e us", round(j["e2e"]["ms_per_step"]*1e3,1))')" >> $OUT
    done
  done
done
cat $OUT
