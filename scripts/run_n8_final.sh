#!/bin/bash
# cfg4 and cfg2 at 8 GPUs on the final build
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29588"
mkdir -p gpurun_out
for cfg in cfg4 cfg2; do
  steps=100; [ $cfg = cfg4 ] && steps=10
  $T bench.py --gpus 8 --config $cfg --steps $steps --warmup 5 > gpurun_out/r02c_bench_${cfg}_n8.json 2> gpurun_out/bench_${cfg}_n8c.err
  python -c "
import json; j=json.load(open('gpurun_out/r02c_bench_${cfg}_n8.json')); print('$cfg', round(j['value'],1), round(j['ms_per_step'],4), 'e2e', round(j['e2e']['value'],1), j['check'].get('planted'), 'frac', round(j['roofline']['frac'],3), j['clocks']['sm_mhz'], j.get('cfg4_efficiency'))"
done
