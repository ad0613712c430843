#!/bin/bash
# One multi-GPU measurement session: bit-equality of the sharded search, then the bench lines of cfg2 / cfg2b16 / cfg4.
N=${1:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555"
mkdir -p gpurun_out
$T tests/mp_sharded_gpu.py > gpurun_out/r02_mp_sharded_${N}gpu_equality.json 2> gpurun_out/mp_${N}.err
echo "equality rc=$?"; python -c "
import json
for l in open('gpurun_out/r02_mp_sharded_${N}gpu_equality.json'):
    if l.startswith('{'):
        j=json.loads(l); print('equality ok=', j['ok'], 'cases', len(j['cases']), 'all equal', all(c['equal'] for c in j['cases']), 'peer_active', {c.get('peer_active') for c in j['cases']})
"
for cfg in cfg2 cfg2b16 cfg4; do
  steps=100; [ $cfg = cfg4 ] && steps=10
  $T bench.py --gpus $N --config $cfg --steps $steps --warmup 5 > gpurun_out/r02_bench_${cfg}_n${N}.json 2> gpurun_out/bench_${cfg}_n${N}.err
  python -c "
import json; j=json.load(open('gpurun_out/r02_bench_${cfg}_n${N}.json')); print('$cfg', round(j['value'],1), round(j['ms_per_step'],4), 'e2e', round(j['e2e']['value'],1), j['check'].get('planted'), j['check'].get('oracle_query'), 'frac', round(j['roofline']['frac'],3), j['roofline']['kernel'], j['clocks'], j.get('cfg4_efficiency'))" || tail -c 600 gpurun_out/bench_${cfg}_n${N}.err
done
