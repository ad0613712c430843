#!/bin/bash
# Round-2 ncu evidence (each capture only after the same command has run clean without ncu):
#  1. launch list of the default bench command (kernel shares of a step)
#  2. --set full capture of the tensor-core search at 4M x 1024 bf16, batch 4096, k = 100 (pre-pass + main launch)
#  3. --set full capture of the batch-1 scan of the default bench
set -x
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-check > gpurun_out/ncu_bench_plain.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_default_bench.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-check > gpurun_out/ncu_bench.log 2>&1
python scripts/time_gemm.py --rows 4000000 --dtype bfloat16 --k 100 --reps 1 > gpurun_out/ncu_gemm_plain.json 2>/dev/null || exit 1
ncu --set full --clock-control none --import-source on -k regex:gemm_topk_kernel -s 3 -c 2 -o gpurun_out/r02_gemm_k100 -f \
    python scripts/time_gemm.py --rows 4000000 --dtype bfloat16 --k 100 --reps 1 > gpurun_out/ncu_gemm.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:scan_tc_kernel -s 2 -c 1 -o gpurun_out/r02_scantc_batch1 -f \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-check > gpurun_out/ncu_scan.log 2>&1
ls -la gpurun_out/*.ncu-rep
