#!/bin/bash
# A/B of the tensor-core search between the in-tree build and an older build of the same ABI (B2FLAT_SO)
OLD=${OLD:-nv_genai_b200/lib/old/libb2flat_r01.so}
ROWS=${ROWS:-10000000}
OUT=${OUT:-gpurun_out/ab_gemm.log}
KS=${KS:-"100 10"}
DTS=${DTS:-"bfloat16 float32"}
mkdir -p gpurun_out
: > $OUT
for dt in $DTS; do
  for k in $KS; do
    echo "== new $dt k=$k" >> $OUT
    B2_GEMM_DEBUG=1 python scripts/time_gemm.py --rows $ROWS --dtype $dt --k $k --reps 3 >> $OUT 2>&1
    if [ -f "$OLD" ]; then
      echo "== old $dt k=$k" >> $OUT
      B2FLAT_SO=$OLD B2_GEMM_DEBUG=1 python scripts/time_gemm.py --rows $ROWS --dtype $dt --k $k --reps 3 >> $OUT 2>&1
    fi
  done
done
grep -E "^==|per MMA warp|^\{" $OUT | cut -c1-420 | awk '!seen[$0]++' | tail -40
