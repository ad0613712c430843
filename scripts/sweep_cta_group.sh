for cg in 1 2; do for mixed in 0 1; do B2_GEMM_CTA_GROUP=$cg B2_GEMM_MIXED=$mixed timeout 300 python scripts/time_gemm.py --dtype float32 --sweep 100:64:2 2>&1 | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('f32 cg=$cg mixed=$mixed', round(d['ms'],1), d['ids_equal_scan_sample'], d['max_rel_score_diff_vs_scan'])"; done; done
for cg in 1 2; do B2_GEMM_CTA_GROUP=$cg timeout 300 python scripts/time_gemm.py --dtype bfloat16 --sweep 100:64:2,100:128:2,100:64:1,100:128:1 2>&1 | python -c "
import json,sys
for l in sys.stdin.read().splitlines():
    if l.startswith('{'):
        d=json.loads(l); print('bf16 cg=$cg sw', d['swizzle'], 'parts', d['products'], round(d['ms'],1), d['ids_equal_scan_sample'])"; done
