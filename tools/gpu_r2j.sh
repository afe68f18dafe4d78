#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_label_batch64.py -x -q -m gpu > $O/r2j_tests.log 2>&1; tail -3 $O/r2j_tests.log
export FTF_CONV_PLAN_FILE=$O/r2j_plan.txt
for c in 1 2 16; do
FTF_BRANCH_POOL_CTAS=$c FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2j_bench_c$c.json 2> $O/r2j_timeline_c$c.txt
python - <<PY
import json
b=json.loads(open('gpurun_out/r2j_bench_c$c.json').read().strip().splitlines()[-1]); print('pool ctas $c: ms_per_step', b['ms_per_step'], 'value', b['value'])
PY
done
wc -l $O/r2j_plan.txt
echo done
