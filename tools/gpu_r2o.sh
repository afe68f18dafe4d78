#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_label_batch64.py -x -q -m gpu -k "pool or network or batch64 or runner" > $O/r2o_tests.log 2>&1; tail -3 $O/r2o_tests.log
run() {  # name, env...
  name=$1; shift
  env "$@" FTF_CONV_PLAN_FILE=$O/r2o_plan.txt FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 30 --warmup 3 > $O/r2o_$name.json 2> $O/r2o_$name.txt
  python - "$name" <<'PY'
import json,sys
b=json.loads(open('gpurun_out/r2o_%s.json'%sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], 'ms_per_step', round(b['ms_per_step'],4), 'value', round(b['value']))
PY
}
run prio A=1
run noprio FTF_CONV_PRIORITY=0
run prio2 A=1
run noprio_cap2 FTF_CONV_PRIORITY=0 FTF_BRANCH_POOL_CTAS=2
run noprio_cap4 FTF_CONV_PRIORITY=0 FTF_BRANCH_POOL_CTAS=4
run prio_cap4 FTF_BRANCH_POOL_CTAS=4
run prio_poolfirst FTF_POOL_FIRST=1
run noprio2 FTF_CONV_PRIORITY=0
echo done
