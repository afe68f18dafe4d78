#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_label_batch64.py tests/test_zz_no_alloc.py -x -q -m gpu > $O/r2n_tests.log 2>&1; tail -3 $O/r2n_tests.log
run() {  # name, env...
  name=$1; shift
  env "$@" FTF_CONV_PLAN_FILE=$O/r2n_plan.txt FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 30 --warmup 3 > $O/r2n_$name.json 2> $O/r2n_$name.txt
  python - "$name" <<'PY'
import json,sys
b=json.loads(open('gpurun_out/r2n_%s.json'%sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], 'ms_per_step', round(b['ms_per_step'],4), 'value', round(b['value']))
PY
}
run base A=1
run base2 A=1
run upfront FTF_POOL_UPFRONT=1
run rev135 FTF_REVERSE=1,3,5
run rev3 FTF_REVERSE=3
run rev36 FTF_REVERSE=3,6
run rev1356 FTF_REVERSE=1,3,5,6
run rev24 FTF_REVERSE=2,4
run base3 A=1
echo done
