#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_label_batch64.py -x -q -m gpu -k "network or batch_64 or layer or stem or fp32 or records" > $O/r2u_tests.log 2>&1; tail -3 $O/r2u_tests.log
run() {  # name, env...
  name=$1; shift
  env FTF_CONV_PLAN_FILE=$O/r2u_plan.txt "$@" FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 30 --warmup 3 > $O/r2u_$name.json 2> $O/r2u_$name.txt
  python - "$name" <<'PY'
import json,sys
b=json.loads(open('gpurun_out/r2u_%s.json'%sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], 'ms_per_step', round(b['ms_per_step'],4), 'value', round(b['value']), 'e2e', round(b['e2e']['value']))
PY
  grep " stem " $O/r2u_$name.txt | head -3 | tr '\n' ';'; echo
}
run c6 A=1
run c4 FTF_STEM_CTAS=4
run c5 FTF_STEM_CTAS=5
run c6b A=1
echo done
