#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
run() {  # name, env...
  name=$1; shift
  env FTF_CONV_PLAN_FILE=$O/r2v_plan.txt "$@" FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 30 --warmup 3 > $O/r2v_$name.json 2> $O/r2v_$name.txt
  python - "$name" <<'PY'
import json,sys
b=json.loads(open('gpurun_out/r2v_%s.json'%sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], 'ms_per_step', round(b['ms_per_step'],4), 'value', round(b['value']))
PY
}
run base A=1
run nostore FTF_EPI_NO_STORE=1
echo done
