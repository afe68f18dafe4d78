#!/bin/bash
# A/B runs of the batch-64 step inside one GPU call: tools/gpu_ab.sh name1 "ENV=.. ENV=.." name2 "..." ...
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
while [ $# -ge 2 ]; do
  name=$1; envs=$2; shift 2
  env FTF_CONV_PLAN_FILE=$O/ab_plan.txt $envs FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 30 --warmup 3 > $O/ab_$name.json 2> $O/ab_$name.txt
  python - "$name" <<'PY'
import json,sys
b=json.loads(open('gpurun_out/ab_%s.json'%sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], 'ms_per_step', round(b['ms_per_step'],4), 'value', round(b['value']))
PY
done
echo done
