#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
export FTF_CONV_PLAN_FILE=$O/r2j_plan.txt
run() { # name, env...
  name=$1; shift
  env "$@" FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2k_bench_$name.json 2> $O/r2k_timeline_$name.txt
  python - <<PY
import json
b=json.loads(open('gpurun_out/r2k_bench_$name.json').read().strip().splitlines()[-1]); print('$name: ms_per_step', b['ms_per_step'])
PY
}
run base A=1
run c2 FTF_BRANCH_POOL_CTAS=2
run c1 FTF_BRANCH_POOL_CTAS=1
run poolfirst FTF_POOL_FIRST=1
run poolfirst_c2 FTF_POOL_FIRST=1 FTF_BRANCH_POOL_CTAS=2
run c4 FTF_BRANCH_POOL_CTAS=4
echo done
