#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
run() {  # name, env...
  name=$1; shift
  env "$@" FTF_CONV_PLAN_FILE=$O/r2p_plan.txt FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 30 --warmup 3 > $O/r2p_$name.json 2> $O/r2p_$name.txt
  python - "$name" <<'PY'
import json,sys
b=json.loads(open('gpurun_out/r2p_%s.json'%sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], 'ms_per_step', round(b['ms_per_step'],4), 'value', round(b['value']))
PY
}
run base A=1
run last FTF_POOL_LAST=1
run last_noprio FTF_POOL_LAST=1 FTF_CONV_PRIORITY=0
run base2 A=1
run last2 FTF_POOL_LAST=1
echo done
