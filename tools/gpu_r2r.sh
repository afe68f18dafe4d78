#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
rm -f $O/r2q_*.ncu-rep
timeout 900 python -m pytest tests/test_gpu_label_batch64.py -x -q -m gpu > $O/r2r_tests.log 2>&1; tail -3 $O/r2r_tests.log
run() {  # name, env...
  name=$1; shift
  env "$@" FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 30 --warmup 3 > $O/r2r_$name.json 2> $O/r2r_$name.txt
  python - "$name" <<'PY'
import json,sys
b=json.loads(open('gpurun_out/r2r_%s.json'%sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], 'ms_per_step', round(b['ms_per_step'],4), 'value', round(b['value']), 'launches', b.get('gpu_launches'))
PY
}
run g16 A=1
run g0 FTF_EARLY_GROUP=0
run g8 FTF_EARLY_GROUP=8
run g32 FTF_EARLY_GROUP=32
run g16b A=1
run g0b FTF_EARLY_GROUP=0
echo done
