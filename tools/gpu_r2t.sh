#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_label_batch64.py tests/test_zz_no_alloc.py -x -q -m gpu -k "pool or network or batch_64 or layer or alloc" > $O/r2t_tests.log 2>&1; tail -3 $O/r2t_tests.log
run() {  # name, env...
  name=$1; shift
  env FTF_CONV_PLAN_FILE=$O/r2t_plan.txt "$@" FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 30 --warmup 3 > $O/r2t_$name.json 2> $O/r2t_$name.txt
  python - "$name" <<'PY'
import json,sys
b=json.loads(open('gpurun_out/r2t_%s.json'%sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], 'ms_per_step', round(b['ms_per_step'],4), 'value', round(b['value']), 'launches', b.get('gpu_launches'))
PY
  grep avgpool $O/r2t_$name.txt | head -9 | tr '\n' ';'; echo
}
run tma A=1
run tma2 A=1
run tma_noprio FTF_CONV_PRIORITY=0
run tma_first FTF_POOL_FIRST=1
run tma_last FTF_POOL_LAST=1
run tma3 A=1
echo done
