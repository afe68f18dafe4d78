#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_conv.py tests/test_gpu_label_batch64.py tests/test_gpu_parity.py -x -q -m gpu > $O/r2e_tests.log 2>&1; tail -3 $O/r2e_tests.log
FTF_CONV_VERBOSE=1 FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2e_bench.json 2> $O/r2e_timeline.txt
grep "^refine" $O/r2e_timeline.txt | tail -40
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2e_bench.json').read().strip().splitlines()[-1]); print('ms_per_step', b['ms_per_step'], 'value', b['value'], 'e2e', b['e2e']['value'])
PY
FTF_NO_REFINE=1 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2e_bench_norefine.json 2>/dev/null
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2e_bench_norefine.json').read().strip().splitlines()[-1]); print('no refine: ms_per_step', b['ms_per_step'])
PY
timeout 300 python bench.py --quick --workload normalize --steps 200 --warmup 5 > $O/r2e_bench_normalize.json 2>/dev/null
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2e_bench_normalize.json').read().strip().splitlines()[-1]); print('normalize ms_per_step', b['ms_per_step'], 'frac', b['roofline']['frac'])
PY
echo done
