#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_label_batch64.py tests/test_c_abi_driver.py -x -q -m gpu > $O/r2c_tests.log 2>&1; tail -3 $O/r2c_tests.log
FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2c_bench.json 2> $O/r2c_timeline.txt; tail -c 300 $O/r2c_bench.json | head -c 10; python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2c_bench.json').read().strip().splitlines()[-1]); print('ms_per_step', b['ms_per_step'], 'value', b['value'], 'e2e', b['e2e']['value'])
PY
for kind in pm wm; do
  FTF_CONV_KIND=$kind FTF_CONV_TUNE=0 FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 10 --warmup 3 > $O/r2c_bench_$kind.json 2> $O/r2c_timeline_$kind.txt
done
FTF_CONV_VERBOSE=1 timeout 300 python bench.py --quick --steps 3 --warmup 3 > /dev/null 2> $O/r2c_plan_choices.txt
echo done
