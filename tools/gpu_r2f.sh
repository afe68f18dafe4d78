#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_conv.py tests/test_gpu_label_batch64.py -x -q -m gpu > $O/r2f_tests.log 2>&1; tail -3 $O/r2f_tests.log
FTF_CONV_VERBOSE=1 FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2f_bench.json 2> $O/r2f_timeline.txt
grep "^refine" $O/r2f_timeline.txt | tail -40
grep -c "wide" $O/r2f_timeline.txt
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2f_bench.json').read().strip().splitlines()[-1]); print('ms_per_step', b['ms_per_step'], 'value', b['value'], 'e2e', b['e2e']['value'])
PY
for kind in pmw pm2; do
FTF_CONV_KIND=$kind FTF_CONV_TUNE=0 FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 10 --warmup 3 > $O/r2f_bench_$kind.json 2> $O/r2f_timeline_$kind.txt
done
echo done
