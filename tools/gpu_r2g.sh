#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r2g_tests.log 2>&1; tail -5 $O/r2g_tests.log
FTF_CONV_VERBOSE=1 FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2g_bench.json 2> $O/r2g_timeline.txt
grep "^refine" $O/r2g_timeline.txt | tail -40
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2g_bench.json').read().strip().splitlines()[-1]); print('ms_per_step', b['ms_per_step'], 'value', b['value'], 'e2e', b['e2e']['value'])
PY
FTF_STEM_DIRECT=1 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2g_bench_stemdirect.json 2>/dev/null
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2g_bench_stemdirect.json').read().strip().splitlines()[-1]); print('stem direct: ms_per_step', b['ms_per_step'])
PY
for kind in pmw pm2; do
FTF_CONV_KIND=$kind FTF_CONV_TUNE=0 FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 10 --warmup 3 > $O/r2g_bench_$kind.json 2> $O/r2g_timeline_$kind.txt
done
echo done
