#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_label_batch64.py -x -q -m gpu > $O/r2d_tests.log 2>&1; tail -3 $O/r2d_tests.log
FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2d_bench.json 2> $O/r2d_timeline.txt
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2d_bench.json').read().strip().splitlines()[-1]); print('ms_per_step', b['ms_per_step'], 'value', b['value'], 'e2e', b['e2e']['value'])
PY
FTF_CONV_KIND=wm FTF_CONV_TUNE=0 FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 10 --warmup 3 > $O/r2d_bench_wm.json 2> $O/r2d_timeline_wm.txt
FTF_CONV_PROFILE=1 timeout 600 python tools/profile_conv.py --json $O/r2d_conv_layers.json > $O/r2d_conv_roles.txt 2>&1
timeout 600 python tools/profile_conv.py --json $O/r2d_conv_layers_noprof.json > $O/r2d_conv_layers_noprof.txt 2>&1
echo done
