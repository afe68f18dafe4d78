#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r2i_tests.log 2>&1; tail -3 $O/r2i_tests.log
timeout 900 python bench.py --steps 20 --warmup 5 > $O/r2i_bench.json 2> $O/r2i_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2i_bench_ref.json 2> $O/r2i_bench_ref.err; echo "ref rc=$?"
FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2i_bench_quick.json 2> $O/r2i_timeline.txt
python - <<'PY'
import json
for f in ('r2i_bench','r2i_bench_quick'):
    b=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1]); print(f,'ms_per_step', b['ms_per_step'], 'value', b['value'], 'e2e', b['e2e']['value'], b['roofline']['frac'])
PY
for wl in half_plus_two normalize jpeg; do timeout 600 python bench.py --workload $wl --steps 50 --warmup 5 > $O/r2i_bench_$wl.json 2> $O/r2i_bench_$wl.err; done
echo done
