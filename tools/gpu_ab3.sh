#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_label_batch64.py tests/test_gpu_jpeg.py tests/test_gpu_parity.py -x -q -m gpu > $O/ab3_tests.log 2>&1; tail -3 $O/ab3_tests.log
for run in 16777216 0; do
FTF_UPLOAD_RUN=$run FTF_CONV_PLAN_FILE=$O/ab3_plan.txt timeout 900 python bench.py > $O/ab3_bench_$run.json 2> $O/ab3_bench.err; echo run=$run rc=$?
python - $run <<'PY'
import json,sys
b=json.loads(open('gpurun_out/ab3_bench_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
print('value', round(b['value']), 'e2e', round(b['e2e']['value']), 'p50', b['e2e']['latency_us_p50'], 'p99', b['e2e']['latency_us_p99'], 'stream', round(b['streaming']['records_per_s']), b['streaming']['latency_us_p50'], 'half-rate p50/p99', b['streaming_half_rate']['latency_us_p50'], b['streaming_half_rate']['latency_us_p99'])
print(' per_record', round(b['per_record']['half_plus_two']['calls_per_s']), round(b['per_record']['inception_label']['calls_per_s']), 'jpeg stream', round(b['streaming_jpeg']['records_per_s']))
PY
done
timeout 600 python bench.py --workload normalize > $O/ab3_norm.json 2>/dev/null; timeout 600 python bench.py --workload half_plus_two > $O/ab3_hpt.json 2>/dev/null
echo done
