#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_label_batch64.py -x -q -m gpu > $O/r2m_tests.log 2>&1; tail -3 $O/r2m_tests.log
FTF_CONV_PLAN_FILE=$O/r2m_plan.txt FTF_CONV_VERBOSE=1 FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2m_bench.json 2> $O/r2m_timeline.txt
grep "^refine" $O/r2m_timeline.txt | tail -20
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2m_bench.json').read().strip().splitlines()[-1]); print('ms_per_step', b['ms_per_step'], 'value', b['value'])
PY
FTF_CONV_PLAN_FILE=$O/r2m_plan.txt timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2m_bench2.json 2>/dev/null
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2m_bench2.json').read().strip().splitlines()[-1]); print('replayed plan: ms_per_step', b['ms_per_step'], 'value', b['value'])
PY
# K1: ncu of the variant that runs at 2^28 (affine_f32_bulk)
export FTF_BENCH_CUDA_PROFILER=1
timeout 300 python bench.py --quick --workload half_plus_two --steps 3 --warmup 3 > $O/r2m_plain_hpt.log 2>&1 &&
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:affine' -c 2 -f -o $O/r2m_k1_bulk python bench.py --quick --workload half_plus_two --steps 3 --warmup 3 > $O/r2m_ncu_k1.log 2>&1
timeout 300 python bench.py --quick --workload normalize --steps 3 --warmup 3 > $O/r2m_plain_norm.log 2>&1 &&
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:resize_norm' -c 2 -f -o $O/r2m_k2 python bench.py --quick --workload normalize --steps 3 --warmup 3 > $O/r2m_ncu_k2.log 2>&1
ls -la $O/*.ncu-rep | tail -3
echo done
