#!/bin/bash
# Round-2 evidence run on one B200: full GPU suite, every bench line, launch list and ncu captures replaying the plain run's kernel plan.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
rm -f $O/*.ncu-rep
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r2w_tests.log 2>&1; tail -3 $O/r2w_tests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r2w_smoke.log 2>&1; tail -1 $O/r2w_smoke.log
FTF_CONV_PLAN_FILE=$O/r2w_plan.txt FTF_TIMELINE=8 timeout 900 python bench.py > $O/r2w_bench_inception.json 2> $O/r2w_timeline.txt; echo "bench rc=$?"
timeout 600 python bench.py --workload half_plus_two > $O/r2w_bench_hpt.json 2>/dev/null; echo "hpt rc=$?"
timeout 600 python bench.py --workload normalize > $O/r2w_bench_normalize.json 2>/dev/null; echo "norm rc=$?"
timeout 600 python bench.py --workload jpeg > $O/r2w_bench_jpeg.json 2>/dev/null; echo "jpeg rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2w_bench_reference.json 2>/dev/null; echo "ref rc=$?"
python - <<'PY'
import json
for n in ['inception','hpt','normalize','jpeg','reference']:
    try:
        b=json.loads(open('gpurun_out/r2w_bench_%s.json'%n).read().strip().splitlines()[-1])
        print(n, b.get('value'), b.get('unit'), 'ms', b.get('ms_per_step'), 'e2e', (b.get('e2e') or {}).get('value'), 'roof', (b.get('roofline') or {}).get('frac'))
    except Exception as e: print(n, 'ERR', e)
PY
export FTF_BENCH_CUDA_PROFILER=1 FTF_NO_GRAPH=1 FTF_CONV_PLAN_FILE=$O/r2w_plan.txt
timeout 300 python bench.py --quick --steps 2 --warmup 3 > $O/r2w_plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv \
    --log-file $O/r2w_launches_traffic.csv python bench.py --quick --steps 2 --warmup 3 > $O/r2w_ncu1.log 2>&1
for win in "A 0 6" "B 20 8" "C 53 8"; do
  set -- $win
  timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:conv_tc_kernel|conv_halo_kernel' -s $2 -c $3 \
      -f -o $O/r2w_conv_full_$1 python bench.py --quick --steps 2 --warmup 3 > $O/r2w_ncu2_$1.log 2>&1
done
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:stem_conv|pool' -c 8 -f -o $O/r2w_stem_pools_full python bench.py --quick --steps 2 --warmup 3 > $O/r2w_ncu3.log 2>&1
ls -la $O/*.ncu-rep; du -sm $O
echo done
