#!/bin/bash
# Round-2 evidence run on one B200: full GPU suite, every bench line, launch list and ncu captures replaying the plain run's kernel plan.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
rm -f $O/*.ncu-rep
timeout 1500 python -m pytest tests -x -q -m gpu > $O/ev_tests.log 2>&1; tail -3 $O/ev_tests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/ev_smoke.log 2>&1; tail -1 $O/ev_smoke.log
FTF_CONV_PLAN_FILE=$O/ev_plan.txt FTF_TIMELINE=8 timeout 900 python bench.py > $O/ev_bench_inception.json 2> $O/ev_timeline.txt; echo "bench rc=$?"
timeout 600 python bench.py --workload half_plus_two > $O/ev_bench_hpt.json 2>/dev/null; echo "hpt rc=$?"
timeout 600 python bench.py --workload normalize > $O/ev_bench_normalize.json 2>/dev/null; echo "norm rc=$?"
timeout 600 python bench.py --workload jpeg > $O/ev_bench_jpeg.json 2>/dev/null; echo "jpeg rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/ev_bench_reference.json 2>/dev/null; echo "ref rc=$?"
python - <<'PY'
import json
for n in ['inception','hpt','normalize','jpeg','reference']:
    try:
        b=json.loads(open('gpurun_out/ev_bench_%s.json'%n).read().strip().splitlines()[-1])
        print(n, b.get('value'), b.get('unit'), 'ms', b.get('ms_per_step'), 'e2e', (b.get('e2e') or {}).get('value'), 'roof', (b.get('roofline') or {}).get('frac'))
    except Exception as e: print(n, 'ERR', e)
PY
export FTF_BENCH_CUDA_PROFILER=1 FTF_NO_GRAPH=1 FTF_CONV_PLAN_FILE=$O/ev_plan.txt
timeout 300 python bench.py --quick --steps 2 --warmup 3 > $O/ev_plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv \
    --log-file $O/ev_launches_traffic.csv python bench.py --quick --steps 2 --warmup 3 > $O/ev_ncu1.log 2>&1
for win in "A 0 6" "B 20 8" "C 53 8"; do
  set -- $win
  timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:conv_tc_kernel|conv_halo_kernel' -s $2 -c $3 \
      -f -o $O/ev_conv_full_$1 python bench.py --quick --steps 2 --warmup 3 > $O/ev_ncu2_$1.log 2>&1
done
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:stem_conv|pool' -c 8 -f -o $O/ev_stem_pools_full python bench.py --quick --steps 2 --warmup 3 > $O/ev_ncu3.log 2>&1
ls -la $O/*.ncu-rep; du -sm $O
echo done
