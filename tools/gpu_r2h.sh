#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_conv.py -x -q -m gpu -k "inception or fp32 or network" > $O/r2h_tests.log 2>&1; tail -3 $O/r2h_tests.log
FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 20 --warmup 3 > $O/r2h_bench.json 2> $O/r2h_timeline.txt
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2h_bench.json').read().strip().splitlines()[-1]); print('ms_per_step', b['ms_per_step'], 'value', b['value'], 'e2e', b['e2e']['value'])
PY
grep -A3 "timeline batch" $O/r2h_timeline.txt | tail -4
export FTF_BENCH_CUDA_PROFILER=1 FTF_NO_GRAPH=1
timeout 300 python bench.py --quick --steps 2 --warmup 3 > $O/r2h_plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv \
    --log-file $O/r2h_launches_traffic.csv python bench.py --quick --steps 2 --warmup 3 > $O/r2h_ncu1.log 2>&1
for win in "A 0 5" "B 20 8" "C 53 6"; do
  set -- $win
  timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:conv_tc_kernel|conv_halo_kernel' -s $2 -c $3 \
      -f -o $O/r2h_conv_full_$1 python bench.py --quick --steps 2 --warmup 3 > $O/r2h_ncu2_$1.log 2>&1
done
timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:stem_conv' -c 1 -f -o $O/r2h_stem_full python bench.py --quick --steps 2 --warmup 3 > $O/r2h_ncu3.log 2>&1
ls -la $O/*.ncu-rep; du -sm $O
echo done
