#!/bin/bash
# Round 2, GPU call B: default bench, kernel timeline, plan choices, role counters, ncu of the shipped conv kernels (CSV exports only).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r2b_bench.json 2> $O/r2b_bench.err; echo "bench rc=$?"
FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 10 --warmup 3 > /dev/null 2> $O/r2b_timeline.txt
FTF_CONV_VERBOSE=1 timeout 300 python bench.py --quick --steps 3 --warmup 3 > /dev/null 2> $O/r2b_plan_choices.txt
for kind in pm pm2 wm; do
  FTF_CONV_KIND=$kind FTF_CONV_TUNE=0 FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 10 --warmup 3 > $O/r2b_bench_$kind.json 2> $O/r2b_timeline_$kind.txt
done
FTF_CONV_PAIR=1 FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 10 --warmup 3 > $O/r2b_bench_pairtuned.json 2> $O/r2b_timeline_pairtuned.txt
FTF_CONV_PROFILE=1 timeout 600 python tools/profile_conv.py --json $O/r2b_conv_layers.json > $O/r2b_conv_roles.txt 2>&1
export FTF_BENCH_CUDA_PROFILER=1 FTF_NO_GRAPH=1
timeout 300 python bench.py --quick --steps 2 --warmup 3 > $O/r2b_plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv \
    --log-file $O/r2b_launches_traffic.csv python bench.py --quick --steps 2 --warmup 3 > $O/r2b_ncu1.log 2>&1
# --set full on the shipped conv kernels, three windows of one timed step: stem + Mixed_5b (9 launches), Mixed_6b (8), Mixed_7b (6)
for win in "A 0 9" "B 20 8" "C 53 6"; do
  set -- $win
  timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:conv_tc_kernel|conv_halo_kernel' -s $2 -c $3 \
      -f -o $O/r2b_conv_full_$1 python bench.py --quick --steps 2 --warmup 3 > $O/r2b_ncu2_$1.log 2>&1
done
ls -la $O/*.ncu-rep
if [ $(du -sm $O | cut -f1) -gt 58 ]; then rm -f $O/r2b_conv_full_C.ncu-rep; fi
du -sh $O
echo done
