#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
rm -f $O/r2h_*.ncu-rep $O/r2m_*.ncu-rep
export FTF_BENCH_CUDA_PROFILER=1 FTF_NO_GRAPH=1
FTF_CONV_PLAN_FILE=$O/r2q_plan.txt timeout 300 python bench.py --quick --steps 2 --warmup 3 > $O/r2q_plain.log 2>&1 &&
FTF_CONV_PLAN_FILE=$O/r2q_plan.txt timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:avgpool3x3|maxpool3x3|global_avgpool' -c 14 \
      -f -o $O/r2q_pools_full python bench.py --quick --steps 2 --warmup 3 > $O/r2q_ncu.log 2>&1
ls -la $O/*.ncu-rep
echo done
