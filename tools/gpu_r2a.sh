#!/bin/bash
# Round 2, first GPU call: full GPU test suite, default bench, kernel timeline, role counters, ncu of the shipped conv kernels.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r2a_gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2a_tests.log
tail -5 gpurun_out/r2a_tests.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
FTF_TIMELINE=8 timeout 300 python bench.py --quick --steps 10 --warmup 3 > /dev/null 2> gpurun_out/r2a_timeline.txt
FTF_CONV_VERBOSE=1 timeout 300 python bench.py --quick --steps 3 --warmup 3 > /dev/null 2> gpurun_out/r2a_plan_choices.txt
FTF_CONV_PROFILE=1 timeout 600 python tools/profile_conv.py --json gpurun_out/r2a_conv_layers.json > gpurun_out/r2a_conv_roles.txt 2>&1
# ncu: DRAM bytes + duration of every kernel of the timed steps (plain stream launches so that every kernel is listed)
export FTF_BENCH_CUDA_PROFILER=1 FTF_NO_GRAPH=1
timeout 300 python bench.py --quick --steps 2 --warmup 3 > gpurun_out/r2a_plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv \
    --log-file gpurun_out/r2a_launches_traffic.csv python bench.py --quick --steps 2 --warmup 3 > gpurun_out/r2a_ncu1.log 2>&1
timeout 1500 ncu --profile-from-start off --set full --clock-control none --import-source on -k 'regex:conv_tc_kernel|conv_halo_kernel' -c 67 \
    -o gpurun_out/r2a_conv_full python bench.py --quick --steps 2 --warmup 3 > gpurun_out/r2a_ncu2.log 2>&1
echo done
