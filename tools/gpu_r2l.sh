#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > $O/r2l_bench_2gpu.json 2> $O/r2l_bench_2gpu.err; echo "rc=$?"
tail -c 600 $O/r2l_bench_2gpu.err
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2l_bench_2gpu.json').read().strip().splitlines()[-1])
print('value', b['value'], 'ms', b['ms_per_step'], 'e2e', b['e2e']['value'], 'match', b.get('outputs_match_rank0'), b.get('logits_bits_match_rank0'))
for k in ['streaming_jpeg','streaming','streaming_half_rate']:
    s=b[k]; print(k, round(s['records_per_s']), s['records_per_s_this_rank'], s['latency_us_p50'], s['latency_us_p99'], s['ranks'])
print(b['host_numa_binding'])
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > $O/r2l_ref_2gpu.json 2>/dev/null; echo "ref rc=$?"; head -c 300 $O/r2l_ref_2gpu.json
echo done
