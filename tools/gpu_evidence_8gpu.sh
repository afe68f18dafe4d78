#!/bin/bash
# 8 GPUs of one box: host -> device bandwidth with all ranks copying at once, then the bench at N = 8 (and 4).
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/ev_topo.txt 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/probe_h2d.py > $O/ev_probe_h2d_8.txt 2>&1; tail -9 $O/ev_probe_h2d_8.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 3 > $O/ev_bench_8gpu.json 2> $O/ev_bench_8gpu.err; echo "bench8 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 > $O/ev_bench_2gpu.json 2> $O/ev_bench_2gpu.err; echo "bench2 rc=$?"
python - <<'PY'
import json
for n in ['8gpu','2gpu']:
    try:
        b=json.loads(open('gpurun_out/ev_bench_%s.json'%n).read().strip().splitlines()[-1])
        print(n, 'value', b.get('value'), 'ms', b.get('ms_per_step'), 'e2e', (b.get('e2e') or {}).get('value'), 'match', b.get('outputs_match_rank0'), b.get('logits_bits_match_rank0'))
        for k in ('streaming_jpeg','streaming','streaming_half_rate','host_numa_binding'):
            v=b.get(k); print('  ',k, (v.get('records_per_s') if isinstance(v,dict) else v))
    except Exception as e: print(n,'ERR',e)
PY
echo done
