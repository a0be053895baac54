#!/bin/bash
# N GPUs (default 2): default bench on HEAD (e2e with the residual read back under the last contractions)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
N=${1:-2}
mkdir -p gpurun_out
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2971$N bench.py --gpus $N --steps 5 --warmup 3 ) > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench N=$N rc=$?"
tail -n 3 gpurun_out/bench_${N}gpu.err
python - $N <<'PY'
import json,sys
n=sys.argv[1]
d=json.loads(open(f'gpurun_out/bench_{n}gpu.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','value_by_w_mode','stage_ms_per_step','stage_ms_slowest_step','clocks','t2new_checksum_vs_stored','barrier_wait_ms_by_rank')})
print(d['e2e'])
print([ (c['N'],c['sampler'][:10],c['ms_per_compression']) for c in d.get('compress_sweep',[])])
PY
