#!/bin/bash
# 2 GPUs: multi-process tests (peer IPC, range-sharded compression, sharded step) and the default bench at N=2 on HEAD
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_peer_ipc_gpu.py tests/test_ranges_gpu.py tests/test_dist_gpu.py tests/test_sharded_gpu.py -m gpu -x -q -p no:cacheprovider ) > gpurun_out/pytest_2gpu.log 2>&1; echo "pytest rc=$?"
tail -n 5 gpurun_out/pytest_2gpu.log
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus 2 --steps 5 --warmup 3 ) > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench N=2 rc=$?"
tail -n 3 gpurun_out/bench_2gpu.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_2gpu.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','value_by_w_mode','stage_ms_per_step','stage_ms_slowest_step','scratch_allocations_in_timed_region','clocks','t2new_checksum_vs_stored','barrier_wait_ms_by_rank')})
print(d['e2e'])
print([ (c['N'],c['sampler'][:10],c['ms_per_compression']) for c in d.get('compress_sweep',[])])
PY
