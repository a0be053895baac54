#!/bin/bash
# 2 GPUs: peer-memory test with the device-side barrier, sharded bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
nvidia-smi -L > gpurun_out/gpus2.txt 2>&1
timeout 900 python -m pytest tests/test_peer_ipc_gpu.py tests/test_dist_gpu.py -q -m gpu -x -p no:cacheprovider > gpurun_out/pytest_peer.log 2>&1; echo "peer tests rc=$?" >> gpurun_out/summary.txt
tail -n 15 gpurun_out/pytest_peer.log
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 ) > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench2 rc=$?" >> gpurun_out/summary.txt
tail -n 8 gpurun_out/bench_n2.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/bench_n2.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','n_gpus','ladder_ms_per_step','stage_ms_per_step','value_by_w_mode','gpu_launches','output_checksums')})
    print(d['roofline']['frac'], d['e2e'], d['clocks'])
    print(d['compress_sweep'])
except Exception as e:
    print("bench parse failed", e)
PY
cat gpurun_out/summary.txt
