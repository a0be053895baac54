#!/bin/bash
# 8 GPUs: 2-process peer test (copy-engine stores), then the sharded bench at N=8 (default, and A/B variants), N=4, N=2
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_peer_ipc_gpu.py -q -m gpu -x -p no:cacheprovider > gpurun_out/pytest_peer.log 2>&1; echo "peer tests rc=$?" >> gpurun_out/summary.txt
tail -n 6 gpurun_out/pytest_peer.log
run() {  # tag n extra...
  tag=$1; n=$2; shift 2
  ( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 5 --warmup 3 "$@" ) > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench_$tag rc=$?" >> gpurun_out/summary.txt
  tail -n 3 gpurun_out/bench_$tag.err
  python - $tag <<'PY'
import json,sys
t=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/bench_{t}.json').read().strip().splitlines()[-1])
    print(t, {k:d.get(k) for k in ('value','n_gpus','ladder_ms_per_step','stage_ms_per_step','value_by_w_mode','gpu_launches','peer_copies','barrier_wait_ms_by_rank')})
    print(d['roofline']['frac'], d['roofline']['avg_launch_ms'], d.get('roofline_planes'), d.get('e2e'), d['clocks'], d.get('t2new_checksum_vs_stored'))
except Exception as e:
    print("bench parse failed", e)
PY
}
run n8 8 --no-sweep
run n8_parts2 8 --no-sweep --no-e2e --last-term-parts 2
run n8_old 8 --no-sweep --no-e2e --push-kernel --no-overlap-planes --last-term-parts 2
run n4 4 --no-sweep
run n2 2 --no-sweep
cat gpurun_out/summary.txt
