#!/bin/bash
# 8 GPUs: A/B of the exchange variants at N=8 (short runs, no e2e / sweep), 2-process peer test first
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_peer_ipc_gpu.py -q -m gpu -x -p no:cacheprovider > gpurun_out/pytest_peer.log 2>&1; echo "peer tests rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/pytest_peer.log
run() {  # tag n extra...
  tag=$1; n=$2; shift 2
  ( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2971$n bench.py --gpus $n --steps 5 --warmup 3 "$@" ) > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench_$tag rc=$?" >> gpurun_out/summary.txt
  tail -n 2 gpurun_out/bench_$tag.err | head -1
  python - $tag <<'PY'
import json,sys
t=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/bench_{t}.json').read().strip().splitlines()[-1])
    print(t, {k:d.get(k) for k in ('value','ladder_ms_per_step','stage_ms_per_step','value_by_w_mode','peer_copies','barrier_wait_ms_by_rank')}, d.get('t2new_checksum_vs_stored'))
except Exception as e:
    print(t, "bench parse failed", e)
PY
}
run ab_default 8 --no-sweep --no-e2e
run ab_carveout 8 --no-sweep --no-e2e --uniform-carveout 100
run ab_ce_fan 8 --no-sweep --no-e2e --push-copy-engine 2 --last-term-parts 1
run ab_overlap 8 --no-sweep --no-e2e --overlap-planes
run ab_carveout_overlap 8 --no-sweep --no-e2e --uniform-carveout 100 --overlap-planes
cat gpurun_out/summary.txt
