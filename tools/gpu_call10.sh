#!/bin/bash
# 8 GPUs: sharded bench at N=8 and N=4
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for n in ${NS:-8}; do
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 ) > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench$n rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/bench_n$n.err
python - $n <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/bench_n{n}.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','n_gpus','ladder_ms_per_step','stage_ms_per_step','value_by_w_mode','gpu_launches','output_checksums')})
    print(d['roofline']['frac'], d['roofline']['avg_launch_ms'], d['roofline_planes']['avg_launch_ms'], d['e2e'], d['clocks'])
    print([ (c['N'],c['sampler'],c['ms_per_compression']) for c in d['compress_sweep']])
except Exception as e:
    print("bench parse failed", e)
PY
done
cat gpurun_out/summary.txt
