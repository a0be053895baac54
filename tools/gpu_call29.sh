#!/bin/bash
# N GPUs: A/B of --overlap-planes (first ladder sub-shard encoded beside the compression, short blocks) against the default
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
N=${1:-2}
mkdir -p gpurun_out
for tag in default overlap; do
  extra=""; [ $tag = overlap ] && extra="--overlap-planes"
  ( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2972$N bench.py --gpus $N --steps 6 --warmup 3 --no-e2e --no-sweep $extra ) > gpurun_out/ab_${tag}_$N.json 2> gpurun_out/ab_${tag}_$N.err; echo "bench $tag rc=$?"
  python - $tag $N <<'PY'
import json,sys
t,n=sys.argv[1:3]
d=json.loads(open(f'gpurun_out/ab_{t}_{n}.json').read().strip().splitlines()[-1])
print(t, {k:d.get(k) for k in ('value','stage_ms_per_step','t2new_checksum_vs_stored')}, (d.get('roofline_planes') or {}).get('avg_launch_ms_beside_the_compression'), (d.get('roofline_planes') or {}).get('avg_launch_ms'))
PY
done
