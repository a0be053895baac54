#!/bin/bash
# 1 GPU: rank-0-of-8 share, event timing vs per-kernel timing (host gaps), then the launch list of the same
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python bench.py --steps 3 --warmup 2 --no-e2e --no-sweep --no-cpu-baseline --as-rank-of 8 > gpurun_out/bench_as8.json 2> gpurun_out/bench_as8.err; echo "as8 rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_as8.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ladder_ms_per_step','stage_ms_per_step','gpu_launches','output_checksums')})
PY
tail -n 3 gpurun_out/bench_as8.err
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_as8.csv \
   python bench.py --steps 1 --warmup 1 --no-e2e --no-sweep --no-cpu-baseline --as-rank-of 8 > gpurun_out/ncu_as8.log 2>&1
echo "ncu rc=$?"
