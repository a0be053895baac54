#!/bin/bash
# full GPU suite + default bench + reference arm
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 1500 python -m pytest tests -q -m gpu -x -p no:cacheprovider > gpurun_out/pytest_all.log 2>&1; echo "all gpu tests rc=$?" >> gpurun_out/summary.txt
tail -n 15 gpurun_out/pytest_all.log
( time timeout 1200 python bench.py ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?" >> gpurun_out/summary.txt
tail -n 6 gpurun_out/bench_default.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/bench_default.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ladder_ms_per_step','stage_ms_per_step','value_by_w_mode','gpu_launches','output_checksums')})
    print(d['roofline']['frac'], d['roofline']['achieved'], d['roofline']['peak'], d['roofline_planes']['avg_launch_ms'], d['e2e'], d['clocks'], d['cpu_baseline'])
    print(d['compress_sweep'])
except Exception as e:
    print("bench parse failed", e)
PY
cat gpurun_out/summary.txt
