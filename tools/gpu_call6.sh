#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_dense_gpu.py tests/test_plugin_tc_gpu.py -q -m gpu -x -p no:cacheprovider > gpurun_out/pytest_tc.log 2>&1; rc=$?
echo "tc tests rc=$rc" >> gpurun_out/summary.txt
tail -n 12 gpurun_out/pytest_tc.log
timeout 600 python tools/bench_kernels.py --planes 6 --reps 3 --skip-two-pass > gpurun_out/bench_kernels6.jsonl 2> gpurun_out/bench_kernels.err; echo "bench_kernels rc=$?" >> gpurun_out/summary.txt
cat gpurun_out/bench_kernels6.jsonl; tail -n 5 gpurun_out/bench_kernels.err
timeout 900 python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline > gpurun_out/bench6.json 2> gpurun_out/bench6.err; echo "bench rc=$?" >> gpurun_out/summary.txt
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/bench6.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ladder_ms_per_step','stage_ms_per_step','value_by_w_mode')})
    print(d['roofline']['frac'], d['roofline']['achieved'], d['roofline']['peak'], d['roofline_planes']['avg_launch_ms'], d['e2e']['value'], d['clocks'])
except Exception as e:
    print("bench parse failed", e)
PY
tail -n 5 gpurun_out/bench6.err
cat gpurun_out/summary.txt
