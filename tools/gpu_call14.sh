#!/bin/bash
# fused pairs: tests, then default bench (N=1)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_plugin_tc_gpu.py tests/test_dense_gpu.py tests/test_device_driver.py -q -m gpu -x -p no:cacheprovider > gpurun_out/pytest_fuse.log 2>&1; echo "tests rc=$?" >> gpurun_out/summary.txt
tail -n 25 gpurun_out/pytest_fuse.log
( time timeout 900 python bench.py --no-sweep --no-cpu-baseline ) > gpurun_out/bench14.json 2> gpurun_out/bench14.err; echo "bench rc=$?" >> gpurun_out/summary.txt
tail -n 5 gpurun_out/bench14.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/bench14.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ladder_ms_per_step','stage_ms_per_step','value_by_w_mode','gpu_launches','output_checksums')})
    print(d['roofline']['frac'], d['e2e']['value'], d.get('t2new_checksum_vs_stored'))
except Exception as e:
    print("bench parse failed", e)
PY
cat gpurun_out/summary.txt
