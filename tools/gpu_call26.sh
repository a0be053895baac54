#!/bin/bash
# 1 GPU: constant-operand planes (test + bench A/B)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_plugin_tc_gpu.py tests/test_device_driver.py tests/test_sharded_gpu.py -m gpu -x -q -p no:cacheprovider ) > gpurun_out/pytest_const.log 2>&1; echo "pytest rc=$?"
tail -n 15 gpurun_out/pytest_const.log
for tag in on off; do
  extra=""; [ $tag = off ] && extra="--no-constant-eris"
  ( time timeout 600 python bench.py --no-cpu-baseline --no-sweep --no-e2e --steps 5 $extra ) > gpurun_out/bench_const_$tag.json 2> gpurun_out/bench_const_$tag.err; echo "bench $tag rc=$?"
  python - $tag <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/bench_const_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(sys.argv[1], {k:d.get(k) for k in ('value','value_by_w_mode','stage_ms_per_step','t2new_checksum_vs_stored')}, d['config']['constant_operands'])
PY
done
