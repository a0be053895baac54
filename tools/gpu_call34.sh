#!/bin/bash
# 1 GPU: prepare_operand -- plugin tests, then the default bench (e2e with the ladder operand encoded during the upload)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests/test_plugin_tc_gpu.py -m gpu -x -q -p no:cacheprovider ) > gpurun_out/pytest_prep.log 2>&1; echo "pytest rc=$?"
tail -n 12 gpurun_out/pytest_prep.log
( time timeout 600 python bench.py --no-cpu-baseline --no-sweep ) > gpurun_out/bench_prep.json 2> gpurun_out/bench_prep.err; echo "bench rc=$?"
tail -n 3 gpurun_out/bench_prep.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_prep.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','value_by_w_mode','t2new_checksum_vs_stored')})
e=d['e2e']; print(e['value'], e.get('prepared_plane_hits'), e['stage_ms_per_step'])
PY
