#!/bin/bash
# 1 GPU: last check of the bench line after the reporting changes (default flags) + smoke
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( time timeout 600 python bench.py --no-cpu-baseline ) > gpurun_out/bench_last.json 2> gpurun_out/bench_last.err; echo "bench rc=$?"
tail -n 3 gpurun_out/bench_last.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_last.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','value_by_w_mode','o3v3_stage_int8_tops','stage_ms_per_step','t2new_checksum_vs_stored','gpu_launches')})
print(d['e2e']['value'], d['roofline']['frac'], d['roofline']['traffic'], d['roofline_planes']['frac'], d['roofline_planes']['traffic'], d['config']['constant_operands'], d['config']['ladder_planes'])
PY
