#!/bin/bash
# 1 GPU: register-resident emit kernel (multi-draw) and 16-byte loads in the map kernel: compression tests, then the sweep
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_compress_gpu.py tests/test_ranges_gpu.py tests/test_parity_mode_gpu.py tests/test_rccsd_energy_gpu.py tests/test_device_driver.py tests/test_parallel_sample_gpu.py -m gpu -x -q -p no:cacheprovider ) > gpurun_out/pytest_compress.log 2>&1; echo "pytest rc=$?"
tail -n 5 gpurun_out/pytest_compress.log
( time timeout 600 python bench.py --no-cpu-baseline --steps 3 ) > gpurun_out/bench_emit.json 2> gpurun_out/bench_emit.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_emit.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','value_by_w_mode','stage_ms_per_step','t2new_checksum_vs_stored')})
print(d['e2e']['value'], d['e2e']['stage_ms_per_step'])
print([ (c['N'],c['sampler'][:10],c['ms_per_compression'],round(c['frac_of_measured_hbm'],3)) for c in d.get('compress_sweep',[])])
PY
