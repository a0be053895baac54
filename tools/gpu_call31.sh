#!/bin/bash
# 1 GPU: final validation of HEAD (pytest -m gpu, smoke, default bench, reference arm) + ncu of the new emission kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
( time timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/summary.txt
tail -n 2 gpurun_out/smoke.log
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/bench_default.err
( time timeout 600 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 ) > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/bench_ref.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"escan_emit|escan_map" -c 4 -f -o gpurun_out/prof_r02_emit python tools/profile_targets.py > gpurun_out/ncu_emit.log 2>&1; echo "ncu emit rc=$?" >> gpurun_out/summary.txt
python - <<'PY'
import json
for f in ('gpurun_out/bench_default.json','gpurun_out/bench_ref.json'):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k:d.get(k) for k in ('value','steps','ladder_ms_per_step','stage_ms_per_step','stage_ms_slowest_step','value_by_w_mode','gpu_launches','t2new_checksum_vs_stored','scratch_allocations_in_timed_region')})
        print('   ', d.get('e2e'), (d.get('roofline') or {}), d.get('roofline_planes'), d.get('clocks'), d.get('cpu_baseline'))
    except Exception as e:
        print(f, "parse failed", e)
PY
cat gpurun_out/summary.txt
