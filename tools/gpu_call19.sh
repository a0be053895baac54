#!/bin/bash
# 1 GPU: validate HEAD (pytest -m gpu, smoke, default bench, reference arm), then ncu --set full of the kernels worth
# a capture (after the same program has exited 0 without ncu) and the launch list of one bench step
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
( time timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/summary.txt
tail -n 6 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/summary.txt
tail -n 2 gpurun_out/smoke.log
( time timeout 900 python bench.py ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/bench_default.err
( time timeout 600 python bench.py --impl reference ) > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/bench_ref.err
timeout 600 python tools/profile_targets.py > gpurun_out/profile_targets_plain.log 2>&1; rc=$?; echo "profile_targets plain rc=$rc" >> gpurun_out/summary.txt
tail -n 3 gpurun_out/profile_targets_plain.log
if [ $rc -eq 0 ]; then
  timeout 1200 ncu --set full --clock-control none --import-source on \
    -k regex:"tc_sign_gemm_pair|planes_fused|planes_predict|escan_emit|escan_map|escan_approx|key_hist|shard_values|shard_candidates" \
    -c 40 -f -o gpurun_out/prof_r02_targets python tools/profile_targets.py > gpurun_out/ncu_targets.log 2>&1; echo "ncu full rc=$?" >> gpurun_out/summary.txt
  tail -n 4 gpurun_out/ncu_targets.log
  ls -la gpurun_out/prof_r02_targets.ncu-rep
fi
( time timeout 900 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-sweep ) > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; rc=$?; echo "bench short rc=$rc" >> gpurun_out/summary.txt
if [ $rc -eq 0 ]; then
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_r02_final.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-sweep > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?" >> gpurun_out/summary.txt
  wc -l gpurun_out/launches_r02_final.csv
fi
python - <<'PY'
import json
for f in ('gpurun_out/bench_default.json','gpurun_out/bench_ref.json','gpurun_out/bench_short.json'):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k:d.get(k) for k in ('value','ladder_ms_per_step','stage_ms_per_step','value_by_w_mode','gpu_launches','t2new_checksum_vs_stored')})
        print('   ', d.get('e2e'), (d.get('roofline') or {}).get('frac'), d.get('roofline_planes'), d.get('clocks'))
    except Exception as e:
        print(f, "parse failed", e)
PY
cat gpurun_out/summary.txt
