#!/bin/bash
# 1 GPU: the whole GPU suite + smoke + default bench with the final library; molecular-shape config on both arms;
# compute-sanitizer memcheck of the hand-rolled mbarrier/TMEM kernels (one tool per call: no ncu here)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_ranges_gpu.py -q -m gpu -x -p no:cacheprovider > gpurun_out/pytest_ranges.log 2>&1; echo "ranges tests rc=$?" >> gpurun_out/summary.txt
tail -n 15 gpurun_out/pytest_ranges.log
timeout 1800 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu_full.log 2>&1; echo "full gpu tests rc=$?" >> gpurun_out/summary.txt
tail -n 12 gpurun_out/pytest_gpu_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/summary.txt
tail -n 3 gpurun_out/smoke.log
( time timeout 900 python bench.py ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/bench_default.err
( time timeout 900 python bench.py --o 21 --v 93 --m 1e6 --steps 5 --warmup 3 ) > gpurun_out/bench_mol.json 2> gpurun_out/bench_mol.err; echo "bench mol rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/bench_mol.err
( time timeout 1500 python bench.py --impl reference --o 21 --v 93 --m 1e6 --steps 1 --warmup 0 ) > gpurun_out/bench_mol_ref.json 2> gpurun_out/bench_mol_ref.err; echo "ref mol rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/bench_mol_ref.err
python - <<'PY'
import json
for f in ('gpurun_out/bench_default.json','gpurun_out/bench_mol.json','gpurun_out/bench_mol_ref.json'):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k:d.get(k) for k in ('value','ms_per_step','stage_ms_per_step','parts_s','gpu_launches','value_by_w_mode','t2new_checksum_vs_stored')})
        print('   ', d.get('e2e'), d.get('cpu_baseline'), (d.get('roofline') or {}).get('frac'), (d.get('roofline_planes') or {}).get('frac'), d.get('clocks'))
        print('   ', d.get('compress_sweep'))
    except Exception as e:
        print(f, "parse failed", e)
PY
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 77 python -m pytest tests/test_dense_gpu.py tests/test_plugin_tc_gpu.py -q -m gpu -x -p no:cacheprovider -k "pair_and_one_cta or predicted_update or single_read or fused_pair or (tc_equals_fp64_and_oracle and 2323)" > gpurun_out/sanitizer_memcheck.log 2>&1; echo "sanitizer rc=$?" >> gpurun_out/summary.txt
tail -n 12 gpurun_out/sanitizer_memcheck.log
cat gpurun_out/summary.txt
