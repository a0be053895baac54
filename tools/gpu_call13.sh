#!/bin/bash
# molecular-shape config on both arms + compute-sanitizer memcheck of the new kernels
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
( time timeout 900 python bench.py --o 21 --v 93 --m 1e6 --steps 5 --warmup 3 ) > gpurun_out/bench_mol.json 2> gpurun_out/bench_mol.err; echo "bench mol rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/bench_mol.err
( time timeout 1500 python bench.py --impl reference --o 21 --v 93 --m 1e6 --steps 1 --warmup 0 ) > gpurun_out/bench_mol_ref.json 2> gpurun_out/bench_mol_ref.err; echo "ref mol rc=$?" >> gpurun_out/summary.txt
tail -n 4 gpurun_out/bench_mol_ref.err
python - <<'PY'
import json
for f in ('gpurun_out/bench_mol.json','gpurun_out/bench_mol_ref.json'):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k:d.get(k) for k in ('value','stage_ms_per_step','parts_s','gpu_launches')}, d.get('e2e'), d.get('cpu_baseline'), d['config'].get('contraction_kernel'), d['config'].get('same_config'))
    except Exception as e:
        print(f, "parse failed", e)
PY
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 77 python -m pytest tests/test_dense_gpu.py tests/test_plugin_tc_gpu.py -q -m gpu -x -p no:cacheprovider -k "pair_and_one_cta or predicted_update or single_read or tc_equals_fp64_and_oracle and 2323" > gpurun_out/sanitizer_memcheck.log 2>&1; echo "sanitizer rc=$?" >> gpurun_out/summary.txt
tail -n 12 gpurun_out/sanitizer_memcheck.log
cat gpurun_out/summary.txt
