#!/bin/bash
# round-2 GPU call 2: smoke, kernel timings after the issue-loop fix, the new bench at N=1, TC tests, fresh ncu of the TC kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/summary.txt
timeout 600 python tools/bench_kernels.py --planes 6,3,2 --reps 3 > gpurun_out/bench_kernels.jsonl 2> gpurun_out/bench_kernels.err; echo "bench_kernels rc=$?" >> gpurun_out/summary.txt
( time timeout 1500 python bench.py --steps 2 --warmup 3 ) > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?" >> gpurun_out/summary.txt
for f in test_plugin_tc_gpu test_dense_gpu test_parity_mode_gpu; do
  timeout 900 python -m pytest tests/$f.py -q -m gpu --maxfail=40 -p no:cacheprovider > gpurun_out/pytest_$f.log 2>&1
  echo "$f rc=$?" >> gpurun_out/summary.txt
done
timeout 300 python tools/bench_kernels.py --planes 6 --reps 1 --skip-two-pass > gpurun_out/bk_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tc_sign_gemm|planes_fused" -c 3 -f -o gpurun_out/prof_tc_r02b \
    python tools/bench_kernels.py --planes 6 --reps 1 --skip-two-pass > gpurun_out/ncu_tc.log 2>&1
echo "ncu rc=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
cat gpurun_out/smoke.log | tail -3
cat gpurun_out/bench_kernels.jsonl
tail -c 3000 gpurun_out/bench_n1.err
head -c 6000 gpurun_out/bench_n1.json
for f in gpurun_out/pytest_*.log; do tail -n 3 $f; done
