#!/bin/bash
# pair kernel: correctness first (bounded), then timings
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_dense_gpu.py -q -m gpu -x -p no:cacheprovider > gpurun_out/pytest_dense.log 2>&1; rc=$?
echo "dense rc=$rc" >> gpurun_out/summary.txt
tail -n 15 gpurun_out/pytest_dense.log
if [ $rc -eq 0 ]; then
  timeout 600 python -m pytest tests/test_plugin_tc_gpu.py -q -m gpu -p no:cacheprovider > gpurun_out/pytest_plugin.log 2>&1; echo "plugin rc=$?" >> gpurun_out/summary.txt
  tail -n 5 gpurun_out/pytest_plugin.log
  timeout 600 python tools/bench_kernels.py --planes 6,3 --reps 3 --skip-two-pass $BK_ARGS > gpurun_out/bench_kernels.jsonl 2> gpurun_out/bench_kernels.err; echo "bench_kernels rc=$?" >> gpurun_out/summary.txt
  cat gpurun_out/bench_kernels.jsonl; tail -n 5 gpurun_out/bench_kernels.err
fi
cat gpurun_out/summary.txt
