#!/bin/bash
# round-2 GPU call 1: new parity/plugin tests (separate processes), old suite, probes, kernel timings, ncu of the TC kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
for f in test_parity_mode_gpu test_rccsd_energy_gpu test_plugin_tc_gpu test_dense_gpu; do
  timeout 900 python -m pytest tests/$f.py -q -m gpu --maxfail=40 -p no:cacheprovider > gpurun_out/pytest_$f.log 2>&1
  echo "$f rc=$?" >> gpurun_out/summary.txt
done
timeout 900 python -m pytest tests -q -m gpu --maxfail=20 -p no:cacheprovider \
   --deselect tests/test_parity_mode_gpu.py --deselect tests/test_rccsd_energy_gpu.py \
   --deselect tests/test_plugin_tc_gpu.py --deselect tests/test_dense_gpu.py > gpurun_out/pytest_rest.log 2>&1
echo "rest rc=$?" >> gpurun_out/summary.txt
timeout 120 tools/_bin/probe_sparse_mma > gpurun_out/probe_sparse_mma.txt 2>&1; echo "probe_sparse rc=$?" >> gpurun_out/summary.txt
timeout 120 tools/_bin/probe_int8_peak > gpurun_out/probe_int8_peak.txt 2>&1; echo "probe_int8 rc=$?" >> gpurun_out/summary.txt
timeout 600 python tools/bench_kernels.py --planes 6,3,2 --reps 3 > gpurun_out/bench_kernels.jsonl 2> gpurun_out/bench_kernels.err; echo "bench_kernels rc=$?" >> gpurun_out/summary.txt
timeout 300 python tools/bench_kernels.py --planes 6 --reps 1 --skip-two-pass > gpurun_out/bk_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_sign_gemm -c 1 -o gpurun_out/prof_tc_r02 \
    python tools/bench_kernels.py --planes 6 --reps 1 --skip-two-pass > gpurun_out/ncu_tc.log 2>&1
echo "ncu rc=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
tail -3 gpurun_out/pytest_*.log
cat gpurun_out/probe_sparse_mma.txt gpurun_out/probe_int8_peak.txt gpurun_out/bench_kernels.jsonl
