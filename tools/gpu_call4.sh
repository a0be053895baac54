#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python tools/bench_kernels.py --planes 6,3 --reps 3 --skip-two-pass --debug > gpurun_out/bench_kernels_dbg.jsonl 2> gpurun_out/bench_kernels.err; echo "rc=$?"
cat gpurun_out/bench_kernels_dbg.jsonl; tail -n 5 gpurun_out/bench_kernels.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tc_sign_gemm_pair" -c 1 -f -o gpurun_out/prof_tc_pair_r02 \
    python tools/bench_kernels.py --planes 6 --reps 1 --skip-two-pass > gpurun_out/ncu_pair.log 2>&1
echo "ncu rc=$?"
