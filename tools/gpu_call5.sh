#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python tools/bench_kernels.py --planes 6 --reps 3 --skip-two-pass > gpurun_out/bench_kernels_sync.jsonl 2> gpurun_out/bench_kernels.err; echo "rc=$?"
cat gpurun_out/bench_kernels_sync.jsonl; tail -n 5 gpurun_out/bench_kernels.err
timeout 900 python bench.py --steps 1 --warmup 1 --no-e2e --no-sweep --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_r02.csv \
   python bench.py --steps 1 --warmup 1 --no-e2e --no-sweep --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
echo "ncu rc=$?"
head -c 1500 gpurun_out/bench_short.json
