#!/bin/bash
# 1 GPU: block cache in front of cudaMallocAsync -- provoked-stall runs again (tracer on), then the full GPU suite
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,clocks.sm,power.draw,clocks_event_reasons.active --format=csv -lms 100 > gpurun_out/ext_smi.csv 2>/dev/null &
SMI=$!
for i in 1 2 3 4; do
  FRICC_TRACE_SLOW_MS=25 timeout 300 python bench.py --no-sweep --no-cpu-baseline --no-e2e --steps 10 --clock-period-ms 50 > gpurun_out/cache_$i.json 2> gpurun_out/cache_$i.err
  python - "$i" <<'PY'
import json,sys
i=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/cache_{i}.json').read().strip().splitlines()[-1])
    w=d['stage_ms_slowest_step']; a=d['stage_ms_per_step']; h=d['stage_host_issue_ms_of_that_step']
    bad={k:(w[k],h[k]) for k in w if k!='compress_x2' and (w[k]>2*a[k]+1 or h[k]>50)}
    print(i,round(d['value'],4), d['scratch_allocations_in_timed_region']['cudaMallocAsync_calls'], d['scratch_allocations_in_timed_region']['served_from_cached_blocks'], 'outliers (device ms, host ms):', bad)
except Exception as e:
    print(i,'parse failed',e)
PY
  grep "fricc slow" gpurun_out/cache_$i.err | grep -v "int8_peak\|device_count\|cudaEventSynchronize" | sort | uniq -c | sort -rn | head -12
done
kill $SMI
( time timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 6 gpurun_out/pytest_gpu.log
( time timeout 600 python bench.py ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','value_by_w_mode','stage_ms_per_step','stage_ms_slowest_step','scratch_allocations_in_timed_region','clocks')})
print(d['e2e'])
PY
