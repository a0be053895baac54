#!/bin/bash
# 1 GPU: provoke the host stalls (fast nvidia-smi sampling, like the driver's own sampler beside ours) with the library's
# slow-call tracer on: which call is the host waiting in?
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,clocks.sm,power.draw,clocks_event_reasons.active --format=csv -lms 100 > gpurun_out/ext_smi.csv 2>/dev/null &
SMI=$!
for i in 1 2 3 4; do
  FRICC_TRACE_SLOW_MS=25 timeout 300 python bench.py --no-sweep --no-cpu-baseline --no-e2e --steps 10 --stall-trace 150 --clock-period-ms 50 > gpurun_out/slow_$i.json 2> gpurun_out/slow_$i.err
  python - "$i" <<'PY'
import json,sys
i=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/slow_{i}.json').read().strip().splitlines()[-1])
    w=d['stage_ms_slowest_step']; a=d['stage_ms_per_step']; h=d['stage_host_issue_ms_of_that_step']
    bad={k:(w[k],h[k]) for k in w if k!='compress_x2' and (w[k]>2*a[k]+1 or h[k]>50)}
    print(i,round(d['value'],4), 'outliers (device ms, host ms):', bad)
except Exception as e:
    print(i,'parse failed',e)
PY
  grep "fricc slow" gpurun_out/slow_$i.err | sort | uniq -c | sort -rn | head -30
done
kill $SMI
wc -l gpurun_out/ext_smi.csv
