#!/bin/bash
# 1 GPU: where does the host stall come from?  default / no sampler / no sampler + gc off, 5 short runs each, with the
# stall watchdog
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
run() { tag=$1; shift
  for i in 1 2 3 4 5; do
    timeout 300 "$@" > gpurun_out/stall_${tag}_$i.json 2> gpurun_out/stall_${tag}_$i.err
    python - "$tag" "$i" <<'PY'
import json,sys
t,i=sys.argv[1:3]
try:
    d=json.loads(open(f'gpurun_out/stall_{t}_{i}.json').read().strip().splitlines()[-1])
    w=d['stage_ms_slowest_step']; a=d['stage_ms_per_step']; h=d['stage_host_issue_ms_of_that_step']
    bad={k:(w[k],h[k]) for k in w if k!='compress_x2' and (w[k]>2*a[k]+1 or h[k]>50)}
    print(t,i,round(d['value'],4), 'outliers (device ms, host ms):', bad)
except Exception as e:
    print(t,i,'parse failed',e)
PY
    grep -A8 "stall-trace" gpurun_out/stall_${tag}_$i.err | head -40
  done
}
run default python bench.py --no-sweep --no-cpu-baseline --no-e2e --steps 6 --stall-trace 100
run noclocks python bench.py --no-sweep --no-cpu-baseline --no-e2e --steps 6 --stall-trace 100 --no-clocks
run nogc env FRICC_BENCH_NOGC=1 python bench.py --no-sweep --no-cpu-baseline --no-e2e --steps 6 --stall-trace 100 --no-clocks
