#!/bin/bash
# 1 GPU: is the slow-stage outlier of the default bench run gone with the clock sampler started before the warm-up?
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for i in 1 2 3; do
  ( time timeout 600 python bench.py --no-sweep --no-cpu-baseline --no-e2e ) > gpurun_out/bench_ab_$i.json 2> gpurun_out/bench_ab_$i.err; echo "rc=$?"
done
( time timeout 600 python bench.py --no-sweep --no-cpu-baseline --no-e2e --steps 8 ) > gpurun_out/bench_ab_4.json 2> gpurun_out/bench_ab_4.err; echo "rc=$?"
python - <<'PY'
import json
for i in (1,2,3,4):
    try:
        d=json.loads(open(f'gpurun_out/bench_ab_{i}.json').read().strip().splitlines()[-1])
        print(i, d['value'], d.get('value_by_w_mode'), d['clocks'])
        print('  avg  ', d['stage_ms_per_step'])
        print('  worst', d['stage_ms_slowest_step'])
        print('  host ', d['stage_host_issue_ms_of_that_step'])
    except Exception as e:
        print(i, 'parse failed', e)
PY
