#!/bin/bash
# usage (GPU box): scripts/handicap_bench.sh <refs> <pct> [<pct> ...]   -- UVG_PIVOT_HANDICAP sweep of the in-tree library
REFS=$1; shift
for h in "$@"; do
  UVG_PIVOT_HANDICAP=$h timeout 180 python bench.py --refs $REFS --steps 3 --warmup 2 --no-e2e --no-cpu > gpurun_out/hc_$h.json 2> gpurun_out/hc_$h.err
  python - "$h" <<'PY'
import json,sys
v=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/hc_{v}.json').read().strip().split('\n')[-1]); r=d['roofline']
    print(f"handicap {v:>3s}%  ms/step {d['ms_per_step']:8.2f}  kernel ms {r['kernel_ms_per_step']:8.2f}  frac {r['frac']:.3f}  checksum {d['result_checksum']}")
except Exception as e:
    print(v,'FAILED',e, open(f'gpurun_out/hc_{v}.err').read()[-300:])
PY
done
