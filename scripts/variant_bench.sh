#!/bin/bash
# usage (on the GPU box): scripts/variant_bench.sh <refs> <variant> [<variant> ...]    ("main" = uvaia_b200/lib)
# one short bench line per kernel-variant library built by `make OUT=../lib_<variant> EXTRA=-D...`
REFS=$1; shift
for v in "$@"; do
  if [ "$v" = "main" ]; then unset UVG_LIB; else export UVG_LIB=$PWD/uvaia_b200/lib_$v/libuvaia_b200.so; fi
  timeout 180 python bench.py --refs $REFS --steps 3 --warmup 2 --no-e2e --no-cpu > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err
  python - "$v" <<'PY'
import json,sys
v=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/var_{v}.json').read().strip().split('\n')[-1])
    r=d['roofline']
    print(f"{v:10s} ms/step {d['ms_per_step']:8.2f}  q/s {d['value']:9.1f}  kernel ms {r['kernel_ms_per_step']:8.2f}  frac {r['frac']:.3f}  executed {r['executed_fraction_of_dense']:.3f}  checksum {d['result_checksum']}")
except Exception as e:
    print(v,'FAILED',e, open(f'gpurun_out/var_{v}.err').read()[-300:])
PY
done
