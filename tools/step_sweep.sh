#!/bin/bash
# Sweep the persistent decode kernel's build-time / environment variants on one B200 and summarise tok/s.
# usage: tools/step_sweep.sh TAG "ENV1=a ENV2=b" "ENV1=c" ...   (each argument = one bench.py run with that environment)
tag=$1; shift
mkdir -p gpurun_out
i=0
for v in "$@"; do
  i=$((i+1))
  out=gpurun_out/${tag}_$i
  env $v python bench.py --steps ${STEPS:-64} --warmup 8 --no-cpu-baseline > $out.json 2> $out.err || { echo "FAILED: $v"; tail -5 $out.err; }
  python - "$v" $out.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    r = d["roofline"]
    print(f"{sys.argv[1]:50s} value {d['value']:8.1f} tok/s  e2e {d['e2e']['value']:8.1f}  frac {r['frac']:.3f}  us/launch {r['us_per_launch']:.1f}  clocks {d['clocks'].get('sm_mhz')} {d['clocks'].get('reasons')}")
except Exception as e:
    print(sys.argv[1], "no result:", e)
PY
done
