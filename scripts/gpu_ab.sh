#!/bin/bash
# A/B runs of the bench line on one cached setup: gpurun -- 'bash scripts/gpu_ab.sh tag "ENV1=.. ENV2=.." "ENV.." ...'
TAG=$1; shift
OUT=gpurun_out; mkdir -p $OUT
PACK=/tmp/bench_ab.ddmpack
python bench.py --pack $PACK --setup-only > $OUT/${TAG}_setup.log 2>&1
i=0
for envs in "$@"; do
  env $envs python bench.py --pack $PACK --no-cpu-baseline > $OUT/${TAG}_ab$i.json 2> $OUT/${TAG}_ab$i.err
  echo "== [$i] $envs (exit $?)"
  python - $OUT/${TAG}_ab$i.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
t=d["timers_ms_per_call"]
print("ms/step %.3f its %d apply %.4f fwd %.4f bwd %.4f restrict %.4f prolong %.4f coarse %.4f spmv %.4f e2e %.2f" % (d["ms_per_step"], d["config"]["iterations"], t["CombinedPreconditioner/apply"], t["Schwarz/local solve fwd"], t["Schwarz/local solve bwd"], t["Schwarz/get defect"], t["Schwarz/add solution"], t["GalerkinPrec/coarse solve"], t["Operator/apply"], d["e2e"]["value"]))
PY
  i=$((i+1))
done
