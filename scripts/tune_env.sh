#!/bin/bash
# Sweep one environment knob over the timed decode steps of bench.py (--tune: decode steps only).
#   scripts/tune_env.sh NAME v1 v2 ...      [STEPS=100]
name=$1; shift
for v in "$@"; do
  out=$(env "$name=$v" python bench.py --steps ${STEPS:-100} --warmup 5 --tune 2>/dev/null | tail -1)
  echo "$name=$v $(python -c "import json,sys; d=json.loads(sys.argv[1]); print('ms/step %.4f frac %.4f median %.4f min %.4f sm %s' % (d['ms_per_step'], d['step_frac'], d['step_ms']['median'], d['step_ms']['min'], d['clocks']['sm_mhz']))" "$out")"
done
