#!/bin/bash
# usage: dev_sweep.sh "WPC:KB WPC:KB ..."
for cfg in $1; do
  wpc=${cfg%%:*}; kb=${cfg##*:}
  GZCUDA_PGS_WPC=$wpc GZCUDA_PGS_SMEM_KB=$kb python bench.py --steps 100 --warmup 30 --no-cpu-baseline 2>&1 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('wpc=$wpc kb=$kb', round(d['value']), 'pgs_ms', round(d['roofline']['avg_launch_ms'],3))"
done
